// sa_api.cu -- C-ABI entry points: text builds, accessors, batched query (see include/sa_cuda.h).
#include <string.h>

#include "sa_index.h"

#define SA_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            sa_set_error(__VA_ARGS__);         \
            return (code);                     \
        }                                      \
    } while (0)

static const uint64_t SA_MAX_N = 0xFFFFFFFFull - 1024ull;

static int prepare_text(sa_index* idx, uint64_t n, uint32_t flags)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    SA_REQUIRE(n <= SA_MAX_N, SA_ERR_LIMIT, "n = %llu exceeds the uint32 position range", (unsigned long long)n);
    SA_REQUIRE(n <= SA_MAX_INDEX_BYTES, SA_ERR_LIMIT,
               "n = %llu: one index holds at most 2^31 bytes (the reference's partition size, suffix_array.pyx:148)",
               (unsigned long long)n);
    SA_CU(cudaSetDevice(idx->device));
    idx->timing = (flags & SA_FLAG_TIMING) != 0;
    idx->n = n;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    // zero the padding (and the alignment slack in front of it)
    const size_t pad_from = (size_t)n & ~(size_t)15;
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + pad_from, 0, bytes - pad_from, idx->stream));
    return SA_OK;
}

extern "C" int sa_build_text(sa_index* idx, const uint8_t* text, uint64_t n, uint32_t flags)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    SA_REQUIRE(text || n == 0, SA_ERR_INVALID, "text is NULL");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    int rc = prepare_text(idx, n, flags);
    if (rc != SA_OK) return rc;
    idx->stats.ms_h2d = 0.f;
    if (n) {
        ScopedPhase ph(idx, &idx->stats.ms_h2d);
        SA_CU(cudaMemcpyAsync(idx->text.p, text, n, cudaMemcpyHostToDevice, idx->stream));
    }
    return sa_build_resident(idx, flags);
}

extern "C" int sa_build_text_device(sa_index* idx, const uint8_t* d_text, uint64_t n, uint32_t flags)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    SA_REQUIRE(d_text || n == 0, SA_ERR_INVALID, "d_text is NULL");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    int rc = prepare_text(idx, n, flags);
    if (rc != SA_OK) return rc;
    idx->stats.ms_h2d = 0.f;
    if (n) SA_CU(cudaMemcpyAsync(idx->text.p, d_text, n, cudaMemcpyDeviceToDevice, idx->stream));
    return sa_build_resident(idx, flags);
}

// ------------------------------------------------------------------------------ accessors
extern "C" uint64_t sa_size(const sa_index* idx) { return idx && idx->built ? idx->n : 0; }
extern "C" uint64_t sa_num_docs(const sa_index* idx) { return idx && idx->built ? (uint64_t)idx->num_nl + 1 : 0; }
extern "C" const uint32_t* sa_device_sa(const sa_index* idx)
{
    if (!idx || !idx->built) return nullptr;
    return idx->csv ? idx->sa_file.as<uint32_t>() : idx->sa_final.as<uint32_t>();
}

static int copy_out(sa_index* idx, const void* dev, void* out, size_t bytes)
{
    if (!bytes) return SA_OK;
    SA_CU(cudaSetDevice(idx->device));
    SA_CU(cudaMemcpyAsync(out, dev, bytes, cudaMemcpyDeviceToHost, idx->stream));
    SA_CU(cudaStreamSynchronize(idx->stream));
    return SA_OK;
}

extern "C" int sa_copy_sa_u32(sa_index* idx, uint32_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(out || idx->n == 0, SA_ERR_INVALID, "out is NULL");
    SA_REQUIRE(cap >= idx->n, SA_ERR_INVALID, "cap %llu < n %llu", (unsigned long long)cap, (unsigned long long)idx->n);
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    return copy_out(idx, sa_device_sa(idx), out, (size_t)idx->n * 4);
}

extern "C" int sa_copy_text_positions_u32(sa_index* idx, uint32_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(cap >= idx->n, SA_ERR_INVALID, "cap too small");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    return copy_out(idx, idx->sa_final.p, out, (size_t)idx->n * 4);
}

extern "C" int sa_copy_sa_i64(sa_index* idx, int64_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(cap >= idx->n, SA_ERR_INVALID, "cap too small");
    if (idx->n == 0) return SA_OK;
    // widen in place from the tail: the u32 copy lands in the upper half of the caller's buffer
    uint32_t* tmp = reinterpret_cast<uint32_t*>(out) + idx->n;
    int rc = sa_copy_sa_u32(idx, tmp, idx->n);
    if (rc != SA_OK) return rc;
    for (uint64_t i = 0; i < idx->n; ++i) out[i] = (int64_t)tmp[i];
    return SA_OK;
}

extern "C" int sa_copy_quoted_bits(sa_index* idx, uint8_t* out_bits, uint64_t cap_bytes)
{
    SA_REQUIRE(idx && idx->built && idx->csv, SA_ERR_INVALID, "no CSV index built");
    const uint64_t need = (idx->n + 7) / 8;
    SA_REQUIRE(cap_bytes >= need, SA_ERR_INVALID, "cap too small");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    return copy_out(idx, idx->qbits.p, out_bits, (size_t)need);
}

extern "C" int sa_copy_text(sa_index* idx, uint8_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(cap >= idx->n, SA_ERR_INVALID, "cap too small");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    return copy_out(idx, idx->text.p, out, (size_t)idx->n);
}

extern "C" int sa_copy_doc_starts(sa_index* idx, uint32_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    const uint64_t nd = (uint64_t)idx->num_nl + 1;
    SA_REQUIRE(cap >= nd, SA_ERR_INVALID, "cap too small");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    out[0] = 0;
    int rc = copy_out(idx, idx->nlpos.p, out + 1, (size_t)idx->num_nl * 4);
    if (rc != SA_OK) return rc;
    for (uint64_t i = 1; i < nd; ++i) out[i] += 1;      // record d starts right after the (d-1)-th '\n'
    return SA_OK;
}

extern "C" int sa_get_stats(const sa_index* idx, sa_stats* out)
{
    SA_REQUIRE(idx && out, SA_ERR_INVALID, "NULL argument");
    *out = idx->stats;
    return SA_OK;
}

extern "C" int sa_debug_copy_lcp(sa_index* idx, uint16_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(cap >= idx->n, SA_ERR_INVALID, "cap too small");
    SA_REQUIRE(idx->lcp.p, SA_ERR_INVALID, "no lcp array");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    return copy_out(idx, idx->lcp.p, out, (size_t)idx->n * 2);
}

extern "C" int sa_debug_set(sa_index* idx, const char* key, int64_t value)
{
    SA_REQUIRE(idx && key, SA_ERR_INVALID, "NULL argument");
    if (!strcmp(key, "max_rounds")) idx->dbg_max_rounds = value;
    else if (!strcmp(key, "skip_finish")) idx->dbg_skip_finish = value;
    else if (!strcmp(key, "force_gbits")) idx->dbg_force_gbits = value;
    else if (!strcmp(key, "code_bits")) idx->dbg_code_bits = value;
    else if (!strcmp(key, "sort0_chars")) idx->dbg_sort0_chars = value;
    else if (!strcmp(key, "csv_scanner")) idx->dbg_csv_scanner = value;
    else if (!strcmp(key, "replicate_memcpy")) idx->dbg_replicate_memcpy = value;
    else if (!strcmp(key, "same_process")) idx->dbg_same_process = value;
    else if (!strcmp(key, "query_kernel")) idx->dbg_query_kernel = value;
    else if (!strcmp(key, "onesweep_variant")) sa_onesweep_variant = (int)value;
    else { sa_set_error("unknown debug key %s", key); return SA_ERR_INVALID; }
    return SA_OK;
}

// ------------------------------------------------------------------------------ query
static int run_ranges(sa_index* idx, const uint8_t* d_q, const uint64_t* d_off, uint64_t nq, uint32_t k,
                      uint32_t* d_start, uint32_t* d_end, bool offsets)
{
    cudaStream_t s = idx->stream;
    {
        ScopedPhase ph(idx, &idx->stats.ms_query);
        SA_CU(sa_launch_query(idx->text.as<uint8_t>(), idx->sa_final.as<uint32_t>(), (uint32_t)idx->n, idx->L, d_q,
                              d_off, (uint32_t)nq, d_start, d_end, (int)idx->dbg_query_kernel, s));
        idx->stats.kernel_launches += 1;
    }
    if (offsets) {
        ScopedPhase ph(idx, &idx->stats.ms_hits);
        SA_CU(idx->qcounts.ensure((size_t)nq * 4));
        SA_CU(idx->hitoff.ensure(((size_t)nq + 1) * 8));
        SA_CU(idx->blocksums.ensure(((size_t)(nq + 1023) / 1024 + 2) * 8));
        SA_CU(sa_launch_hit_counts(d_start, d_end, (uint32_t)nq, k, idx->qcounts.as<uint32_t>(), s));
        SA_CU(sa_launch_scan_u32_to_u64(idx->qcounts.as<uint32_t>(), (uint32_t)nq, idx->hitoff.as<uint64_t>(),
                                        idx->blocksums.as<uint32_t>(), s));
        idx->stats.kernel_launches += 4;
    }
    return SA_OK;
}

extern "C" int sa_query_ranges(sa_index* idx, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq, uint32_t k,
                               uint32_t* start, uint32_t* end, uint64_t* hit_off)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(qoff || nq == 0, SA_ERR_INVALID, "qoff is NULL");
    SA_REQUIRE(nq < (1ull << 28), SA_ERR_LIMIT, "too many queries in one batch");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    sa_invalidate_results(idx);
    idx->stats.ms_query = idx->stats.ms_hits = 0.f;
    idx->stats.kernel_launches = 0;
    if (hit_off) hit_off[0] = 0;
    idx->last_nq = nq;
    idx->last_total_hits = 0;
    if (nq == 0) { idx->have_ranges = true; return SA_OK; }
    if (idx->n == 0) {       // empty index: every query misses (engine.c:896-899 sentinel)
        for (uint64_t i = 0; i < nq; ++i) {
            if (start) start[i] = 0xFFFFFFFFu;
            if (end) end[i] = 0xFFFFFFFFu;
            if (hit_off) hit_off[i + 1] = 0;
        }
        idx->last_nq = 0;
        idx->have_ranges = true;
        return SA_OK;
    }
    cudaStream_t s = idx->stream;
    const uint64_t qlen = qoff[nq];
    SA_REQUIRE(qbytes || qlen == 0, SA_ERR_INVALID, "qbytes is NULL");
    SA_CU(idx->qbytes.ensure((size_t)qlen + 64));
    SA_CU(idx->qoff.ensure(((size_t)nq + 1) * 8));
    SA_CU(idx->qstart.ensure((size_t)nq * 4));
    SA_CU(idx->qend.ensure((size_t)nq * 4));
    if (qlen) SA_CU(cudaMemcpyAsync(idx->qbytes.p, qbytes, qlen, cudaMemcpyHostToDevice, s));
    SA_CU(cudaMemcpyAsync(idx->qoff.p, qoff, (nq + 1) * 8, cudaMemcpyHostToDevice, s));
    int rc = run_ranges(idx, idx->qbytes.as<uint8_t>(), idx->qoff.as<uint64_t>(), nq, k, idx->qstart.as<uint32_t>(),
                        idx->qend.as<uint32_t>(), true);
    if (rc != SA_OK) return rc;
    if (start) SA_CU(cudaMemcpyAsync(start, idx->qstart.p, nq * 4, cudaMemcpyDeviceToHost, s));
    if (end) SA_CU(cudaMemcpyAsync(end, idx->qend.p, nq * 4, cudaMemcpyDeviceToHost, s));
    if (hit_off) SA_CU(cudaMemcpyAsync(hit_off, idx->hitoff.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, s));
    uint64_t* h_total = reinterpret_cast<uint64_t*>(idx->h_small + 32);
    SA_CU(cudaMemcpyAsync(h_total, idx->hitoff.as<uint64_t>() + nq, 8, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    idx->last_total_hits = *h_total;
    idx->have_ranges = true;
    sa_time_collect(idx);
    return SA_OK;
}

extern "C" int sa_query_hits(sa_index* idx, uint32_t* hit_pos, uint32_t* hit_doc, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built && idx->have_ranges, SA_ERR_INVALID, "no sa_query_ranges result to expand");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    const uint64_t total = idx->last_total_hits;
    SA_REQUIRE(cap >= total, SA_ERR_INVALID, "cap %llu < %llu hits", (unsigned long long)cap, (unsigned long long)total);
    if (total == 0 || idx->last_nq == 0) return SA_OK;
    cudaStream_t s = idx->stream;
    if (hit_pos) SA_CU(idx->hitpos.ensure((size_t)total * 4));
    if (hit_doc) SA_CU(idx->hitdoc.ensure((size_t)total * 4));
    {
        ScopedPhase ph(idx, &idx->stats.ms_hits);
        // text index: position in the text, record = '\n' bytes in front of it; CSV index: offset in the
        // FILE (what the remapped array holds, engine.c:647-651), record = row of the file (header = 0)
        const bool rows = idx->csv && idx->rowbits.p && idx->rowrank.p;
        SA_CU(sa_launch_hits(rows ? idx->sa_file.as<uint32_t>() : idx->sa_final.as<uint32_t>(), idx->qstart.as<uint32_t>(),
                             idx->hitoff.as<uint64_t>(), (uint32_t)idx->last_nq,
                             rows ? idx->rowbits.as<unsigned long long>() : idx->nlbits.as<unsigned long long>(),
                             rows ? idx->rowrank.as<uint32_t>() : idx->nlrank.as<uint32_t>(),
                             hit_pos ? idx->hitpos.as<uint32_t>() : nullptr, hit_doc ? idx->hitdoc.as<uint32_t>() : nullptr,
                             s));
        idx->stats.kernel_launches += 1;
    }
    if (hit_pos) SA_CU(cudaMemcpyAsync(hit_pos, idx->hitpos.p, total * 4, cudaMemcpyDeviceToHost, s));
    if (hit_doc) SA_CU(cudaMemcpyAsync(hit_doc, idx->hitdoc.p, total * 4, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    sa_time_collect(idx);
    return SA_OK;
}

extern "C" int sa_query_ranges_device(sa_index* idx, const uint8_t* d_qbytes, const uint64_t* d_qoff, uint64_t nq,
                                      uint64_t qbytes_len, uint32_t k, uint32_t* d_start, uint32_t* d_end)
{
    (void)qbytes_len; (void)k;
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(idx->n > 0, SA_ERR_INVALID, "empty index");
    SA_REQUIRE(d_qbytes && d_qoff && d_start && d_end, SA_ERR_INVALID, "NULL device pointer");
    SA_REQUIRE(nq < (1ull << 28), SA_ERR_LIMIT, "too many queries in one batch");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    idx->stats.ms_query = idx->stats.ms_hits = 0.f;
    idx->stats.kernel_launches = 0;
    int rc = run_ranges(idx, d_qbytes, d_qoff, nq, k, d_start, d_end, false);
    if (rc != SA_OK) return rc;
    SA_CU(cudaStreamSynchronize(idx->stream));
    sa_time_collect(idx);
    return SA_OK;
}
