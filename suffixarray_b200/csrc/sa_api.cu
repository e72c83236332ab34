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
static const uint64_t SA_MAX_TAIL = SA_TEXT_PAD - 64;      // the zero padding behind it keeps every window read in bounds

// Device text = n indexed bytes + `tail` look-ahead bytes + zero padding.
static int prepare_text(sa_index* idx, uint64_t n, uint32_t flags, uint32_t tail = 0)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    SA_REQUIRE(n <= SA_MAX_N, SA_ERR_LIMIT, "n = %llu exceeds the uint32 position range", (unsigned long long)n);
    SA_REQUIRE(n <= SA_MAX_INDEX_BYTES, SA_ERR_LIMIT,
               "n = %llu: one index holds at most 2^31 bytes (the reference's partition size, suffix_array.pyx:148)",
               (unsigned long long)n);
    SA_CU(cudaSetDevice(idx->device));
    idx->timing = (flags & SA_FLAG_TIMING) != 0;
    idx->n = n;
    idx->tail = tail;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    // zero the padding (and the alignment slack in front of it)
    const size_t pad_from = ((size_t)n + tail) & ~(size_t)15;
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + pad_from, 0, bytes - pad_from, idx->stream));
    return SA_OK;
}

extern "C" int sa_build_text(sa_index* idx, const uint8_t* text, uint64_t n, uint32_t flags)
{
    return sa_build_text_tail(idx, text, n, 0, flags);
}

// The reference hands construct_truncated_suffix_array `text + idx * 2 GiB` of ONE contiguous buffer
// (suffix_array.pyx:176-180): the suffixes within max_suffix_length + 17 bytes of a partition's end are
// sorted -- and later compared by the bisection -- on the first bytes of the NEXT partition.  `tail`
// such look-ahead bytes follow the n indexed ones in `text` (at most SA_MAX_TAIL are used).
extern "C" int sa_build_text_tail(sa_index* idx, const uint8_t* text, uint64_t n, uint64_t tail, uint32_t flags)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    SA_REQUIRE(text || n == 0, SA_ERR_INVALID, "text is NULL");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    if (n == 0) tail = 0;
    if (tail > SA_MAX_TAIL) tail = SA_MAX_TAIL;
    int rc = prepare_text(idx, n, flags, (uint32_t)tail);
    if (rc != SA_OK) return rc;
    idx->stats.ms_h2d = 0.f;
    if (n) {
        ScopedPhase ph(idx, &idx->stats.ms_h2d);
        SA_CU(cudaMemcpyAsync(idx->text.p, text, n + tail, cudaMemcpyHostToDevice, idx->stream));
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
    if (!idx || !idx->built || idx->sharded_sa) return nullptr;     // a sharded array has no single device pointer
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
    if (idx->sharded_sa) {                       // gather the key ranges from their owners (peer memory)
        SA_CU(cudaSetDevice(idx->device));
        for (int g = 0; g < idx->view.n; ++g) {
            const uint64_t a = idx->view.bound[g], b = idx->view.bound[g + 1];
            if (b > a) SA_CU(cudaMemcpyAsync(out + a, idx->view.base[g] + a, (size_t)(b - a) * 4, cudaMemcpyDefault, idx->stream));
        }
        SA_CU(cudaStreamSynchronize(idx->stream));
        return SA_OK;
    }
    return copy_out(idx, sa_device_sa(idx), out, (size_t)idx->n * 4);
}

extern "C" int sa_copy_text_positions_u32(sa_index* idx, uint32_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && idx->built, SA_ERR_INVALID, "index not built");
    if (idx->sharded_sa) return sa_copy_sa_u32(idx, out, cap);
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
    else if (!strcmp(key, "timeline")) idx->dbg_timeline = value;
    else if (!strcmp(key, "replicate_ctas")) idx->dbg_replicate_ctas = value;
    else if (!strcmp(key, "finish_chunks")) idx->dbg_finish_chunks = value;
    else if (!strcmp(key, "text_push")) idx->dbg_text_push = value;
    else if (!strcmp(key, "ext_fill")) idx->dbg_ext_fill = value;
    else if (!strcmp(key, "replicate_fused")) idx->dbg_replicate_fused = value;
    else if (!strcmp(key, "split_gamma_x100")) idx->dbg_split_gamma_x100 = value;
    else if (!strcmp(key, "same_process")) idx->dbg_same_process = value;
    else if (!strcmp(key, "query_kernel")) idx->dbg_query_kernel = value;
    else if (!strcmp(key, "qtop_levels")) { idx->dbg_qtop_levels = value; idx->qtop_ready = false; }
    else { sa_set_error("unknown debug key %s", key); return SA_ERR_INVALID; }
    return SA_OK;
}

