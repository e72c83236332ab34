// sa_query.cu -- C-ABI query entry points (see include/sa_cuda.h): rank ranges, hit lists, segmented
// dedupe, record materialisation, the CSV row table.  All per-call state lives in an sa_result, so
// several threads can query one built index at once; the index is only read (shared lock).
#include <string.h>

#include <new>

#include "sa_index.h"

#define SA_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            sa_set_error(__VA_ARGS__);         \
            return (code);                     \
        }                                      \
    } while (0)

// ------------------------------------------------------------------------------ result objects
extern "C" int sa_result_create(sa_index* idx, sa_result** out)
{
    SA_REQUIRE(idx && out, SA_ERR_INVALID, "sa_result_create: NULL argument");
    *out = nullptr;
    SA_CU(cudaSetDevice(idx->device));
    sa_result* r = new (std::nothrow) sa_result();
    SA_REQUIRE(r, SA_ERR_OOM, "host allocation failed");
    r->idx = idx;
    r->device = idx->device;
    cudaError_t e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
    for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreate(&r->ev[i]);
    if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&r->h_pin), 8 * sizeof(uint64_t));
    if (e != cudaSuccess) { sa_result_destroy(r); return sa_cuda_fail(e, "sa_result_create"); }
    *out = r;
    return SA_OK;
}

extern "C" void sa_result_destroy(sa_result* r)
{
    if (!r) return;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    DevBuf* all[] = {&r->qbytes, &r->qoff, &r->qstart, &r->qend, &r->qcounts, &r->hitoff, &r->blocksums, &r->hitpos,
                     &r->hitdoc, &r->ucount, &r->uoff, &r->upos, &r->udoc, &r->table, &r->reclen, &r->recoff,
                     &r->recbytes, &r->recsrc};
    for (DevBuf* b : all) b->release();
    for (cudaEvent_t e : r->ev)
        if (e) cudaEventDestroy(e);
    if (r->h_pin) cudaFreeHost(r->h_pin);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
}

static sa_result* default_result(sa_index* idx)
{
    std::lock_guard<std::mutex> g(idx->res_mu);
    if (!idx->default_res && sa_result_create(idx, &idx->default_res) != SA_OK) return nullptr;
    return idx->default_res;
}

static float elapsed(cudaEvent_t a, cudaEvent_t b)
{
    float ms = 0.f;
    return cudaEventElapsedTime(&ms, a, b) == cudaSuccess ? ms : 0.f;
}

// ------------------------------------------------------------------------------ probe cache
// The first thread-per-query batch after a build fills the table (about 1 M gathers); later batches and
// other threads find it ready.  Small batches (8 lanes per query) do not use it.
static int ensure_qtop(sa_index* idx, cudaStream_t s)
{
    std::lock_guard<std::mutex> g(idx->qtop_mu);
    if (idx->qtop_ready) return SA_OK;
    int levels = 0;
    if (idx->dbg_qtop_levels != 0 && idx->n >= 2) {
        while (levels < 20 && (2ull << levels) <= idx->n) ++levels;      // 2^levels <= n: every cached level is full-ish
        if (idx->dbg_qtop_levels > 0 && idx->dbg_qtop_levels < levels) levels = (int)idx->dbg_qtop_levels;
    }
    idx->qtop_levels = levels;
    if (levels > 0) {
        SA_CU(idx->qtop.ensure(((size_t)1 << levels) * 16));
        if (idx->sharded_sa)
            SA_CU(sa_launch_qtop_build_sharded(idx->text.as<uint8_t>(), idx->view, (uint32_t)idx->n, levels, idx->qtop.p, s));
        else
            SA_CU(sa_launch_qtop_build(idx->text.as<uint8_t>(), idx->sa_final.as<uint32_t>(), (uint32_t)idx->n, levels, idx->qtop.p, s));
        SA_CU(cudaStreamSynchronize(s));
    }
    idx->qtop_ready = true;
    return SA_OK;
}

// ------------------------------------------------------------------------------ ranges
static int ranges_locked(sa_index* idx, sa_result* r, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq, uint32_t k,
                         uint32_t flags, uint32_t* start, uint32_t* end, uint64_t* hit_off)
{
    SA_CU(cudaSetDevice(idx->device));
    const bool file_mode = (flags & SA_QUERY_FILE_SEMANTICS) != 0;
    if (file_mode) {
        SA_REQUIRE(idx->csv, SA_ERR_INVALID, "SA_QUERY_FILE_SEMANTICS needs a CSV index");
        SA_REQUIRE(idx->csv_file.p && idx->csv_file_bytes, SA_ERR_INVALID,
                   "the CSV file is not resident on the device (index restored by sa_load: call sa_attach_csv_file)");
    }
    // first batch since the build / load: whatever the index stream still has queued comes first
    if (r->generation != idx->generation) SA_CU(cudaStreamSynchronize(idx->stream));
    r->have_ranges = r->have_hits = r->have_records = false;
    r->ms_query = r->ms_hits = 0.f;
    r->launches = 0;
    r->nq = nq;
    r->k = k;
    r->range_flags = flags;
    r->total_hits = 0;
    r->generation = idx->generation;
    if (hit_off) hit_off[0] = 0;
    if (nq == 0) { r->have_ranges = true; return SA_OK; }
    if (idx->n == 0) {       // empty index: every query misses (engine.c:896-899 sentinel)
        for (uint64_t i = 0; i < nq; ++i) {
            if (start) start[i] = 0xFFFFFFFFu;
            if (end) end[i] = 0xFFFFFFFFu;
            if (hit_off) hit_off[i + 1] = 0;
        }
        r->nq = 0;
        r->have_ranges = true;
        return SA_OK;
    }
    cudaStream_t s = r->stream;
    const uint64_t qlen = qoff[nq];
    SA_REQUIRE(qbytes || qlen == 0, SA_ERR_INVALID, "qbytes is NULL");
    SA_CU(r->qbytes.ensure((size_t)qlen + 64));
    SA_CU(r->qoff.ensure(((size_t)nq + 1) * 8));
    SA_CU(r->qstart.ensure((size_t)nq * 4));
    SA_CU(r->qend.ensure((size_t)nq * 4));
    SA_CU(r->qcounts.ensure((size_t)nq * 4));
    SA_CU(r->hitoff.ensure(((size_t)nq + 1) * 8));
    SA_CU(r->blocksums.ensure(((size_t)(nq + 1023) / 1024 + 2) * 8));
    if (qlen) SA_CU(cudaMemcpyAsync(r->qbytes.p, qbytes, qlen, cudaMemcpyHostToDevice, s));
    SA_CU(cudaMemcpyAsync(r->qoff.p, qoff, (nq + 1) * 8, cudaMemcpyHostToDevice, s));
    // an array whose key ranges live on several GPUs is always queried one thread per query (with the probe cache)
    const bool thread_kernel = !file_mode && (idx->sharded_sa || idx->dbg_query_kernel == 1 || (idx->dbg_query_kernel == 0 && nq >= 4096));
    if (thread_kernel) {
        int rc = ensure_qtop(idx, s);
        if (rc != SA_OK) return rc;
    }
    SA_CU(cudaEventRecord(r->ev[0], s));
    if (idx->sharded_sa && !file_mode)
        SA_CU(sa_launch_query_sharded(idx->text.as<uint8_t>(), idx->view, (uint32_t)idx->n, idx->L, r->qbytes.as<uint8_t>(),
                                      r->qoff.as<uint64_t>(), (uint32_t)nq, r->qstart.as<uint32_t>(), r->qend.as<uint32_t>(), s,
                                      idx->qtop_levels ? idx->qtop.p : nullptr, idx->qtop_levels));
    else if (file_mode)
        SA_CU(sa_launch_query_file(idx->csv_file.as<uint8_t>(), idx->csv_file_bytes, idx->sa_file.as<uint32_t>(),
                                   (uint32_t)idx->n, idx->L, r->qbytes.as<uint8_t>(), r->qoff.as<uint64_t>(), (uint32_t)nq,
                                   r->qstart.as<uint32_t>(), r->qend.as<uint32_t>(), s));
    else
        SA_CU(sa_launch_query(idx->text.as<uint8_t>(), idx->sa_final.as<uint32_t>(), (uint32_t)idx->n, idx->L,
                              r->qbytes.as<uint8_t>(), r->qoff.as<uint64_t>(), (uint32_t)nq, r->qstart.as<uint32_t>(),
                              r->qend.as<uint32_t>(), (int)idx->dbg_query_kernel, s,
                              thread_kernel && idx->qtop_levels ? idx->qtop.p : nullptr, idx->qtop_levels));
    SA_CU(cudaEventRecord(r->ev[1], s));
    SA_CU(cudaEventRecord(r->ev[2], s));
    SA_CU(sa_launch_hit_counts(r->qstart.as<uint32_t>(), r->qend.as<uint32_t>(), (uint32_t)nq, k, (uint32_t)idx->n,
                               r->qcounts.as<uint32_t>(), s));
    SA_CU(sa_launch_scan_u32_to_u64(r->qcounts.as<uint32_t>(), (uint32_t)nq, r->hitoff.as<uint64_t>(),
                                    r->blocksums.as<uint32_t>(), s));
    SA_CU(cudaEventRecord(r->ev[3], s));
    r->launches += 5;
    if (start) SA_CU(cudaMemcpyAsync(start, r->qstart.p, nq * 4, cudaMemcpyDeviceToHost, s));
    if (end) SA_CU(cudaMemcpyAsync(end, r->qend.p, nq * 4, cudaMemcpyDeviceToHost, s));
    if (hit_off) SA_CU(cudaMemcpyAsync(hit_off, r->hitoff.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaMemcpyAsync(r->h_pin, r->hitoff.as<uint64_t>() + nq, 8, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    r->total_hits = r->h_pin[0];
    r->ms_query = elapsed(r->ev[0], r->ev[1]);
    r->ms_hits = elapsed(r->ev[2], r->ev[3]);
    r->have_ranges = true;
    return SA_OK;
}

extern "C" int sa_query_ranges_r(sa_index* idx, sa_result* res, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq,
                                 uint32_t k, uint32_t flags, uint32_t* start, uint32_t* end, uint64_t* hit_off)
{
    SA_REQUIRE(idx && res && res->idx == idx, SA_ERR_INVALID, "result object does not belong to this index");
    SA_REQUIRE(qoff || nq == 0, SA_ERR_INVALID, "qoff is NULL");
    SA_REQUIRE(nq < (1ull << 28), SA_ERR_LIMIT, "too many queries in one batch");
    SA_REQUIRE((flags & ~SA_QUERY_FILE_SEMANTICS) == 0, SA_ERR_INVALID, "sa_query_ranges_r: unknown flag");
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    SA_REQUIRE(idx->built, SA_ERR_INVALID, "index not built");
    std::lock_guard<std::mutex> g(res->mu);
    return ranges_locked(idx, res, qbytes, qoff, nq, k, flags, start, end, hit_off);
}

// ------------------------------------------------------------------------------ hits (+ dedupe)
// Device part: hit positions / record ids of res's ranges in res->hitpos / hitdoc (or, with dedupe,
// compacted in upos / udoc with offsets uoff).  Events 2..3 are (re-)recorded around it.
static int expand_locked(sa_index* idx, sa_result* r, uint32_t flags, bool want_pos)
{
    SA_REQUIRE(r->have_ranges, SA_ERR_INVALID, "no sa_query_ranges result to expand");
    SA_REQUIRE(r->generation == idx->generation, SA_ERR_INVALID, "the index was rebuilt since these ranges were computed");
    SA_CU(cudaSetDevice(idx->device));
    const bool dedupe = (flags & SA_QUERY_DEDUPE) != 0;
    r->have_hits = r->have_records = false;
    r->deduped = dedupe;
    r->out_hits = 0;
    const uint64_t total = r->total_hits;
    if (total == 0 || r->nq == 0) { r->have_hits = true; return SA_OK; }
    cudaStream_t s = r->stream;
    const uint32_t nq = (uint32_t)r->nq;
    if (want_pos) SA_CU(r->hitpos.ensure((size_t)total * 4));
    SA_CU(r->hitdoc.ensure((size_t)total * 4));
    SA_CU(cudaEventRecord(r->ev[2], s));
    // text index: position in the text, record = '\n' bytes in front of it; CSV index: offset in the
    // FILE (what the remapped array holds, engine.c:647-651), record = row of the file (header = 0)
    const bool rows = idx->csv && idx->rowbits.p && idx->rowrank.p;
    uint32_t* d_pos = want_pos ? r->hitpos.as<uint32_t>() : nullptr;
    if (idx->sharded_sa && !rows)
        SA_CU(sa_launch_hits_sharded(idx->view, r->qstart.as<uint32_t>(), r->hitoff.as<uint64_t>(), nq,
                                     idx->nlbits.as<unsigned long long>(), idx->nlrank.as<uint32_t>(), d_pos,
                                     r->hitdoc.as<uint32_t>(), s));
    else
    SA_CU(sa_launch_hits(rows ? idx->sa_file.as<uint32_t>() : idx->sa_final.as<uint32_t>(), r->qstart.as<uint32_t>(),
                         r->hitoff.as<uint64_t>(), nq,
                         rows ? idx->rowbits.as<unsigned long long>() : idx->nlbits.as<unsigned long long>(),
                         rows ? idx->rowrank.as<uint32_t>() : idx->nlrank.as<uint32_t>(), d_pos, r->hitdoc.as<uint32_t>(), s));
    r->launches += 1;
    if (dedupe) {
        SA_CU(r->ucount.ensure((size_t)nq * 4));
        SA_CU(r->uoff.ensure(((size_t)nq + 1) * 8));
        if (want_pos) SA_CU(r->upos.ensure((size_t)total * 4));
        SA_CU(r->udoc.ensure((size_t)total * 4));
        if (r->k > 1024) SA_CU(r->table.ensure((size_t)total * 16));     // lists longer than the shared-memory table
        SA_CU(sa_launch_dedupe_mark(r->hitoff.as<uint64_t>(), nq, d_pos, r->hitdoc.as<uint32_t>(), r->ucount.as<uint32_t>(),
                                    r->table.as<uint32_t>(), s));
        SA_CU(sa_launch_scan_u32_to_u64(r->ucount.as<uint32_t>(), nq, r->uoff.as<uint64_t>(), r->blocksums.as<uint32_t>(), s));
        SA_CU(sa_launch_segment_compact(r->hitoff.as<uint64_t>(), r->uoff.as<uint64_t>(), nq, d_pos, r->hitdoc.as<uint32_t>(),
                                        want_pos ? r->upos.as<uint32_t>() : nullptr, r->udoc.as<uint32_t>(), s));
        r->launches += 5;
        SA_CU(cudaMemcpyAsync(r->h_pin + 1, r->uoff.as<uint64_t>() + nq, 8, cudaMemcpyDeviceToHost, s));
    }
    SA_CU(cudaEventRecord(r->ev[3], s));
    if (dedupe) {
        SA_CU(cudaStreamSynchronize(s));
        r->out_hits = r->h_pin[1];
    } else {
        r->out_hits = total;
    }
    r->have_hits = true;
    return SA_OK;
}

extern "C" int sa_query_hits_r(sa_index* idx, sa_result* res, uint32_t flags, uint32_t* hit_pos, uint32_t* hit_doc,
                               uint64_t cap, uint64_t* hit_off_out)
{
    SA_REQUIRE(idx && res && res->idx == idx, SA_ERR_INVALID, "result object does not belong to this index");
    SA_REQUIRE((flags & ~SA_QUERY_DEDUPE) == 0, SA_ERR_INVALID, "sa_query_hits_r: unknown flag");
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    SA_REQUIRE(idx->built, SA_ERR_INVALID, "index not built");
    std::lock_guard<std::mutex> g(res->mu);
    SA_REQUIRE(res->have_ranges, SA_ERR_INVALID, "no sa_query_ranges result to expand");
    const bool dedupe = (flags & SA_QUERY_DEDUPE) != 0;
    SA_REQUIRE(!dedupe || hit_off_out, SA_ERR_INVALID, "SA_QUERY_DEDUPE needs hit_off_out");
    SA_REQUIRE(cap >= res->total_hits, SA_ERR_INVALID, "cap %llu < %llu hits", (unsigned long long)cap,
               (unsigned long long)res->total_hits);
    int rc = expand_locked(idx, res, flags, hit_pos != nullptr);
    if (rc != SA_OK) return rc;
    const uint64_t nq = res->nq;
    if (res->total_hits == 0 || nq == 0) {
        if (hit_off_out) memset(hit_off_out, 0, (size_t)(res->nq + 1) * 8);
        return SA_OK;
    }
    cudaStream_t s = res->stream;
    const uint64_t out = res->out_hits;
    if (hit_pos && out) SA_CU(cudaMemcpyAsync(hit_pos, dedupe ? res->upos.p : res->hitpos.p, out * 4, cudaMemcpyDeviceToHost, s));
    if (hit_doc && out) SA_CU(cudaMemcpyAsync(hit_doc, dedupe ? res->udoc.p : res->hitdoc.p, out * 4, cudaMemcpyDeviceToHost, s));
    if (hit_off_out) SA_CU(cudaMemcpyAsync(hit_off_out, dedupe ? res->uoff.p : res->hitoff.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    res->ms_hits += elapsed(res->ev[2], res->ev[3]);
    return SA_OK;
}

// ------------------------------------------------------------------------------ CSV row table
// positions of the record ends of the file, from the rank structure (built once, on first use)
static int ensure_rowpos(sa_index* idx, cudaStream_t s)
{
    std::lock_guard<std::mutex> g(idx->rowpos_mu);
    if (idx->rowpos_ready) return SA_OK;
    SA_REQUIRE(idx->csv && idx->rowbits.p && idx->rowrank.p && idx->csv_file_bytes, SA_ERR_INVALID, "no CSV row table");
    const uint64_t nwords = (idx->csv_file_bytes >> 6) + 1;
    unsigned long long last_bits = 0;
    uint32_t last_rank = 0;
    SA_CU(cudaMemcpyAsync(&last_bits, idx->rowbits.as<unsigned long long>() + (nwords - 1), 8, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaMemcpyAsync(&last_rank, idx->rowrank.as<uint32_t>() + (nwords - 1), 4, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    idx->num_row_ends = (uint64_t)last_rank + (uint64_t)__builtin_popcountll(last_bits);
    SA_CU(idx->rowpos.ensure((size_t)(idx->num_row_ends + 1) * 4));
    SA_CU(sa_launch_rank_positions(idx->rowbits.as<unsigned long long>(), idx->rowrank.as<uint32_t>(), nwords,
                                   idx->rowpos.as<uint32_t>(), s));
    SA_CU(cudaStreamSynchronize(s));
    idx->rowpos_ready = true;
    return SA_OK;
}

extern "C" uint64_t sa_num_rows(const sa_index* cidx)
{
    sa_index* idx = const_cast<sa_index*>(cidx);
    if (!idx) return 0;
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    if (!idx->built || !idx->csv) return 0;
    if (cudaSetDevice(idx->device) != cudaSuccess) return 0;
    if (ensure_rowpos(idx, idx->stream) != SA_OK) return 0;
    return idx->num_row_ends + 1;
}

extern "C" int sa_copy_row_starts(sa_index* idx, uint64_t* out, uint64_t cap)
{
    SA_REQUIRE(idx && out, SA_ERR_INVALID, "NULL argument");
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    SA_REQUIRE(idx->built && idx->csv, SA_ERR_INVALID, "no CSV index built");
    SA_CU(cudaSetDevice(idx->device));
    int rc = ensure_rowpos(idx, idx->stream);
    if (rc != SA_OK) return rc;
    const uint64_t ends = idx->num_row_ends;
    SA_REQUIRE(cap >= ends + 2, SA_ERR_INVALID, "cap %llu < num_rows + 1 = %llu", (unsigned long long)cap,
               (unsigned long long)(ends + 2));
    // the u32 positions land in the upper half of the caller's buffer and are widened in place
    uint32_t* tmp = reinterpret_cast<uint32_t*>(out + 1) + ends;
    if (ends) {
        std::lock_guard<std::mutex> g(idx->rowpos_mu);       // idx->stream is shared by the accessors
        SA_CU(cudaMemcpyAsync(tmp, idx->rowpos.p, ends * 4, cudaMemcpyDeviceToHost, idx->stream));
        SA_CU(cudaStreamSynchronize(idx->stream));
    }
    out[0] = 0;
    for (uint64_t i = 0; i < ends; ++i) out[i + 1] = (uint64_t)tmp[i] + 1;
    out[ends + 1] = idx->csv_file_bytes;
    return SA_OK;
}

// ------------------------------------------------------------------------------ records
extern "C" int sa_query_records_size(sa_index* idx, sa_result* res, uint32_t flags, uint64_t* num_records,
                                     uint64_t* num_bytes)
{
    SA_REQUIRE(idx && res && res->idx == idx, SA_ERR_INVALID, "result object does not belong to this index");
    SA_REQUIRE((flags & ~(SA_QUERY_DEDUPE | SA_RECORDS_REFERENCE_COMPAT)) == 0, SA_ERR_INVALID, "sa_query_records_size: unknown flag");
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    SA_REQUIRE(idx->built, SA_ERR_INVALID, "index not built");
    std::lock_guard<std::mutex> g(res->mu);
    if (idx->csv)
        SA_REQUIRE(idx->csv_file.p && idx->csv_file_bytes, SA_ERR_INVALID,
                   "the CSV file is not resident on the device (index restored by sa_load: call sa_attach_csv_file)");
    int rc = expand_locked(idx, res, flags & SA_QUERY_DEDUPE, true);
    if (rc != SA_OK) return rc;
    const uint64_t num = res->out_hits;
    res->rec_bytes = 0;
    if (num_records) *num_records = num;
    if (num_bytes) *num_bytes = 0;
    if (num == 0) { res->have_records = true; return SA_OK; }
    SA_REQUIRE(num < 0xFFFFFFFFull, SA_ERR_LIMIT, "%llu records in one batch: split the batch", (unsigned long long)num);
    cudaStream_t s = res->stream;
    if (idx->csv) {
        rc = ensure_rowpos(idx, s);
        if (rc != SA_OK) return rc;
    }
    SA_CU(res->reclen.ensure((size_t)num * 4));
    SA_CU(res->recsrc.ensure((size_t)num * 8));
    SA_CU(res->recoff.ensure(((size_t)num + 1) * 8));
    SA_CU(res->blocksums.ensure(((size_t)(num + 1023) / 1024 + 2) * 8));
    const bool dd = res->deduped;
    const uint32_t* d_pos = dd ? res->upos.as<uint32_t>() : res->hitpos.as<uint32_t>();
    const uint32_t* d_doc = dd ? res->udoc.as<uint32_t>() : res->hitdoc.as<uint32_t>();
    SA_CU(cudaEventRecord(res->ev[2], s));
    if (idx->csv)
        SA_CU(sa_launch_record_bounds(d_pos, d_doc, num, idx->rowpos.as<uint32_t>(), idx->num_row_ends, idx->csv_file_bytes,
                                      (flags & SA_RECORDS_REFERENCE_COMPAT) != 0, res->recsrc.as<unsigned long long>(),
                                      res->reclen.as<uint32_t>(), s));
    else
        SA_CU(sa_launch_record_bounds(d_pos, d_doc, num, idx->nlpos.as<uint32_t>(), idx->num_nl, idx->n, false,
                                      res->recsrc.as<unsigned long long>(), res->reclen.as<uint32_t>(), s));
    SA_CU(sa_launch_scan_u32_to_u64(res->reclen.as<uint32_t>(), (uint32_t)num, res->recoff.as<uint64_t>(),
                                    res->blocksums.as<uint32_t>(), s));
    SA_CU(cudaEventRecord(res->ev[3], s));
    res->launches += 4;
    SA_CU(cudaMemcpyAsync(res->h_pin + 2, res->recoff.as<uint64_t>() + num, 8, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    res->rec_bytes = res->h_pin[2];
    res->ms_hits += elapsed(res->ev[2], res->ev[3]);
    if (num_bytes) *num_bytes = res->rec_bytes;
    res->have_records = true;
    return SA_OK;
}

extern "C" int sa_query_records(sa_index* idx, sa_result* res, uint8_t* out, uint64_t cap_bytes, uint64_t* rec_off,
                                uint64_t cap_records, uint64_t* hit_off_out, uint32_t* hit_doc)
{
    SA_REQUIRE(idx && res && res->idx == idx, SA_ERR_INVALID, "result object does not belong to this index");
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    SA_REQUIRE(idx->built, SA_ERR_INVALID, "index not built");
    std::lock_guard<std::mutex> g(res->mu);
    SA_REQUIRE(res->have_records && res->generation == idx->generation, SA_ERR_INVALID,
               "call sa_query_records_size first");
    const uint64_t num = res->out_hits, bytes = res->rec_bytes, nq = res->nq;
    SA_REQUIRE(cap_bytes >= bytes, SA_ERR_INVALID, "cap_bytes %llu < %llu", (unsigned long long)cap_bytes, (unsigned long long)bytes);
    SA_REQUIRE(!rec_off || cap_records >= num, SA_ERR_INVALID, "cap_records %llu < %llu", (unsigned long long)cap_records,
               (unsigned long long)num);
    SA_CU(cudaSetDevice(idx->device));
    if (num == 0) {
        if (rec_off) rec_off[0] = 0;
        if (hit_off_out) memset(hit_off_out, 0, (size_t)(nq + 1) * 8);
        return SA_OK;
    }
    cudaStream_t s = res->stream;
    const bool dd = res->deduped;
    if (bytes) {
        SA_REQUIRE(out, SA_ERR_INVALID, "out is NULL");
        SA_CU(res->recbytes.ensure((size_t)bytes));
        SA_CU(cudaEventRecord(res->ev[2], s));
        SA_CU(sa_launch_record_gather(idx->csv ? idx->csv_file.as<uint8_t>() : idx->text.as<uint8_t>(),
                                      res->recsrc.as<unsigned long long>(), res->recoff.as<uint64_t>(), num,
                                      res->recbytes.as<uint8_t>(), s));
        SA_CU(cudaEventRecord(res->ev[3], s));
        res->launches += 1;
        SA_CU(cudaMemcpyAsync(out, res->recbytes.p, bytes, cudaMemcpyDeviceToHost, s));
    }
    if (rec_off) SA_CU(cudaMemcpyAsync(rec_off, res->recoff.p, (num + 1) * 8, cudaMemcpyDeviceToHost, s));
    if (hit_off_out) SA_CU(cudaMemcpyAsync(hit_off_out, dd ? res->uoff.p : res->hitoff.p, (nq + 1) * 8, cudaMemcpyDeviceToHost, s));
    if (hit_doc) SA_CU(cudaMemcpyAsync(hit_doc, dd ? res->udoc.p : res->hitdoc.p, num * 4, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    if (bytes) res->ms_hits += elapsed(res->ev[2], res->ev[3]);
    return SA_OK;
}

extern "C" int sa_result_stats(const sa_result* res, float* ms_query, float* ms_hits, uint64_t* kernel_launches)
{
    SA_REQUIRE(res, SA_ERR_INVALID, "NULL argument");
    if (ms_query) *ms_query = res->ms_query;
    if (ms_hits) *ms_hits = res->ms_hits;
    if (kernel_launches) *kernel_launches = res->launches;
    return SA_OK;
}

// ------------------------------------------------------------------------------ one batch at a time per index
static void mirror_stats(sa_index* idx, const sa_result* r)
{
    idx->stats.ms_query = r->ms_query;
    idx->stats.ms_hits = r->ms_hits;
    idx->stats.kernel_launches = r->launches;
}

extern "C" int sa_query_ranges(sa_index* idx, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq, uint32_t k,
                               uint32_t* start, uint32_t* end, uint64_t* hit_off)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    sa_result* r = default_result(idx);
    if (!r) return SA_ERR_CUDA;
    int rc = sa_query_ranges_r(idx, r, qbytes, qoff, nq, k, 0, start, end, hit_off);
    if (rc == SA_OK) mirror_stats(idx, r);
    return rc;
}

extern "C" int sa_query_hits(sa_index* idx, uint32_t* hit_pos, uint32_t* hit_doc, uint64_t cap)
{
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    sa_result* r = default_result(idx);
    if (!r) return SA_ERR_CUDA;
    int rc = sa_query_hits_r(idx, r, 0, hit_pos, hit_doc, cap, nullptr);
    if (rc == SA_OK) mirror_stats(idx, r);
    return rc;
}

// Device-resident variant for callers that keep batches in HBM (bench `value` path): ranges only.
extern "C" int sa_query_ranges_device(sa_index* idx, const uint8_t* d_qbytes, const uint64_t* d_qoff, uint64_t nq,
                                      uint64_t qbytes_len, uint32_t k, uint32_t* d_start, uint32_t* d_end)
{
    (void)qbytes_len; (void)k;
    SA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    SA_REQUIRE(d_qbytes && d_qoff && d_start && d_end, SA_ERR_INVALID, "NULL device pointer");
    SA_REQUIRE(nq < (1ull << 28), SA_ERR_LIMIT, "too many queries in one batch");
    sa_result* r = default_result(idx);
    if (!r) return SA_ERR_CUDA;
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    SA_REQUIRE(idx->built, SA_ERR_INVALID, "index not built");
    SA_REQUIRE(idx->n > 0, SA_ERR_INVALID, "empty index");
    SA_REQUIRE(!idx->sharded_sa, SA_ERR_INVALID, "sa_query_ranges_device: not available on a SA_FLAG_SHARDED_SA index");
    std::lock_guard<std::mutex> g(r->mu);
    SA_CU(cudaSetDevice(idx->device));
    r->have_ranges = r->have_hits = r->have_records = false;
    r->launches = 1;
    const bool thread_kernel = idx->dbg_query_kernel == 1 || (idx->dbg_query_kernel == 0 && nq >= 4096);
    if (thread_kernel) {
        int rc = ensure_qtop(idx, r->stream);
        if (rc != SA_OK) return rc;
    }
    SA_CU(cudaEventRecord(r->ev[0], r->stream));
    SA_CU(sa_launch_query(idx->text.as<uint8_t>(), idx->sa_final.as<uint32_t>(), (uint32_t)idx->n, idx->L, d_qbytes, d_qoff,
                          (uint32_t)nq, d_start, d_end, (int)idx->dbg_query_kernel, r->stream,
                          thread_kernel && idx->qtop_levels ? idx->qtop.p : nullptr, idx->qtop_levels));
    SA_CU(cudaEventRecord(r->ev[1], r->stream));
    SA_CU(cudaStreamSynchronize(r->stream));
    r->ms_query = elapsed(r->ev[0], r->ev[1]);
    r->ms_hits = 0.f;
    mirror_stats(idx, r);
    return SA_OK;
}
