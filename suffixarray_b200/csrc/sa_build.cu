// sa_build.cu -- lifecycle and the build pipeline of libsa_cuda.so.
//
// Replaces construct_truncated_suffix_array (engine.c:837-866): identity fill + recursive
// bucket sort become   pack -> onesweep LSD sort (round 0) -> tie refinement rounds -> finish.
#include <stdarg.h>
#include <string.h>

#include "sa_index.h"

static thread_local char g_err[512] = "";

void sa_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int sa_cuda_fail(cudaError_t e, const char* what)
{
    sa_set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return e == cudaErrorMemoryAllocation ? SA_ERR_OOM : SA_ERR_CUDA;
}

// ------------------------------------------------------------------------------ timing
cudaEvent_t sa_take_event(sa_index* idx)
{
    if (idx->events_used == idx->event_pool.size()) {
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        idx->event_pool.push_back(e);
    }
    return idx->event_pool[idx->events_used++];
}

void sa_time_collect(sa_index* idx)
{
    if (!idx->timing) return;
    cudaStreamSynchronize(idx->stream);
    for (const PhaseTimer& t : idx->timers) {
        if (!t.target) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) *t.target += ms;
    }
    idx->timers.clear();
    idx->events_used = 0;
}

// ------------------------------------------------------------------------------ lifecycle
extern "C" const char* sa_last_error(void) { return g_err; }
extern "C" int sa_abi_version(void) { return SA_CUDA_ABI_VERSION; }

extern "C" int sa_create(sa_index** out, uint32_t max_suffix_length, int device)
{
    if (!out) { sa_set_error("sa_create: out is NULL"); return SA_ERR_INVALID; }
    *out = nullptr;
    if (max_suffix_length > 4096) { sa_set_error("max_suffix_length %u > 4096", max_suffix_length); return SA_ERR_LIMIT; }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        sa_set_error("no CUDA device (%s): libsa_cuda has no CPU fallback", cudaGetErrorString(e));
        return SA_ERR_CUDA;
    }
    if (device < 0 || device >= count) { sa_set_error("device %d out of range (%d devices)", device, count); return SA_ERR_INVALID; }
    SA_CU(cudaSetDevice(device));
    sa_index* idx = new (std::nothrow) sa_index();
    if (!idx) { sa_set_error("host allocation failed"); return SA_ERR_OOM; }
    idx->device = device;
    idx->L = max_suffix_length;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) idx->num_sms = prop.multiProcessorCount;
    e = cudaStreamCreateWithFlags(&idx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&idx->h_small), H_SMALL_WORDS * sizeof(uint32_t));
    if (e == cudaSuccess) e = idx->small.ensure(SM_WORDS * sizeof(uint32_t));
    if (e != cudaSuccess) { sa_destroy(idx); return sa_cuda_fail(e, "sa_create"); }
    *out = idx;
    return SA_OK;
}

extern "C" void sa_destroy(sa_index* idx)
{
    if (!idx) return;
    cudaSetDevice(idx->device);
    if (idx->stream) cudaStreamSynchronize(idx->stream);
    if (idx->default_res) { sa_result_destroy(idx->default_res); idx->default_res = nullptr; }
    DevBuf* all[] = {&idx->text, &idx->sa_final, &idx->nlpos, &idx->sa_file, &idx->qbits, &idx->map,
                     &idx->quoted_text, &idx->keys[0], &idx->keys[1], &idx->vals[0], &idx->vals[1],
                     &idx->aux[0], &idx->aux[1], &idx->lcp, &idx->tie, &idx->amask, &idx->hmask, &idx->counts,
                     &idx->status, &idx->small, &idx->nl_counts, &idx->codes, &idx->rowpos, &idx->qtop, &idx->pack_keys, &idx->pack_pos, &idx->hist16,
                     &idx->ext_sa, &idx->ext_lcp, &idx->ext_out, &idx->split_table, &idx->split_offsets,
                     &idx->csv_file, &idx->csv_tq, &idx->csv_td, &idx->csv_te, &idx->csv_tr, &idx->rowbits, &idx->rowrank,
                     &idx->nlbits, &idx->nlrank};
    for (DevBuf* b : all) b->release();
    for (int i = 0; i < 16; ++i) {
        if (idx->copy_streams[i]) cudaStreamDestroy(idx->copy_streams[i]);
        if (idx->chunk_ev[i]) cudaEventDestroy(idx->chunk_ev[i]);
        if (idx->ext_ev[i]) cudaEventDestroy(idx->ext_ev[i]);
    }
    if (idx->ext_ev[16]) cudaEventDestroy(idx->ext_ev[16]);
    for (int i = 0; i < 2; ++i) {
        if (idx->h_ring[i]) cudaFreeHost(idx->h_ring[i]);
        if (idx->ring_ev[i]) cudaEventDestroy(idx->ring_ev[i]);
    }
    for (int g = 0; g < 16; ++g)
        if (idx->peer_mapped[g]) { cudaIpcCloseMemHandle(idx->peer_keys[g]); cudaIpcCloseMemHandle(idx->peer_pos[g]); }
    for (int g = 0; g < 16; ++g)
        if (idx->peer_final_mapped[g]) cudaIpcCloseMemHandle(idx->peer_final[g]);
    for (int g = 0; g < 16; ++g) {
        if (idx->peer_text[g].mapped) cudaIpcCloseMemHandle(idx->peer_text[g].ptr);
        if (idx->peer_nlbits[g].mapped) cudaIpcCloseMemHandle(idx->peer_nlbits[g].ptr);
        if (idx->peer_nlrank[g].mapped) cudaIpcCloseMemHandle(idx->peer_nlrank[g].ptr);
    }
    if (idx->h_big) cudaFreeHost(idx->h_big);
    if (idx->push_stream) cudaStreamDestroy(idx->push_stream);
    for (cudaEvent_t e : idx->event_pool) cudaEventDestroy(e);
    if (idx->h_small) cudaFreeHost(idx->h_small);
    if (idx->stream) cudaStreamDestroy(idx->stream);
    delete idx;
}

// ------------------------------------------------------------------------------ radix driver
struct SortBufs {
    uint64_t* k[2];
    uint32_t* v[2];
    int cur;
};

// Stable LSD sort of m records on key bits [low_bit, low_bit + nbits).  identity: values are the
// record indices (generated by the first executed pass).  Passes whose digit is constant are
// skipped (one tiny read-back).
static int radix_sort(sa_index* idx, SortBufs& sb, uint32_t m, int low_bit, int nbits, bool identity)
{
    cudaStream_t s = idx->stream;
    uint32_t* sm = idx->small.as<uint32_t>();
    const int npasses = (nbits + 7) / 8;
    if (npasses < 1 || npasses > 8) { sa_set_error("radix_sort: bad pass count %d", npasses); return SA_ERR_INTERNAL; }
    {
        ScopedPhase ph(idx, &idx->stats.ms_hist);
        SA_CU(sa_launch_hist(sb.k[sb.cur], m, low_bit, npasses, sm + SM_HIST, idx->num_sms, s));
        SA_CU(sa_launch_scan_hist(sm + SM_HIST, m, npasses, sm + SM_BASE, sm + SM_SKIP, s));
        idx->stats.kernel_launches += 2;
    }
    SA_CU(cudaMemcpyAsync(idx->h_small, sm + SM_SKIP, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    uint32_t skip[8];
    memcpy(skip, idx->h_small, sizeof skip);
    for (int p = 0; p < npasses; ++p) {
        if (skip[p]) continue;
        {
            ScopedPhase ph(idx, &idx->stats.ms_onesweep_sum);
            SA_CU(sa_launch_onesweep(sb.k[sb.cur], sb.v[sb.cur], sb.k[sb.cur ^ 1], sb.v[sb.cur ^ 1], m,
                                     low_bit + 8 * p, sm + SM_BASE + p * 256, idx->status.as<uint32_t>(),
                                     sm + SM_TILECTR, sm + SM_ERR, identity, s));
        }
        identity = false;
        sb.cur ^= 1;
        idx->stats.radix_passes += 1;
        idx->stats.kernel_launches += 1;
        idx->stats.onesweep_elems += m;
    }
    if (identity) {   // every pass skipped: values were never generated
        SA_CU(sa_launch_iota(sb.v[sb.cur], m, s));
        idx->stats.kernel_launches += 1;
    }
    return SA_OK;
}

static int check_device_error(sa_index* idx)
{
    uint32_t* sm = idx->small.as<uint32_t>();
    SA_CU(cudaMemcpyAsync(idx->h_small, sm + SM_ERR, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, idx->stream));
    SA_CU(cudaStreamSynchronize(idx->stream));
    const uint32_t err = idx->h_small[0];
    unsigned long long runs;
    memcpy(&runs, idx->h_small + 1, sizeof runs);
    idx->stats.finish_runs = runs;
    if (err) {
        sa_set_error("device-side guard tripped (flags 0x%x): %s%s", err,
                     (err & SA_DEVERR_LOOKBACK_TIMEOUT) ? "onesweep look-back timed out; " : "",
                     (err & SA_DEVERR_RUN_TOO_LONG) ? "finish saw a bucket longer than 32" : "");
        return SA_ERR_INTERNAL;
    }
    return SA_OK;
}

// ------------------------------------------------------------------------------ build stages
// Stage 1 (text resident in idx->text): lower-casing, record table, alphabet / code array.
// Leaves idx->code, idx->cb0, idx->max_chars set.
int sa_stage_prepare(sa_index* idx, uint32_t flags)
{
    cudaStream_t s = idx->stream;
    const uint32_t n = (uint32_t)idx->n;
    sa_stats& st = idx->stats;
    uint32_t* sm = idx->small.as<uint32_t>();
    SA_CU(cudaMemsetAsync(sm + SM_ERR, 0, 4 * sizeof(uint32_t), s));       // err + run counter

    if (flags & SA_FLAG_LOWERCASE) {
        SA_CU(sa_launch_lowercase(idx->text.as<uint8_t>(), (uint64_t)n + idx->tail, s));
        st.kernel_launches += 1;
    }
    // document table: positions of '\n' (the reference scans for them per hit, engine.c:1189-1191)
    {
        ScopedPhase ph(idx, &st.ms_docs);
        const uint32_t tiles = sa_newline_tiles(n);
        SA_CU(idx->nl_counts.ensure((size_t)tiles * sizeof(uint32_t)));
        SA_CU(sa_launch_newline_count(idx->text.as<uint8_t>(), n, idx->nl_counts.as<uint32_t>(), s));
        SA_CU(sa_launch_scan_u32(idx->nl_counts.as<uint32_t>(), tiles, sm + SM_NLTOTAL, s));
        SA_CU(cudaMemcpyAsync(idx->h_small + 8, sm + SM_NLTOTAL, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        // which byte values occur (decides the code width); read back with the newline total
        SA_CU(sa_launch_byte_hist(idx->text.as<uint8_t>(), (uint64_t)n + idx->tail, sm + SM_BYTEHIST, idx->num_sms, s));
        SA_CU(cudaMemcpyAsync(idx->h_small + 256, sm + SM_BYTEHIST, 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        st.kernel_launches += 1;
        SA_CU(cudaStreamSynchronize(s));
        idx->num_nl = idx->h_small[8];
        SA_CU(idx->nlpos.ensure((size_t)idx->num_nl * sizeof(uint32_t)));
        SA_CU(idx->nlbits.ensure(((size_t)(n >> 6) + 1) * sizeof(uint64_t)));
        SA_CU(idx->nlrank.ensure(((size_t)(n >> 6) + 1) * sizeof(uint32_t)));
        SA_CU(sa_launch_newline_write(idx->text.as<uint8_t>(), n, idx->nl_counts.as<uint32_t>(),
                                      idx->nlpos.as<uint32_t>(), idx->nlbits.as<unsigned long long>(),
                                      idx->nlrank.as<uint32_t>(), s));
        st.kernel_launches += 3;
        st.num_docs = (uint64_t)idx->num_nl + 1;
    }
    return sa_stage_alphabet(idx, idx->h_small + 256);
}

// Alphabet: dense order-preserving codes when <= 128 byte values occur (NUL is always code 0).
// present[256]: which byte values occur in the text (host).  Encodes idx->text into idx->codes and
// leaves idx->code, idx->cb0, idx->max_chars set.
int sa_stage_alphabet(sa_index* idx, const uint32_t* present, uint64_t enc_lo, uint64_t enc_hi)
{
    cudaStream_t s = idx->stream;
    const uint32_t n = (uint32_t)idx->n;
    sa_stats& st = idx->stats;
    uint32_t* sm = idx->small.as<uint32_t>();
    sa::CodeInfo ci{8, 0x100u, 0x100u};
    if (idx->lprime) {
        ScopedPhase ph(idx, &st.ms_pack);
        uint8_t* table = reinterpret_cast<uint8_t*>(idx->h_small + 512);
        uint32_t next = 1;
        table[0] = 0;
        for (int b = 1; b < 256; ++b) table[b] = present[b] ? (uint8_t)(next++) : 0;
        int bits = 1;
        while ((1u << bits) < next) ++bits;
        if (idx->dbg_code_bits >= 1 && idx->dbg_code_bits > bits) bits = (int)idx->dbg_code_bits;
        if (bits > 8) bits = 8;
        if (bits < 8) {
            ci.bits = bits;
            ci.nl = present['\n'] ? table['\n'] : 0x100u;
            ci.ff = present[255] ? table[255] : 0x100u;
            const size_t tbytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
            SA_CU(idx->codes.ensure(tbytes));
            uint8_t* d_table = reinterpret_cast<uint8_t*>(sm + SM_TABLE);
            SA_CU(cudaMemcpyAsync(d_table, table, 256, cudaMemcpyHostToDevice, s));
            // [enc_lo, enc_hi) (16-byte aligned; default: everything) -- the sharded build encodes its own
            // range first and the rest once the peers' slices have landed (sa_stage_encode_range)
            if (enc_hi > tbytes) enc_hi = tbytes;
            if (enc_lo < enc_hi) {
                SA_CU(sa_launch_encode(idx->text.as<uint8_t>() + enc_lo, enc_hi - enc_lo, d_table, idx->codes.as<uint8_t>() + enc_lo, s));
                st.kernel_launches += 1;
            }
        }
    }
    idx->code = ci;
    st.code_bits = (uint32_t)ci.bits;
    idx->max_chars = ci.bits == 8 ? 8 : (63 / ci.bits > 16 ? 16 : 63 / ci.bits);
    idx->cb0 = idx->lprime < (uint32_t)idx->max_chars ? (int)idx->lprime : idx->max_chars;
    if (idx->dbg_sort0_chars >= 1 && idx->dbg_sort0_chars < idx->cb0) idx->cb0 = (int)idx->dbg_sort0_chars;
    return SA_OK;
}

// Encode text[lo, hi) (16-byte aligned bounds, hi clipped to the padded size) with the table sa_stage_alphabet left
// on the device; a no-op for 8-bit alphabets (the text is its own key source).
int sa_stage_encode_range(sa_index* idx, uint64_t lo, uint64_t hi)
{
    if (idx->code.bits == 8 || !idx->lprime) return SA_OK;
    const size_t tbytes = ((size_t)idx->n + SA_TEXT_PAD + 255) & ~(size_t)255;
    if (hi > tbytes) hi = tbytes;
    if (lo >= hi) return SA_OK;
    const uint8_t* d_table = reinterpret_cast<const uint8_t*>(idx->small.as<uint32_t>() + SM_TABLE);
    SA_CU(sa_launch_encode(idx->text.as<uint8_t>() + lo, hi - lo, d_table, idx->codes.as<uint8_t>() + lo, idx->stream));
    idx->stats.kernel_launches += 1;
    return SA_OK;
}

const uint8_t* sa_key_source(const sa_index* idx)
{
    return idx->code.bits == 8 ? idx->text.as<uint8_t>() : idx->codes.as<uint8_t>();
}

// scratch for sorting / refining m records
int sa_ensure_scratch(sa_index* idx, uint32_t m)
{
    const size_t m8 = (size_t)m * 8, m4 = (size_t)m * 4;
    for (int i = 0; i < 2; ++i) {
        SA_CU(idx->keys[i].ensure(m8));
        SA_CU(idx->vals[i].ensure(m4));
        SA_CU(idx->aux[i].ensure(m4));
    }
    SA_CU(idx->lcp.ensure((size_t)m * 2 + 64));
    const uint32_t nwords_data = (m + 31) / 32;
    SA_CU(idx->tie.ensure((size_t)(nwords_data + 2) * 4));
    SA_CU(idx->amask.ensure((size_t)nwords_data * 4));
    SA_CU(idx->hmask.ensure((size_t)nwords_data * 4));
    const uint32_t scan_tiles = (nwords_data + SA_SCAN_TILE / 32 - 1) / (SA_SCAN_TILE / 32);
    SA_CU(idx->counts.ensure((size_t)scan_tiles * 2 * sizeof(uint32_t)));
    const uint32_t os_tiles = (m + sa_onesweep_tile_items() - 1) / sa_onesweep_tile_items();
    SA_CU(idx->status.ensure((size_t)os_tiles * 256 * sizeof(uint32_t)));
    return SA_OK;
}

// Stage 2: m records (round-0 keys in keys[0]; positions in vals[0] unless `identity`, in which
// case record i is suffix i) -> sorted by the masked L'-byte prefix as far as ORDER-E needs it:
// *sa_sorted = positions in prefix order, idx->lcp = adjacent LCPs (u16, bytes).
int sa_stage_sort_refine(sa_index* idx, uint32_t m, bool identity, uint32_t** sa_sorted)
{
    cudaStream_t s = idx->stream;
    const uint32_t lprime = idx->lprime;
    sa_stats& st = idx->stats;
    uint32_t* sm = idx->small.as<uint32_t>();
    const sa::CodeInfo ci = idx->code;
    const int cb0 = idx->cb0, max_chars = idx->max_chars;
    const uint8_t* src = sa_key_source(idx);
    const uint32_t nwords_data = (m + 31) / 32;
    const uint32_t tie_words = nwords_data + 2;
    const uint32_t scan_tiles = (nwords_data + SA_SCAN_TILE / 32 - 1) / (SA_SCAN_TILE / 32);

    SortBufs sb0{{idx->keys[0].as<uint64_t>(), idx->keys[1].as<uint64_t>()},
                 {idx->vals[0].as<uint32_t>(), idx->vals[1].as<uint32_t>()}, 0};
    {
        ScopedPhase ph(idx, &st.ms_sort0);
        int rc = radix_sort(idx, sb0, m, 64 - ci.bits * cb0, ci.bits * cb0, identity);
        if (rc != SA_OK) return rc;
    }
    uint32_t* sa = sb0.v[sb0.cur];
    uint16_t* lcp = idx->lcp.as<uint16_t>();
    uint32_t depth = (uint32_t)cb0;
    {
        ScopedPhase ph(idx, &st.ms_refine);
        SA_CU(sa_launch_lcp0(sb0.k[sb0.cur], m, cb0, ci.bits, lprime, lcp, idx->tie.as<uint32_t>(), tie_words, s));
        st.kernel_launches += 1;

        // ---- refinement rounds: tied groups of more than 32 members, next cb bytes
        uint32_t* cv_a = sb0.v[sb0.cur ^ 1];
        while (depth < lprime && (idx->dbg_max_rounds < 0 || (int64_t)st.rounds < idx->dbg_max_rounds)) {
            // the tie mask for this depth was left by lcp0 / the previous round's write-back
            SA_CU(sa_launch_group_flags(idx->tie.as<uint32_t>(), nwords_data, idx->amask.as<uint32_t>(),
                                        idx->hmask.as<uint32_t>(), idx->counts.as<uint32_t>(), s));
            SA_CU(sa_launch_scan_pairs(idx->counts.as<uint32_t>(), scan_tiles, sm + SM_TOTALS, s));
            SA_CU(cudaMemcpyAsync(idx->h_small + 16, sm + SM_TOTALS, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
            SA_CU(cudaStreamSynchronize(s));
            st.kernel_launches += 2;
            const uint32_t ma = idx->h_small[16], g = idx->h_small[17];
            if (ma == 0) break;
            int gbits = 1;
            while (gbits < 32 && (1ull << gbits) < (uint64_t)g) ++gbits;
            if (idx->dbg_force_gbits > gbits) gbits = (int)idx->dbg_force_gbits;
            int cb = ci.bits == 8 ? (64 - gbits) / 8 : (63 - gbits) / ci.bits;
            if (cb > max_chars) cb = max_chars;
            if ((uint32_t)cb > lprime - depth) cb = (int)(lprime - depth);
            if (cb < 1) { sa_set_error("refine: no key bits left (gbits %d)", gbits); return SA_ERR_INTERNAL; }
            SortBufs sb{{idx->keys[0].as<uint64_t>(), idx->keys[1].as<uint64_t>()},
                        {cv_a, idx->aux[0].as<uint32_t>()}, 0};
            uint32_t* slots = idx->aux[1].as<uint32_t>();
            SA_CU(sa_launch_compact_gather(idx->amask.as<uint32_t>(), idx->hmask.as<uint32_t>(), nwords_data,
                                           idx->counts.as<uint32_t>(), sa, src, depth, cb, gbits, ci, slots, sb.k[0],
                                           sb.v[0], s));
            st.kernel_launches += 1;
            int rc = radix_sort(idx, sb, ma, 64 - gbits - ci.bits * cb, gbits + ci.bits * cb, false);
            if (rc != SA_OK) return rc;
            SA_CU(sa_launch_writeback(slots, sb.k[sb.cur], sb.v[sb.cur], ma, depth, cb, gbits, ci.bits, lprime, sa, lcp,
                                      idx->tie.as<uint32_t>(), tie_words, s));
            st.kernel_launches += 1;
            st.refined_elems += ma;
            st.rounds += 1;
            depth += (uint32_t)cb;
        }
    }
    *sa_sorted = sa;
    return SA_OK;
}

// Stage 3: <= 32 buckets, 17-byte window order.  Ranks [r_begin, r_end) of the m-entry arrays are
// placed (the rest are halo records of a shard, see sa_shard.cu).
int sa_stage_finish(sa_index* idx, const uint32_t* sa, const uint16_t* lcp, uint32_t m, uint32_t r_begin,
                    uint32_t r_end, uint32_t* out)
{
    sa::FinishOut fo{};
    fo.out[0] = out;
    fo.n = 1;
    fo.lo = 0;
    fo.hi = m;
    return sa_stage_finish_to(idx, sa, lcp, m, r_begin, r_end, fo);
}

int sa_stage_finish_to(sa_index* idx, const uint32_t* sa, const uint16_t* lcp, uint32_t m, uint32_t r_begin,
                       uint32_t r_end, const sa::FinishOut& outs, uint32_t block0, uint32_t nblocks)
{
    uint32_t* out = outs.out[0];
    cudaStream_t s = idx->stream;
    sa_stats& st = idx->stats;
    uint32_t* sm = idx->small.as<uint32_t>();
    ScopedPhase ph(idx, &st.ms_finish);
    if (idx->dbg_skip_finish) {
        SA_CU(cudaMemcpyAsync(out, sa, (size_t)m * 4, cudaMemcpyDeviceToDevice, s));
    } else {
        SA_CU(sa_launch_finish(sa, lcp, idx->text.as<uint8_t>(), m, idx->lprime, r_begin, r_end, outs,
                               reinterpret_cast<unsigned long long*>(sm + SM_RUNCTR), sm + SM_ERR, s, block0, nblocks));
        st.kernel_launches += 1;
    }
    return SA_OK;
}

static int build_pipeline(sa_index* idx, uint32_t flags)
{
    cudaStream_t s = idx->stream;
    const uint32_t n = (uint32_t)idx->n;
    sa_stats& st = idx->stats;
    int rc = sa_stage_prepare(idx, flags);
    if (rc != SA_OK) return rc;
    SA_CU(idx->sa_final.ensure((size_t)n * sizeof(uint32_t)));
    if (idx->lprime == 0) {       // depth 0 == max depth: the identity (engine.c:706)
        SA_CU(sa_launch_iota(idx->sa_final.as<uint32_t>(), n, s));
        return SA_OK;
    }
    rc = sa_ensure_scratch(idx, n);
    if (rc != SA_OK) return rc;
    // ---- round 0 keys: every suffix, bytes [0, cb0)
    {
        ScopedPhase ph(idx, &st.ms_pack);
        SA_CU(sa_launch_pack0(sa_key_source(idx), n, idx->cb0, idx->code, idx->keys[0].as<uint64_t>(), s));
        st.kernel_launches += 1;
    }
    uint32_t* sa = nullptr;
    rc = sa_stage_sort_refine(idx, n, true, &sa);
    if (rc != SA_OK) return rc;
    return sa_stage_finish(idx, sa, idx->lcp.as<uint16_t>(), n, 0, n, idx->sa_final.as<uint32_t>());
}

int sa_check_device_error(sa_index* idx) { return check_device_error(idx); }

// idx->text holds n bytes followed by zero padding.
int sa_build_resident(sa_index* idx, uint32_t flags)
{
    const uint32_t n = (uint32_t)idx->n;
    sa_stats& st = idx->stats;
    const float h2d = st.ms_h2d;
    memset(&st, 0, sizeof st);
    st.ms_h2d = h2d;
    idx->lprime = idx->L < n ? idx->L : n;                    // engine.c:841
    st.n = n; st.max_suffix_length = idx->L; st.lprime = idx->lprime;
    idx->built = false;
    sa_invalidate_results(idx);
    idx->csv = false;
    idx->shard_prepared = false;
    idx->shard_csv = false;
    if (n == 0) { idx->num_nl = 0; st.num_docs = 1; idx->built = true; return SA_OK; }
    int rc;
    {
        ScopedPhase total(idx, &st.ms_total);
        rc = build_pipeline(idx, flags);
    }
    if (rc == SA_OK) rc = check_device_error(idx);
    sa_time_collect(idx);
    if (rc != SA_OK) return rc;
    idx->built = true;
    return SA_OK;
}
