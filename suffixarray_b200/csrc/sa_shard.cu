// sa_shard.cu -- building ONE suffix array over several GPUs as a sample sort.
//
// One sa_index per GPU (one process per GPU); the text is replicated.  Rank g packs the round-0
// keys of its slice of text positions, the ranks agree on splitters from the all-reduced
// histogram of the keys' top 16 bits (so equal keys -- and therefore whole tied groups -- always
// land on one rank), records are split by destination (stable, so every run stays in ascending
// position order), exchanged with one all-to-all over NVLink (NCCL; done by the host side between
// sa_shard_split and sa_shard_sort), and every rank then runs the ordinary pipeline on what it
// received: LSD radix sort, tie refinement, and the <= 32 bucket finish -- the latter on its
// records plus a 128-record halo from each neighbour, because ORDER-E's bucket depths d* depend on
// the 32 ranks either side (sa_stage_finish places only the ranks it owns).  The final array is
// the concatenation of the per-rank slices.
//
// The reference has no counterpart: its only split is the sequential 2 GiB partitioning into
// independent indexes (suffix_array.pyx:148-180).
#include <string.h>

#include "sa_index.h"

namespace sa {

__global__ void iota_offset_kernel(uint32_t* __restrict__ out, uint32_t m, uint32_t first)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) out[i] = first + i;
}

// top-16-bit histogram of a sample of the keys (global reductions: the bins are far too many for shared
// memory).  Every 16th key of small slices, every 64th of slices of 16 M records and more (1.5 M samples
// of a 100 M-record slice still balance the shards to a fraction of a percent); lanes that hit the same
// bin are merged into one atomic (text keys are skewed: a handful of bins take most of the samples).
__global__ void hist16_kernel(const uint64_t* __restrict__ keys, uint32_t m, uint32_t stride_keys, uint32_t* __restrict__ hist)
{
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t ms = (m + stride_keys - 1) / stride_keys;
    const uint32_t rounds = (ms + stride - 1) / stride;
    for (uint32_t it = 0; it < rounds; ++it) {
        const uint32_t i = it * stride + blockIdx.x * blockDim.x + threadIdx.x;
        const bool ok = i < ms;
        const uint32_t bin = ok ? (uint32_t)(__ldg(keys + (size_t)i * stride_keys) >> 48) : 0xFFFFFFFFu;
        const uint32_t peers = __match_any_sync(0xffffffffu, bin);
        if (ok && (threadIdx.x & 31) == (uint32_t)(__ffs((int)peers) - 1)) atomicAdd(hist + bin, (uint32_t)__popc(peers));
    }
}

constexpr int SPLIT_THREADS = 256, SPLIT_IPT = 8, SPLIT_TILE = SPLIT_THREADS * SPLIT_IPT, SPLIT_MAXR = 16;

// counts[g * tiles + tile] = records of the tile whose destination is rank g
__global__ void __launch_bounds__(SPLIT_THREADS)
split_count_kernel(const uint64_t* __restrict__ keys, uint32_t m, const uint8_t* __restrict__ bin2rank, int nranks,
                   uint32_t tiles, uint32_t* __restrict__ counts)
{
    __shared__ uint32_t s_cnt[SPLIT_MAXR];
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < SPLIT_MAXR) s_cnt[tid] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SPLIT_TILE + (tid >> 5) * (32 * SPLIT_IPT) + lane;
    uint32_t cnt[SPLIT_MAXR];
#pragma unroll
    for (int g = 0; g < SPLIT_MAXR; ++g) cnt[g] = 0;
#pragma unroll
    for (int i = 0; i < SPLIT_IPT; ++i) {
        const uint32_t idx = base + i * 32;
        const uint32_t d = idx < m ? bin2rank[(uint32_t)(ld_stream_u64(keys + idx) >> 48)] : 0xFFu;
#pragma unroll
        for (int g = 0; g < SPLIT_MAXR; ++g)
            if (g < nranks) cnt[g] += __popc(__ballot_sync(0xffffffffu, d == (uint32_t)g));
    }
    if (lane == 0) {
#pragma unroll
        for (int g = 0; g < SPLIT_MAXR; ++g)
            if (g < nranks && cnt[g]) atomicAdd(&s_cnt[g], cnt[g]);
    }
    __syncthreads();
    if (tid < nranks) counts[(size_t)tid * tiles + blockIdx.x] = s_cnt[tid];
}

// offsets = exclusive scan of counts in (rank-major, tile-minor) order: stable scatter
__global__ void __launch_bounds__(SPLIT_THREADS)
split_scatter_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ pos, uint32_t m,
                     const uint8_t* __restrict__ bin2rank, int nranks, uint32_t tiles,
                     const uint32_t* __restrict__ offsets, uint64_t* __restrict__ keys_out,
                     uint32_t* __restrict__ pos_out)
{
    constexpr int WARPS = SPLIT_THREADS / 32;
    __shared__ uint32_t s_wcnt[WARPS][SPLIT_MAXR];
    __shared__ uint32_t s_wbase[WARPS][SPLIT_MAXR];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = lanemask_lt();
    const uint32_t base = blockIdx.x * SPLIT_TILE + warp * (32 * SPLIT_IPT) + lane;
    uint64_t key[SPLIT_IPT];
    uint32_t d[SPLIT_IPT], rk[SPLIT_IPT];
    uint32_t cnt[SPLIT_MAXR];
#pragma unroll
    for (int g = 0; g < SPLIT_MAXR; ++g) cnt[g] = 0;
#pragma unroll
    for (int i = 0; i < SPLIT_IPT; ++i) {
        const uint32_t idx = base + i * 32;
        key[i] = idx < m ? ld_stream_u64(keys + idx) : 0ull;
        d[i] = idx < m ? bin2rank[(uint32_t)(key[i] >> 48)] : 0xFFu;
        rk[i] = 0;
#pragma unroll
        for (int g = 0; g < SPLIT_MAXR; ++g) {
            if (g < nranks) {
                const uint32_t b = __ballot_sync(0xffffffffu, d[i] == (uint32_t)g);
                if (d[i] == (uint32_t)g) rk[i] = cnt[g] + __popc(b & lt);
                cnt[g] += __popc(b);
            }
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int g = 0; g < SPLIT_MAXR; ++g) s_wcnt[warp][g] = cnt[g];
    }
    __syncthreads();
    if (tid < nranks) {
        uint32_t run = offsets[(size_t)tid * tiles + blockIdx.x];
        for (int w = 0; w < WARPS; ++w) { s_wbase[w][tid] = run; run += s_wcnt[w][tid]; }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SPLIT_IPT; ++i) {
        const uint32_t idx = base + i * 32;
        if (idx < m) {
            const uint32_t o = s_wbase[warp][d[i]] + rk[i];
            keys_out[o] = key[i];
            pos_out[o] = ld_stream_u32(pos + idx);
        }
    }
}

// Fused split + exchange: the same stable split, but every record is stored straight into its
// destination rank's receive buffer (that rank's keys[0] / vals[0], mapped through CUDA IPC), so the
// transfer over NVLink happens inside the scatter, tile by tile.  dst[g] = receive-buffer pointers
// of rank g already advanced to where this rank's run starts there.
struct PeerTable {
    uint64_t* keys[SPLIT_MAXR];
    uint32_t* pos[SPLIT_MAXR];
    uint32_t first[SPLIT_MAXR];      // offsets[g * tiles] of the local scan (start of the rank-g run)
};

__global__ void __launch_bounds__(SPLIT_THREADS)
split_scatter_p2p_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ pos, uint32_t m,
                         const uint8_t* __restrict__ bin2rank, int nranks, uint32_t tiles,
                         const uint32_t* __restrict__ offsets, PeerTable pt)
{
    constexpr int WARPS = SPLIT_THREADS / 32;
    __shared__ uint32_t s_wcnt[WARPS][SPLIT_MAXR];
    __shared__ uint32_t s_wbase[WARPS][SPLIT_MAXR];        // slot of the warp's first record for a destination
    __shared__ uint32_t s_tstart[SPLIT_MAXR + 1];           // first slot of each destination inside the tile
    __shared__ uint32_t s_run[SPLIT_MAXR];                  // index in the destination run of the tile's first record
    __shared__ uint64_t* s_k[SPLIT_MAXR];
    __shared__ uint32_t* s_p[SPLIT_MAXR];
    __shared__ uint64_t s_keys[SPLIT_TILE];                 // tile re-ordered by destination: peer stores become
    __shared__ uint32_t s_pos[SPLIT_TILE];                  // long coalesced runs instead of 32-byte fragments
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < SPLIT_MAXR) { s_k[tid] = pt.keys[tid]; s_p[tid] = pt.pos[tid]; }
    const uint32_t lt = lanemask_lt();
    const uint32_t tile_base = blockIdx.x * SPLIT_TILE;
    const uint32_t tile_n = (m - tile_base) < (uint32_t)SPLIT_TILE ? (m - tile_base) : (uint32_t)SPLIT_TILE;
    const uint32_t base = tile_base + warp * (32 * SPLIT_IPT) + lane;
    uint64_t key[SPLIT_IPT];
    uint32_t val[SPLIT_IPT];
    uint32_t d[SPLIT_IPT], rk[SPLIT_IPT];
    uint32_t cnt[SPLIT_MAXR];
#pragma unroll
    for (int g = 0; g < SPLIT_MAXR; ++g) cnt[g] = 0;
#pragma unroll
    for (int i = 0; i < SPLIT_IPT; ++i) {
        const uint32_t idx = base + i * 32;
        key[i] = idx < m ? ld_stream_u64(keys + idx) : 0ull;
        val[i] = idx < m ? ld_stream_u32(pos + idx) : 0u;
    }
#pragma unroll
    for (int i = 0; i < SPLIT_IPT; ++i) {
        const uint32_t idx = base + i * 32;
        d[i] = idx < m ? bin2rank[(uint32_t)(key[i] >> 48)] : 0xFFu;
        rk[i] = 0;
#pragma unroll
        for (int g = 0; g < SPLIT_MAXR; ++g) {
            if (g < nranks) {
                const uint32_t b = __ballot_sync(0xffffffffu, d[i] == (uint32_t)g);
                if (d[i] == (uint32_t)g) rk[i] = cnt[g] + __popc(b & lt);
                cnt[g] += __popc(b);
            }
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int g = 0; g < SPLIT_MAXR; ++g) s_wcnt[warp][g] = cnt[g];
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t run = 0;
        for (int g = 0; g < nranks; ++g) {
            s_tstart[g] = run;
            for (int w = 0; w < WARPS; ++w) { s_wbase[w][g] = run; run += s_wcnt[w][g]; }
        }
        for (int g = nranks; g <= SPLIT_MAXR; ++g) s_tstart[g] = run;
    }
    if (tid < nranks) s_run[tid] = offsets[(size_t)tid * tiles + blockIdx.x] - pt.first[tid];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SPLIT_IPT; ++i) {
        const uint32_t idx = base + i * 32;
        if (idx < m) {
            const uint32_t slot = s_wbase[warp][d[i]] + rk[i];
            s_keys[slot] = key[i];
            s_pos[slot] = val[i];
        }
    }
    __syncthreads();
    for (uint32_t j = tid; j < tile_n; j += SPLIT_THREADS) {
        int g = 0;
#pragma unroll
        for (int t = 1; t < SPLIT_MAXR; ++t) g += (t < nranks && j >= s_tstart[t]) ? 1 : 0;
        const uint32_t o = s_run[g] + (j - s_tstart[g]);
        s_k[g][o] = s_keys[j];
        s_p[g][o] = s_pos[j];
    }
}

// ext arrays for the finish: [left halo | own | right halo]
__global__ void ext_fill_kernel(const uint32_t* __restrict__ own_sa, const uint16_t* __restrict__ own_lcp, uint32_t m,
                                const uint32_t* __restrict__ halo_pos, const uint16_t* __restrict__ halo_lcp,
                                uint32_t nleft, uint32_t nright, uint32_t lcp_first, uint32_t* __restrict__ ext_sa,
                                uint16_t* __restrict__ ext_lcp)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t total = nleft + m + nright;
    if (i >= total) return;
    if (i < nleft) {
        ext_sa[i] = halo_pos[i];
        ext_lcp[i] = halo_lcp[i];
    } else if (i < nleft + m) {
        const uint32_t j = i - nleft;
        ext_sa[i] = own_sa[j];
        ext_lcp[i] = j == 0 ? (uint16_t)lcp_first : own_lcp[j];
    } else {
        const uint32_t j = i - nleft - m;
        ext_sa[i] = halo_pos[nleft + j];
        ext_lcp[i] = halo_lcp[nleft + j];
    }
}

// The halos alone (the shard itself is staged piece by piece with device copies, sa_shard_finish_bcast)
__global__ void ext_halo_kernel(const uint32_t* __restrict__ halo_pos, const uint16_t* __restrict__ halo_lcp, uint32_t m,
                                uint32_t nleft, uint32_t nright, uint32_t* __restrict__ ext_sa, uint16_t* __restrict__ ext_lcp)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nleft + nright) return;
    const uint32_t x = i < nleft ? i : i + m;
    ext_sa[x] = halo_pos[i];
    ext_lcp[x] = halo_lcp[i];
}

__global__ void set_u16_kernel(uint16_t* p, uint32_t v) { *p = (uint16_t)v; }

}  // namespace sa

#define SH_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            sa_set_error(__VA_ARGS__);         \
            return (code);                     \
        }                                      \
    } while (0)

// Upload the (replicated) text and run the preparation stage.  Replaces nothing in the reference.
extern "C" int sa_shard_begin(sa_index* idx, const uint8_t* text, uint64_t n, uint32_t flags)
{
    SH_REQUIRE(idx && (text || n == 0), SA_ERR_INVALID, "sa_shard_begin: NULL argument");
    SH_REQUIRE(n > 0 && n < 0xFFFFFFFFull - 1024ull, SA_ERR_LIMIT, "sa_shard_begin: n out of range");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    idx->timing = (flags & SA_FLAG_TIMING) != 0;
    idx->n = n; idx->tail = 0;
    idx->built = false;
    sa_invalidate_results(idx);
    idx->csv = false;
    idx->shard_prepared = false;
    idx->shard_csv = false;
    memset(&idx->stats, 0, sizeof idx->stats);
    idx->lprime = idx->L < n ? idx->L : (uint32_t)n;
    idx->stats.n = n; idx->stats.max_suffix_length = idx->L; idx->stats.lprime = idx->lprime;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    const size_t pad_from = (size_t)n & ~(size_t)15;
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + pad_from, 0, bytes - pad_from, s));
    {
        ScopedPhase ph(idx, &idx->stats.ms_h2d);
        SA_CU(cudaMemcpyAsync(idx->text.p, text, n,
                              (flags & SA_FLAG_DEVICE_TEXT) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
    }
    int rc = sa_stage_prepare(idx, flags);
    if (rc != SA_OK) return rc;
    SH_REQUIRE(idx->lprime > 0, SA_ERR_INVALID, "sa_shard_begin: max_suffix_length 0 needs no sharded build");
    idx->shard_prepared = true;
    return SA_OK;
}

// Round-0 keys + positions of the suffixes in [pos_begin, pos_end) into caller-owned device
// buffers; optionally accumulates the top-16-bit key histogram (65536 u32, caller zeroes it).
extern "C" int sa_shard_pack(sa_index* idx, uint64_t pos_begin, uint64_t pos_end, uint64_t* d_keys, uint32_t* d_pos,
                             uint32_t* d_hist16)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_pack: call sa_shard_begin first");
    SH_REQUIRE(pos_begin <= pos_end && pos_end <= idx->n, SA_ERR_INVALID, "sa_shard_pack: bad slice");
    SH_REQUIRE((pos_begin & 15) == 0, SA_ERR_INVALID, "sa_shard_pack: pos_begin must be a multiple of 16");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = (uint32_t)(pos_end - pos_begin);
    if (m == 0) return SA_OK;
    SH_REQUIRE(d_keys && d_pos, SA_ERR_INVALID, "sa_shard_pack: NULL buffer");
    {
        ScopedPhase ph(idx, &idx->stats.ms_pack);
        SA_CU(sa_launch_pack0(sa_key_source(idx) + pos_begin, m, idx->cb0, idx->code, d_keys, s));
        sa::iota_offset_kernel<<<(m + 255) / 256, 256, 0, s>>>(d_pos, m, (uint32_t)pos_begin);
        if (d_hist16) sa::hist16_kernel<<<idx->num_sms * 8, 256, 0, s>>>(d_keys, m, m >= (16u << 20) ? 64u : 16u, d_hist16);
        SA_CU(cudaGetLastError());
        idx->stats.kernel_launches += d_hist16 ? 3 : 2;
    }
    SA_CU(cudaStreamSynchronize(s));
    return SA_OK;
}

// Stable split of m records by destination rank = bin2rank[key >> 48] (host table, 65536 bytes).
// Output: records grouped by destination in rank order; counts_out[g] = records for rank g.
extern "C" int sa_shard_split(sa_index* idx, const uint64_t* d_keys, const uint32_t* d_pos, uint64_t m64,
                              const uint8_t* bin2rank, int nranks, uint64_t* d_keys_out, uint32_t* d_pos_out,
                              uint64_t* counts_out)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_split: call sa_shard_begin first");
    SH_REQUIRE(bin2rank && counts_out, SA_ERR_INVALID, "sa_shard_split: NULL argument");
    SH_REQUIRE(nranks >= 1 && nranks <= sa::SPLIT_MAXR, SA_ERR_LIMIT, "sa_shard_split: 1..16 ranks supported");
    SH_REQUIRE(m64 < (1ull << 32), SA_ERR_LIMIT, "sa_shard_split: too many records");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = (uint32_t)m64;
    for (int g = 0; g < nranks; ++g) counts_out[g] = 0;
    if (m == 0) return SA_OK;
    const uint32_t tiles = (m + sa::SPLIT_TILE - 1) / sa::SPLIT_TILE;
    // scratch: table (64 KiB) + counts/offsets (nranks * tiles + 1 words), in the aux buffers
    SA_CU(idx->aux[0].ensure(65536));
    SA_CU(idx->aux[1].ensure(((size_t)nranks * tiles + 2) * sizeof(uint32_t)));
    uint8_t* d_table = idx->aux[0].as<uint8_t>();
    uint32_t* d_counts = idx->aux[1].as<uint32_t>();
    uint32_t* sm = idx->small.as<uint32_t>();
    SA_CU(cudaMemcpyAsync(d_table, bin2rank, 65536, cudaMemcpyHostToDevice, s));
    sa::split_count_kernel<<<tiles, sa::SPLIT_THREADS, 0, s>>>(d_keys, m, d_table, nranks, tiles, d_counts);
    SA_CU(cudaGetLastError());
    SA_CU(sa_launch_scan_u32(d_counts, (uint32_t)nranks * tiles, sm + SM_NLTOTAL, s));
    sa::split_scatter_kernel<<<tiles, sa::SPLIT_THREADS, 0, s>>>(d_keys, d_pos, m, d_table, nranks, tiles, d_counts,
                                                                 d_keys_out, d_pos_out);
    SA_CU(cudaGetLastError());
    idx->stats.kernel_launches += 3;
    // per-rank totals = differences of the scanned offsets at tile 0 of each rank
    std::vector<uint32_t> first((size_t)nranks + 1, 0);
    for (int g = 0; g < nranks; ++g)
        SA_CU(cudaMemcpyAsync(&first[g], d_counts + (size_t)g * tiles, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    first[nranks] = m;
    for (int g = 0; g < nranks; ++g) counts_out[g] = first[g + 1] - first[g];
    return SA_OK;
}

// Sort + refine the m records this rank received (device buffers, runs in ascending position
// order per source rank, sources in rank order).
extern "C" int sa_shard_sort(sa_index* idx, const uint64_t* d_keys, const uint32_t* d_pos, uint64_t m64)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_sort: call sa_shard_begin first");
    SH_REQUIRE(m64 <= SA_MAX_INDEX_BYTES, SA_ERR_LIMIT, "sa_shard_sort: %llu records exceed the 2^31 per-GPU limit",
               (unsigned long long)m64);
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = (uint32_t)m64;
    idx->shard_m = m;
    idx->shard_sa = nullptr;
    if (m == 0) return SA_OK;
    int rc = sa_ensure_scratch(idx, m);
    if (rc != SA_OK) return rc;
    SA_CU(cudaMemcpyAsync(idx->keys[0].p, d_keys, (size_t)m * 8, cudaMemcpyDeviceToDevice, s));
    SA_CU(cudaMemcpyAsync(idx->vals[0].p, d_pos, (size_t)m * 4, cudaMemcpyDeviceToDevice, s));
    rc = sa_stage_sort_refine(idx, m, false, &idx->shard_sa);
    if (rc != SA_OK) return rc;
    SA_CU(cudaStreamSynchronize(s));
    return SA_OK;
}

// First and last min(edge, m) records (position + adjacent LCP) of the sorted shard, to the host.
extern "C" int sa_shard_edges(sa_index* idx, uint32_t edge, uint32_t* head_pos, uint16_t* head_lcp,
                              uint32_t* tail_pos, uint16_t* tail_lcp)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_edges: no sorted shard");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = idx->shard_m, k = edge < m ? edge : m;
    if (k == 0) return SA_OK;
    const uint16_t* lcp = idx->lcp.as<uint16_t>();
    SA_CU(cudaMemcpyAsync(head_pos, idx->shard_sa, (size_t)k * 4, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaMemcpyAsync(head_lcp, lcp, (size_t)k * 2, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaMemcpyAsync(tail_pos, idx->shard_sa + (m - k), (size_t)k * 4, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaMemcpyAsync(tail_lcp, lcp + (m - k), (size_t)k * 2, cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    return SA_OK;
}

// Finish the shard with its halos.  left_*: the nleft records that precede this shard in the
// global order (left_lcp[0] is ignored), lcp_first = LCP(last left record, first own record);
// right_*: the nright records that follow (right_lcp[0] = LCP(last own record, first right record)).
// The m placed positions go to d_sa_out (device).
// Shared body of the two finish entry points: halo records -> extended arrays -> finish_kernel.
// outs == nullptr: results go to idx->ext_out (extended rank space).
static int shard_finish_common(sa_index* idx, const uint32_t* left_pos, const uint16_t* left_lcp, uint32_t nleft,
                               uint32_t lcp_first, const uint32_t* right_pos, const uint16_t* right_lcp,
                               uint32_t nright, const sa::FinishOut* outs, bool bulk = true)
{
    cudaStream_t s = idx->stream;
    const uint32_t m = idx->shard_m;
    const uint32_t total = nleft + m + nright;
    SA_CU(idx->ext_sa.ensure((size_t)total * 4));
    SA_CU(idx->ext_lcp.ensure((size_t)total * 2 + 64));
    if (!outs) SA_CU(idx->ext_out.ensure((size_t)total * 4));
    // halo records: pinned staging is overkill for <= 8192 records; a pageable copy is synchronous
    const uint32_t nh = nleft + nright;
    SA_CU(idx->aux[0].ensure((size_t)(nh + 1) * 4));
    SA_CU(idx->aux[1].ensure((size_t)(nh + 1) * 2));
    if (nleft) {
        SA_CU(cudaMemcpyAsync(idx->aux[0].p, left_pos, (size_t)nleft * 4, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->aux[1].p, left_lcp, (size_t)nleft * 2, cudaMemcpyHostToDevice, s));
    }
    if (nright) {
        SA_CU(cudaMemcpyAsync(idx->aux[0].as<uint32_t>() + nleft, right_pos, (size_t)nright * 4, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->aux[1].as<uint16_t>() + nleft, right_lcp, (size_t)nright * 2, cudaMemcpyHostToDevice, s));
    }
    if (bulk)
        sa::ext_fill_kernel<<<(total + 255) / 256, 256, 0, s>>>(idx->shard_sa, idx->lcp.as<uint16_t>(), m,
                                                                idx->aux[0].as<uint32_t>(), idx->aux[1].as<uint16_t>(), nleft,
                                                                nright, lcp_first, idx->ext_sa.as<uint32_t>(),
                                                                idx->ext_lcp.as<uint16_t>());
    else if (nh)
        sa::ext_halo_kernel<<<(nh + 255) / 256, 256, 0, s>>>(idx->aux[0].as<uint32_t>(), idx->aux[1].as<uint16_t>(), m, nleft, nright,
                                                             idx->ext_sa.as<uint32_t>(), idx->ext_lcp.as<uint16_t>());
    SA_CU(cudaGetLastError());
    idx->stats.kernel_launches += 1;
    const uint32_t r_begin = nleft - (nleft < 32 ? nleft : 32);
    const uint32_t r_end = nleft + m + (nright < 32 ? nright : 32);
    if (outs) return SA_OK;                   // the caller launches the finishing kernels itself (chunked)
    return sa_stage_finish(idx, idx->ext_sa.as<uint32_t>(), idx->ext_lcp.as<uint16_t>(), total, r_begin, r_end,
                           idx->ext_out.as<uint32_t>());
}

extern "C" int sa_shard_finish(sa_index* idx, const uint32_t* left_pos, const uint16_t* left_lcp, uint32_t nleft,
                               uint32_t lcp_first, const uint32_t* right_pos, const uint16_t* right_lcp,
                               uint32_t nright, uint32_t* d_sa_out)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_finish: no sorted shard");
    SH_REQUIRE(nleft <= 4096 && nright <= 4096, SA_ERR_LIMIT, "sa_shard_finish: halo too large");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = idx->shard_m;
    if (m == 0) return SA_OK;
    SH_REQUIRE(d_sa_out, SA_ERR_INVALID, "sa_shard_finish: d_sa_out is NULL");
    int rc = shard_finish_common(idx, left_pos, left_lcp, nleft, lcp_first, right_pos, right_lcp, nright, nullptr);
    if (rc != SA_OK) return rc;
    SA_CU(cudaMemcpyAsync(d_sa_out, idx->ext_out.as<uint32_t>() + nleft, (size_t)m * 4, cudaMemcpyDeviceToDevice, s));
    rc = sa_check_device_error(idx);
    sa_time_collect(idx);
    return rc;
}

// Push `count` u32 entries of the local final array to the same offset of every peer's array: one
// read of local HBM, one 16-byte store per peer over NVLink (head / tail to 16-byte alignment with
// 4-byte stores; all arrays share the alignment of the offset).
namespace sa {
struct ReplicaTable {
    uint32_t* dst[15];
    int n;
};

__global__ void __launch_bounds__(512) replicate_kernel(const uint32_t* __restrict__ src, uint64_t count, ReplicaTable t)
{
    const uint64_t head = min(count, (uint64_t)((16u - ((uintptr_t)src & 15u)) & 15u) / 4u);
    const uint64_t nvec = (count - head) / 4u;
    const uint64_t tail0 = head + nvec * 4u;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint4* sv = reinterpret_cast<const uint4*>(src + head);
    for (uint64_t i = tid; i < nvec; i += nthreads) {
        const uint4 v = __ldcs(sv + i);
#pragma unroll 1
        for (int g = 0; g < t.n; ++g) reinterpret_cast<uint4*>(t.dst[g] + head)[i] = v;
    }
    for (uint64_t i = tid; i < head + (count - tail0); i += nthreads) {
        const uint64_t j = i < head ? i : tail0 + (i - head);
        const uint32_t v = src[j];
        for (int g = 0; g < t.n; ++g) t.dst[g][j] = v;
    }
}

// The same push as a kernel that fits BESIDE the finishing kernel instead of competing with it for block
// slots: six resident finish_kernel blocks (256 threads x 40 registers, 33 KB of shared memory each) leave
// 512 threads and 4096 registers of an SM free -- exactly one block of 128 threads x 32 registers.  Four
// 16-byte loads per thread are in flight before their 4 x n peer stores are issued (stores are posted, so
// only the local load latency is exposed).
constexpr int REP_THREADS = 128, REP_UNROLL = 4;
__global__ void __launch_bounds__(REP_THREADS, 16) replicate_small_kernel(const uint32_t* __restrict__ src, uint64_t count,
                                                                          ReplicaTable t)
{
    const uint64_t head = min(count, (uint64_t)((16u - ((uintptr_t)src & 15u)) & 15u) / 4u);
    const uint64_t nvec = (count - head) / 4u;
    const uint64_t tail0 = head + nvec * 4u;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint4* sv = reinterpret_cast<const uint4*>(src + head);
    const uint32_t nv = (uint32_t)nvec, nt = (uint32_t)nthreads;          // count <= n < 2^32 entries -> nvec < 2^30
    for (uint32_t i0 = (uint32_t)tid; i0 < nv; i0 += nt * REP_UNROLL) {
        uint4 v[REP_UNROLL];
#pragma unroll
        for (int u = 0; u < REP_UNROLL; ++u) {
            const uint32_t i = i0 + (uint32_t)u * nt;
            if (i < nv) v[u] = __ldcs(sv + i);
        }
#pragma unroll 1
        for (int g = 0; g < t.n; ++g) {
            uint4* d = reinterpret_cast<uint4*>(t.dst[g] + head);
#pragma unroll
            for (int u = 0; u < REP_UNROLL; ++u) {
                const uint32_t i = i0 + (uint32_t)u * nt;
                if (i < nv) d[i] = v[u];
            }
        }
    }
    for (uint64_t i = tid; i < head + (count - tail0); i += nthreads) {
        const uint64_t j = i < head ? i : tail0 + (i - head);
        const uint32_t v = src[j];
        for (int g = 0; g < t.n; ++g) t.dst[g][j] = v;
    }
}
}  // namespace sa

// ------------------------------------------------------------------------------ finish + replicate, fused
// Instead of "finish into a local slice, all-gather the slices, copy into place", the shard is finished straight
// into this rank's final array AND into every peer's (CUDA IPC mappings, NVLink): finish_kernel<STAGED> collects the
// positions of a block's 1024 ranks in shared memory and stores them 16 bytes at a time into all replicas, so the
// gather / adopt passes and the re-read of the finished array disappear.  (Round 2 measured the alternatives:
// single 4-byte peer stores from the unstaged kernel reach ~260 GB/s at 8 GPUs; finishing locally in chunks and
// pushing finished chunks with replicate_kernel on a high-priority stream overlaps, but kernel and push slow each
// other -- profiles/r02_multi_gpu_phases.md; sa_debug_set("replicate_fused", 0) selects that path.)
// The shard itself is staged into the [left halo | shard | right halo] arrays in pieces beside the finishing kernels.
// 1. sa_shard_final_prepare: allocate this rank's final array (n entries) and export its IPC handle.
extern "C" int sa_shard_final_prepare(sa_index* idx, uint8_t* handle_final)
{
    SH_REQUIRE(idx && idx->shard_prepared && handle_final, SA_ERR_INVALID, "sa_shard_final_prepare: bad argument");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    idx->built = false;
    SA_CU(idx->sa_final.ensure((size_t)(idx->n ? idx->n : 1) * 4));
    memset(handle_final, 0, 64);
    if (idx->dbg_same_process) {
        memcpy(handle_final, "SAMEPROC", 8);
        memcpy(handle_final + 8, &idx->sa_final.p, sizeof(void*));
        return SA_OK;
    }
    cudaIpcMemHandle_t h;
    SA_CU(cudaIpcGetMemHandle(&h, idx->sa_final.p));
    memcpy(handle_final, &h, 64);
    return SA_OK;
}

// 2. sa_shard_finish_bcast: like sa_shard_finish, but the m placed positions of this shard end up in
//    final[global_offset .. global_offset + m) of every rank (handles: 64 bytes per rank, from
//    sa_shard_final_prepare).  After a barrier over all ranks call sa_shard_mark_built.
extern "C" int sa_shard_finish_bcast(sa_index* idx, const uint32_t* left_pos, const uint16_t* left_lcp, uint32_t nleft,
                                     uint32_t lcp_first, const uint32_t* right_pos, const uint16_t* right_lcp,
                                     uint32_t nright, int nranks, int my_rank, const uint8_t* handles,
                                     uint64_t global_offset)
{
    SH_REQUIRE(idx && idx->shard_prepared && handles, SA_ERR_INVALID, "sa_shard_finish_bcast: bad argument");
    SH_REQUIRE(nleft <= 4096 && nright <= 4096, SA_ERR_LIMIT, "sa_shard_finish_bcast: halo too large");
    SH_REQUIRE(nranks >= 1 && nranks <= 16 && my_rank >= 0 && my_rank < nranks, SA_ERR_INVALID, "sa_shard_finish_bcast: bad rank");
    SH_REQUIRE(global_offset + idx->shard_m <= idx->n, SA_ERR_INVALID, "sa_shard_finish_bcast: slice outside the array");
    SH_REQUIRE(idx->sa_final.p, SA_ERR_INVALID, "sa_shard_finish_bcast: call sa_shard_final_prepare first");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    const uint32_t m = idx->shard_m;
    uint32_t* peer_base[16] = {};
    for (int g = 0; g < nranks; ++g) {
        void* base = nullptr;
        const uint8_t* h = handles + (size_t)g * 64;
        if (g == my_rank) {
            base = idx->sa_final.p;
        } else if (memcmp(h, "SAMEPROC", 8) == 0) {
            memcpy(&base, h + 8, sizeof base);
        } else {
            if (!idx->peer_final_mapped[g] || memcmp(&idx->peer_hf[g], h, 64) != 0) {
                if (idx->peer_final_mapped[g]) { cudaIpcCloseMemHandle(idx->peer_final[g]); idx->peer_final_mapped[g] = false; }
                memcpy(&idx->peer_hf[g], h, 64);
                SA_CU(cudaIpcOpenMemHandle(&idx->peer_final[g], idx->peer_hf[g], cudaIpcMemLazyEnablePeerAccess));
                idx->peer_final_mapped[g] = true;
            }
            base = idx->peer_final[g];
        }
        SH_REQUIRE(base, SA_ERR_INVALID, "sa_shard_finish_bcast: rank %d has no final array", g);
        peer_base[g] = static_cast<uint32_t*>(base);
        idx->shard_peer_base[g] = peer_base[g];
    }
    const bool replicate = !idx->shard_no_replicate;             // SA_FLAG_SHARDED_SA: the shard stays where it is
    if (m == 0) return SA_OK;
    cudaStream_t s = idx->stream;
    // replication runs beside the finishing kernels: highest stream priority, so that its blocks are
    // scheduled as soon as SM resources free up instead of queueing behind the finish kernel's grid
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    for (int g = 0; g < nranks; ++g)
        if (g != my_rank && !idx->copy_streams[g])
            SA_CU(cudaStreamCreateWithPriority(&idx->copy_streams[g], cudaStreamNonBlocking, prio_hi));
    constexpr int MAX_CHUNKS = 16;
    const int CHUNKS = idx->dbg_finish_chunks >= 1 && idx->dbg_finish_chunks <= MAX_CHUNKS ? (int)idx->dbg_finish_chunks : 8;
    for (int c = 0; c < CHUNKS; ++c)
        if (!idx->chunk_ev[c]) SA_CU(cudaEventCreateWithFlags(&idx->chunk_ev[c], cudaEventDisableTiming));

    // extended rank x (>= nleft) lives at global rank global_offset + x - nleft of the final arrays
    sa::FinishOut fo{};
    fo.n = 1;
    fo.lo = nleft;
    fo.hi = nleft + m;
    fo.out[0] = peer_base[my_rank] + global_offset - nleft;
    // fused (default): finish_kernel<STAGED> stores every finished block into all replicas itself -- 16-byte stores
    // over NVLink from the block's shared-memory slots -- so nothing is re-read and no copy kernel competes with the
    // finishing blocks; sa_debug_set("replicate_fused", 0) = finish locally, push finished chunks with replicate_kernel
    const bool fused = replicate && nranks > 1 && idx->dbg_replicate_fused != 0;
    if (fused)
        for (int k = 1; k < nranks; ++k) fo.out[fo.n++] = peer_base[(my_rank + k) % nranks] + global_offset - nleft;
    // the halos go into the [left halo | shard | right halo] arrays with a small kernel; the shard itself follows in
    // pieces -- device copies on a second stream, piece c + 1 beside the finishing kernels of chunk c
    const bool staged = idx->dbg_ext_fill == 0;
    int rc = shard_finish_common(idx, left_pos, left_lcp, nleft, lcp_first, right_pos, right_lcp, nright, &fo, !staged);
    if (rc != SA_OK) return rc;
    const uint32_t total = nleft + m + nright;
    const uint32_t r_begin = nleft - (nleft < 32 ? nleft : 32);
    const uint32_t r_end = nleft + m + (nright < 32 ? nright : 32);
    const uint32_t blocks = sa_finish_blocks(total), FTR = sa_finish_block_ranks();
    const uint32_t per = (blocks + CHUNKS - 1) / CHUNKS;
    if (staged) {
        if (!idx->push_stream) SA_CU(cudaStreamCreateWithPriority(&idx->push_stream, cudaStreamNonBlocking, prio_hi));
        for (int c = 0; c <= CHUNKS; ++c)
            if (!idx->ext_ev[c]) SA_CU(cudaEventCreateWithFlags(&idx->ext_ev[c], cudaEventDisableTiming));
        cudaStream_t ps = idx->push_stream;
        SA_CU(cudaEventRecord(idx->ext_ev[CHUNKS], s));          // sorted shard + halos are in place
        SA_CU(cudaStreamWaitEvent(ps, idx->ext_ev[CHUNKS], 0));
        uint64_t done = 0;                                        // records of the shard staged so far
        for (int c = 0; c < CHUNKS && (uint64_t)c * per < blocks; ++c) {
            // chunk c reads LCPs up to 96 ranks past its last block
            const bool last = (uint64_t)(c + 1) * per >= blocks;
            uint64_t need = last ? (uint64_t)total : (uint64_t)((uint64_t)c * per + per) * FTR + 128;
            if (need > total) need = total;
            uint64_t hi = need > nleft ? need - nleft : 0;
            if (hi > m) hi = m;
            if (hi > done) {
                SA_CU(cudaMemcpyAsync(idx->ext_sa.as<uint32_t>() + nleft + done, idx->shard_sa + done, (size_t)(hi - done) * 4,
                                      cudaMemcpyDeviceToDevice, ps));
                SA_CU(cudaMemcpyAsync(idx->ext_lcp.as<uint16_t>() + nleft + done, idx->lcp.as<uint16_t>() + done, (size_t)(hi - done) * 2,
                                      cudaMemcpyDeviceToDevice, ps));
                if (done == 0) {                                  // the first record's LCP is the junction with the left halo
                    sa::set_u16_kernel<<<1, 1, 0, ps>>>(idx->ext_lcp.as<uint16_t>() + nleft, lcp_first);
                    SA_CU(cudaGetLastError());
                }
                done = hi;
            }
            SA_CU(cudaEventRecord(idx->ext_ev[c], ps));
        }
    }
    uint32_t copied_to = nleft;               // extended rank up to which the peers are up to date
    // sa_debug_set("timeline", 1): when each chunk was finished / pushed, in ms after the first launch
    const bool tl = idx->dbg_timeline != 0;
    cudaEvent_t tl_ev[1 + 3 * MAX_CHUNKS] = {};
    if (tl) {
        for (auto& e : tl_ev) SA_CU(cudaEventCreate(&e));
        SA_CU(cudaEventRecord(tl_ev[0], s));
    }
    // sa_debug_set("replicate_ctas", k): k > 0 = push kernel with k blocks of 512 threads (default 64, measured best at
    // 4 GPUs: profiles/r02_multi_gpu_phases.md); k < 0 = the 128-thread kernel that fits beside the finishing blocks
    const unsigned rep_ctas = idx->dbg_replicate_ctas > 0 ? (unsigned)idx->dbg_replicate_ctas : (idx->dbg_replicate_ctas < 0 ? 0u : 64u);
    for (int c = 0; c < CHUNKS && (uint64_t)c * per < blocks; ++c) {
        const uint32_t b0 = (uint32_t)c * per;
        if (staged) SA_CU(cudaStreamWaitEvent(s, idx->ext_ev[c], 0));
        rc = sa_stage_finish_to(idx, idx->ext_sa.as<uint32_t>(), idx->ext_lcp.as<uint16_t>(), total, r_begin, r_end, fo, b0, per);
        if (rc != SA_OK) return rc;
        SA_CU(cudaEventRecord(idx->chunk_ev[c], s));
        if (tl) SA_CU(cudaEventRecord(tl_ev[1 + 3 * c], s));
        // a bucket may reach 32 ranks back into the previous chunk: ranks below hi_c - 32 are final now
        const bool last = (uint64_t)(c + 1) * per >= blocks;
        uint64_t upto = last ? (uint64_t)nleft + m : (uint64_t)(b0 + per) * FTR;
        if (!last) upto = upto > 32 ? upto - 32 : 0;
        if (upto > (uint64_t)nleft + m) upto = (uint64_t)nleft + m;
        if (upto <= copied_to || !replicate || fused) continue;
        const size_t off = (size_t)global_offset + copied_to - nleft, cnt = (size_t)(upto - copied_to);
        if (idx->dbg_replicate_memcpy) {      // copy engines instead of the store kernel (slower at 8 GPUs)
            for (int k = 1; k < nranks; ++k) {
                const int g = (my_rank + k) % nranks;
                SA_CU(cudaStreamWaitEvent(idx->copy_streams[g], idx->chunk_ev[c], 0));
                SA_CU(cudaMemcpyAsync(peer_base[g] + off, peer_base[my_rank] + off, cnt * 4, cudaMemcpyDeviceToDevice,
                                      idx->copy_streams[g]));
            }
        } else if (nranks > 1) {
            sa::ReplicaTable rt{};
            for (int k = 1; k < nranks; ++k) rt.dst[rt.n++] = peer_base[(my_rank + k) % nranks] + off;
            cudaStream_t cs = idx->copy_streams[(my_rank + 1) % nranks];
            SA_CU(cudaStreamWaitEvent(cs, idx->chunk_ev[c], 0));
            if (tl) SA_CU(cudaEventRecord(tl_ev[2 + 3 * c], cs));
            if (rep_ctas)
                sa::replicate_kernel<<<rep_ctas, 512, 0, cs>>>(peer_base[my_rank] + off, (uint64_t)cnt, rt);
            else
                sa::replicate_small_kernel<<<idx->num_sms * 2, sa::REP_THREADS, 0, cs>>>(peer_base[my_rank] + off, (uint64_t)cnt, rt);
            SA_CU(cudaGetLastError());
            if (tl) SA_CU(cudaEventRecord(tl_ev[3 + 3 * c], cs));
            idx->stats.kernel_launches += 1;
        }
        copied_to = (uint32_t)upto;
    }
    for (int g = 0; g < nranks; ++g)
        if (g != my_rank) SA_CU(cudaStreamSynchronize(idx->copy_streams[g]));
    rc = sa_check_device_error(idx);          // synchronises the build stream
    if (tl) {
        char line[2048];
        int at = snprintf(line, sizeof line, "[timeline rank %d] chunk: finished | push start..end (ms)", my_rank);
        for (int c = 0; c < CHUNKS; ++c) {
            float f = -1.f, a = -1.f, b = -1.f;
            cudaEventElapsedTime(&f, tl_ev[0], tl_ev[1 + 3 * c]);
            cudaEventElapsedTime(&a, tl_ev[0], tl_ev[2 + 3 * c]);
            cudaEventElapsedTime(&b, tl_ev[0], tl_ev[3 + 3 * c]);
            at += snprintf(line + at, sizeof line - (size_t)at, "  %d: %.2f | %.2f..%.2f", c, f, a, b);
        }
        cudaGetLastError();                   // events of skipped chunks were never recorded
        fprintf(stderr, "%s\n", line);
        for (auto& e : tl_ev) cudaEventDestroy(e);
    }
    sa_time_collect(idx);
    return rc;
}

// 3. every rank's finish has completed (barrier): the local final array is complete.
extern "C" int sa_shard_mark_built(sa_index* idx)
{
    SH_REQUIRE(idx && idx->shard_prepared && idx->sa_final.p, SA_ERR_INVALID, "sa_shard_mark_built: nothing to adopt");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    if (idx->shard_csv) {
        SA_CU(cudaSetDevice(idx->device));
        int rc = sa_csv_finalize_shard(idx);
        if (rc != SA_OK) return rc;
    }
    idx->built = true;
    return SA_OK;
}

// Text the shard build works on (after sa_shard_begin / sa_shard_begin_csv): its size, and a copy of
// any byte range of it for the host side of the driver (a few boundary LCPs are computed there).
extern "C" uint64_t sa_shard_text_size(const sa_index* idx) { return idx && idx->shard_prepared ? idx->n : 0; }

extern "C" int sa_shard_copy_text(sa_index* idx, uint64_t offset, uint64_t len, uint8_t* out)
{
    SH_REQUIRE(idx && idx->shard_prepared && (out || len == 0), SA_ERR_INVALID, "sa_shard_copy_text: bad argument");
    SH_REQUIRE(offset <= idx->n && len <= idx->n - offset, SA_ERR_INVALID, "sa_shard_copy_text: range outside the text");
    if (len == 0) return SA_OK;
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    SA_CU(cudaMemcpyAsync(out, idx->text.as<uint8_t>() + offset, (size_t)len, cudaMemcpyDeviceToHost, idx->stream));
    SA_CU(cudaStreamSynchronize(idx->stream));
    return SA_OK;
}

// Install a complete suffix array (n entries, e.g. the all-gathered shard outputs) so that the
// index answers queries: every rank then holds text + SA and takes a share of each query batch.
extern "C" int sa_adopt_sa(sa_index* idx, const uint32_t* d_sa_full, uint64_t n)
{
    SH_REQUIRE(idx && d_sa_full, SA_ERR_INVALID, "sa_adopt_sa: NULL argument");
    SH_REQUIRE(idx->shard_prepared && n == idx->n, SA_ERR_INVALID, "sa_adopt_sa: the index holds no text of %llu bytes",
               (unsigned long long)n);
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    SA_CU(idx->sa_final.ensure((size_t)n * 4));
    SA_CU(cudaMemcpyAsync(idx->sa_final.p, d_sa_full, (size_t)n * 4, cudaMemcpyDeviceToDevice, idx->stream));
    SA_CU(cudaStreamSynchronize(idx->stream));
    if (idx->shard_csv) {
        int rc = sa_csv_finalize_shard(idx);
        if (rc != SA_OK) return rc;
    }
    idx->built = true;
    return SA_OK;
}

// ------------------------------------------------------------------------------ fused split + exchange (CUDA IPC)
// 1. sa_shard_count: how many local records go to each rank (counts_out[nranks]); the scanned
//    per-tile offsets stay on the device for step 3.
extern "C" int sa_shard_count(sa_index* idx, const uint64_t* d_keys, uint64_t m64, const uint8_t* bin2rank, int nranks,
                              uint64_t* counts_out)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_count: call sa_shard_begin first");
    SH_REQUIRE(bin2rank && counts_out, SA_ERR_INVALID, "sa_shard_count: NULL argument");
    SH_REQUIRE(nranks >= 1 && nranks <= sa::SPLIT_MAXR, SA_ERR_LIMIT, "sa_shard_count: 1..16 ranks supported");
    SH_REQUIRE(m64 < (1ull << 32), SA_ERR_LIMIT, "sa_shard_count: too many records");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = (uint32_t)m64;
    for (int g = 0; g < nranks; ++g) counts_out[g] = 0;
    if (m == 0) return SA_OK;
    const uint32_t tiles = (m + sa::SPLIT_TILE - 1) / sa::SPLIT_TILE;
    SA_CU(idx->split_table.ensure(65536));
    SA_CU(idx->split_offsets.ensure(((size_t)nranks * tiles + 2) * sizeof(uint32_t)));
    uint8_t* d_table = idx->split_table.as<uint8_t>();
    uint32_t* d_counts = idx->split_offsets.as<uint32_t>();
    uint32_t* sm = idx->small.as<uint32_t>();
    SA_CU(cudaMemcpyAsync(d_table, bin2rank, 65536, cudaMemcpyHostToDevice, s));
    sa::split_count_kernel<<<tiles, sa::SPLIT_THREADS, 0, s>>>(d_keys, m, d_table, nranks, tiles, d_counts);
    SA_CU(cudaGetLastError());
    SA_CU(sa_launch_scan_u32(d_counts, (uint32_t)nranks * tiles, sm + SM_NLTOTAL, s));
    idx->stats.kernel_launches += 2;
    uint32_t* first = idx->h_small + 64;                       // pinned, [64, 64 + nranks]
    for (int g = 0; g < nranks; ++g)
        SA_CU(cudaMemcpyAsync(first + g, d_counts + (size_t)g * tiles, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    first[nranks] = m;
    for (int g = 0; g < nranks; ++g) counts_out[g] = first[g + 1] - first[g];
    return SA_OK;
}

// 2. sa_shard_recv_prepare: size this rank's receive buffers (its sort buffers keys[0] / vals[0])
//    for m_recv records and export their CUDA IPC handles (64 bytes each).
extern "C" int sa_shard_recv_prepare(sa_index* idx, uint64_t m_recv, uint8_t* handle_keys, uint8_t* handle_pos)
{
    SH_REQUIRE(idx && idx->shard_prepared && handle_keys && handle_pos, SA_ERR_INVALID, "sa_shard_recv_prepare: bad argument");
    SH_REQUIRE(m_recv <= SA_MAX_INDEX_BYTES, SA_ERR_LIMIT, "sa_shard_recv_prepare: %llu records exceed the 2^31 per-GPU limit",
               (unsigned long long)m_recv);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    SA_CU(cudaStreamSynchronize(idx->stream));                  // nothing of an earlier build still uses them
    int rc = sa_ensure_scratch(idx, m_recv ? (uint32_t)m_recv : 1u);
    if (rc != SA_OK) return rc;
    if (idx->dbg_same_process) {                                // emulated ranks share one process: raw pointers
        memset(handle_keys, 0, 64);
        memset(handle_pos, 0, 64);
        memcpy(handle_keys, "SAMEPROC", 8);
        memcpy(handle_pos, "SAMEPROC", 8);
        memcpy(handle_keys + 8, &idx->keys[0].p, sizeof(void*));
        memcpy(handle_pos + 8, &idx->vals[0].p, sizeof(void*));
        return SA_OK;
    }
    cudaIpcMemHandle_t hk, hp;
    SA_CU(cudaIpcGetMemHandle(&hk, idx->keys[0].p));
    SA_CU(cudaIpcGetMemHandle(&hp, idx->vals[0].p));
    memcpy(handle_keys, &hk, 64);
    memcpy(handle_pos, &hp, 64);
    return SA_OK;
}

// 3. sa_shard_scatter_p2p: handles[g] = the two handles rank g exported (128 bytes per rank, keys
//    then positions); dst_offset[g] = record index in rank g's buffers where this rank's run starts
//    (sum of the counts of lower ranks for destination g).  Stores go over NVLink / the local HBM.
extern "C" int sa_shard_scatter_p2p(sa_index* idx, const uint64_t* d_keys, const uint32_t* d_pos, uint64_t m64,
                                    int nranks, int my_rank, const uint8_t* handles, const uint64_t* dst_offset)
{
    SH_REQUIRE(idx && idx->shard_prepared && handles && dst_offset, SA_ERR_INVALID, "sa_shard_scatter_p2p: bad argument");
    SH_REQUIRE(nranks >= 1 && nranks <= sa::SPLIT_MAXR && my_rank >= 0 && my_rank < nranks, SA_ERR_INVALID,
               "sa_shard_scatter_p2p: bad rank");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    cudaStream_t s = idx->stream;
    const uint32_t m = (uint32_t)m64;
    sa::PeerTable pt{};
    for (int g = 0; g < nranks; ++g) {
        void *pk = nullptr, *pp = nullptr;
        if (g == my_rank) {
            pk = idx->keys[0].p; pp = idx->vals[0].p;
        } else {
            const uint8_t* h = handles + (size_t)g * 128;
            if (memcmp(h, "SAMEPROC", 8) == 0) {               // single-process emulation: raw device pointers
                memcpy(&pk, h + 8, sizeof pk);
                memcpy(&pp, h + 64 + 8, sizeof pp);
                pt.keys[g] = static_cast<uint64_t*>(pk) + dst_offset[g];
                pt.pos[g] = static_cast<uint32_t*>(pp) + dst_offset[g];
                continue;
            }
            if (!idx->peer_mapped[g] || memcmp(&idx->peer_hk[g], h, 64) != 0 || memcmp(&idx->peer_hp[g], h + 64, 64) != 0) {
                if (idx->peer_mapped[g]) {
                    cudaIpcCloseMemHandle(idx->peer_keys[g]);
                    cudaIpcCloseMemHandle(idx->peer_pos[g]);
                    idx->peer_mapped[g] = false;
                }
                memcpy(&idx->peer_hk[g], h, 64);
                memcpy(&idx->peer_hp[g], h + 64, 64);
                SA_CU(cudaIpcOpenMemHandle(&idx->peer_keys[g], idx->peer_hk[g], cudaIpcMemLazyEnablePeerAccess));
                SA_CU(cudaIpcOpenMemHandle(&idx->peer_pos[g], idx->peer_hp[g], cudaIpcMemLazyEnablePeerAccess));
                idx->peer_mapped[g] = true;
            }
            pk = idx->peer_keys[g]; pp = idx->peer_pos[g];
        }
        pt.keys[g] = static_cast<uint64_t*>(pk) + dst_offset[g];
        pt.pos[g] = static_cast<uint32_t*>(pp) + dst_offset[g];
    }
    if (m == 0) return SA_OK;
    const uint32_t tiles = (m + sa::SPLIT_TILE - 1) / sa::SPLIT_TILE;
    const uint32_t* first = idx->h_small + 64;                  // left there by sa_shard_count
    for (int g = 0; g < nranks; ++g) pt.first[g] = first[g];
    sa::split_scatter_p2p_kernel<<<tiles, sa::SPLIT_THREADS, 0, s>>>(d_keys, d_pos, m, idx->split_table.as<uint8_t>(),
                                                                     nranks, tiles, idx->split_offsets.as<uint32_t>(), pt);
    SA_CU(cudaGetLastError());
    idx->stats.kernel_launches += 1;
    SA_CU(cudaStreamSynchronize(s));
    return SA_OK;
}

// 4. sa_shard_sort_recv: after a barrier (every rank's scatter has completed), sort + refine the
//    m records that now sit in this rank's keys[0] / vals[0].
extern "C" int sa_shard_sort_recv(sa_index* idx, uint64_t m64)
{
    SH_REQUIRE(idx && idx->shard_prepared, SA_ERR_INVALID, "sa_shard_sort_recv: call sa_shard_begin first");
    SH_REQUIRE(m64 <= SA_MAX_INDEX_BYTES, SA_ERR_LIMIT, "sa_shard_sort_recv: too many records");
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    const uint32_t m = (uint32_t)m64;
    idx->shard_m = m;
    idx->shard_sa = nullptr;
    if (m == 0) return SA_OK;
    int rc = sa_stage_sort_refine(idx, m, false, &idx->shard_sa);
    if (rc != SA_OK) return rc;
    SA_CU(cudaStreamSynchronize(idx->stream));
    return SA_OK;
}

// Same, from the padded all-gather layout: part g holds counts[g] entries at d_parts + g * stride.
extern "C" int sa_adopt_sa_parts(sa_index* idx, const uint32_t* d_parts, uint64_t stride, const uint64_t* counts,
                                 int nparts)
{
    SH_REQUIRE(idx && d_parts && counts && nparts >= 1, SA_ERR_INVALID, "sa_adopt_sa_parts: bad argument");
    SH_REQUIRE(idx->shard_prepared, SA_ERR_INVALID, "sa_adopt_sa_parts: the index holds no text");
    uint64_t total = 0;
    for (int g = 0; g < nparts; ++g) total += counts[g];
    SH_REQUIRE(total == idx->n, SA_ERR_INVALID, "sa_adopt_sa_parts: parts hold %llu entries, text has %llu bytes",
               (unsigned long long)total, (unsigned long long)idx->n);
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    SA_CU(idx->sa_final.ensure((size_t)idx->n * 4));
    uint64_t off = 0;
    for (int g = 0; g < nparts; ++g) {
        if (counts[g])
            SA_CU(cudaMemcpyAsync(idx->sa_final.as<uint32_t>() + off, d_parts + (size_t)g * stride, (size_t)counts[g] * 4,
                                  cudaMemcpyDeviceToDevice, idx->stream));
        off += counts[g];
    }
    SA_CU(cudaStreamSynchronize(idx->stream));
    if (idx->shard_csv) {
        int rc = sa_csv_finalize_shard(idx);
        if (rc != SA_OK) return rc;
    }
    idx->built = true;
    return SA_OK;
}
