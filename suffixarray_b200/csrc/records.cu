// records.cu -- what follows the hit lists (sm_100a): segmented dedupe by record id, the row table of
// a CSV file as positions, and record materialisation (the record copies of get_matching_records[_file],
// engine.c:1207-1210 / :1369-1387, batched: one packed byte buffer + offsets instead of one malloc per hit).
#include "common.cuh"
#include "kernels.h"

namespace sa {

// ------------------------------------------------------------------------------ segmented dedupe
// One warp per query.  Record ids of the query's hits go into an open-addressing table
// (key = id + 1, value = smallest hit index with that id); a hit survives when it is that smallest
// index.  Survivors are compacted to the front of the query's segment, order kept.
// Up to DD_SMALL hits the table lives in shared memory (covers the default k = 1000); longer lists
// use 4 words per hit of global scratch.
constexpr int DD_SMALL = 1024;
constexpr int DD_SLOTS = 2048;
constexpr int DD_WARPS = 2;

__device__ __forceinline__ uint32_t dd_hash(uint32_t x)
{
    x *= 0x9E3779B1u;
    return x ^ (x >> 15);
}

__global__ void __launch_bounds__(DD_WARPS * 32)
dedupe_mark_kernel(const uint64_t* __restrict__ hit_off, uint32_t nq, uint32_t* __restrict__ hit_pos,
                   uint32_t* __restrict__ hit_doc, uint32_t* __restrict__ ucount, uint32_t* __restrict__ gtable)
{
    __shared__ uint32_t s_key[DD_WARPS][DD_SLOTS];
    __shared__ uint32_t s_val[DD_WARPS][DD_SLOTS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t q = blockIdx.x * DD_WARPS + warp;
    if (q >= nq) return;
    const uint64_t o = hit_off[q];
    const uint32_t take = (uint32_t)(hit_off[q + 1] - o);
    if (take == 0) { if (lane == 0) ucount[q] = 0; return; }
    uint32_t *key, *val, slots;
    if (take <= DD_SMALL) {
        key = s_key[warp]; val = s_val[warp]; slots = DD_SLOTS;
    } else {
        key = gtable + 4 * o; slots = 2 * take; val = key + slots;
    }
    for (uint32_t i = lane; i < slots; i += 32) { key[i] = 0u; val[i] = 0xFFFFFFFFu; }
    __syncwarp();
    for (uint32_t j = lane; j < take; j += 32) {
        const uint32_t d = hit_doc[o + j] + 1u;                   // record ids are < 2^32 - 1
        uint32_t slot = dd_hash(d) % slots;
        for (;;) {
            const uint32_t prev = atomicCAS(&key[slot], 0u, d);
            if (prev == 0u || prev == d) { atomicMin(&val[slot], j); break; }
            slot = slot + 1 == slots ? 0 : slot + 1;
        }
    }
    __threadfence_block();
    __syncwarp();
    uint32_t written = 0;
    for (uint32_t base = 0; base < take; base += 32) {
        const uint32_t j = base + lane;
        bool keep = false;
        uint32_t p = 0, d = 0;
        if (j < take) {
            p = hit_pos ? hit_pos[o + j] : 0u;
            d = hit_doc[o + j];
            uint32_t slot = dd_hash(d + 1u) % slots;
            while (key[slot] != d + 1u) slot = slot + 1 == slots ? 0 : slot + 1;
            keep = val[slot] == j;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        __syncwarp();                                              // all reads of this chunk precede its writes
        if (keep) {
            const uint32_t u = written + __popc(bal & ((1u << lane) - 1u));      // u <= j: never ahead of the reads
            if (hit_pos) hit_pos[o + u] = p;
            hit_doc[o + u] = d;
        }
        written += __popc(bal);
        __syncwarp();
    }
    if (lane == 0) ucount[q] = written;
}

__global__ void __launch_bounds__(256)
segment_compact_kernel(const uint64_t* __restrict__ hit_off, const uint64_t* __restrict__ uoff, uint32_t nq,
                       const uint32_t* __restrict__ hit_pos, const uint32_t* __restrict__ hit_doc,
                       uint32_t* __restrict__ upos, uint32_t* __restrict__ udoc)
{
    const uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= nq) return;
    const uint64_t o = hit_off[q], u = uoff[q];
    const uint32_t cnt = (uint32_t)(uoff[q + 1] - u);
    for (uint32_t j = threadIdx.x & 31; j < cnt; j += 32) {
        if (upos) upos[u + j] = hit_pos[o + j];
        udoc[u + j] = hit_doc[o + j];
    }
}

// ------------------------------------------------------------------------------ rank structure -> positions
__global__ void __launch_bounds__(256)
rank_positions_kernel(const unsigned long long* __restrict__ bits, const uint32_t* __restrict__ rank, uint64_t nwords,
                      uint32_t* __restrict__ pos)
{
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= nwords) return;
    unsigned long long b = bits[w];
    uint32_t r = rank[w];
    while (b) {
        const int i = __ffsll((long long)b) - 1;
        pos[r++] = (uint32_t)(w * 64 + i);
        b &= b - 1;
    }
}

// ------------------------------------------------------------------------------ records
__global__ void __launch_bounds__(256)
record_bounds_kernel(const uint32_t* __restrict__ hit_pos, const uint32_t* __restrict__ hit_doc, uint64_t num,
                     const uint32_t* __restrict__ ends, uint64_t num_ends, uint64_t total, bool compat,
                     unsigned long long* __restrict__ src_off, uint32_t* __restrict__ len)
{
    const uint64_t h = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= num) return;
    const uint32_t d = hit_doc[h];
    uint64_t a = d ? (uint64_t)ends[d - 1] + 1 : 0;
    uint64_t b = d < num_ends ? (uint64_t)ends[d] : total;
    if (compat) {
        // engine.c:1314: the byte in front of the '\n' is dropped; :1262/:1277: for hits at offsets
        // >= 4096 the backward scan starts one byte early and the record loses its first byte too
        if (b > a) b -= 1;
        if (hit_pos[h] >= 4096u && a < b) a += 1;
    }
    src_off[h] = a;
    len[h] = (uint32_t)(b - a);
}

__global__ void __launch_bounds__(256)
record_gather_kernel(const uint8_t* __restrict__ src, const unsigned long long* __restrict__ src_off,
                     const uint64_t* __restrict__ rec_off, uint64_t num, uint8_t* __restrict__ out)
{
    const uint64_t h = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (h >= num) return;
    const uint64_t o = rec_off[h];
    const uint32_t len = (uint32_t)(rec_off[h + 1] - o);
    const uint8_t* s = src + src_off[h];
    for (uint32_t i = threadIdx.x & 31; i < len; i += 32) out[o + i] = s[i];
}

// sa_load: how many entries of a[0, n) are >= limit (a corrupt container must not reach the query kernels)
__global__ void __launch_bounds__(256)
count_out_of_range_kernel(const uint32_t* __restrict__ a, uint64_t n, uint64_t limit, unsigned int* __restrict__ bad)
{
    uint32_t c = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        c += (uint64_t)a[i] >= limit;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(bad, c);
}

}  // namespace sa

cudaError_t sa_launch_count_out_of_range(const uint32_t* a, uint64_t n, uint64_t limit, unsigned int* bad, cudaStream_t s)
{
    if (!n) return cudaSuccess;
    const uint64_t blocks = (n + 255) / 256;
    sa::count_out_of_range_kernel<<<(unsigned)(blocks < 2368 ? blocks : 2368), 256, 0, s>>>(a, n, limit, bad);
    return cudaGetLastError();
}

cudaError_t sa_launch_dedupe_mark(const uint64_t* hit_off, uint32_t nq, uint32_t* hit_pos, uint32_t* hit_doc,
                                  uint32_t* ucount, uint32_t* table, cudaStream_t s)
{
    if (!nq) return cudaSuccess;
    sa::dedupe_mark_kernel<<<(nq + sa::DD_WARPS - 1) / sa::DD_WARPS, sa::DD_WARPS * 32, 0, s>>>(hit_off, nq, hit_pos, hit_doc,
                                                                                              ucount, table);
    return cudaGetLastError();
}

cudaError_t sa_launch_segment_compact(const uint64_t* hit_off, const uint64_t* uoff, uint32_t nq, const uint32_t* hit_pos,
                                      const uint32_t* hit_doc, uint32_t* upos, uint32_t* udoc, cudaStream_t s)
{
    if (!nq) return cudaSuccess;
    const uint64_t threads = (uint64_t)nq * 32;
    sa::segment_compact_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(hit_off, uoff, nq, hit_pos, hit_doc, upos, udoc);
    return cudaGetLastError();
}

cudaError_t sa_launch_rank_positions(const unsigned long long* bits, const uint32_t* rank, uint64_t nwords, uint32_t* pos,
                                     cudaStream_t s)
{
    if (!nwords) return cudaSuccess;
    sa::rank_positions_kernel<<<(unsigned)((nwords + 255) / 256), 256, 0, s>>>(bits, rank, nwords, pos);
    return cudaGetLastError();
}

cudaError_t sa_launch_record_bounds(const uint32_t* hit_pos, const uint32_t* hit_doc, uint64_t num, const uint32_t* ends,
                                    uint64_t num_ends, uint64_t total, bool compat, unsigned long long* src_off,
                                    uint32_t* len, cudaStream_t s)
{
    if (!num) return cudaSuccess;
    sa::record_bounds_kernel<<<(unsigned)((num + 255) / 256), 256, 0, s>>>(hit_pos, hit_doc, num, ends, num_ends, total, compat,
                                                                          src_off, len);
    return cudaGetLastError();
}

cudaError_t sa_launch_record_gather(const uint8_t* src, const unsigned long long* src_off, const uint64_t* rec_off,
                                    uint64_t num, uint8_t* out, cudaStream_t s)
{
    if (!num) return cudaSuccess;
    const uint64_t threads = num * 32;
    sa::record_gather_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(src, src_off, rec_off, num, out);
    return cudaGetLastError();
}
