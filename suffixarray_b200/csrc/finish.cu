// finish.cu -- the "<= 32 bucket" finishing step of ORDER-E, data-parallel.
//
// recursive_bucket_sort stops splitting a bucket as soon as it has <= 32 members and finishes
// it with a stable insertion sort under strncmp_128 (engine.c:710-729, :664-677).  On the
// sorted-by-prefix array those buckets are contiguous, and the depth at which rank r's bucket
// first has <= 32 members is
//     d*(r) = 1 + max over the 33-rank windows containing r of (min adjacent LCP inside it)
// (0 when n <= 32).  Each block takes 1024 ranks (+ halo):
//   A. adjacent LCPs -> shared memory
//   B. 32-wide sliding minimum, then 33-wide sliding maximum, both by the block-prefix /
//      block-suffix decomposition (van Herk / Gil-Werman) with warp shuffles: 2 barriers
//   C. per rank: d*, bucket-head flag (lcp < d*), and -- when the bucket gets re-sorted -- the
//      17-byte window key at text[pos + d*] (one 32-byte aligned read per rank, the byte in
//      front of the window tells whether the shared prefix ends in a terminator)
//   D. bucket bounds from the head bit masks (clz / ffs), rank-by-counting inside the bucket
// A bucket is left alone when d* >= L' (engine.c:706), when it has one member, or when the
// shared d*-byte prefix ends in '\n' / 0xFF (such buckets are never entered, engine.c:771-773;
// a terminator can only sit at byte d*-1: an earlier one would have frozen a larger bucket).
#include "common.cuh"
#include "kernels.h"

namespace sa {

constexpr int FT = 1024;             // ranks per block
constexpr int FEXT = FT + 64;        // ranks with keys: [b0-32, b0+FT+32)            (34 warps' worth)
constexpr int FLCP = FT + 160;       // lcp entries: ranks [b0-64, b0+FT+96)          (37 * 32)
constexpr int FTHREADS = 256;
constexpr int FWARPS = FTHREADS / 32;

__device__ __forceinline__ int warp_prefix_min(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v = min(v, y); }
    return v;
}
__device__ __forceinline__ int warp_suffix_min(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_down_sync(0xffffffffu, v, o); if (lane + o < 32) v = min(v, y); }
    return v;
}
__device__ __forceinline__ int warp_prefix_max(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v = max(v, y); }
    return v;
}
__device__ __forceinline__ int warp_suffix_max(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_down_sync(0xffffffffu, v, o); if (lane + o < 32) v = max(v, y); }
    return v;
}

// (h2, l2, t2) < (hi, lo, tp) in lexicographic order = the borrow out of the 192-bit subtraction;
// returns 0xFFFFFFFF when less, else 0.
__device__ __forceinline__ uint32_t less192(uint64_t h2, uint64_t l2, uint64_t t2, uint64_t hi, uint64_t lo, uint64_t tp)
{
    uint32_t r;
    asm("{\n"
        ".reg .u32 t;\n"
        "sub.cc.u32 t, %1, %7;\n"
        "subc.cc.u32 t, %2, %8;\n"
        "subc.cc.u32 t, %3, %9;\n"
        "subc.cc.u32 t, %4, %10;\n"
        "subc.cc.u32 t, %5, %11;\n"
        "subc.cc.u32 t, %6, %12;\n"
        "subc.u32 %0, 0, 0;\n"
        "}"
        : "=r"(r)
        : "r"((uint32_t)t2), "r"((uint32_t)(t2 >> 32)), "r"((uint32_t)l2), "r"((uint32_t)(l2 >> 32)), "r"((uint32_t)h2),
          "r"((uint32_t)(h2 >> 32)), "r"((uint32_t)tp), "r"((uint32_t)(tp >> 32)), "r"((uint32_t)lo),
          "r"((uint32_t)(lo >> 32)), "r"((uint32_t)hi), "r"((uint32_t)(hi >> 32)));
    return r;
}

// text gathers are the only re-used stream of this kernel: ask L2 to keep them (evict_last), the
// rank-ordered streams go through with evict-first hints (__ldcs / __stcs)
__device__ __forceinline__ uint64_t ld_keep_u64(const uint64_t* p, uint64_t policy)
{
    uint64_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(policy));
    return v;
}

// One placed position goes to every replica of the output (1 locally; all ranks' arrays through
// peer stores over NVLink in a sharded build, where only ranks [lo, hi) belong to this shard).
__device__ __forceinline__ void put(const FinishOut& o, int64_t x, uint32_t pos)
{
    if (x < (int64_t)o.lo || x >= (int64_t)o.hi) return;
#pragma unroll 1
    for (int g = 0; g < o.n; ++g) __stcs(o.out[g] + x, pos);
}

// STAGED (a sharded build that replicates: outs.n > 1): positions that land inside the block's own 1024 ranks are
// collected in shared memory and leave as 16-byte stores -- one per replica, i.e. the push of the finished ranks
// into every peer's array over NVLink happens inside this kernel, block by block, instead of in a copy kernel that
// re-reads the local array afterwards.  Positions that land in a neighbouring block's ranks (a bucket may reach 31
// ranks across the boundary) go out as single stores right away; slots nobody wrote keep the sentinel and are
// skipped (a neighbour, or nobody -- ranks outside [lo, hi) -- owns them).
constexpr uint32_t FSENT = 0xFFFFFFFFu;      // not a text position: n < 2^32 - 1024

template <bool STAGED>
__device__ __forceinline__ void place(const FinishOut& o, uint32_t* s_out, int64_t b0, int64_t x, uint32_t pos)
{
    if (!STAGED) { put(o, x, pos); return; }
    if (x < (int64_t)o.lo || x >= (int64_t)o.hi) return;
    const int64_t j = x - b0;
    if (j >= 0 && j < FT) { s_out[j] = pos; return; }
#pragma unroll 1
    for (int g = 0; g < o.n; ++g) __stcs(o.out[g] + x, pos);
}

template <bool STAGED>
__global__ void __launch_bounds__(FTHREADS)
finish_kernel(const uint32_t* __restrict__ sa_in, const uint16_t* __restrict__ lcp,
              const uint8_t* __restrict__ text, uint32_t n, uint32_t lprime, uint32_t r_begin, uint32_t r_end,
              const FinishOut outs, unsigned long long* __restrict__ run_counter, uint32_t* __restrict__ err,
              uint32_t block0)
{
    __shared__ __align__(16) int16_t s_lps[3 * FLCP];
    int16_t* const s_l = s_lps;              // adjacent LCPs, -1 outside [1, n)
    int16_t* const s_p = s_lps + FLCP;       // block prefix min of l, later block prefix max of m
    int16_t* const s_s = s_lps + 2 * FLCP;   // block suffix min of l, later block suffix max of m
    uint32_t* const s_out = reinterpret_cast<uint32_t*>(s_lps);   // STAGED: the block's FT output slots (phase D on: s_l/s_p/s_s are dead)
    static_assert(3 * FLCP * sizeof(int16_t) >= FT * sizeof(uint32_t), "output slots fit into the LCP arrays");
    __shared__ uint64_t s_hi[FEXT], s_lo[FEXT], s_tp[FEXT];   // window key words; tp = byte16 << 32 | pos
    __shared__ uint32_t s_head[FEXT / 32 + 2], s_sort[FEXT / 32 + 2];
    __shared__ uint32_t s_runs;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t b0 = (int64_t)(blockIdx.x + block0) * FT;
    uint64_t keep;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
    if (tid == 0) s_runs = 0;
    for (int j = tid; j < FLCP; j += FTHREADS) {
        const int64_t r = b0 - 64 + j;
        s_l[j] = (r >= 1 && r < (int64_t)n) ? (int16_t)__ldcs(lcp + r) : (int16_t)-1;
    }
    __syncthreads();
    // B1: per 32-block prefix / suffix minima of l
    for (int blk = warp; blk < FLCP / 32; blk += FWARPS) {
        const int j = blk * 32 + lane;
        const int v = s_l[j];
        s_p[j] = (int16_t)warp_prefix_min(v, lane);
        s_s[j] = (int16_t)warp_suffix_min(v, lane);
    }
    __syncthreads();
    // B2: m[j] = min(l[j+1 .. j+32]) = min(suffix[j+1], prefix[j+32]); per 32-block prefix / suffix maxima of m
    constexpr int MIT = (FLCP / 32 + FWARPS - 1) / FWARPS;
    int mreg[MIT];
#pragma unroll
    for (int it = 0; it < MIT; ++it) {
        const int j = (warp + it * FWARPS) * 32 + lane;
        mreg[it] = (j + 32 < FLCP) ? min((int)s_s[j + 1], (int)s_p[j + 32]) : -1;
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < MIT; ++it) {
        const int blk = warp + it * FWARPS;
        if (blk < FLCP / 32) {                                   // warp-uniform
            const int j = blk * 32 + lane;
            s_p[j] = (int16_t)warp_prefix_max(mreg[it], lane);
            s_s[j] = (int16_t)warp_suffix_max(mreg[it], lane);
        }
    }
    __syncthreads();

    // C: keys for ranks [b0-32, b0+FT+32), one warp-iteration per 32 consecutive ranks
    const int lim = lprime < 17u ? (int)lprime : 17;
    for (int eb = warp; eb < FEXT / 32; eb += FWARPS) {
        const int e = eb * 32 + lane;
        const int64_t r = b0 - 32 + e;
        const int j = e + 32;                                    // index into the lcp-space arrays
        uint64_t hi = 0, lo = 0, tp = 0;
        bool head = true, resort = false;
        if (r >= 0 && r < (int64_t)n) {
            // max(m[j-32 .. j]) = max(block suffix max at j-32, block prefix max at j)
            const int d = max((int)s_s[j - 32], (int)s_p[j]) + 1;
            const uint32_t pos = __ldcs(sa_in + r);
            const int lj = s_l[j];
            head = lj < d;
            tp = pos;
            const bool has_nb = (lj >= d) || ((int)s_l[j + 1] >= d);
            if ((uint32_t)d < lprime && has_nb) {
                // 32 aligned bytes that hold text[pos+d-1 .. pos+d+16]
                const uint64_t a = (uint64_t)pos + (uint32_t)d;
                const uint64_t a0 = d ? a - 1 : a;
                const uint64_t* wp = reinterpret_cast<const uint64_t*>(text + (a0 & ~7ull));
                const uint64_t w0 = ld_keep_u64(wp, keep), w1 = ld_keep_u64(wp + 1, keep), w2 = ld_keep_u64(wp + 2, keep),
                               w3 = ld_keep_u64(wp + 3, keep);
                uint32_t o = (uint32_t)(a - (a0 & ~7ull));       // 0..8: offset of the window in the 32 bytes
                bool term = false;
                if (d) {
                    const uint32_t ob = o - 1;
                    const uint32_t pb = (uint32_t)((ob < 8 ? w0 : w1) >> (8 * (ob & 7))) & 0xFFu;
                    term = pb == 0x0Au || pb == 0xFFu;
                }
                if (!term) {
                    uint64_t x0 = w0, x1 = w1, x2 = w2;
                    if (o == 8) { x0 = w1; x1 = w2; x2 = w3; o = 0; }
                    const uint32_t sh = o * 8;
                    uint64_t v0 = x0, v1 = x1;
                    if (sh) { v0 = (x0 >> sh) | (x1 << (64 - sh)); v1 = (x1 >> sh) | (x2 << (64 - sh)); }
                    hi = bswap64(v0);
                    lo = bswap64(v1);
                    uint32_t b16 = (uint32_t)(x2 >> sh) & 0xFFu;
                    // bytes after the first NUL of the 16-byte window do not take part (pcmpistri)
                    uint64_t z = zero_bytes(hi);
                    if (z) {
                        const int k = __clzll((long long)z) >> 3;
                        hi = k ? (hi & (~0ull << (64 - 8 * k))) : 0ull;
                        lo = 0;
                    } else if ((z = zero_bytes(lo)) != 0) {
                        const int k = __clzll((long long)z) >> 3;
                        lo = k ? (lo & (~0ull << (64 - 8 * k))) : 0ull;
                    }
                    // the idx < max_depth gate: only the first min(L', 17) bytes count
                    if (lim < 17) b16 = 0;
                    if (lim < 16) lo = (lim > 8) ? keep_top_bytes(lo, lim - 8) : 0ull;
                    if (lim < 8) hi = keep_top_bytes(hi, lim);
                    tp |= (uint64_t)b16 << 32;
                    resort = true;
                }
            }
        }
        s_hi[e] = hi; s_lo[e] = lo; s_tp[e] = tp;
        const uint32_t hb = __ballot_sync(0xffffffffu, head), rb = __ballot_sync(0xffffffffu, resort);
        if (lane == 0) { s_head[eb] = hb; s_sort[eb] = rb; }
    }
    if (tid == 0) { s_head[FEXT / 32] = 0xFFFFFFFFu; s_sort[FEXT / 32] = 0; }
    __syncthreads();
    if (STAGED) {
        for (int j = tid; j < FT; j += FTHREADS) s_out[j] = FSENT;
        __syncthreads();
    }

    // D: rank inside the bucket
    uint32_t my_runs = 0;
    for (int t = tid; t < FT; t += FTHREADS) {
        const int64_t r = b0 + t;
        if (r >= (int64_t)n) break;
        if (r < (int64_t)r_begin || r >= (int64_t)r_end) continue;    // halo record of a shard
        const int e = t + 32, q = e >> 5, i = e & 31;
        const uint64_t tp = s_tp[e];
        const uint32_t pos = (uint32_t)tp;
        if (!((s_sort[q] >> i) & 1u)) { place<STAGED>(outs, s_out, b0, r, pos); continue; }
        // bucket = [first, last]: last head at or before e, next head after e
        const uint64_t X = (((uint64_t)s_head[q] << 32) | s_head[q - 1]) & (~0ull >> (31 - i));
        const int first = (q - 1) * 32 + (63 - __clzll((long long)X));
        const uint64_t Y = ((((uint64_t)s_head[q + 1] << 32) | s_head[q]) >> i) >> 1;
        const int last = Y ? e + __ffsll((long long)Y) - 1 : e + 32;
        if (X == 0 || last - first + 1 > SA_SMALL) { atomicOr(err, SA_DEVERR_RUN_TOO_LONG); place<STAGED>(outs, s_out, b0, r, pos); continue; }
        const uint64_t hi = s_hi[e], lo = s_lo[e];
        uint32_t rank = 0;
        for (int x = first; x <= last; ++x)                 // branch-free (key, byte 16, position) compare
            rank -= less192(s_hi[x], s_lo[x], s_tp[x], hi, lo, tp);
        place<STAGED>(outs, s_out, b0, r - (e - first) + rank, pos);
        if (e == first) ++my_runs;
    }
    if (my_runs) atomicAdd(&s_runs, my_runs);
    __syncthreads();
    if (tid == 0 && s_runs) atomicAdd(run_counter, (unsigned long long)s_runs);
    if (STAGED) {
        // the block's slots, in groups of four aligned to 16 bytes in the output arrays (all replicas share the alignment:
        // same offset into 256-byte aligned allocations); a group with an empty slot or at the ragged ends goes out singly
        const int mis = (int)((reinterpret_cast<uintptr_t>(outs.out[0] + b0) >> 2) & 3u);
        for (int q = tid; q <= FT / 4; q += FTHREADS) {
            const int j0 = 4 * q - mis;
            uint32_t v[4];
            bool all = true;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u;
                v[u] = (j >= 0 && j < FT) ? s_out[j] : FSENT;
                all = all && v[u] != FSENT;
            }
            if (all) {
                const uint4 w = make_uint4(v[0], v[1], v[2], v[3]);
#pragma unroll 1
                for (int g = 0; g < outs.n; ++g) __stcs(reinterpret_cast<uint4*>(outs.out[g] + (b0 + j0)), w);
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (v[u] == FSENT) continue;
#pragma unroll 1
                    for (int g = 0; g < outs.n; ++g) __stcs(outs.out[g] + (b0 + j0 + u), v[u]);
                }
            }
        }
    }
}

}  // namespace sa

cudaError_t sa_launch_finish(const uint32_t* sa_in, const uint16_t* lcp, const uint8_t* text, uint32_t n,
                             uint32_t lprime, uint32_t r_begin, uint32_t r_end, const sa::FinishOut& outs,
                             unsigned long long* run_counter, uint32_t* err, cudaStream_t s, uint32_t block0,
                             uint32_t nblocks)
{
    if (!n) return cudaSuccess;
    const uint32_t blocks = sa_finish_blocks(n);
    if (block0 >= blocks) return cudaSuccess;
    if (nblocks > blocks - block0) nblocks = blocks - block0;
    // several replicas (a sharded build that replicates): the push into the peers' arrays is fused into the kernel
    if (outs.n > 1)
        sa::finish_kernel<true><<<nblocks, sa::FTHREADS, 0, s>>>(sa_in, lcp, text, n, lprime, r_begin, r_end, outs, run_counter,
                                                                 err, block0);
    else
        sa::finish_kernel<false><<<nblocks, sa::FTHREADS, 0, s>>>(sa_in, lcp, text, n, lprime, r_begin, r_end, outs, run_counter,
                                                                  err, block0);
    return cudaGetLastError();
}

uint32_t sa_finish_blocks(uint32_t n) { return (n + sa::FT - 1) / sa::FT; }
uint32_t sa_finish_block_ranks() { return sa::FT; }
