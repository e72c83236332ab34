// finish.cu -- the "<= 32 bucket" finishing step of ORDER-E, data-parallel.
//
// recursive_bucket_sort stops splitting a bucket as soon as it has <= 32 members and finishes
// it with a stable insertion sort under strncmp_128 (engine.c:710-729, :664-677).  On the
// sorted-by-prefix array those buckets are contiguous, and the depth at which rank r's bucket
// first has <= 32 members is
//     d*(r) = 1 + max over the 33-rank windows containing r of (min adjacent LCP inside it)
// (0 when n <= 32).  Each block takes 1024 ranks (+32 halo each side): sliding min/max over the
// LCPs in shared memory, a 17-byte window key per member (gathered from the L2-resident text),
// then every rank counts the members of its own bucket that sort before it.
//   bucket is left alone when d* >= L' (engine.c:706), when it has one member, or when the
//   shared d*-byte prefix holds '\n' / 0xFF (such buckets are never entered, engine.c:771-773).
#include "common.cuh"
#include "kernels.h"

namespace sa {

constexpr int FT = 1024;             // ranks per block
constexpr int FEXT = FT + 64;        // ranks with keys: [b0-32, b0+FT+32)
constexpr int FLCP = FT + 129;       // lcp entries: ranks [b0-64, b0+FT+65)
constexpr int FTHREADS = 256;

__global__ void __launch_bounds__(FTHREADS)
finish_kernel(const uint32_t* __restrict__ sa_in, const uint16_t* __restrict__ lcp,
              const uint8_t* __restrict__ text, uint32_t n, uint32_t lprime, uint32_t* __restrict__ sa_out,
              unsigned long long* __restrict__ run_counter, uint32_t* __restrict__ err)
{
    __shared__ int16_t s_l[FLCP];        // adjacent LCPs, -1 outside [1, n)
    __shared__ int16_t s_a[FLCP];
    __shared__ int16_t s_b[FLCP];
    __shared__ int16_t s_m[FLCP];        // window minima
    __shared__ uint64_t s_hi[FEXT], s_lo[FEXT];
    __shared__ uint32_t s_pos[FEXT];
    __shared__ uint32_t s_meta[FEXT];    // bits 0-7: 17th byte, bit 8: bucket gets re-sorted
    __shared__ int16_t s_d[FEXT];
    __shared__ uint32_t s_runs;

    const int tid = threadIdx.x;
    const int64_t b0 = (int64_t)blockIdx.x * FT;
    if (tid == 0) s_runs = 0;
    for (int j = tid; j < FLCP; j += FTHREADS) {
        const int64_t r = b0 - 64 + j;
        s_l[j] = (r >= 1 && r < (int64_t)n) ? (int16_t)lcp[r] : (int16_t)-1;
    }
    __syncthreads();
    // m[j] = min(l[j+1 .. j+32]) by doubling: widths 2,4,8,16,32
    for (int j = tid; j < FLCP; j += FTHREADS) s_a[j] = (j + 2 < FLCP) ? min(s_l[j + 1], s_l[j + 2]) : (int16_t)-1;
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_b[j] = (j + 2 < FLCP) ? min(s_a[j], s_a[j + 2]) : (int16_t)-1;
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_a[j] = (j + 4 < FLCP) ? min(s_b[j], s_b[j + 4]) : (int16_t)-1;
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_b[j] = (j + 8 < FLCP) ? min(s_a[j], s_a[j + 8]) : (int16_t)-1;
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_m[j] = (j + 16 < FLCP) ? min(s_b[j], s_b[j + 16]) : (int16_t)-1;
    __syncthreads();
    // M[j] = max(m[j-32 .. j]): widths 2,4,8,16,32 then one more element
    for (int j = tid; j < FLCP; j += FTHREADS) s_a[j] = (j >= 1) ? max(s_m[j], s_m[j - 1]) : s_m[j];
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_b[j] = (j >= 2) ? max(s_a[j], s_a[j - 2]) : s_a[j];
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_a[j] = (j >= 4) ? max(s_b[j], s_b[j - 4]) : s_b[j];
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_b[j] = (j >= 8) ? max(s_a[j], s_a[j - 8]) : s_a[j];
    __syncthreads();
    for (int j = tid; j < FLCP; j += FTHREADS) s_a[j] = (j >= 16) ? max(s_b[j], s_b[j - 16]) : s_b[j];
    __syncthreads();

    // keys for ranks [b0-32, b0+FT+32)
    const int lim = lprime < 17u ? (int)lprime : 17;
    for (int e = tid; e < FEXT; e += FTHREADS) {
        const int64_t r = b0 - 32 + e;
        const int j = e + 32;                                    // index into the lcp-space arrays
        uint32_t meta = 0, pos = 0;
        uint64_t hi = 0, lo = 0;
        int16_t d = 0;
        if (r >= 0 && r < (int64_t)n) {
            d = (int16_t)(max(s_a[j], s_m[j - 32]) + 1);
            pos = sa_in[r];
            const bool has_nb = (s_l[j] >= d) || (s_l[j + 1] >= d);
            if ((uint32_t)d < lprime && has_nb) {
                bool term = false;
                for (int w = 0; w < d; w += 8) {
                    uint64_t be = load_be64(text, (uint64_t)pos + w);
                    const int kb = (d - w) < 8 ? (d - w) : 8;
                    term |= terminator_bytes(keep_top_bytes(be, kb)) != 0;
                }
                if (!term) {
                    const uint64_t q = (uint64_t)pos + (uint32_t)d;
                    hi = load_be64(text, q);
                    lo = load_be64(text, q + 8);
                    uint32_t b16 = text[q + 16];
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
                    meta = b16 | 0x100u;
                }
            }
        }
        s_hi[e] = hi; s_lo[e] = lo; s_pos[e] = pos; s_meta[e] = meta; s_d[e] = d;
    }
    __syncthreads();

    uint32_t my_runs = 0;
    for (int t = tid; t < FT; t += FTHREADS) {
        const int64_t r = b0 + t;
        if (r >= (int64_t)n) break;
        const int e = t + 32, j = t + 64;
        const uint32_t pos = s_pos[e], meta = s_meta[e];
        if (!(meta & 0x100u)) { sa_out[r] = pos; continue; }
        const int d = s_d[e];
        int back = 0, fwd = 0;
        while (back < SA_SMALL && s_l[j - back] >= d) ++back;      // members before r
        while (fwd < SA_SMALL && s_l[j + 1 + fwd] >= d) ++fwd;      // members after r
        if (back + fwd + 1 > SA_SMALL) { atomicOr(err, SA_DEVERR_RUN_TOO_LONG); sa_out[r] = pos; continue; }
        const uint64_t hi = s_hi[e], lo = s_lo[e];
        const uint32_t b16 = meta & 0xFFu;
        uint32_t rank = 0;
        for (int x = e - back; x <= e + fwd; ++x) {
            const uint64_t h2 = s_hi[x], l2 = s_lo[x];
            const uint32_t c2 = s_meta[x] & 0xFFu, p2 = s_pos[x];
            bool less;
            if (h2 != hi) less = h2 < hi;
            else if (l2 != lo) less = l2 < lo;
            else if (c2 != b16) less = c2 < b16;
            else less = p2 < pos;
            rank += less;
        }
        sa_out[r - back + rank] = pos;
        if (back == 0) ++my_runs;
    }
    if (my_runs) atomicAdd(&s_runs, my_runs);
    __syncthreads();
    if (tid == 0 && s_runs) atomicAdd(run_counter, (unsigned long long)s_runs);
}

}  // namespace sa

cudaError_t sa_launch_finish(const uint32_t* sa_in, const uint16_t* lcp, const uint8_t* text, uint32_t n,
                             uint32_t lprime, uint32_t* sa_out, unsigned long long* run_counter, uint32_t* err,
                             cudaStream_t s)
{
    if (!n) return cudaSuccess;
    const uint32_t blocks = (n + sa::FT - 1) / sa::FT;
    sa::finish_kernel<<<blocks, sa::FTHREADS, 0, s>>>(sa_in, lcp, text, n, lprime, sa_out, run_counter, err);
    return cudaGetLastError();
}
