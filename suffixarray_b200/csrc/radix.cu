// radix.cu -- onesweep-style LSD radix sort of (u64 key, u32 value) records for sm_100a.
//
// Replaces the per-byte counting sorts of recursive_bucket_sort (engine.c:734-758): instead
// of one pass over a bucket per key byte with random text reads, the suffix keys are packed
// once and sorted 8 bits at a time with stable scatter passes.
//
//   hist_kernel      one read of the keys, all digit-place histograms (shared-memory atomics)
//   scan_hist_kernel exclusive scan of each 256-bin histogram + "single bin" skip flags
//   onesweep_kernel  one stable scatter pass: warp-level match/ballot ranking into per-warp
//                    digit counters, per-tile digit counts chained between tiles by decoupled
//                    look-back (one launch per pass, keys read once), tile reordered in shared
//                    memory so global stores are coalesced runs.
//
// Algorithmic traffic of one pass over m records: read 12 m, write 12 m bytes (HBM bound).
#include "common.cuh"
#include "kernels.h"
#include "onesweep.cuh"

namespace sa {

// ------------------------------------------------------------------------------ histograms
__global__ void __launch_bounds__(512) hist_kernel(const uint64_t* __restrict__ keys, uint32_t n, int shift0,
                                                    int npasses, uint32_t* __restrict__ hist)
{
    __shared__ uint32_t h[8 * 256];
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t k = ld_stream_u64(keys + i);
        k >>= shift0;
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            if (p < npasses) atomicAdd(&h[p * 256 + (uint32_t)(k & 0xFF)], 1u);
            k >>= 8;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npasses * 256; i += blockDim.x) {
        uint32_t v = h[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

// one block of 256 threads per digit place: base = exclusive scan, skip[p] = 1 when one bin
// holds all n records (the pass would be the identity permutation).
__global__ void __launch_bounds__(256) scan_hist_kernel(const uint32_t* __restrict__ hist, uint32_t n,
                                                         uint32_t* __restrict__ base, uint32_t* __restrict__ skip)
{
    __shared__ uint32_t wsum[8];
    __shared__ uint32_t any_full;
    const int p = blockIdx.x, t = threadIdx.x;
    if (t == 0) any_full = 0;
    __syncthreads();
    uint32_t v = hist[p * 256 + t];
    if (v == n) any_full = 1;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if ((t & 31) >= o) inc += y;
    }
    if ((t & 31) == 31) wsum[t >> 5] = inc;
    __syncthreads();
    uint32_t pre = 0;
    for (int w = 0; w < (t >> 5); ++w) pre += wsum[w];
    base[p * 256 + t] = pre + inc - v;
    if (t == 0) skip[p] = any_full;
}

}  // namespace sa

// ------------------------------------------------------------------------------ launchers
namespace {
// Configuration chosen on B200 with scripts/bench_onesweep.cu (profiles/r01_onesweep_variants.md):
// 8 ranking warps x 12 records per thread (3072-record tiles) + 8 look-back warps, 2 persistent
// CTAs per SM, match.any ranking, 4-wide look-back window, 24 / 104 registers per thread.
constexpr int OS_IPT = 12, OS_LBW = 4, OS_LB_REGS = 24;
constexpr int OS_TILE = 256 * OS_IPT;
}  // namespace

uint32_t sa_onesweep_tile_items() { return OS_TILE; }   // smallest tile of any variant: sizes the status words

cudaError_t sa_launch_hist(const uint64_t* keys, uint32_t n, int shift0, int npasses, uint32_t* hist,
                           int num_sms, cudaStream_t s)
{
    cudaError_t e = cudaMemsetAsync(hist, 0, 8 * 256 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    if (n == 0) return cudaSuccess;
    uint32_t blocks = (n + 511) / 512;
    uint32_t cap = (uint32_t)num_sms * 4;
    if (blocks > cap) blocks = cap;
    sa::hist_kernel<<<blocks, 512, 0, s>>>(keys, n, shift0, npasses, hist);
    return cudaGetLastError();
}

cudaError_t sa_launch_scan_hist(const uint32_t* hist, uint32_t n, int npasses, uint32_t* base, uint32_t* skip,
                                cudaStream_t s)
{
    sa::scan_hist_kernel<<<npasses, 256, 0, s>>>(hist, n, base, skip);
    return cudaGetLastError();
}

cudaError_t sa_launch_onesweep(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                               uint32_t* vals_out, uint32_t n, int shift, const uint32_t* digit_base,
                               uint32_t* status, uint32_t* tile_counter, uint32_t* err, bool identity_vals,
                               cudaStream_t s)
{
    return sa::launch_onesweep4_cfg<OS_IPT, 2, OS_LBW, OS_LB_REGS>(keys_in, vals_in, keys_out, vals_out, n, shift, digit_base,
                                                                   status, tile_counter, err, identity_vals, s);
}
