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

// ------------------------------------------------------------------------------ onesweep
constexpr uint32_t ST_PARTIAL = 1u << 30;
constexpr uint32_t ST_INCLUSIVE = 2u << 30;
constexpr uint32_t ST_MASK = (1u << 30) - 1u;
constexpr uint32_t LOOKBACK_SPIN_LIMIT = 1u << 26;   // a hung look-back sets d_err and exits

template <int THREADS, int IPT, bool IDENTITY_VALS>
__global__ void __launch_bounds__(THREADS)
onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift,
                const uint32_t* __restrict__ digit_base, volatile uint32_t* status, uint32_t* tile_counter,
                uint32_t* err)
{
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * IPT;
    static_assert(THREADS >= 256, "one thread per digit");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* keys_s = reinterpret_cast<uint64_t*>(smem_raw);                 // TILE
    uint32_t* vals_s = reinterpret_cast<uint32_t*>(keys_s + TILE);            // TILE
    uint32_t* whist = vals_s + TILE;                                           // WARPS*256
    uint32_t* tile_start = whist + WARPS * 256;                                // 256
    uint32_t* gofs = tile_start + 256;                                         // 256
    uint32_t* wsum = gofs + 256;                                               // 8
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < WARPS * 256; i += THREADS) whist[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t base64 = (uint64_t)tile * TILE;
    if (base64 >= n) return;
    const uint32_t base = (uint32_t)base64;
    const uint32_t tile_n = (n - base) < (uint32_t)TILE ? (n - base) : (uint32_t)TILE;
    const uint32_t wbase = base + warp * (32 * IPT) + lane;

    uint64_t key[IPT];
    uint32_t rank[IPT];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        uint32_t idx = wbase + i * 32;
        key[i] = (idx < n) ? ld_stream_u64(keys_in + idx) : ~0ull;
    }
    const uint32_t lt = lanemask_lt();
    uint32_t* myhist = whist + warp * 256;
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const bool valid = (wbase + i * 32) < n;
        const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
        const uint32_t mval = valid ? d : (256u + lane);
        const uint32_t peers = __match_any_sync(0xffffffffu, mval);
        const int leader = __ffs(peers) - 1;
        uint32_t b = 0;
        if (lane == leader && valid) {
            b = myhist[d];
            myhist[d] = b + __popc(peers);
        }
        b = __shfl_sync(0xffffffffu, b, leader);
        rank[i] = b + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();

    // per digit: exclusive scan over warps, tile count, decoupled look-back over earlier tiles
    uint32_t count = 0;
    if (tid < 256) {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            uint32_t c = whist[w * 256 + tid];
            whist[w * 256 + tid] = run;
            run += c;
        }
        count = run;
        volatile uint32_t* mine = status + (uint64_t)tile * 256 + tid;
        uint32_t excl = 0;
        if (tile == 0) {
            *mine = ST_INCLUSIVE | count;
        } else {
            *mine = ST_PARTIAL | count;
            uint32_t spins = 0;
            for (int64_t p = (int64_t)tile - 1; p >= 0; --p) {
                volatile uint32_t* st = status + (uint64_t)p * 256 + tid;
                uint32_t v;
                while (((v = *st) & (ST_PARTIAL | ST_INCLUSIVE)) == 0) {
                    if (++spins > LOOKBACK_SPIN_LIMIT) { atomicOr(err, SA_DEVERR_LOOKBACK_TIMEOUT); v = ST_INCLUSIVE; break; }
                }
                excl += v & ST_MASK;
                if (v & ST_INCLUSIVE) break;
            }
            *mine = ST_INCLUSIVE | (excl + count);
        }
        gofs[tid] = digit_base[tid] + excl;     // global start of this tile's digit run
    }
    // block exclusive scan of the 256 digit counts -> start of each digit inside the tile
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (tid < 256 && lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (tid < 256) {
        uint32_t pre = 0;
        for (int w = 0; w < warp; ++w) pre += wsum[w];
        const uint32_t ts = pre + inc - count;
        tile_start[tid] = ts;
        gofs[tid] -= ts;                        // global index = gofs[d] + position in tile
    }
    __syncthreads();

    // reorder the tile in shared memory (stable: digit, then warp, then item order)
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const uint32_t idx = wbase + i * 32;
        if (idx < n) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
            const uint32_t lp = tile_start[d] + myhist[d] + rank[i];
            keys_s[lp] = key[i];
            vals_s[lp] = IDENTITY_VALS ? idx : ld_stream_u32(vals_in + idx);
        }
    }
    __syncthreads();
#pragma unroll 4
    for (uint32_t j = tid; j < tile_n; j += THREADS) {
        const uint64_t k = keys_s[j];
        const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
        keys_out[g] = k;
        vals_out[g] = vals_s[j];
    }
}

}  // namespace sa

// ------------------------------------------------------------------------------ launchers
namespace {
constexpr int OS_THREADS = 256;
constexpr int OS_IPT = 16;
constexpr int OS_TILE = OS_THREADS * OS_IPT;
constexpr size_t OS_SMEM = (size_t)OS_TILE * 12 + (OS_THREADS / 32) * 256 * 4 + 256 * 4 * 2 + 64;
}  // namespace

uint32_t sa_onesweep_tile_items() { return OS_TILE; }

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
    if (n == 0) return cudaSuccess;
    const uint32_t tiles = (n + OS_TILE - 1) / OS_TILE;
    cudaError_t e = cudaMemsetAsync(status, 0, (size_t)tiles * 256 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(tile_counter, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(sa::onesweep_kernel<OS_THREADS, OS_IPT, true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OS_SMEM);
        cudaFuncSetAttribute(sa::onesweep_kernel<OS_THREADS, OS_IPT, false>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OS_SMEM);
        attr_set = true;
    }
    if (identity_vals)
        sa::onesweep_kernel<OS_THREADS, OS_IPT, true><<<tiles, OS_THREADS, OS_SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, digit_base, status, tile_counter, err);
    else
        sa::onesweep_kernel<OS_THREADS, OS_IPT, false><<<tiles, OS_THREADS, OS_SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, digit_base, status, tile_counter, err);
    return cudaGetLastError();
}
