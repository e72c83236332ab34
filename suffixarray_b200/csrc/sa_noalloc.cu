// sa_noalloc.cu -- ONE entry point that allocates nothing: sa_hit_offsets_noalloc, the hit-offset scan of a
// query batch (per query min(k, end - start + 1) hits, engine.c:1181-1185, exclusive prefix sum in u64) for
// callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid allocation.  It is a variant of the hit_counts / block_sum / scan / offsets kernels of query.cu
// behind sa_query_ranges_r (which keep their scratch -- per-query counts and per-block sums -- in buffers
// grown with cudaMalloc): here the counts are recomputed instead of stored and the per-block sums live in a
// file-scope __device__ array of fixed size, so every GPU that loads the library holds
// (SA_NOALLOC_MAX_QUERIES / 1024 + 1) * 8 bytes = 128 KiB + 8 B of it until the process exits.
//
// No counterpart in the reference.
//
// Rules of the scratch array:
//  * host code never uses the array's NAME as a pointer (in host code it is a host-side stand-in, not a
//    device address): at set-up, once per device and with that device current, cudaGetSymbolAddress gives
//    the device address, which is what the kernels receive;
//  * every call shares the array and it keeps the last call's contents.  No word of it is read before the
//    same call has written it (block b's sum is stored by the first kernel before the scan kernel reads it;
//    the scan kernel stores the total behind the last block), so nothing has to start at zero -- the slot
//    of the total is cleared all the same, so that a batch of zero blocks cannot see a stale one;
//  * calls that overlap on the GPU (two streams, two host threads on two streams) RACE on the array: the
//    caller serialises them (one stream, or events between streams).  Calls on one stream are ordered.
#include <atomic>

#include "sa_index.h"

#define NA_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            sa_set_error(__VA_ARGS__);         \
            return (code);                     \
        }                                      \
    } while (0)

namespace {

constexpr uint32_t NA_BLOCK = 1024;                                          // queries per block sum
constexpr uint32_t NA_MAX_BLOCKS = (uint32_t)(SA_NOALLOC_MAX_QUERIES / NA_BLOCK);
constexpr int NA_MAX_DEVICES = 64;

// the scratch: one u64 per block of 1024 queries + the total
static __device__ __align__(256) unsigned long long g_noalloc_sums[NA_MAX_BLOCKS + 1];

// device address of g_noalloc_sums on each device (set up on the first call that uses the device)
std::atomic<unsigned long long*> g_sums_addr[NA_MAX_DEVICES];

// engine.c:1181-1185 (same clamps as hit_counts_kernel, query.cu)
__device__ __forceinline__ uint32_t na_hit_count(uint32_t s, uint32_t e, uint32_t k, uint32_t n)
{
    uint32_t c = 0;
    if (s != 0xFFFFFFFFu) {
        c = e - s + 1u;
        if (c > k) c = k;
        if (s >= n) c = 0;
        else if (c > n - s) c = n - s;
    }
    return c;
}

// sum over a 1024-thread block; the result is valid in every thread
__device__ __forceinline__ unsigned long long na_block_sum(unsigned long long v, unsigned long long* ws)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long t = ws[threadIdx.x & 31];
#pragma unroll
    for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    __syncthreads();
    return t;
}

// exclusive scan over a 1024-thread block; *total receives the block's sum (valid in every thread)
__device__ __forceinline__ unsigned long long na_block_exclusive(unsigned long long a, unsigned long long* ws,
                                                                 unsigned long long* total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long ia = a;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, ia, o);
        if (lane >= o) ia += y;
    }
    if (lane == 31) ws[warp] = ia;
    __syncthreads();
    if (warp == 0) {
        const unsigned long long v = ws[lane];
        unsigned long long s2 = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, s2, o);
            if (lane >= o) s2 += y;
        }
        ws[lane] = s2 - v;
        if (lane == 31) ws[32] = s2;
    }
    __syncthreads();
    const unsigned long long ex = ws[warp] + ia - a;
    *total = ws[32];
    __syncthreads();
    return ex;
}

__global__ void __launch_bounds__(NA_BLOCK)
na_block_sums_kernel(const uint32_t* __restrict__ start, const uint32_t* __restrict__ end, uint32_t nq, uint32_t k,
                     uint32_t n, unsigned long long* __restrict__ sums)
{
    __shared__ unsigned long long ws[33];
    const uint32_t i = blockIdx.x * NA_BLOCK + threadIdx.x;
    const unsigned long long c = i < nq ? na_hit_count(start[i], end[i], k, n) : 0u;
    const unsigned long long t = na_block_sum(c, ws);
    if (threadIdx.x == 0) sums[blockIdx.x] = t;
}

// in place: sums[b] -> sum of the blocks in front of b; sums[nb] = the total
__global__ void __launch_bounds__(NA_BLOCK) na_scan_sums_kernel(unsigned long long* __restrict__ sums, uint32_t nb)
{
    __shared__ unsigned long long ws[33];
    unsigned long long carry = 0;
    for (uint32_t base = 0; base < nb; base += NA_BLOCK) {
        const uint32_t i = base + threadIdx.x;
        const unsigned long long a = i < nb ? sums[i] : 0ull;
        unsigned long long total;
        const unsigned long long ex = na_block_exclusive(a, ws, &total);
        if (i < nb) sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) sums[nb] = carry;
}

__global__ void __launch_bounds__(NA_BLOCK)
na_offsets_kernel(const uint32_t* __restrict__ start, const uint32_t* __restrict__ end, uint32_t nq, uint32_t k,
                  uint32_t n, const unsigned long long* __restrict__ sums, uint32_t nb,
                  unsigned long long* __restrict__ hit_off)
{
    __shared__ unsigned long long ws[33];
    const uint32_t i = blockIdx.x * NA_BLOCK + threadIdx.x;
    const unsigned long long c = i < nq ? na_hit_count(start[i], end[i], k, n) : 0u;
    unsigned long long total;
    const unsigned long long ex = na_block_exclusive(c, ws, &total);
    if (i < nq) hit_off[i] = sums[blockIdx.x] + ex;
    if (i + 1 == nq || (nq == 0 && i == 0)) hit_off[nq] = sums[nb];     // the total (0 for an empty batch)
}

}  // namespace

// set-up, once per device: the device address of the scratch array (the device must be current)
static int na_scratch_address(int device, unsigned long long** out)
{
    NA_REQUIRE(device >= 0 && device < NA_MAX_DEVICES, SA_ERR_LIMIT, "sa_hit_offsets_noalloc: device %d >= %d", device,
               NA_MAX_DEVICES);
    unsigned long long* p = g_sums_addr[device].load(std::memory_order_acquire);
    if (!p) {
        void* sym = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_sums);
        if (e != cudaSuccess) return sa_cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_sums)");
        NA_REQUIRE(sym, SA_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_sums) returned NULL");
        p = static_cast<unsigned long long*>(sym);
        g_sums_addr[device].store(p, std::memory_order_release);     // two racing threads store the same value
    }
    *out = p;
    return SA_OK;
}

// include/sa_cuda.h: no cudaMalloc / cudaFree / allocating routine on this path, no synchronisation either:
// three kernels are queued on `stream` and the call returns.
extern "C" int sa_hit_offsets_noalloc(sa_index* idx, const uint32_t* d_start, const uint32_t* d_end, uint64_t nq,
                                      uint32_t k, uint64_t* d_hit_off, void* stream)
{
    NA_REQUIRE(idx, SA_ERR_INVALID, "index is NULL");
    NA_REQUIRE(d_hit_off, SA_ERR_INVALID, "sa_hit_offsets_noalloc: d_hit_off is NULL");
    NA_REQUIRE(nq <= SA_NOALLOC_MAX_QUERIES, SA_ERR_LIMIT,
               "sa_hit_offsets_noalloc: %llu queries exceed the fixed scratch (SA_NOALLOC_MAX_QUERIES = %llu); split the batch "
               "or use sa_query_ranges_r",
               (unsigned long long)nq, (unsigned long long)SA_NOALLOC_MAX_QUERIES);
    NA_REQUIRE(nq == 0 || (d_start && d_end), SA_ERR_INVALID, "sa_hit_offsets_noalloc: NULL device pointer");
    std::shared_lock<std::shared_mutex> rd(idx->mu);
    NA_REQUIRE(idx->built, SA_ERR_INVALID, "index not built");
    SA_CU(cudaSetDevice(idx->device));
    unsigned long long* sums = nullptr;
    int rc = na_scratch_address(idx->device, &sums);
    if (rc != SA_OK) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const uint32_t n32 = (uint32_t)idx->n;
    const uint32_t nb = (uint32_t)((nq + NA_BLOCK - 1) / NA_BLOCK);
    unsigned long long* off = reinterpret_cast<unsigned long long*>(d_hit_off);
    if (nb) na_block_sums_kernel<<<nb, NA_BLOCK, 0, s>>>(d_start, d_end, (uint32_t)nq, k, n32, sums);
    na_scan_sums_kernel<<<1, NA_BLOCK, 0, s>>>(sums, nb);            // nb == 0: clears the slot of the total
    if (nb) na_offsets_kernel<<<nb, NA_BLOCK, 0, s>>>(d_start, d_end, (uint32_t)nq, k, n32, sums, nb, off);
    else na_offsets_kernel<<<1, 32, 0, s>>>(nullptr, nullptr, 0, k, n32, sums, 0, off);
    SA_CU(cudaGetLastError());
    return SA_OK;
}
