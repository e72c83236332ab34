// onesweep_variants.cuh -- the earlier forms of the onesweep pass (v1: one CTA per tile; v2: persistent
// CTAs with in-line windowed look-back, optionally with a bulk-copy prefetch of the next tile), kept
// for scripts/bench_onesweep.cu only.  libsa_cuda.so builds the warp-specialised v4 pass alone
// (suffixarray_b200/csrc/onesweep.cuh); profiles/r01_onesweep_variants.md has the comparison.
#pragma once
#include "../suffixarray_b200/csrc/onesweep.cuh"

namespace sa {

template <int THREADS, int IPT>
constexpr size_t onesweep_smem_bytes()
{
    return (size_t)THREADS * IPT * 12 + (size_t)(THREADS / 32) * 256 * 4 + 256 * 4 * 2 + 64;
}

template <int THREADS, int IPT, int MINB, bool IDENTITY_VALS, int RANK_MODE>
__global__ void __launch_bounds__(THREADS, MINB)
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
    const bool full = tile_n == (uint32_t)TILE;

    uint64_t key[IPT];
    uint32_t val[IPT];
    uint32_t rank[IPT];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const uint32_t idx = wbase + i * 32;
        key[i] = (full || idx < n) ? ld_stream_u64(keys_in + idx) : ~0ull;
    }
    if (!IDENTITY_VALS) {
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const uint32_t idx = wbase + i * 32;
            val[i] = (full || idx < n) ? ld_stream_u32(vals_in + idx) : 0u;
        }
    }
    const uint32_t lt = lanemask_lt();
    uint32_t* myhist = whist + warp * 256;
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const bool valid = full || (wbase + i * 32) < n;
        const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
        const uint32_t peers = digit_peers<RANK_MODE>(d, valid, lane);
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
            *mine = st_partial(count);
            uint32_t spins = 0;
            for (int64_t p = (int64_t)tile - 1; p >= 0; --p) {
                volatile uint32_t* st = status + (uint64_t)p * 256 + tid;
                uint32_t v;
                while ((v = *st) == 0) {
                    if (++spins > LOOKBACK_SPIN_LIMIT) { atomicOr(err, SA_DEVERR_LOOKBACK_TIMEOUT); v = ST_INCLUSIVE; break; }
                }
                excl += st_value(v);
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
        if (full || idx < n) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
            const uint32_t lp = tile_start[d] + myhist[d] + rank[i];
            keys_s[lp] = key[i];
            vals_s[lp] = IDENTITY_VALS ? idx : val[i];
        }
    }
    __syncthreads();
    if (full) {
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const uint32_t j = tid + i * THREADS;
            const uint64_t k = keys_s[j];
            const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
            keys_out[g] = k;
            vals_out[g] = vals_s[j];
        }
    } else {
        for (uint32_t j = tid; j < tile_n; j += THREADS) {
            const uint64_t k = keys_s[j];
            const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
            keys_out[g] = k;
            vals_out[g] = vals_s[j];
        }
    }
}

// --- bulk-copy (TMA, 1-D) prefetch of the next tile into a shared-memory stage, completion on an mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra W_%=;\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

template <int THREADS, int IPT, bool PF>
constexpr size_t onesweep2_smem_bytes()
{
    return onesweep_smem_bytes<THREADS, IPT>() + (PF ? (size_t)THREADS * IPT * 12 + 16 : 0);
}

// PF: the next tile's keys / values are bulk-copied into a second shared-memory stage while the
// current tile is scanned, re-ordered and written out (needs 16-byte aligned inputs).
template <int THREADS, int IPT, int MINB, bool IDENTITY_VALS, int RANK_MODE, int LBW, bool PF>
__global__ void __launch_bounds__(THREADS, MINB)
onesweep2_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift,
                 const uint32_t* __restrict__ digit_base, uint32_t* status, uint32_t* tile_counter,
                 uint32_t* err, uint32_t num_tiles)
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
    uint32_t* wsum = gofs + 256;                                               // 16
    uint64_t* bar = reinterpret_cast<uint64_t*>(wsum + 16);                    // PF: mbarrier
    uint64_t* stage_k = bar + 2;                                               // PF: TILE keys
    uint32_t* stage_v = reinterpret_cast<uint32_t*>(stage_k + TILE);           // PF: TILE values
    __shared__ uint32_t s_tile[2];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t lt = lanemask_lt();
    uint32_t* myhist = whist + warp * 256;
    const uint32_t full_tiles = n / (uint32_t)TILE;          // tiles below this index are complete
    uint32_t phase = 0, next_tile = 0;
    if (tid == 0) {
        const uint32_t t0 = atomicAdd(tile_counter, 1u);
        s_tile[0] = t0;
        if (PF) {
            mbar_init(bar, 1);
            if (t0 < full_tiles) {
                mbar_expect_tx(bar, (uint32_t)TILE * (IDENTITY_VALS ? 8u : 12u));
                bulk_g2s(stage_k, keys_in + (uint64_t)t0 * TILE, (uint32_t)TILE * 8u, bar);
                if (!IDENTITY_VALS) bulk_g2s(stage_v, vals_in + (uint64_t)t0 * TILE, (uint32_t)TILE * 4u, bar);
            }
        }
    }
    int buf = 0;
    OS_CLK_INIT;
    for (;;) {
        for (int i = tid; i < WARPS * 256; i += THREADS) whist[i] = 0;
        __syncthreads();                                     // s_tile[buf] + zeroed counters visible
        OS_CLK(0);
        const uint32_t tile = s_tile[buf];
        if (tile >= num_tiles) return;
        if (tid == 0) {
            next_tile = atomicAdd(tile_counter, 1u);         // next tile, latency hidden
            s_tile[buf ^ 1] = next_tile;
        }
        buf ^= 1;
        const uint32_t base = tile * (uint32_t)TILE;
        const uint32_t tile_n = (n - base) < (uint32_t)TILE ? (n - base) : (uint32_t)TILE;
        const uint32_t wbase = base + warp * (32 * IPT) + lane;
        const bool full = tile_n == (uint32_t)TILE;

        uint64_t key[IPT];
        uint32_t val[IPT];
        uint32_t rank[IPT];
        if (PF && full) {
            mbar_wait(bar, phase);
            phase ^= 1u;
            const int sb = warp * (32 * IPT) + lane;
#pragma unroll
            for (int i = 0; i < IPT; ++i) key[i] = stage_k[sb + i * 32];
            if (!IDENTITY_VALS) {
#pragma unroll
                for (int i = 0; i < IPT; ++i) val[i] = stage_v[sb + i * 32];
            }
        } else {
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                const uint32_t idx = wbase + i * 32;
                key[i] = (full || idx < n) ? ld_stream_u64(keys_in + idx) : ~0ull;
            }
            if (!IDENTITY_VALS) {
#pragma unroll
                for (int i = 0; i < IPT; ++i) {
                    const uint32_t idx = wbase + i * 32;
                    val[i] = (full || idx < n) ? ld_stream_u32(vals_in + idx) : 0u;
                }
            }
        }
#ifdef OS_PHASE_CLOCKS
        if (tid == 0 && key[IPT - 1] == 0x123456789ull) clk_ += 1;   // wait for the loads
#endif
        OS_CLK(1);
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const bool valid = full || (wbase + i * 32) < n;
            const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
            const uint32_t peers = digit_peers<RANK_MODE>(d, valid, lane);
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
        OS_CLK(2);
        if (PF && tid == 0 && next_tile < full_tiles) {       // the stage has been consumed: refill it
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(bar, (uint32_t)TILE * (IDENTITY_VALS ? 8u : 12u));
            bulk_g2s(stage_k, keys_in + (uint64_t)next_tile * TILE, (uint32_t)TILE * 8u, bar);
            if (!IDENTITY_VALS) bulk_g2s(stage_v, vals_in + (uint64_t)next_tile * TILE, (uint32_t)TILE * 4u, bar);
        }

        // per digit: exclusive scan over warps; publish this tile's count at once
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
            st_relaxed_gpu(status + (uint64_t)tile * 256 + tid, tile == 0 ? (ST_INCLUSIVE | count) : st_partial(count));
        }
        uint32_t inc = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += y;
        }
        if (tid < 256 && lane == 31) wsum[warp] = inc;
        __syncthreads();
        uint32_t ts = 0;
        if (tid < 256) {
            uint32_t pre = 0;
            for (int w = 0; w < warp; ++w) pre += wsum[w];
            ts = pre + inc - count;
            tile_start[tid] = ts;
        }
        __syncthreads();
        OS_CLK(3);
        // re-order the tile in shared memory (stable: digit, then warp, then item order)
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const uint32_t idx = wbase + i * 32;
            if (full || idx < n) {
                const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
                const uint32_t lp = tile_start[d] + myhist[d] + rank[i];
                keys_s[lp] = key[i];
                vals_s[lp] = IDENTITY_VALS ? idx : val[i];
            }
        }
        OS_CLK(4);
        // windowed decoupled look-back, as late as possible
        if (tid < 256) {
            uint32_t excl = 0;
#ifdef OS_FAKE_LOOKBACK   // timing experiment only (wrong output): no waiting on predecessors
            if (true) {
                const uint32_t tot = (tid == 255 ? n : digit_base[tid + 1]) - digit_base[tid];
                excl = (uint32_t)((uint64_t)tile * tot / num_tiles);
                if (excl + count > tot) excl = tot > count ? tot - count : 0;
            } else
#endif
            if (tile != 0) {
                uint32_t spins = 0;
                int64_t p = (int64_t)tile - 1;
                bool done = false;
                while (!done) {
                    uint32_t v[LBW];
#pragma unroll
                    for (int j = 0; j < LBW; ++j)
                        v[j] = (p - j >= 0) ? ld_relaxed_gpu(status + (uint64_t)(p - j) * 256 + tid) : ST_INCLUSIVE;
                    int used = 0;
#pragma unroll
                    for (int j = 0; j < LBW; ++j) {
                        if (done || used != j) continue;
                        if (v[j] == 0) continue;                                      // not published yet: re-poll from here
                        excl += st_value(v[j]);
                        used = j + 1;
                        if (v[j] & ST_INCLUSIVE) done = true;
                    }
                    p -= used;
                    if (used == 0 && ++spins > LOOKBACK_SPIN_LIMIT) { atomicOr(err, SA_DEVERR_LOOKBACK_TIMEOUT); done = true; }
                }
                st_relaxed_gpu(status + (uint64_t)tile * 256 + tid, ST_INCLUSIVE | (excl + count));
            }
            gofs[tid] = digit_base[tid] + excl - ts;     // global index = gofs[d] + position in tile
        }
        OS_CLK(5);
        __syncthreads();
        OS_CLK(6);
        if (full) {
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                const uint32_t j = tid + i * THREADS;
                const uint64_t k = keys_s[j];
                const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
                keys_out[g] = k;
                vals_out[g] = vals_s[j];
            }
        } else {
            for (uint32_t j = tid; j < tile_n; j += THREADS) {
                const uint64_t k = keys_s[j];
                const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
                keys_out[g] = k;
                vals_out[g] = vals_s[j];
            }
        }
        OS_CLK(7);
        __syncthreads();       // keys_s / gofs / whist are reused by the next tile
    }
}

template <int THREADS, int IPT, int MINB, int RANK_MODE, int LBW, bool PF = false>
cudaError_t launch_onesweep2_cfg(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                 uint32_t* vals_out, uint32_t n, int shift, const uint32_t* digit_base,
                                 uint32_t* status, uint32_t* tile_counter, uint32_t* err, bool identity_vals,
                                 cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    constexpr uint32_t TILE = THREADS * IPT;
    constexpr size_t SMEM = onesweep2_smem_bytes<THREADS, IPT, PF>();
    if (PF && ((((uintptr_t)keys_in) | ((uintptr_t)vals_in)) & 15u))        // bulk copies need 16-byte alignment
        return launch_onesweep2_cfg<THREADS, IPT, MINB, RANK_MODE, LBW, false>(keys_in, vals_in, keys_out, vals_out, n, shift,
                                                                               digit_base, status, tile_counter, err,
                                                                               identity_vals, s);
    const uint32_t tiles = (n + TILE - 1) / TILE;
    cudaError_t e = cudaMemsetAsync(status, 0, (size_t)tiles * 256 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(tile_counter, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    static PerDeviceInt grid_cache;     // per device: the attribute below is per device too
    int grid = grid_cache.get();
    if (!grid) {
        auto k1 = onesweep2_kernel<THREADS, IPT, MINB, true, RANK_MODE, LBW, PF>;
        auto k0 = onesweep2_kernel<THREADS, IPT, MINB, false, RANK_MODE, LBW, PF>;
        cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        int dev = 0, sms = 148, per_sm = 1;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k0, THREADS, SMEM);
        if (per_sm < 1) per_sm = 1;
        grid = sms * per_sm;
        grid_cache.set(grid);
    }
    const uint32_t g = tiles < (uint32_t)grid ? tiles : (uint32_t)grid;
    if (identity_vals)
        onesweep2_kernel<THREADS, IPT, MINB, true, RANK_MODE, LBW, PF><<<g, THREADS, SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, digit_base, status, tile_counter, err, tiles);
    else
        onesweep2_kernel<THREADS, IPT, MINB, false, RANK_MODE, LBW, PF><<<g, THREADS, SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, digit_base, status, tile_counter, err, tiles);
    return cudaGetLastError();
}

// Launch one pass with a compile-time configuration.  status must hold tiles*256 words.
template <int THREADS, int IPT, int MINB, int RANK_MODE>
cudaError_t launch_onesweep_cfg(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                uint32_t* vals_out, uint32_t n, int shift, const uint32_t* digit_base,
                                uint32_t* status, uint32_t* tile_counter, uint32_t* err, bool identity_vals,
                                cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    constexpr uint32_t TILE = THREADS * IPT;
    constexpr size_t SMEM = onesweep_smem_bytes<THREADS, IPT>();
    const uint32_t tiles = (n + TILE - 1) / TILE;
    cudaError_t e = cudaMemsetAsync(status, 0, (size_t)tiles * 256 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(tile_counter, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    static PerDeviceInt attr_set;
    if (!attr_set.get()) {
        cudaFuncSetAttribute(onesweep_kernel<THREADS, IPT, MINB, true, RANK_MODE>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        cudaFuncSetAttribute(onesweep_kernel<THREADS, IPT, MINB, false, RANK_MODE>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        attr_set.set(1);
    }
    if (identity_vals)
        onesweep_kernel<THREADS, IPT, MINB, true, RANK_MODE><<<tiles, THREADS, SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, digit_base, status, tile_counter, err);
    else
        onesweep_kernel<THREADS, IPT, MINB, false, RANK_MODE><<<tiles, THREADS, SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, digit_base, status, tile_counter, err);
    return cudaGetLastError();
}

}  // namespace sa
