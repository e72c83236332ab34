// onesweep.cuh -- one stable 8-bit LSD scatter pass over (u64 key, u32 value) records (sm_100a).
//
// One CTA per tile of THREADS*IPT records, tiles taken in order from an atomic counter (so a
// tile's predecessors are always resident or done -> the look-back cannot deadlock):
//   1. coalesced streaming loads of the tile's keys (and values) into registers
//   2. warp-level multisplit: peers with the same digit (match.any or 8 ballots), rank inside
//      the warp, per-warp digit counters in shared memory
//   3. per digit: exclusive scan over warps, tile total published, decoupled look-back over the
//      earlier tiles' totals -> global base of this tile's run of that digit
//   4. records re-ordered in shared memory (digit-major, stable), then written out as digit runs
// Algorithmic traffic: read 12 B + write 12 B per record (8 + 12 when the values are generated).
#pragma once
#include "common.cuh"

#include <atomic>

namespace sa {

// One cached int per CUDA device (launch configuration derived from per-device attributes:
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and occupancy are properties of the device a
// kernel runs on, not of the process).  0 = not computed yet; a race between two threads computes
// the same value twice, which is harmless.
struct PerDeviceInt {
    std::atomic<int> v[64] = {};
    static int dev()
    {
        int d = 0;
        cudaGetDevice(&d);
        return (d >= 0 && d < 64) ? d : 0;
    }
    int get() const { return v[dev()].load(std::memory_order_acquire); }
    void set(int x) { v[dev()].store(x, std::memory_order_release); }
};

// Look-back status word of one (tile, digit): 0 = not published; bit 31 clear = the tile's own
// count + 1; bit 31 set = inclusive prefix over tiles 0..t in the low 31 bits (a prefix is at most
// n - 1 < 2^31 whenever the pass runs at all: a digit that holds every record skips the pass).
constexpr uint32_t ST_INCLUSIVE = 1u << 31;
constexpr uint32_t ST_MASK = ST_INCLUSIVE - 1u;
__device__ __forceinline__ uint32_t st_partial(uint32_t count) { return count + 1u; }
__device__ __forceinline__ uint32_t st_value(uint32_t v) { return (v & ST_INCLUSIVE) ? (v & ST_MASK) : v - 1u; }
constexpr uint32_t LOOKBACK_SPIN_LIMIT = 1u << 26;   // a hung look-back sets d_err and exits

// RANK_MODE 0: match.any.sync; 1: eight ballots (CUB-style warp multisplit)
template <int RANK_MODE>
__device__ __forceinline__ uint32_t digit_peers(uint32_t d, bool valid, uint32_t lane)
{
    if (RANK_MODE == 0) {
        return __match_any_sync(0xffffffffu, valid ? d : (256u + lane));
    } else {
        uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            peers &= bit ? bal : ~bal;
        }
        return valid ? peers : (1u << lane);
    }
}

// status traffic uses relaxed gpu-scope accesses (the word carries flag + count together)
__device__ __forceinline__ uint32_t ld_relaxed_gpu(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu(uint32_t* p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Per-phase clock accounting for scripts/bench_onesweep.cu (compiled in only with -DOS_PHASE_CLOCKS)
#ifdef OS_PHASE_CLOCKS
__device__ unsigned long long os_phase_cycles[8];
#define OS_CLK(k)                                                                   \
    do {                                                                            \
        if (tid == 0) { const long long c_ = clock64(); atomicAdd(&os_phase_cycles[k], (unsigned long long)(c_ - clk_)); clk_ = c_; } \
    } while (0)
#define OS_CLK_INIT long long clk_ = clock64()
#else
#define OS_CLK(k) do { } while (0)
#define OS_CLK_INIT do { } while (0)
#endif

// ------------------------------------------------------------------------------------------------
// v4: warp-specialised pass.  Each persistent CTA has 8 ranking warps and one look-back warp:
//   * ranking warps: rank + scan + shared-memory re-order of tile i into buffer i&1, then issue the
//     global loads of tile i+1 into registers, then write tile i-1 out of buffer (i-1)&1
//   * look-back warp: as soon as tile i's digit counts exist it runs the windowed decoupled
//     look-back for all 256 digits (8 per lane) and publishes the inclusive prefixes; the result
//     (global offset per digit) is only needed one tile later, so the chain of L2 round trips the
//     look-back costs (~3-5 per tile at B200 tile rates) is off the ranking warps' critical path.
// Named barriers: 1 = ranking warps only; 2+b = "counts of buffer b ready" (ranking warps arrive,
// look-back warp waits); 4+b = "offsets of buffer b ready" (look-back warp arrives, ranking warps wait).
__device__ __forceinline__ void named_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void named_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

template <int IPT>
constexpr size_t onesweep4_smem_bytes()
{
    return (size_t)256 * IPT * 12 * 2 + 8 * 256 * 4 + 4 * 256 * 4 + 64;
}

template <int IPT, int MINB, bool IDENTITY_VALS, int LBW, int LB_REGS>
__global__ void __launch_bounds__(512, MINB)
onesweep4_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift,
                 const uint32_t* __restrict__ digit_base, uint32_t* status, uint32_t* tile_counter,
                 uint32_t* err, uint32_t num_tiles)
{
    constexpr int MAIN = 256, WARPS = 8, ALL = 512;
    constexpr int MAIN_REGS = 128 - LB_REGS;      // 256 * (LB + MAIN) = the CTA's half of the register file
    static_assert(IPT % 2 == 0, "ranks are packed in pairs");
    constexpr int TILE = MAIN * IPT;
    constexpr uint32_t NO_TILE = 0xFFFFFFFFu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* keys_s = reinterpret_cast<uint64_t*>(smem_raw);                 // 2 * TILE
    uint32_t* vals_s = reinterpret_cast<uint32_t*>(keys_s + 2 * TILE);        // 2 * TILE
    uint32_t* whist = vals_s + 2 * TILE;                                       // WARPS * 256
    uint32_t* tst = whist + WARPS * 256;                                       // 2 * 256: digit start inside the tile
    uint32_t* cnt = tst + 512;                                                 // 2 * 256: digit count, later global offset
    uint32_t* wsum = cnt + 512;                                                // 8
    uint32_t* s_tile = wsum + 8;                                               // 4: ring of tile ids
    uint32_t* s_lbtile = s_tile + 4;                                           // 2: tile of buffer b for the look-back warp

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (warp >= WARPS) {
        // ------------------------------------------------------------------ look-back warps (one digit per thread)
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(LB_REGS));
        const int d = tid - MAIN;
        const uint32_t dbase = digit_base[d];
        for (uint32_t it = 0;; ++it) {
            const int b = it & 1;
            named_sync(2 + b, ALL);
            const uint32_t tile = s_lbtile[b];
            if (tile == NO_TILE) return;
            const uint32_t count = cnt[b * 256 + d];
            uint32_t excl = 0;
            if (tile != 0) {
                uint32_t spins = 0;
                int64_t p = (int64_t)tile - 1;
                bool done = false;
                while (!done) {
                    uint32_t v[LBW];
#pragma unroll
                    for (int j = 0; j < LBW; ++j)
                        v[j] = (p - j >= 0) ? ld_relaxed_gpu(status + (uint64_t)(p - j) * 256 + d) : ST_INCLUSIVE;
                    int used = 0;
#pragma unroll
                    for (int j = 0; j < LBW; ++j) {
                        if (done || used != j) continue;
                        if (v[j] == 0) continue;
                        excl += st_value(v[j]);
                        used = j + 1;
                        if (v[j] & ST_INCLUSIVE) done = true;
                    }
                    p -= used;
                    if (used == 0 && ++spins > LOOKBACK_SPIN_LIMIT) { atomicOr(err, SA_DEVERR_LOOKBACK_TIMEOUT); done = true; }
                }
                st_relaxed_gpu(status + (uint64_t)tile * 256 + d, ST_INCLUSIVE | (excl + count));
            }
            cnt[b * 256 + d] = dbase + excl - tst[b * 256 + d];   // global index = offset + position in tile
            named_arrive(4 + b, ALL);
        }
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MAIN_REGS));

    // ---------------------------------------------------------------------- ranking warps
    const uint32_t lt = lanemask_lt();
    uint32_t* myhist = whist + warp * 256;
    uint64_t key[IPT];
    uint32_t val[IPT];
    uint32_t rank2[IPT / 2];                                 // two 16-bit ranks per register
    if (tid == 0) s_tile[0] = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < WARPS * 256; i += MAIN) whist[i] = 0;
    named_sync(1, MAIN);
    {
        const uint32_t t0 = s_tile[0];
        const uint32_t wb = t0 * (uint32_t)TILE + warp * (32 * IPT) + lane;
        if (t0 < num_tiles) {
#pragma unroll
            for (int i = 0; i < IPT; ++i) key[i] = (wb + i * 32 < n) ? ld_stream_u64(keys_in + wb + i * 32) : ~0ull;
            if (!IDENTITY_VALS) {
#pragma unroll
                for (int i = 0; i < IPT; ++i) val[i] = (wb + i * 32 < n) ? ld_stream_u32(vals_in + wb + i * 32) : 0u;
            }
        }
    }
    uint32_t prev_n = 0;
    OS_CLK_INIT;
    for (uint32_t it = 0;; ++it) {
        const int b = it & 1;
        // s_tile[it & 3] was written one iteration ago (barriers in between); the per-warp counters
        // were re-zeroed by their own warp after the previous re-order
        OS_CLK(0);
        const uint32_t tile = s_tile[it & 3];
        const bool valid = tile < num_tiles;
        uint32_t tile_n = 0;
        if (valid) {
            if (tid == 0) s_tile[(it + 1) & 3] = atomicAdd(tile_counter, 1u);
            const uint32_t base = tile * (uint32_t)TILE;
            tile_n = (n - base) < (uint32_t)TILE ? (n - base) : (uint32_t)TILE;
            const uint32_t wbase = base + warp * (32 * IPT) + lane;
            const bool full = tile_n == (uint32_t)TILE;
#ifdef OS_PHASE_CLOCKS
            if (tid == 0 && key[IPT - 1] == 0x123456789ull) clk_ += 1;   // wait for the loads
#endif
            OS_CLK(1);
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                const bool ok = full || (wbase + i * 32) < n;
                const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
                const uint32_t peers = __match_any_sync(0xffffffffu, ok ? d : (256u + lane));
                const int leader = __ffs(peers) - 1;
                uint32_t bb = 0;
                if (lane == leader && ok) {
                    bb = myhist[d];
                    myhist[d] = bb + __popc(peers);
                }
                bb = __shfl_sync(0xffffffffu, bb, leader);
                const uint32_t r = bb + __popc(peers & lt);   // < 32 * IPT
                rank2[i >> 1] = (i & 1) ? (rank2[i >> 1] | (r << 16)) : r;
                __syncwarp();
            }
            named_sync(1, MAIN);
            OS_CLK(2);
            // per digit (one thread each): exclusive scan over warps, publish the tile's count at once
            uint32_t count = 0;
            {
                uint32_t run = 0;
#pragma unroll
                for (int w = 0; w < WARPS; ++w) {
                    const uint32_t c = whist[w * 256 + tid];
                    whist[w * 256 + tid] = run;
                    run += c;
                }
                count = run;
                st_relaxed_gpu(status + (uint64_t)tile * 256 + tid, tile == 0 ? (ST_INCLUSIVE | count) : st_partial(count));
            }
            uint32_t inc = count;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += y;
            }
            if (lane == 31) wsum[warp] = inc;
            named_sync(1, MAIN);
            {
                uint32_t pre = 0;
                for (int w = 0; w < warp; ++w) pre += wsum[w];
                tst[b * 256 + tid] = pre + inc - count;
                cnt[b * 256 + tid] = count;
                if (tid == 0) s_lbtile[b] = tile;
            }
            named_arrive(2 + b, ALL);                         // the look-back warp may start on this tile
            named_sync(1, MAIN);
            OS_CLK(3);
            // re-order the tile in shared memory (stable: digit, then warp, then item order)
            uint64_t* kb = keys_s + b * TILE;
            uint32_t* vb = vals_s + b * TILE;
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                const uint32_t idx = wbase + i * 32;
                if (full || idx < n) {
                    const uint32_t d = (uint32_t)(key[i] >> shift) & 0xFFu;
                    const uint32_t lp = tst[b * 256 + d] + myhist[d] + ((rank2[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu);
                    kb[lp] = key[i];
                    vb[lp] = IDENTITY_VALS ? idx : val[i];
                }
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i) myhist[i * 32 + lane] = 0;     // own counters, ready for the next tile
            OS_CLK(4);
            // the registers are free again: start the loads of the next tile
            const uint32_t nt = s_tile[(it + 1) & 3];
            if (nt < num_tiles) {
                const uint32_t wb = nt * (uint32_t)TILE + warp * (32 * IPT) + lane;
#pragma unroll
                for (int i = 0; i < IPT; ++i) key[i] = (wb + i * 32 < n) ? ld_stream_u64(keys_in + wb + i * 32) : ~0ull;
                if (!IDENTITY_VALS) {
#pragma unroll
                    for (int i = 0; i < IPT; ++i) val[i] = (wb + i * 32 < n) ? ld_stream_u32(vals_in + wb + i * 32) : 0u;
                }
            }
        } else {
            if (tid == 0) s_lbtile[b] = NO_TILE;
            named_arrive(2 + b, ALL);                         // tells the look-back warp to leave
        }
        if (it > 0) {
            // write the previous tile out of the other buffer
            const int pb = b ^ 1;
            OS_CLK(5);
            named_sync(4 + pb, ALL);
            OS_CLK(6);
            const uint64_t* kb = keys_s + pb * TILE;
            const uint32_t* vb = vals_s + pb * TILE;
            const uint32_t* gofs = cnt + pb * 256;
            if (prev_n == (uint32_t)TILE) {
#pragma unroll
                for (int i = 0; i < IPT; ++i) {
                    const uint32_t j = tid + i * MAIN;
                    const uint64_t k = kb[j];
                    const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
                    keys_out[g] = k;
                    vals_out[g] = vb[j];
                }
            } else {
                for (uint32_t j = tid; j < prev_n; j += MAIN) {
                    const uint64_t k = kb[j];
                    const uint32_t g = gofs[(uint32_t)(k >> shift) & 0xFFu] + j;
                    keys_out[g] = k;
                    vals_out[g] = vb[j];
                }
            }
            OS_CLK(7);
        }
        if (!valid) return;
        prev_n = tile_n;
    }
}

template <int IPT, int MINB, int LBW, int LB_REGS>
cudaError_t launch_onesweep4_cfg(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                 uint32_t* vals_out, uint32_t n, int shift, const uint32_t* digit_base,
                                 uint32_t* status, uint32_t* tile_counter, uint32_t* err, bool identity_vals,
                                 cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    constexpr uint32_t TILE = 256 * IPT;
    constexpr size_t SMEM = onesweep4_smem_bytes<IPT>();
    const uint32_t tiles = (n + TILE - 1) / TILE;
    cudaError_t e = cudaMemsetAsync(status, 0, (size_t)tiles * 256 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(tile_counter, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    static PerDeviceInt grid_cache;     // per device: the attribute below is per device too
    int grid = grid_cache.get();
    if (!grid) {
        auto k1 = onesweep4_kernel<IPT, MINB, true, LBW, LB_REGS>;
        auto k0 = onesweep4_kernel<IPT, MINB, false, LBW, LB_REGS>;
        cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        int dev = 0, sms = 148, per_sm = 1;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k0, 512, SMEM);
        if (per_sm < 1) per_sm = 1;
        grid = sms * per_sm;
        grid_cache.set(grid);
    }
    const uint32_t g = tiles < (uint32_t)grid ? tiles : (uint32_t)grid;
    if (identity_vals)
        onesweep4_kernel<IPT, MINB, true, LBW, LB_REGS><<<g, 512, SMEM, s>>>(keys_in, vals_in, keys_out, vals_out, n, shift,
                                                                   digit_base, status, tile_counter, err, tiles);
    else
        onesweep4_kernel<IPT, MINB, false, LBW, LB_REGS><<<g, 512, SMEM, s>>>(keys_in, vals_in, keys_out, vals_out, n, shift,
                                                                    digit_base, status, tile_counter, err, tiles);
    return cudaGetLastError();
}


}  // namespace sa
