// onesweep_wide.cuh -- EXPERIMENT, not part of libsa_cuda.so: the warp-specialised onesweep pass of
// csrc/onesweep.cuh (onesweep4_kernel) generalised from 8-bit to RB-bit digits (RB = 8, 9, 10), so
// that the 60-bit keys of the first sort take 7 or 6 passes instead of 8 (profiles/r01_digit_model.md).
// Only scripts/bench_onesweep_wide.cu includes it.  Written without GPU access: compiled and checked
// with ptxas here, never run -- the bench verifies order and stability of every variant it times.
//
// Differences from onesweep4_kernel:
//   * BINS = 1 << RB digit values; the digit of a pass is (key >> shift) & mask with mask < BINS
//   * per-warp digit counters are 16-bit (a warp ranks 32 * IPT <= 768 records, a tile < 65536)
//   * a ranking thread owns DPT = BINS / 256 consecutive digits in the per-digit step, a look-back
//     thread DPT digits 256 apart (coalesced status rows) whose windows are polled together
//   * the status row of a tile is BINS words
#pragma once
#include "../suffixarray_b200/csrc/onesweep.cuh"

namespace sa {

template <int RB, int IPT>
constexpr size_t onesweepw_smem_bytes()
{
    return (size_t)256 * IPT * 12 * 2 + (size_t)4 * (1 << RB) * 4 + 64 + (size_t)8 * (1 << RB) * 2;
}

template <int RB, int IPT, int MINB, bool IDENTITY_VALS, int LBW, int LB_REGS>
__global__ void __launch_bounds__(512, MINB)
onesweepw_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift, uint32_t mask,
                 const uint32_t* __restrict__ digit_base, uint32_t* status, uint32_t* tile_counter,
                 uint32_t* err, uint32_t num_tiles)
{
    constexpr int MAIN = 256, WARPS = 8, ALL = 512;
    constexpr int BINS = 1 << RB, DPT = BINS / 256;
    constexpr int MAIN_REGS = MINB >= 2 ? 128 - LB_REGS : (256 - LB_REGS > 232 ? 232 : 256 - LB_REGS);
    static_assert(RB >= 8 && RB <= 10, "256, 512 or 1024 digit values");
    static_assert(IPT % 2 == 0, "ranks are packed in pairs");
    static_assert(MAIN_REGS % 8 == 0 && LB_REGS % 8 == 0, "setmaxnreg takes multiples of 8");
    constexpr int TILE = MAIN * IPT;
    constexpr uint32_t NO_TILE = 0xFFFFFFFFu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* keys_s = reinterpret_cast<uint64_t*>(smem_raw);                 // 2 * TILE
    uint32_t* vals_s = reinterpret_cast<uint32_t*>(keys_s + 2 * TILE);        // 2 * TILE
    uint32_t* tst = vals_s + 2 * TILE;                                         // 2 * BINS: digit start inside the tile
    uint32_t* cnt = tst + 2 * BINS;                                            // 2 * BINS: digit count, later global offset
    uint32_t* wsum = cnt + 2 * BINS;                                           // 8
    uint32_t* s_tile = wsum + 8;                                               // 4: ring of tile ids
    uint32_t* s_lbtile = s_tile + 4;                                           // 2 (+2 pad)
    uint16_t* whist = reinterpret_cast<uint16_t*>(s_lbtile + 4);               // WARPS * BINS

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (warp >= WARPS) {
        // ------------------------------------------------------------------ look-back warps
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(LB_REGS));
        const int t = tid - MAIN;
        for (uint32_t it = 0;; ++it) {
            const int b = it & 1;
            named_sync(2 + b, ALL);
            const uint32_t tile = s_lbtile[b];
            if (tile == NO_TILE) return;
            uint32_t excl[DPT];
#pragma unroll
            for (int j = 0; j < DPT; ++j) excl[j] = 0;
            if (tile != 0) {
                int p[DPT];
#pragma unroll
                for (int j = 0; j < DPT; ++j) p[j] = (int)tile - 1;
                uint32_t pending = (1u << DPT) - 1u, spins = 0;
                while (pending) {
                    uint32_t v[DPT][LBW];
#pragma unroll
                    for (int j = 0; j < DPT; ++j)
#pragma unroll
                        for (int w = 0; w < LBW; ++w)
                            v[j][w] = (((pending >> j) & 1u) && p[j] - w >= 0)
                                          ? ld_relaxed_gpu(status + (uint64_t)(p[j] - w) * BINS + (t + 256 * j))
                                          : ST_INCLUSIVE;
                    bool progressed = false;
#pragma unroll
                    for (int j = 0; j < DPT; ++j) {
                        if (!((pending >> j) & 1u)) continue;
                        int used = 0;
                        bool done = false;
#pragma unroll
                        for (int w = 0; w < LBW; ++w) {
                            if (done || used != w) continue;
                            if (v[j][w] == 0) continue;
                            excl[j] += st_value(v[j][w]);
                            used = w + 1;
                            if (v[j][w] & ST_INCLUSIVE) done = true;
                        }
                        p[j] -= used;
                        if (used) progressed = true;
                        if (done) pending &= ~(1u << j);
                    }
                    if (!progressed && ++spins > LOOKBACK_SPIN_LIMIT) { atomicOr(err, SA_DEVERR_LOOKBACK_TIMEOUT); pending = 0; }
                }
#pragma unroll
                for (int j = 0; j < DPT; ++j)
                    st_relaxed_gpu(status + (uint64_t)tile * BINS + (t + 256 * j),
                                   ST_INCLUSIVE | (excl[j] + cnt[b * BINS + t + 256 * j]));
            }
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int d = t + 256 * j;
                cnt[b * BINS + d] = digit_base[d] + excl[j] - tst[b * BINS + d];   // global index = offset + position in tile
            }
            named_arrive(4 + b, ALL);
        }
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MAIN_REGS));

    // ---------------------------------------------------------------------- ranking warps
    const uint32_t lt = lanemask_lt();
    uint16_t* myhist = whist + warp * BINS;
    uint32_t* myhist32 = reinterpret_cast<uint32_t*>(myhist);
    uint64_t key[IPT];
    uint32_t val[IPT];
    uint32_t rank2[IPT / 2];                                 // two 16-bit ranks per register
    if (tid == 0) s_tile[0] = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < WARPS * BINS / 2; i += MAIN) reinterpret_cast<uint32_t*>(whist)[i] = 0;
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
    for (uint32_t it = 0;; ++it) {
        const int b = it & 1;
        const uint32_t tile = s_tile[it & 3];
        const bool valid = tile < num_tiles;
        uint32_t tile_n = 0;
        if (valid) {
            if (tid == 0) s_tile[(it + 1) & 3] = atomicAdd(tile_counter, 1u);
            const uint32_t base = tile * (uint32_t)TILE;
            tile_n = (n - base) < (uint32_t)TILE ? (n - base) : (uint32_t)TILE;
            const uint32_t wbase = base + warp * (32 * IPT) + lane;
            const bool full = tile_n == (uint32_t)TILE;
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                const bool ok = full || (wbase + i * 32) < n;
                const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
                const uint32_t peers = __match_any_sync(0xffffffffu, ok ? d : ((uint32_t)BINS + lane));
                const int leader = __ffs(peers) - 1;
                uint32_t bb = 0;
                if (lane == leader && ok) {
                    bb = myhist[d];
                    myhist[d] = (uint16_t)(bb + __popc(peers));
                }
                bb = __shfl_sync(0xffffffffu, bb, leader);
                const uint32_t r = bb + __popc(peers & lt);   // < 32 * IPT
                rank2[i >> 1] = (i & 1) ? (rank2[i >> 1] | (r << 16)) : r;
                __syncwarp();
            }
            named_sync(1, MAIN);
            // per digit (DPT consecutive digits per thread): exclusive scan over warps, publish the tile's counts
            uint32_t cj[DPT], tot = 0;
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int d = tid * DPT + j;
                uint32_t run = 0;
#pragma unroll
                for (int w = 0; w < WARPS; ++w) {
                    const uint32_t c = whist[w * BINS + d];
                    whist[w * BINS + d] = (uint16_t)run;
                    run += c;
                }
                cj[j] = run;
                tot += run;
                st_relaxed_gpu(status + (uint64_t)tile * BINS + d, tile == 0 ? (ST_INCLUSIVE | run) : st_partial(run));
            }
            uint32_t inc = tot;
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
                uint32_t at = pre + inc - tot;
#pragma unroll
                for (int j = 0; j < DPT; ++j) {
                    tst[b * BINS + tid * DPT + j] = at;
                    cnt[b * BINS + tid * DPT + j] = cj[j];
                    at += cj[j];
                }
                if (tid == 0) s_lbtile[b] = tile;
            }
            named_arrive(2 + b, ALL);                         // the look-back warps may start on this tile
            named_sync(1, MAIN);
            // re-order the tile in shared memory (stable: digit, then warp, then item order)
            uint64_t* kb = keys_s + b * TILE;
            uint32_t* vb = vals_s + b * TILE;
#pragma unroll
            for (int i = 0; i < IPT; ++i) {
                const uint32_t idx = wbase + i * 32;
                if (full || idx < n) {
                    const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
                    const uint32_t lp = tst[b * BINS + d] + myhist[d] + ((rank2[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu);
                    kb[lp] = key[i];
                    vb[lp] = IDENTITY_VALS ? idx : val[i];
                }
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < BINS / 64; ++i) myhist32[i * 32 + lane] = 0;   // own counters, ready for the next tile
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
            named_arrive(2 + b, ALL);                         // tells the look-back warps to leave
        }
        if (it > 0) {
            // write the previous tile out of the other buffer
            const int pb = b ^ 1;
            named_sync(4 + pb, ALL);
            const uint64_t* kb = keys_s + pb * TILE;
            const uint32_t* vb = vals_s + pb * TILE;
            const uint32_t* gofs = cnt + pb * BINS;
            if (prev_n == (uint32_t)TILE) {
#pragma unroll
                for (int i = 0; i < IPT; ++i) {
                    const uint32_t j = tid + i * MAIN;
                    const uint64_t k = kb[j];
                    const uint32_t g = gofs[(uint32_t)(k >> shift) & mask] + j;
                    keys_out[g] = k;
                    vals_out[g] = vb[j];
                }
            } else {
                for (uint32_t j = tid; j < prev_n; j += MAIN) {
                    const uint64_t k = kb[j];
                    const uint32_t g = gofs[(uint32_t)(k >> shift) & mask] + j;
                    keys_out[g] = k;
                    vals_out[g] = vb[j];
                }
            }
        }
        if (!valid) return;
        prev_n = tile_n;
    }
}

// One pass on key bits [shift, shift + nbits), nbits <= RB.  digit_base: BINS exclusive digit
// offsets of this pass; status: tiles * BINS words.
template <int RB, int IPT, int MINB, int LBW, int LB_REGS>
cudaError_t launch_onesweepw_cfg(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                 uint32_t* vals_out, uint32_t n, int shift, int nbits, const uint32_t* digit_base,
                                 uint32_t* status, uint32_t* tile_counter, uint32_t* err, bool identity_vals,
                                 cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    if (nbits < 1 || nbits > RB) return cudaErrorInvalidValue;
    constexpr uint32_t TILE = 256 * IPT;
    constexpr uint32_t BINS = 1u << RB;
    constexpr size_t SMEM = onesweepw_smem_bytes<RB, IPT>();
    const uint32_t tiles = (n + TILE - 1) / TILE;
    const uint32_t mask = (1u << nbits) - 1u;
    cudaError_t e = cudaMemsetAsync(status, 0, (size_t)tiles * BINS * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(tile_counter, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    static int grid = 0;
    if (!grid) {
        auto k1 = onesweepw_kernel<RB, IPT, MINB, true, LBW, LB_REGS>;
        auto k0 = onesweepw_kernel<RB, IPT, MINB, false, LBW, LB_REGS>;
        cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
        int dev = 0, sms = 148, per_sm = 1;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k0, 512, SMEM);
        if (per_sm < 1) per_sm = 1;
        grid = sms * per_sm;
    }
    const uint32_t g = tiles < (uint32_t)grid ? tiles : (uint32_t)grid;
    if (identity_vals)
        onesweepw_kernel<RB, IPT, MINB, true, LBW, LB_REGS><<<g, 512, SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, mask, digit_base, status, tile_counter, err, tiles);
    else
        onesweepw_kernel<RB, IPT, MINB, false, LBW, LB_REGS><<<g, 512, SMEM, s>>>(
            keys_in, vals_in, keys_out, vals_out, n, shift, mask, digit_base, status, tile_counter, err, tiles);
    return cudaGetLastError();
}

}  // namespace sa
