// refine.cu -- key packing and tie refinement ("prefix extension up to max_suffix_length").
//
// Round 0 sorts every suffix by the masked word of bytes [0,8).  Afterwards only groups that
// (a) are still tied on all `depth` bytes sorted so far, (b) have more than 32 members and
// (c) did not run into a terminator are re-sorted by the next cb bytes -- exactly the buckets
// recursive_bucket_sort would keep splitting (engine.c:706-710, :771-811).  A round is:
//   tie mask -> group flags/counts -> scan -> compact + gather composite keys
//   (dense group id | next cb bytes) -> onesweep passes -> write back + LCP update.
// The adjacent-LCP array (u16, bytes) is what finish.cu turns into the <=32 bucket depths.
#include "common.cuh"
#include "kernels.h"

namespace sa {

// suffix_array.pyx:103-107 -- ASCII A-Z -> a-z, 16 bytes per thread.
__global__ void lowercase_kernel(uint4* __restrict__ text16, uint64_t nvec)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nvec) return;
    uint4 v = text16[i];
    uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t x = w[j];
        // per byte: is it in 'A'..'Z'?  (x + 0x3F) has bit7 set iff byte >= 0x41 (for bytes < 0x80)
        uint32_t hi = x & 0x80808080u;
        uint32_t ge_a = ((x & 0x7F7F7F7Fu) + 0x3F3F3F3Fu);     // bit7 set iff (byte&0x7f) >= 0x41
        uint32_t gt_z = ((x & 0x7F7F7F7Fu) + 0x25252525u);     // bit7 set iff (byte&0x7f) >= 0x5B
        uint32_t is_up = ge_a & ~gt_z & ~hi & 0x80808080u;
        w[j] = x | (is_up >> 2);                                 // +0x20
    }
    text16[i] = v;
}

__global__ void iota_kernel(uint32_t* __restrict__ out, uint32_t n)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}

// 256-bin histogram of the text bytes (which byte values occur decides the code width).
__global__ void __launch_bounds__(256) byte_hist_kernel(const uint4* __restrict__ text16, uint64_t nvec, uint64_t n,
                                                        uint32_t* __restrict__ hist)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const uint4 v = text16[i];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (i * 16 + j < n) {
                // presence is all that matters: a plain store avoids the atomic
                h[(w[j >> 2] >> (8 * (j & 3))) & 0xFFu] = 1;
            }
        }
    }
    __syncthreads();
    if (h[threadIdx.x]) hist[threadIdx.x] = 1;
}

// codes[i] = table[text[i]], 16 bytes per thread (the pad after n is written as code 0).
__global__ void __launch_bounds__(256) encode_kernel(const uint4* __restrict__ text16, uint64_t nvec,
                                                     const uint8_t* __restrict__ table, uint4* __restrict__ codes16)
{
    __shared__ uint8_t tab[256];
    tab[threadIdx.x] = table[threadIdx.x];
    __syncthreads();
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nvec) return;
    uint4 v = text16[i];
    uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t x = w[j];
        w[j] = (uint32_t)tab[x & 0xFFu] | ((uint32_t)tab[(x >> 8) & 0xFFu] << 8) |
               ((uint32_t)tab[(x >> 16) & 0xFFu] << 16) | ((uint32_t)tab[x >> 24] << 24);
    }
    codes16[i] = v;
}

// Round-0 keys of every suffix: masked word of bytes [0, cb) (bits == 8) or cb packed codes.
__global__ void pack0_kernel(const uint8_t* __restrict__ src, uint32_t n, int cb, CodeInfo ci,
                             uint64_t* __restrict__ keys)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    keys[p] = mask_word(load_be64(src, p), cb);
}

// Compressed alphabet: one thread packs the keys of 8 consecutive suffixes from 32 code bytes.
// F_k = the cb codes starting at k (unmasked) slides by one code per suffix; the masked key keeps
// the fields up to and including the first terminator at or after k.
__global__ void __launch_bounds__(256)
pack0_codes_kernel(const uint8_t* __restrict__ codes, uint32_t n, int cb, CodeInfo ci, uint64_t* __restrict__ keys)
{
    const uint32_t p0 = (blockIdx.x * blockDim.x + threadIdx.x) * 8u;
    if (p0 >= n) return;
    // 8 + 16 = 24 code bytes needed (p0 is a multiple of 8; the allocation is padded)
    const uint64_t* c8 = reinterpret_cast<const uint64_t*>(codes + p0);
    const uint64_t W0 = c8[0], W1 = c8[1], W2 = c8[2];
    auto code_at = [&](int i) -> uint32_t {
        const uint64_t ww = i < 8 ? W0 : (i < 16 ? W1 : W2);
        return (uint32_t)(ww >> (8 * (i & 7))) & 0xFFu;
    };
    // bit i of termmask: code i is a terminator
    uint32_t termmask = 0;
#pragma unroll
    for (int i = 0; i < 24; ++i) {
        const uint32_t c = code_at(i);
        termmask |= (uint32_t)(c == ci.nl || c == ci.ff) << i;
    }
    const int used = cb * ci.bits;
    const uint64_t fmask = used >= 64 ? ~0ull : ((1ull << used) - 1ull);
    uint64_t F = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (i < cb) F = (F << ci.bits) | code_at(i);
    uint64_t out[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t tm = (termmask >> k) & ((cb >= 32) ? 0xFFFFFFFFu : ((1u << cb) - 1u));
        uint64_t key = F;
        if (tm) {
            const int t = __ffs((int)tm) - 1;                     // first terminator inside the window
            key &= ~((1ull << (ci.bits * (cb - 1 - t))) - 1ull);
        }
        key <<= (64 - used);
        if (tm) key |= 1ull << (63 - used);
        out[k] = key;
        // slide: drop the oldest code, append code k + cb
        F = ((F << ci.bits) & fmask) | code_at(k + cb);
    }
    if (p0 + 8 <= n) {
        uint4* o = reinterpret_cast<uint4*>(keys + p0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint4 v;
            v.x = (uint32_t)out[2 * k]; v.y = (uint32_t)(out[2 * k] >> 32);
            v.z = (uint32_t)out[2 * k + 1]; v.w = (uint32_t)(out[2 * k + 1] >> 32);
            o[k] = v;
        }
    } else {
        for (int k = 0; k < 8 && p0 + k < n; ++k) keys[p0 + k] = out[k];
    }
}

// Adjacent LCPs after round 0 and, in the same pass, the tie mask of the first refinement round
// (bit r = lcp[r] == cb; layout as in tie_mask_kernel; the caller zeroes the mask first).
__global__ void __launch_bounds__(256)
lcp0_kernel(const uint64_t* __restrict__ k, uint32_t n, int cb, int bits, uint32_t lprime,
            uint16_t* __restrict__ lcp, uint32_t* __restrict__ tie)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t v = 0;
    if (r < n) {
        if (r) v = bits == 8 ? pair_lcp(k[r - 1], k[r], 0, cb, lprime) : pair_lcp_codes(k[r - 1], k[r], 0, cb, bits, lprime);
        lcp[r] = (uint16_t)v;
    }
    const uint32_t b = __ballot_sync(0xffffffffu, r < n && r > 0 && v == (uint32_t)cb);
    if ((threadIdx.x & 31) == 0 && (r >> 5) < ((n + 31) >> 5)) tie[(r >> 5) + 1] = b;
}

// word w (w >= 1) holds ranks [32(w-1), 32w); word 0 is a zero pad.
__global__ void tie_mask_kernel(const uint16_t* __restrict__ lcp, uint32_t n, uint32_t depth,
                                uint32_t* __restrict__ tie, uint32_t num_words)
{
    uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;      // global warp = data word index
    if (gw + 1 >= num_words) return;
    uint32_t r = gw * 32 + (threadIdx.x & 31);
    bool t = (r < n) && (r > 0) && (lcp[r] == depth);
    uint32_t b = __ballot_sync(0xffffffffu, t);
    if ((threadIdx.x & 31) == 0) {
        tie[gw + 1] = b;
        if (gw == 0) tie[0] = 0;
    }
}

// One thread per 32 ranks.  amask: ranks belonging to a tied group of more than 32 members;
// hmask: first rank of such a group.  counts[2*tile], counts[2*tile+1]: per-tile popcounts.
__global__ void __launch_bounds__(SA_SCAN_TILE / 32)
group_flags_kernel(const uint32_t* __restrict__ tie, uint32_t nwords_data, uint32_t* __restrict__ amask,
                   uint32_t* __restrict__ hmask, uint32_t* __restrict__ counts)
{
    __shared__ uint32_t sa_[SA_SCAN_TILE / 32 / 32], sh_[SA_SCAN_TILE / 32 / 32];
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;       // data word (ranks 32w ..)
    uint32_t am = 0, hm = 0;
    if (w < nwords_data) {
        const uint32_t prev = tie[w], cur = tie[w + 1], next = tie[w + 2];
        const uint64_t X = ((uint64_t)cur << 32) | prev;             // rank 32w+i at bit i+32
        const uint64_t Y = ((uint64_t)next << 32) | cur;             // rank 32w+j at bit j
#pragma unroll 8
        for (int i = 0; i < 32; ++i) {
            const uint32_t wb = (uint32_t)(X >> (i + 1));            // bit 31 = tie[r] ... bit 0 = tie[r-31]
            const uint32_t wf = (uint32_t)(Y >> (i + 1));            // bit 0 = tie[r+1] ... bit 31 = tie[r+32]
            const uint32_t back = __clz((int)~wb);                   // ~wb == 0 -> 32
            const uint32_t nf = ~wf;
            const uint32_t fwd = nf ? (uint32_t)(__ffs((int)nf) - 1) : 32u;
            const bool active = (back + fwd) >= (uint32_t)SA_SMALL;
            am |= (uint32_t)active << i;
            hm |= (uint32_t)(active && !((cur >> i) & 1u)) << i;
        }
        amask[w] = am;
        hmask[w] = hm;
    }
    uint32_t ca = __popc(am), ch = __popc(hm);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        ca += __shfl_xor_sync(0xffffffffu, ca, o);
        ch += __shfl_xor_sync(0xffffffffu, ch, o);
    }
    if ((threadIdx.x & 31) == 0) { sa_[threadIdx.x >> 5] = ca; sh_[threadIdx.x >> 5] = ch; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t ta = 0, th = 0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { ta += sa_[i]; th += sh_[i]; }
        counts[2 * blockIdx.x] = ta;
        counts[2 * blockIdx.x + 1] = th;
    }
}

// Single block: in-place exclusive scan of `num` (a, b) pairs; totals[0], totals[1] = sums.
__global__ void __launch_bounds__(1024) scan_pairs_kernel(uint32_t* __restrict__ c, uint32_t num,
                                                          uint32_t* __restrict__ totals)
{
    __shared__ uint32_t wa[32], wb[32];
    __shared__ uint32_t carry_a, carry_b;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) { carry_a = 0; carry_b = 0; }
    __syncthreads();
    for (uint32_t base = 0; base < num; base += 1024) {
        const uint32_t i = base + t;
        uint32_t a = 0, b = 0;
        if (i < num) { a = c[2 * i]; b = c[2 * i + 1]; }
        uint32_t ia = a, ib = b;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t ya = __shfl_up_sync(0xffffffffu, ia, o), yb = __shfl_up_sync(0xffffffffu, ib, o);
            if (lane >= o) { ia += ya; ib += yb; }
        }
        if (lane == 31) { wa[warp] = ia; wb[warp] = ib; }
        __syncthreads();
        if (warp == 0) {
            uint32_t va = wa[lane], vb = wb[lane];
            uint32_t sa2 = va, sb2 = vb;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t ya = __shfl_up_sync(0xffffffffu, sa2, o), yb = __shfl_up_sync(0xffffffffu, sb2, o);
                if (lane >= o) { sa2 += ya; sb2 += yb; }
            }
            wa[lane] = sa2 - va;
            wb[lane] = sb2 - vb;
        }
        __syncthreads();
        const uint32_t ea = carry_a + wa[warp] + ia - a, eb = carry_b + wb[warp] + ib - b;
        if (i < num) { c[2 * i] = ea; c[2 * i + 1] = eb; }
        __syncthreads();
        if (t == 1023) { carry_a = ea + a; carry_b = eb + b; }
        __syncthreads();
    }
    if (t == 0) { totals[0] = carry_a; totals[1] = carry_b; }
}

// One block per tile of SA_SCAN_TILE ranks: list the active ranks (ascending) with their dense
// group ids in shared memory, then gather keys cooperatively so global traffic is coalesced.
__global__ void __launch_bounds__(SA_SCAN_TILE / 32)
compact_gather_kernel(const uint32_t* __restrict__ amask, const uint32_t* __restrict__ hmask,
                      uint32_t nwords_data, const uint32_t* __restrict__ tile_offsets,
                      const uint32_t* __restrict__ sa, const uint8_t* __restrict__ text, uint32_t depth, int cb,
                      int gbits, CodeInfo ci, uint32_t* __restrict__ slots, uint64_t* __restrict__ ckeys,
                      uint32_t* __restrict__ cvals)
{
    constexpr int T = SA_SCAN_TILE / 32;
    __shared__ uint32_t s_rank[SA_SCAN_TILE];
    __shared__ uint32_t s_gid[SA_SCAN_TILE];
    __shared__ uint32_t wsa[T / 32], wsh[T / 32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t w = blockIdx.x * T + t;
    uint32_t am = 0, hm = 0;
    if (w < nwords_data) { am = amask[w]; hm = hmask[w]; }
    const uint32_t ca = __popc(am), ch = __popc(hm);
    uint32_t ia = ca, ih = ch;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t ya = __shfl_up_sync(0xffffffffu, ia, o), yh = __shfl_up_sync(0xffffffffu, ih, o);
        if (lane >= o) { ia += ya; ih += yh; }
    }
    if (lane == 31) { wsa[warp] = ia; wsh[warp] = ih; }
    __syncthreads();
    uint32_t pa = 0, ph = 0;
    for (int i = 0; i < warp; ++i) { pa += wsa[i]; ph += wsh[i]; }
    uint32_t tot = 0;
    for (int i = 0; i < T / 32; ++i) tot += wsa[i];
    uint32_t lo = pa + ia - ca;                           // first local slot of this word
    const uint32_t gid0 = tile_offsets[2 * blockIdx.x + 1] + ph + ih - ch;   // heads before this word
    uint32_t m = am;
    while (m) {
        const int b = __ffs((int)m) - 1;
        m &= m - 1;
        s_rank[lo] = w * 32 + b;
        s_gid[lo] = gid0 + __popc(hm & (0xFFFFFFFFu >> (31 - b))) - 1u;   // heads at or before b
        ++lo;
    }
    __syncthreads();
    const uint32_t out0 = tile_offsets[2 * blockIdx.x];
    // two dependent random reads per record (sa[r], then the text at sa[r] + depth): four records
    // per thread are kept in flight so the gather is not one DRAM latency per record
    constexpr int B = 4;
    for (uint32_t j0 = t; j0 < tot; j0 += T * B) {
        uint32_t r[B], p[B];
        uint64_t w0[B], w1[B];
#pragma unroll
        for (int u = 0; u < B; ++u) {
            const uint32_t j = j0 + u * T;
            r[u] = j < tot ? s_rank[j] : 0u;
        }
#pragma unroll
        for (int u = 0; u < B; ++u) p[u] = (j0 + u * T < tot) ? __ldg(sa + r[u]) : 0u;
#pragma unroll
        for (int u = 0; u < B; ++u) {
            w0[u] = w1[u] = 0;
            if (j0 + u * T < tot) {
                if (ci.bits == 8) w0[u] = load_be64(text, (uint64_t)p[u] + depth);
                else load_be128(text, (uint64_t)p[u] + depth, w0[u], w1[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < B; ++u) {
            const uint32_t j = j0 + u * T;
            if (j >= tot) continue;
            const uint64_t wd = ci.bits == 8 ? mask_word(w0[u], cb) : pack_chars(w0[u], w1[u], cb, ci);
            slots[out0 + j] = r[u];
            cvals[out0 + j] = p[u];
            ckeys[out0 + j] = ((uint64_t)s_gid[j] << (64 - gbits)) | (wd >> gbits);
        }
    }
}

__global__ void writeback_kernel(const uint32_t* __restrict__ slots, const uint64_t* __restrict__ ckeys,
                                 const uint32_t* __restrict__ cvals, uint32_t m, uint32_t depth, int cb, int gbits,
                                 int bits, uint32_t lprime, uint32_t* __restrict__ sa, uint16_t* __restrict__ lcp,
                                 uint32_t* __restrict__ tie_next)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t r = slots[i];
    sa[r] = cvals[i];
    if (i == 0) return;
    const uint64_t a = ckeys[i - 1], b = ckeys[i];
    if ((a >> (64 - gbits)) != (b >> (64 - gbits))) return;          // first member of its group
    const uint32_t v = bits == 8 ? pair_lcp(a << gbits, b << gbits, depth, cb, lprime)
                                 : pair_lcp_codes(a << gbits, b << gbits, depth, cb, bits, lprime);
    lcp[r] = (uint16_t)v;
    // still tied through the bytes sorted so far: candidate of the next round (mask zeroed by the caller)
    if (v == depth + (uint32_t)cb) atomicOr(tie_next + (r >> 5) + 1, 1u << (r & 31));
}

}  // namespace sa

// ------------------------------------------------------------------------------ launchers
cudaError_t sa_launch_iota(uint32_t* out, uint32_t n, cudaStream_t s)
{
    if (!n) return cudaSuccess;
    sa::iota_kernel<<<(n + 255) / 256, 256, 0, s>>>(out, n);
    return cudaGetLastError();
}

cudaError_t sa_launch_lowercase(uint8_t* text, uint64_t n, cudaStream_t s)
{
    uint64_t nvec = (n + 15) / 16;
    if (!nvec) return cudaSuccess;
    sa::lowercase_kernel<<<(unsigned)((nvec + 255) / 256), 256, 0, s>>>(reinterpret_cast<uint4*>(text), nvec);
    return cudaGetLastError();
}

cudaError_t sa_launch_byte_hist(const uint8_t* text, uint64_t n, uint32_t* hist256, int num_sms, cudaStream_t s)
{
    cudaError_t e = cudaMemsetAsync(hist256, 0, 256 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    const uint64_t nvec = (n + 15) / 16;
    if (!nvec) return cudaSuccess;
    uint64_t blocks = (nvec + 255) / 256;
    if (blocks > (uint64_t)num_sms * 8) blocks = (uint64_t)num_sms * 8;
    sa::byte_hist_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const uint4*>(text), nvec, n, hist256);
    return cudaGetLastError();
}

cudaError_t sa_launch_encode(const uint8_t* text, uint64_t nbytes_padded, const uint8_t* table, uint8_t* codes,
                             cudaStream_t s)
{
    const uint64_t nvec = nbytes_padded / 16;
    if (!nvec) return cudaSuccess;
    sa::encode_kernel<<<(unsigned)((nvec + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint4*>(text), nvec, table,
                                                                     reinterpret_cast<uint4*>(codes));
    return cudaGetLastError();
}

cudaError_t sa_launch_pack0(const uint8_t* src, uint32_t n, int cb, sa::CodeInfo ci, uint64_t* keys, cudaStream_t s)
{
    if (!n) return cudaSuccess;
    if (ci.bits == 8) {
        sa::pack0_kernel<<<(n + 255) / 256, 256, 0, s>>>(src, n, cb, ci, keys);
    } else {
        const uint32_t threads = (n + 7) / 8;
        sa::pack0_codes_kernel<<<(threads + 255) / 256, 256, 0, s>>>(src, n, cb, ci, keys);
    }
    return cudaGetLastError();
}

cudaError_t sa_launch_lcp0(const uint64_t* k, uint32_t n, int cb, int bits, uint32_t lprime, uint16_t* lcp,
                           uint32_t* tie, uint32_t tie_words, cudaStream_t s)
{
    if (!n) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(tie, 0, (size_t)tie_words * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    sa::lcp0_kernel<<<(n + 255) / 256, 256, 0, s>>>(k, n, cb, bits, lprime, lcp, tie);
    return cudaGetLastError();
}

cudaError_t sa_launch_tie_mask(const uint16_t* lcp, uint32_t n, uint32_t depth, uint32_t* tie, uint32_t num_words,
                               cudaStream_t s)
{
    // one warp per data word 0 .. num_words-2
    uint64_t threads = (uint64_t)(num_words - 1) * 32;
    sa::tie_mask_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(lcp, n, depth, tie, num_words);
    return cudaGetLastError();
}

cudaError_t sa_launch_group_flags(const uint32_t* tie, uint32_t nwords_data, uint32_t* amask, uint32_t* hmask,
                                  uint32_t* counts, cudaStream_t s)
{
    constexpr uint32_t T = SA_SCAN_TILE / 32;
    const uint32_t tiles = (nwords_data + T - 1) / T;
    if (!tiles) return cudaSuccess;
    sa::group_flags_kernel<<<tiles, T, 0, s>>>(tie, nwords_data, amask, hmask, counts);
    return cudaGetLastError();
}

cudaError_t sa_launch_scan_pairs(uint32_t* counts, uint32_t num, uint32_t* totals, cudaStream_t s)
{
    sa::scan_pairs_kernel<<<1, 1024, 0, s>>>(counts, num, totals);
    return cudaGetLastError();
}

cudaError_t sa_launch_compact_gather(const uint32_t* amask, const uint32_t* hmask, uint32_t nwords_data,
                                     const uint32_t* tile_offsets, const uint32_t* sa_, const uint8_t* text,
                                     uint32_t depth, int cb, int gbits, sa::CodeInfo ci, uint32_t* slots,
                                     uint64_t* ckeys, uint32_t* cvals, cudaStream_t s)
{
    constexpr uint32_t T = SA_SCAN_TILE / 32;
    const uint32_t tiles = (nwords_data + T - 1) / T;
    if (!tiles) return cudaSuccess;
    sa::compact_gather_kernel<<<tiles, T, 0, s>>>(amask, hmask, nwords_data, tile_offsets, sa_, text, depth, cb,
                                                  gbits, ci, slots, ckeys, cvals);
    return cudaGetLastError();
}

cudaError_t sa_launch_writeback(const uint32_t* slots, const uint64_t* ckeys, const uint32_t* cvals, uint32_t m,
                                uint32_t depth, int cb, int gbits, int bits, uint32_t lprime, uint32_t* sa_,
                                uint16_t* lcp, uint32_t* tie_next, uint32_t tie_words, cudaStream_t s)
{
    if (!m) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(tie_next, 0, (size_t)tie_words * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    sa::writeback_kernel<<<(m + 255) / 256, 256, 0, s>>>(slots, ckeys, cvals, m, depth, cb, gbits, bits, lprime, sa_,
                                                         lcp, tie_next);
    return cudaGetLastError();
}
