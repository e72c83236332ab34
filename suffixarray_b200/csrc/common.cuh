// common.cuh -- shared device helpers for libsa_cuda (sm_100a).
//
// Key representation used everywhere: a "masked word" is a big-endian u64 whose TOP cb bytes
// are bytes [depth, depth+cb) of a suffix, with every byte after the first terminator
// ('\n' or 0xFF -- the buckets the reference never refines, engine.c:771-773,795-796) set to
// zero, and the low 8-cb bytes zero.  Unsigned integer order on masked words == the order of
// the reference's per-byte counting sorts over those bytes (engine.c:734-758).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SA_SMALL 32            // insertion-sort cut-off, engine.c:710
#define SA_TEXT_PAD 1024       // zero bytes kept after the device text

namespace sa {

__device__ __forceinline__ uint64_t bswap64(uint64_t v)
{
    uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
    lo = __byte_perm(lo, 0, 0x0123);
    hi = __byte_perm(hi, 0, 0x0123);
    return ((uint64_t)lo << 32) | hi;
}

// 8 text bytes starting at byte offset `a` (any alignment), first byte in the top byte.
// The text allocation is 8-byte aligned and zero padded, so the two aligned loads are safe.
__device__ __forceinline__ uint64_t load_be64(const uint8_t* __restrict__ text, uint64_t a)
{
    const uint64_t* w = reinterpret_cast<const uint64_t*>(text + (a & ~7ull));
    uint64_t lo = __ldg(w);
    unsigned sh = (unsigned)(a & 7ull) * 8u;
    uint64_t v = lo;
    if (sh) {
        uint64_t hi = __ldg(w + 1);
        v = (lo >> sh) | (hi << (64u - sh));
    }
    return bswap64(v);
}

// 0x80 in every byte of v that is zero, exact (no borrow artefacts).
__device__ __forceinline__ uint64_t zero_bytes(uint64_t v)
{
    const uint64_t k7f = 0x7F7F7F7F7F7F7F7Full;
    return ~(((v & k7f) + k7f) | v | k7f);
}

// 0x80 in every byte of be that is '\n' (0x0A) or 0xFF.
__device__ __forceinline__ uint64_t terminator_bytes(uint64_t be)
{
    return zero_bytes(be ^ 0x0A0A0A0A0A0A0A0Aull) | zero_bytes(~be);
}

// keep the top cb bytes (1 <= cb <= 8)
__device__ __forceinline__ uint64_t keep_top_bytes(uint64_t be, int cb)
{
    return cb >= 8 ? be : (be & (~0ull << (8 * (8 - cb))));
}

// masked word: top cb bytes, zero after the first terminator.
__device__ __forceinline__ uint64_t mask_word(uint64_t be, int cb)
{
    be = keep_top_bytes(be, cb);
    uint64_t t = terminator_bytes(be);
    if (t) {
        int j = __clzll((long long)t) >> 3;           // index of the first terminator byte
        be &= ~0ull << (8 * (7 - j));
    }
    return be;
}

__device__ __forceinline__ bool word_has_terminator(uint64_t masked) { return terminator_bytes(masked) != 0; }

// LCP contribution of two masked words sorted adjacent in one group: returns the new lcp value
// (bytes), lprime when they are tied through a terminator or to the truncation.
__device__ __forceinline__ uint32_t pair_lcp(uint64_t a, uint64_t b, uint32_t depth, int cb, uint32_t lprime)
{
    uint64_t x = a ^ b;
    uint32_t v;
    if (x == 0) {
        v = word_has_terminator(b) ? lprime : depth + (uint32_t)cb;
    } else {
        uint32_t cm = (uint32_t)__clzll((long long)x) >> 3;
        if (cm > (uint32_t)cb) cm = (uint32_t)cb;
        v = depth + cm;
    }
    return v < lprime ? v : lprime;
}

// ---------------------------------------------------------------------------------------------
// Compressed alphabet (code_bits < 8).  When the text uses at most 128 distinct byte values the
// build works on a code array (codes[i] = order-preserving dense rank of text[i], NUL always
// code 0) and packs `nchars` codes of `bits` bits into a word: chars top-aligned, then ONE flag bit
// ("a terminator is among these chars") right below them, so nchars * bits <= 63.  Chars after
// the first terminator are code 0, exactly like masked words.  More suffix bytes per 64-bit key
// means fewer radix passes per byte and fewer refinement rounds.
struct CodeInfo {
    int bits;        // 1..8; 8 = identity (keys are masked words of raw bytes)
    uint32_t nl;     // code of '\n' (0x100 when absent)
    uint32_t ff;     // code of 0xFF (0x100 when absent)
};

// 16 bytes starting at byte offset a (any alignment) as two big-endian words.
__device__ __forceinline__ void load_be128(const uint8_t* __restrict__ t, uint64_t a, uint64_t& w0, uint64_t& w1)
{
    const uint64_t* w = reinterpret_cast<const uint64_t*>(t + (a & ~7ull));
    const uint64_t x0 = __ldg(w), x1 = __ldg(w + 1);
    const unsigned sh = (unsigned)(a & 7ull) * 8u;
    uint64_t lo = x0, hi = x1;
    if (sh) {
        const uint64_t x2 = __ldg(w + 2);
        lo = (x0 >> sh) | (x1 << (64u - sh));
        hi = (x1 >> sh) | (x2 << (64u - sh));
    }
    w0 = bswap64(lo);
    w1 = bswap64(hi);
}

// nchars (<= 16, nchars * bits <= 63) codes from the two big-endian code words.
__device__ __forceinline__ uint64_t pack_chars(uint64_t w0, uint64_t w1, int nchars, CodeInfo ci)
{
    uint64_t key = 0;
    bool term = false;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        if (j < nchars) {
            uint32_t c = (uint32_t)((j < 8 ? (w0 >> (56 - 8 * j)) : (w1 >> (56 - 8 * (j - 8)))) & 0xFFull);
            if (term) c = 0;
            key = (key << ci.bits) | c;
            term = term || c == ci.nl || c == ci.ff;
        }
    }
    const int used = nchars * ci.bits;
    key <<= (64 - used);
    if (term) key |= 1ull << (63 - used);
    return key;
}

// LCP contribution of two packed-code words (same nchars) sorted adjacent in one group.
__device__ __forceinline__ uint32_t pair_lcp_codes(uint64_t a, uint64_t b, uint32_t depth, int nchars, int bits,
                                                    uint32_t lprime)
{
    const uint64_t x = a ^ b;
    uint32_t v;
    if (x == 0) {
        v = ((b >> (63 - nchars * bits)) & 1ull) ? lprime : depth + (uint32_t)nchars;
    } else {
        uint32_t cm = (uint32_t)__clzll((long long)x) / (uint32_t)bits;
        if (cm > (uint32_t)nchars) cm = (uint32_t)nchars;
        v = depth + cm;
    }
    return v < lprime ? v : lprime;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t lanemask_lt()
{
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// streaming (evict-first) 8/4-byte loads and stores for the sort passes
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t* p)
{
    uint64_t v;
    asm volatile("ld.global.cs.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.global.cs.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

}  // namespace sa

// error flag bits written by device-side guards (sa_index::d_err)
namespace sa {
// where finish_kernel puts a placed position: out[g] + rank for every replica g, ranks [lo, hi) only
struct FinishOut {
    uint32_t* out[16];
    int n;
    uint32_t lo, hi;
};
}  // namespace sa

namespace sa {
// The complete array as a set of key ranges that live on different GPUs (a sharded build without
// replication): global rank r with bound[g] <= r < bound[g + 1] is base[g][r] -- base[g] is rank g's final
// array (its own memory, or a peer's mapped through CUDA IPC and read over NVLink).
struct SaView {
    const uint32_t* base[16];
    uint32_t bound[17];
    int n;
};
__device__ __forceinline__ uint32_t sa_at(const SaView& v, uint64_t r)
{
    int g = 0;
#pragma unroll
    for (int t = 1; t < 16; ++t) g += (t < v.n && r >= (uint64_t)v.bound[t]) ? 1 : 0;
    return __ldg(v.base[g] + r);
}
}  // namespace sa

#define SA_DEVERR_LOOKBACK_TIMEOUT 1u
#define SA_DEVERR_RUN_TOO_LONG 2u
// csv_scan.cu: the file is outside the cases where the tile-parallel scan equals the reference's state machine
#define SA_CSV_IRREGULAR_NUL 1u
#define SA_CSV_IRREGULAR_QUOTE 2u
