// query.cu -- batched substring query, hit expansion, document table, CSV remap (sm_100a).
//
// query_kernel restates get_substring_positions (engine.c:869-918) for a batch: 8 lanes per
// query, each lane compares 8 bytes of the <= L-byte window per probe; both bisections follow
// the reference's probe sequence mid = (first + last) / 2 exactly (the SA is only "sorted
// enough", so an equivalent lower_bound would not be bit-exact) and run interleaved so two
// dependent load chains are in flight per query.
#include "common.cuh"
#include "kernels.h"

namespace sa {

// ------------------------------------------------------------------------------ scans
// single block, in-place exclusive scan of u32; *total = sum
__global__ void __launch_bounds__(1024) scan_u32_kernel(uint32_t* __restrict__ c, uint32_t num,
                                                        uint32_t* __restrict__ total)
{
    __shared__ uint32_t ws[32];
    __shared__ uint32_t carry;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < num; base += 1024) {
        const uint32_t i = base + t;
        const uint32_t a = (i < num) ? c[i] : 0u;
        uint32_t ia = a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, ia, o);
            if (lane >= o) ia += y;
        }
        if (lane == 31) ws[warp] = ia;
        __syncthreads();
        if (warp == 0) {
            uint32_t v = ws[lane], s2 = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, s2, o);
                if (lane >= o) s2 += y;
            }
            ws[lane] = s2 - v;
        }
        __syncthreads();
        const uint32_t ea = carry + ws[warp] + ia - a;
        if (i < num) c[i] = ea;
        __syncthreads();
        if (t == 1023) carry = ea + a;
        __syncthreads();
    }
    if (t == 0) *total = carry;
}

// u32 counts -> u64 exclusive offsets, three small kernels
__global__ void __launch_bounds__(1024) block_sum_kernel(const uint32_t* __restrict__ c, uint32_t num,
                                                         unsigned long long* __restrict__ sums)
{
    __shared__ unsigned long long ws[32];
    const uint32_t i = blockIdx.x * 1024 + threadIdx.x;
    unsigned long long v = (i < num) ? c[i] : 0ull;
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = ws[threadIdx.x];
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) sums[blockIdx.x] = v;
    }
}

// exclusive scan of the num block sums in place (one block, 1024 per step); sums[num] = total
__global__ void __launch_bounds__(1024) scan_u64_block_kernel(unsigned long long* __restrict__ sums, uint32_t num)
{
    __shared__ unsigned long long ws[32];
    __shared__ unsigned long long carry;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < num; base += 1024) {
        const uint32_t i = base + t;
        const unsigned long long a = (i < num) ? sums[i] : 0ull;
        unsigned long long ia = a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, ia, o);
            if (lane >= o) ia += y;
        }
        if (lane == 31) ws[warp] = ia;
        __syncthreads();
        if (warp == 0) {
            unsigned long long v = ws[lane], s2 = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long y = __shfl_up_sync(0xffffffffu, s2, o);
                if (lane >= o) s2 += y;
            }
            ws[lane] = s2 - v;
        }
        __syncthreads();
        const unsigned long long ea = carry + ws[warp] + ia - a;
        if (i < num) sums[i] = ea;
        __syncthreads();
        if (t == 1023) carry = ea + a;
        __syncthreads();
    }
    if (t == 0) sums[num] = carry;
}

__global__ void __launch_bounds__(1024) offsets_kernel(const uint32_t* __restrict__ c, uint32_t num,
                                                       const unsigned long long* __restrict__ sums,
                                                       uint32_t nblocks, uint64_t* __restrict__ off)
{
    __shared__ unsigned long long ws[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t i = blockIdx.x * 1024 + t;
    const unsigned long long a = (i < num) ? c[i] : 0ull;
    unsigned long long ia = a;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long y = __shfl_up_sync(0xffffffffu, ia, o);
        if (lane >= o) ia += y;
    }
    if (lane == 31) ws[warp] = ia;
    __syncthreads();
    if (warp == 0) {
        unsigned long long v = ws[lane], s2 = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, s2, o);
            if (lane >= o) s2 += y;
        }
        ws[lane] = s2 - v;
    }
    __syncthreads();
    if (i < num) off[i] = sums[blockIdx.x] + ws[warp] + ia - a;
    if (i == 0) off[num] = sums[nblocks];
}

// ------------------------------------------------------------------------------ document table
constexpr int NL_TILE = 4096;   // bytes per block, 16 per thread

__device__ __forceinline__ uint32_t count_nl16(uint4 v)
{
    const uint64_t a = ((uint64_t)v.y << 32) | v.x, b = ((uint64_t)v.w << 32) | v.z;
    return __popcll(zero_bytes(a ^ 0x0A0A0A0A0A0A0A0Aull)) + __popcll(zero_bytes(b ^ 0x0A0A0A0A0A0A0A0Aull));
}

__global__ void __launch_bounds__(256) newline_count_kernel(const uint4* __restrict__ text16, uint32_t n,
                                                            uint32_t* __restrict__ counts)
{
    __shared__ uint32_t ws[8];
    const uint64_t byte0 = (uint64_t)blockIdx.x * NL_TILE + threadIdx.x * 16;
    uint32_t c = 0;
    if (byte0 < n) c = count_nl16(text16[byte0 >> 4]);     // bytes past n are zero padding
#pragma unroll
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < 8; ++i) t += ws[i];
        counts[blockIdx.x] = t;
    }
}

// 8-bit mask of the bytes of w that equal '\n' (bit i = byte i)
__device__ __forceinline__ uint32_t nl_mask8(uint64_t w)
{
    const uint64_t z = zero_bytes(w ^ 0x0A0A0A0A0A0A0A0Aull) >> 7;          // 0x01 per matching byte
    return (uint32_t)((z * 0x0102040810204080ull) >> 56);
}

// positions of the '\n' bytes (nlpos, ascending) and a rank structure over them: one bit per text
// byte (nlbits, 64 per word) + the number of '\n' in front of every 64-byte block (nlrank), so
// "record index of text position p" is two independent loads instead of a bisection over nlpos.
// count_base / pos_base: the kernel may run on a slice of a longer text (sharded preparation): text16,
// nlbits and nlrank then point at the slice, the counts start at count_base (newlines in front of the
// slice) and positions at pos_base; nlpos may be NULL (built afterwards from the rank structure).
__global__ void __launch_bounds__(256) newline_write_kernel(const uint4* __restrict__ text16, uint32_t n,
                                                            const uint32_t* __restrict__ tile_offsets,
                                                            uint32_t* __restrict__ nlpos,
                                                            unsigned long long* __restrict__ nlbits,
                                                            uint32_t* __restrict__ nlrank, uint32_t count_base,
                                                            uint32_t pos_base)
{
    __shared__ uint32_t ws[8];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint64_t byte0 = (uint64_t)blockIdx.x * NL_TILE + t * 16;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (byte0 < n) v = text16[byte0 >> 4];
    const uint32_t c = count_nl16(v);
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    uint32_t pre = tile_offsets[blockIdx.x] + count_base;
    for (int w = 0; w < warp; ++w) pre += ws[w];
    uint32_t o = pre + inc - c;
    {
        const uint32_t m16 = nl_mask8(((uint64_t)v.y << 32) | v.x) | (nl_mask8(((uint64_t)v.w << 32) | v.z) << 8);
        const uint32_t m1 = __shfl_down_sync(0xffffffffu, m16, 1), m2 = __shfl_down_sync(0xffffffffu, m16, 2),
                       m3 = __shfl_down_sync(0xffffffffu, m16, 3);
        const uint64_t blk = byte0 >> 6;
        if ((lane & 3) == 0 && blk <= (uint64_t)(n >> 6)) {
            nlbits[blk] = (unsigned long long)m16 | ((unsigned long long)m1 << 16) | ((unsigned long long)m2 << 32) |
                          ((unsigned long long)m3 << 48);
            nlrank[blk] = o;
        }
    }
    if (c && nlpos) {
        const uint8_t* b = reinterpret_cast<const uint8_t*>(&v);
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (b[i] == '\n') nlpos[o++] = pos_base + (uint32_t)(byte0 + i);
    }
}

// ------------------------------------------------------------------------------ query
__device__ __forceinline__ uint64_t load_query_word(const uint8_t* __restrict__ q, uint32_t off, uint32_t c)
{
    // bytes [off, off+8) of the query, big-endian, zero past c.  Unconditional (clamped) byte loads
    // followed by one mask: the per-byte predicated form gave wrong words inside the probe loop of
    // query_thread_kernel for c = 8k + 1 > 32 with nvcc 12.9 (equal windows compared as "less").
    if (off >= c) return 0ull;
    const uint32_t last = c - 1u;
    uint64_t w = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint32_t idx = off + i;
        w = (w << 8) | (uint64_t)q[idx < last ? idx : last];
    }
    const uint32_t nb = (c - off) < 8u ? (c - off) : 8u;
    return keep_top_bytes(w, (int)nb);
}

// sign of strncmp(text + pos, q, c) computed by the 8 lanes of a group.
__device__ __forceinline__ int group_compare(const uint8_t* __restrict__ text, uint32_t pos,
                                             const uint8_t* __restrict__ q, uint32_t c, uint64_t qw0,
                                             uint32_t gmask, int sub, int gshift)
{
    for (uint32_t cbase = 0; cbase < c; cbase += 64) {
        const uint32_t off = cbase + 8u * sub;
        uint64_t tw = 0, qw = 0;
        if (off < c) {
            const uint32_t nb = (c - off) < 8u ? (c - off) : 8u;
            tw = keep_top_bytes(load_be64(text, (uint64_t)pos + off), (int)nb);
            qw = cbase ? load_query_word(q, off, c) : qw0;
        }
        const uint32_t bal = (__ballot_sync(gmask, tw != qw) >> gshift) & 0xFFu;
        if (bal) {
            const int src = gshift + __ffs((int)bal) - 1;
            const uint64_t t2 = __shfl_sync(gmask, tw, src), q2 = __shfl_sync(gmask, qw, src);
            return t2 < q2 ? -1 : 1;
        }
    }
    return 0;
}

__global__ void __launch_bounds__(256)
query_kernel(const uint8_t* __restrict__ text, const uint32_t* __restrict__ sa, uint32_t n, uint32_t L,
             const uint8_t* __restrict__ qbytes, const uint64_t* __restrict__ qoff, uint32_t nq,
             uint32_t* __restrict__ start_out, uint32_t* __restrict__ end_out)
{
    const uint32_t gq = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    if (gq >= nq) return;
    const int lane = threadIdx.x & 31, sub = lane & 7, gshift = lane & 24;
    const uint32_t gmask = 0xFFu << gshift;
    const uint64_t qb = qoff[gq];
    const uint32_t len = (uint32_t)(qoff[gq + 1] - qb);
    const uint8_t* q = qbytes + qb;
    // strlen (engine.c:874): cut at the first NUL
    uint32_t m = len;
    for (uint32_t cbase = 0; cbase < len; cbase += 64) {
        const uint32_t off = cbase + 8u * sub;
        uint32_t first_zero = 8;
#pragma unroll
        for (int i = 7; i >= 0; --i)
            if (off + i < len && q[off + i] == 0) first_zero = i;
        const uint32_t bal = (__ballot_sync(gmask, first_zero < 8) >> gshift) & 0xFFu;
        if (bal) {
            const int src = gshift + __ffs((int)bal) - 1;
            const uint32_t fz = __shfl_sync(gmask, first_zero, src);
            m = cbase + 8u * (uint32_t)(src - gshift) + fz;
            break;
        }
    }
    const uint32_t c = m < L ? m : L;                               // engine.c:881 (L, not L')
    const uint64_t qw0 = load_query_word(q, 8u * sub, c);

    uint32_t start = 0xFFFFFFFFu, end = 0xFFFFFFFFu;
    int64_t f1 = 0, l1 = (int64_t)n - 1, f2 = 0, l2 = (int64_t)n - 1;
    while (f1 <= l1 || f2 <= l2) {
        const bool a1 = f1 <= l1, a2 = f2 <= l2;
        const int64_t m1 = (f1 + l1) / 2, m2 = (f2 + l2) / 2;
        const uint32_t p1 = a1 ? __ldg(sa + m1) : 0u;
        const uint32_t p2 = a2 ? ((a1 && m2 == m1) ? p1 : __ldg(sa + m2)) : 0u;
        if (a1) {
            const int cmp = group_compare(text, p1, q, c, qw0, gmask, sub, gshift);
            if (cmp < 0) f1 = m1 + 1;                               // engine.c:887-893
            else { l1 = m1 - 1; start = (uint32_t)m1; }
            if (a2 && m2 == m1) {
                if (cmp > 0) l2 = m2 - 1; else { f2 = m2 + 1; end = (uint32_t)m2; }
                continue;
            }
        }
        if (a2) {
            const int cmp = group_compare(text, p2, q, c, qw0, gmask, sub, gshift);
            if (cmp > 0) l2 = m2 - 1;                               // engine.c:904-913
            else { f2 = m2 + 1; end = (uint32_t)m2; }
        }
    }
    if (start == 0xFFFFFFFFu) end = 0xFFFFFFFFu;                    // engine.c:896-899
    if (sub == 0) { start_out[gq] = start; end_out[gq] = end; }
}

// Thread-per-query variant: the bisections are chains of dependent loads (SA[mid], then the text
// window), so throughput is set by how many chains are in flight; one query per lane keeps 32 per
// warp in flight instead of 4.  The first 32 query bytes live in registers, text words are
// fetched lazily (most probes differ in the first 8 bytes).  Same probe sequence as above.
//
// Probe cache ("qtop"): the probe sequence mid = (first + last) / 2 does not depend on the query, so the
// nodes of the top T levels of the bisection tree are the same for every query.  qtop[node] (heap order:
// root 1, "go left" 2i, "go right" 2i + 1) holds the first 16 text bytes of the suffix at that node's
// mid: a probe there costs ONE 16-byte read of a 16 MB, L2-resident table instead of the two dependent
// random reads SA[mid] -> text.  Only when the query and the suffix agree on all 16 bytes (and the query
// is longer) does the probe go to SA + text for bytes 16... -- the comparison result, hence the path and
// the ranges, are exactly those of strncmp(text + SA[mid], q, c).
__device__ __forceinline__ int thread_compare(const uint8_t* __restrict__ text, uint32_t pos, const uint8_t* q,
                                              uint32_t c, uint64_t q0, uint64_t q1, uint64_t q2, uint64_t q3,
                                              uint32_t from = 0)
{
    for (uint32_t off = from; off < c; off += 8) {
        const uint32_t nb = (c - off) < 8u ? (c - off) : 8u;
        const uint64_t tw = keep_top_bytes(load_be64(text, (uint64_t)pos + off), (int)nb);
        const uint32_t j = off >> 3;
        // words past the first 32 bytes only exist when max_suffix_length > 32
        const uint64_t w = j == 0 ? q0 : (j == 1 ? q1 : (j == 2 ? q2 : (j == 3 ? q3 : load_query_word(q, off, c))));
        if (tw != w) return tw < w ? -1 : 1;
    }
    return 0;
}

// SA[r]: a plain array, or key ranges spread over several GPUs (SaView, common.cuh)
struct SaPlain {
    const uint32_t* sa;
    __device__ __forceinline__ uint32_t operator()(uint64_t r) const { return __ldg(sa + r); }
};
struct SaSharded {
    SaView v;
    __device__ __forceinline__ uint32_t operator()(uint64_t r) const { return sa_at(v, r); }
};

// one thread per node of the top `levels` levels: walk from the root along the bits of the node id
template <class SA>
__global__ void __launch_bounds__(256)
qtop_build_kernel(const uint8_t* __restrict__ text, const SA sa, uint32_t n, int levels,
                  ulonglong2* __restrict__ qtop)
{
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node == 0 || node >= (1u << levels)) return;
    const int depth = 31 - __clz((int)node);
    int64_t first = 0, last = (int64_t)n - 1;
    for (int b = depth - 1; b >= 0 && first <= last; --b) {
        const int64_t mid = (first + last) / 2;
        if ((node >> b) & 1u) first = mid + 1; else last = mid - 1;
    }
    if (first > last) return;                       // no query ever reaches this node
    const uint32_t p = sa((uint64_t)((first + last) / 2));
    uint64_t w0, w1;
    load_be128(text, p, w0, w1);
    qtop[node] = make_ulonglong2(w0, w1);
}

// sign of strncmp(text + SA[mid], q, c) for a probe at `node` (< 2^levels): from the table, falling
// back to SA + text past byte 16
template <class SA>
__device__ __forceinline__ int qtop_compare(const ulonglong2* __restrict__ qtop, uint32_t node, const uint8_t* __restrict__ text,
                                            const SA& sa, int64_t mid, const uint8_t* q, uint32_t c,
                                            uint64_t q0, uint64_t q1, uint64_t q2, uint64_t q3)
{
    if (c == 0u) return 0;
    const ulonglong2 k = __ldg(qtop + node);
    const uint64_t t0 = keep_top_bytes(k.x, (int)(c < 8u ? c : 8u));
    if (t0 != q0) return t0 < q0 ? -1 : 1;
    if (c <= 8u) return 0;
    const uint64_t t1 = keep_top_bytes(k.y, (int)(c - 8u < 8u ? c - 8u : 8u));
    if (t1 != q1) return t1 < q1 ? -1 : 1;
    if (c <= 16u) return 0;
    return thread_compare(text, sa((uint64_t)mid), q, c, q0, q1, q2, q3, 16u);
}

template <class SA>
__global__ void __launch_bounds__(128)
query_thread_kernel(const uint8_t* __restrict__ text, const SA sa, uint32_t n, uint32_t L,
                    const uint8_t* __restrict__ qbytes, const uint64_t* __restrict__ qoff, uint32_t nq,
                    uint32_t* __restrict__ start_out, uint32_t* __restrict__ end_out,
                    const ulonglong2* __restrict__ qtop, int levels)
{
    const uint32_t gq = blockIdx.x * blockDim.x + threadIdx.x;
    if (gq >= nq) return;
    const uint64_t qb = qoff[gq];
    const uint32_t len = (uint32_t)(qoff[gq + 1] - qb);
    const uint8_t* q = qbytes + qb;
    // c = min(strlen(q), L) (engine.c:874, :881): a NUL past byte L cannot matter
    uint32_t c = len < L ? len : L;
    for (uint32_t i = 0; i < c; ++i)
        if (q[i] == 0) { c = i; break; }
    const uint64_t q0 = load_query_word(q, 0, c), q1 = load_query_word(q, 8, c), q2 = load_query_word(q, 16, c),
                   q3 = load_query_word(q, 24, c);
    const uint32_t top = qtop ? (1u << levels) : 0u;

    uint32_t start = 0xFFFFFFFFu, end = 0xFFFFFFFFu;
    int64_t f1 = 0, l1 = (int64_t)n - 1, f2 = 0, l2 = (int64_t)n - 1;
    uint32_t n1 = 1, n2 = 1;                                        // nodes of the two searches (heap order)
    while (f1 <= l1 || f2 <= l2) {
        const bool a1 = f1 <= l1, a2 = f2 <= l2;
        const int64_t m1 = (f1 + l1) / 2, m2 = (f2 + l2) / 2;
        const bool same = a1 && a2 && m1 == m2;
        int c1 = 0, c2 = 0;
        if (n1 < top || n2 < top) {
            // top of the tree: one table read per probe
            if (a1) c1 = n1 < top ? qtop_compare(qtop, n1, text, sa, m1, q, c, q0, q1, q2, q3)
                                  : thread_compare(text, sa((uint64_t)m1), q, c, q0, q1, q2, q3);
            if (a2) c2 = same ? c1 : (n2 < top ? qtop_compare(qtop, n2, text, sa, m2, q, c, q0, q1, q2, q3)
                                               : thread_compare(text, sa((uint64_t)m2), q, c, q0, q1, q2, q3));
        } else {
            // both SA entries first (independent loads), then the text windows
            const uint32_t p1 = a1 ? sa((uint64_t)m1) : 0u;
            const uint32_t p2 = (a2 && !same) ? sa((uint64_t)m2) : p1;
            if (a1) c1 = thread_compare(text, p1, q, c, q0, q1, q2, q3);
            if (a2) c2 = same ? c1 : thread_compare(text, p2, q, c, q0, q1, q2, q3);
        }
        if (a1) {
            if (c1 < 0) { f1 = m1 + 1; n1 = n1 < top ? 2 * n1 + 1 : top; }      // engine.c:887-893
            else { l1 = m1 - 1; start = (uint32_t)m1; n1 = n1 < top ? 2 * n1 : top; }
        }
        if (a2) {
            if (c2 > 0) { l2 = m2 - 1; n2 = n2 < top ? 2 * n2 : top; }          // engine.c:904-913
            else { f2 = m2 + 1; end = (uint32_t)m2; n2 = n2 < top ? 2 * n2 + 1 : top; }
        }
    }
    if (start == 0xFFFFFFFFu) end = 0xFFFFFFFFu;                    // engine.c:896-899
    start_out[gq] = start;
    end_out[gq] = end;
}

// ------------------------------------------------------------------------------ file semantics
// get_substring_positions_file (engine.c:920-999), the bisection behind the reference's only
// Python-reachable query (suffix_array.pyx:221-247 -> get_matching_records_file).  Differences to
// get_substring_positions that make it a different function of (array, query):
//   * each probe compares the RAW FILE bytes at offset SA[mid] (fseek + fread + tolower, :939-947):
//     past the end of a cell it sees `,next cell` / `"` / the rest of the line, not '\n' + next row
//   * start / end are only recorded when cmp == 0 (:953-957, :987-991)
//   * the second bisection starts at first = start (:969)
//   * `line` is one malloc'd buffer for all probes: fread near the end of the file fills fewer than
//     cmp_length bytes and the tail keeps the (lower-cased) bytes of earlier probes -- reproduced
//     from the list of this query's probe offsets (zero before the first write: the checker builds
//     the reference with zero-filled malloc, oracle/ref_prelude.h)
// The mid == 0 wrap of `last = mid - 1` is fenced as in the in-memory kernels (signed bounds).
__device__ __forceinline__ uint64_t lower8(uint64_t x)
{
    const uint64_t hi = x & 0x8080808080808080ull, lo = x & 0x7F7F7F7F7F7F7F7Full;
    const uint64_t ge_a = lo + 0x3F3F3F3F3F3F3F3Full;      // bit 7 set iff (byte & 0x7f) >= 'A'
    const uint64_t gt_z = lo + 0x2525252525252525ull;      // bit 7 set iff (byte & 0x7f) >  'Z'
    return x | ((ge_a & ~gt_z & ~hi & 0x8080808080808080ull) >> 2);
}

__device__ __forceinline__ uint32_t lower1(uint32_t b) { return (b >= 'A' && b <= 'Z') ? b + 32u : b; }

constexpr int QF_MAX_PROBES = 72;      // 2 * (32 + 1) probes for n < 2^32, rounded up

__global__ void __launch_bounds__(128)
query_file_kernel(const uint8_t* __restrict__ file, uint64_t fsize, const uint32_t* __restrict__ sa, uint32_t n,
                  uint32_t L, const uint8_t* __restrict__ qbytes, const uint64_t* __restrict__ qoff, uint32_t nq,
                  uint32_t* __restrict__ start_out, uint32_t* __restrict__ end_out)
{
    const uint32_t gq = blockIdx.x * blockDim.x + threadIdx.x;
    if (gq >= nq) return;
    const uint64_t qb = qoff[gq];
    const uint32_t len = (uint32_t)(qoff[gq + 1] - qb);
    const uint8_t* q = qbytes + qb;
    uint32_t c = len < L ? len : L;                                  // :932 min(strlen, L)
    for (uint32_t i = 0; i < c; ++i)
        if (q[i] == 0) { c = i; break; }
    const uint64_t q0 = load_query_word(q, 0, c), q1 = load_query_word(q, 8, c), q2 = load_query_word(q, 16, c),
                   q3 = load_query_word(q, 24, c);
    uint32_t probes[QF_MAX_PROBES];
    int np = 0;

    auto compare = [&](uint32_t pos) -> int {
        int r = 0;
        if ((uint64_t)pos + c <= fsize) {
            for (uint32_t off = 0; off < c; off += 8) {
                const uint32_t nb = (c - off) < 8u ? (c - off) : 8u;
                const uint64_t tw = keep_top_bytes(lower8(load_be64(file, (uint64_t)pos + off)), (int)nb);
                const uint32_t j = off >> 3;
                const uint64_t w = j == 0 ? q0 : (j == 1 ? q1 : (j == 2 ? q2 : (j == 3 ? q3 : load_query_word(q, off, c))));
                if (tw != w) { r = tw < w ? -1 : 1; break; }
            }
        } else {
            // the read crosses the end of the file: byte i comes from the latest probe that filled line[i]
            for (uint32_t i = 0; i < c; ++i) {
                uint32_t b = 0;
                if ((uint64_t)pos + i < fsize) b = lower1(file[(uint64_t)pos + i]);
                else
                    for (int h = np - 1; h >= 0; --h)
                        if ((uint64_t)probes[h] + i < fsize) { b = lower1(file[(uint64_t)probes[h] + i]); break; }
                const uint32_t qc = q[i];
                if (b != qc) { r = b < qc ? -1 : 1; break; }
            }
        }
        if (np < QF_MAX_PROBES) probes[np++] = pos;
        return r;
    };

    uint32_t start = 0xFFFFFFFFu, end = 0xFFFFFFFFu;
    int64_t first = 0, last = (int64_t)n - 1;
    while (first <= last) {                                          // :937-960
        const int64_t mid = (first + last) / 2;
        const int cmp = compare(__ldg(sa + mid));
        if (cmp < 0) first = mid + 1;
        else { last = mid - 1; if (cmp == 0) start = (uint32_t)mid; }
    }
    if (start != 0xFFFFFFFFu) {                                      // :962-966
        first = start; last = (int64_t)n - 1;                        // :969
        while (first <= last) {                                      // :971-993
            const int64_t mid = (first + last) / 2;
            const int cmp = compare(__ldg(sa + mid));
            if (cmp > 0) last = mid - 1;
            else { first = mid + 1; if (cmp == 0) end = (uint32_t)mid; }
        }
    }
    start_out[gq] = start;
    end_out[gq] = end;
}

// engine.c:1181-1185: first == UINT32_MAX -> 0 hits, else min(k, second - first + 1) in u32
__global__ void hit_counts_kernel(const uint32_t* __restrict__ start, const uint32_t* __restrict__ end,
                                  uint32_t nq, uint32_t k, uint32_t n, uint32_t* __restrict__ counts)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    uint32_t s = start[i], c = 0;
    if (s != 0xFFFFFFFFu) {
        c = end[i] - s + 1u;
        if (c > k) c = k;
        if (s >= n) c = 0;                       // memory safety only: the reference would read past its array here
        else if (c > n - s) c = n - s;
    }
    counts[i] = c;
}

// one warp per query: rank -> text position -> record index (# of '\n' before the position)
template <class SA>
__global__ void __launch_bounds__(256)
hits_kernel(const SA sa, const uint32_t* __restrict__ start,
            const uint64_t* __restrict__ hit_off, uint32_t nq, const unsigned long long* __restrict__ nlbits,
            const uint32_t* __restrict__ nlrank, uint32_t* __restrict__ hit_pos, uint32_t* __restrict__ hit_doc)
{
    const uint32_t wq = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wq >= nq) return;
    const uint64_t o = hit_off[wq];
    const uint32_t take = (uint32_t)(hit_off[wq + 1] - o);
    const uint32_t s = start[wq];
    for (uint32_t j = threadIdx.x & 31; j < take; j += 32) {
        const uint32_t p = sa((uint64_t)s + j);      // s + j <= end < n by construction of the counts
        if (hit_pos) hit_pos[o + j] = p;
        if (hit_doc) {
            // record index = number of '\n' in front of p
            const unsigned long long bits = __ldg(nlbits + (p >> 6));
            hit_doc[o + j] = __ldg(nlrank + (p >> 6)) + (uint32_t)__popcll(bits & ((1ull << (p & 63u)) - 1ull));
        }
    }
}

// engine.c:632-651: quoted bit of rank i = flag of text byte SA[i]; SA[i] = map[SA[i]]
__global__ void remap_kernel(const uint32_t* __restrict__ sa, uint32_t n, const uint32_t* __restrict__ map,
                             const uint8_t* __restrict__ quoted_by_text, uint32_t* __restrict__ sa_out,
                             uint32_t* __restrict__ qbits)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool qd = false;
    if (i < n) {
        const uint32_t p = sa[i];
        // copy_buffer_bit (engine.c:256-261) copies only n / 8 whole bytes of the flags, so the
        // last n % 8 text bytes read as "not quoted" in the reference's permutation loop
        // the device column scanner packs the flag into bit 31 of the offset (files are < 2 GiB)
        const uint32_t m = map[p];
        qd = (p < (n & ~7u)) && (quoted_by_text ? quoted_by_text[p] != 0 : (m >> 31) != 0);
        sa_out[i] = quoted_by_text ? m : (m & 0x7FFFFFFFu);
    }
    const uint32_t b = __ballot_sync(0xffffffffu, qd);
    if ((threadIdx.x & 31) == 0 && (i >> 5) < ((n + 31) >> 5)) qbits[i >> 5] = b;
}

}  // namespace sa

// ------------------------------------------------------------------------------ launchers
cudaError_t sa_launch_newline_count(const uint8_t* text, uint32_t n, uint32_t* counts, cudaStream_t s)
{
    const uint32_t tiles = sa_newline_tiles(n);
    if (!tiles) return cudaSuccess;
    sa::newline_count_kernel<<<tiles, 256, 0, s>>>(reinterpret_cast<const uint4*>(text), n, counts);
    return cudaGetLastError();
}

// 64-bit arithmetic: n may be as large as 2^32 - 1025 (sharded indexes), where n + 4095 wraps in 32 bits
uint32_t sa_newline_tiles(uint32_t n) { return (uint32_t)(((uint64_t)n + sa::NL_TILE - 1) / sa::NL_TILE); }

cudaError_t sa_launch_scan_u32(uint32_t* counts, uint32_t num, uint32_t* total, cudaStream_t s)
{
    sa::scan_u32_kernel<<<1, 1024, 0, s>>>(counts, num, total);
    return cudaGetLastError();
}

cudaError_t sa_launch_newline_write(const uint8_t* text, uint32_t n, const uint32_t* tile_offsets, uint32_t* nlpos,
                                    unsigned long long* nlbits, uint32_t* nlrank, cudaStream_t s, uint32_t count_base,
                                    uint32_t pos_base)
{
    const uint32_t tiles = sa_newline_tiles(n);
    if (!tiles) return cudaSuccess;
    sa::newline_write_kernel<<<tiles, 256, 0, s>>>(reinterpret_cast<const uint4*>(text), n, tile_offsets, nlpos, nlbits,
                                                   nlrank, count_base, pos_base);
    return cudaGetLastError();
}

cudaError_t sa_launch_query(const uint8_t* text, const uint32_t* sa_, uint32_t n, uint32_t max_suffix_length,
                            const uint8_t* qbytes, const uint64_t* qoff, uint32_t nq, uint32_t* start,
                            uint32_t* end, int mode, cudaStream_t s, const void* qtop, int qtop_levels)
{
    if (!nq) return cudaSuccess;
    if (mode == 1 || (mode == 0 && nq >= 4096)) {        // enough queries to fill the GPU with one query per lane
        sa::query_thread_kernel<<<(nq + 127) / 128, 128, 0, s>>>(text, sa::SaPlain{sa_}, n, max_suffix_length, qbytes, qoff, nq,
                                                                 start, end, static_cast<const ulonglong2*>(qtop), qtop_levels);
    } else {                 // small batches: 8 lanes per query shorten each query's critical path
        const uint64_t threads = (uint64_t)nq * 8;
        sa::query_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(text, sa_, n, max_suffix_length, qbytes,
                                                                          qoff, nq, start, end);
    }
    return cudaGetLastError();
}

// levels: top levels of the bisection tree to cache; qtop holds 2^levels entries of 16 bytes
cudaError_t sa_launch_qtop_build(const uint8_t* text, const uint32_t* sa_, uint32_t n, int levels, void* qtop, cudaStream_t s)
{
    const uint32_t nodes = 1u << levels;
    sa::qtop_build_kernel<<<(nodes + 255) / 256, 256, 0, s>>>(text, sa::SaPlain{sa_}, n, levels, static_cast<ulonglong2*>(qtop));
    return cudaGetLastError();
}

// ---- the same three kernels over an array whose key ranges live on several GPUs (no replication)
cudaError_t sa_launch_query_sharded(const uint8_t* text, const sa::SaView& view, uint32_t n, uint32_t max_suffix_length,
                                    const uint8_t* qbytes, const uint64_t* qoff, uint32_t nq, uint32_t* start, uint32_t* end,
                                    cudaStream_t s, const void* qtop, int qtop_levels)
{
    if (!nq) return cudaSuccess;
    sa::query_thread_kernel<<<(nq + 127) / 128, 128, 0, s>>>(text, sa::SaSharded{view}, n, max_suffix_length, qbytes, qoff, nq,
                                                             start, end, static_cast<const ulonglong2*>(qtop), qtop_levels);
    return cudaGetLastError();
}

cudaError_t sa_launch_qtop_build_sharded(const uint8_t* text, const sa::SaView& view, uint32_t n, int levels, void* qtop,
                                         cudaStream_t s)
{
    const uint32_t nodes = 1u << levels;
    sa::qtop_build_kernel<<<(nodes + 255) / 256, 256, 0, s>>>(text, sa::SaSharded{view}, n, levels, static_cast<ulonglong2*>(qtop));
    return cudaGetLastError();
}

cudaError_t sa_launch_hits_sharded(const sa::SaView& view, const uint32_t* start, const uint64_t* hit_off, uint32_t nq,
                                   const unsigned long long* nlbits, const uint32_t* nlrank, uint32_t* hit_pos,
                                   uint32_t* hit_doc, cudaStream_t s)
{
    if (!nq) return cudaSuccess;
    const uint64_t threads = (uint64_t)nq * 32;
    sa::hits_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(sa::SaSharded{view}, start, hit_off, nq, nlbits, nlrank,
                                                                     hit_pos, hit_doc);
    return cudaGetLastError();
}

cudaError_t sa_launch_hit_counts(const uint32_t* start, const uint32_t* end, uint32_t nq, uint32_t k, uint32_t n,
                                 uint32_t* counts, cudaStream_t s)
{
    if (!nq) return cudaSuccess;
    sa::hit_counts_kernel<<<(nq + 255) / 256, 256, 0, s>>>(start, end, nq, k, n, counts);
    return cudaGetLastError();
}

// block_sums must hold ceil(num/1024) + 1 u64 entries
cudaError_t sa_launch_scan_u32_to_u64(const uint32_t* counts, uint32_t num, uint64_t* off, uint32_t* block_sums,
                                      cudaStream_t s)
{
    unsigned long long* sums = reinterpret_cast<unsigned long long*>(block_sums);
    const uint32_t nb = (num + 1023) / 1024;
    if (!num) return cudaMemsetAsync(off, 0, sizeof(uint64_t), s);
    sa::block_sum_kernel<<<nb, 1024, 0, s>>>(counts, num, sums);
    sa::scan_u64_block_kernel<<<1, 1024, 0, s>>>(sums, nb);
    sa::offsets_kernel<<<nb, 1024, 0, s>>>(counts, num, sums, nb, off);
    return cudaGetLastError();
}

cudaError_t sa_launch_hits(const uint32_t* sa_, const uint32_t* start, const uint64_t* hit_off, uint32_t nq,
                           const unsigned long long* nlbits, const uint32_t* nlrank, uint32_t* hit_pos,
                           uint32_t* hit_doc, cudaStream_t s)
{
    if (!nq) return cudaSuccess;
    const uint64_t threads = (uint64_t)nq * 32;
    sa::hits_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(sa::SaPlain{sa_}, start, hit_off, nq, nlbits, nlrank,
                                                                     hit_pos, hit_doc);
    return cudaGetLastError();
}

cudaError_t sa_launch_remap(const uint32_t* sa_, uint32_t n, const uint32_t* map, const uint8_t* quoted_by_text,
                            uint32_t* sa_out, uint32_t* quoted_bits_by_rank, cudaStream_t s)
{
    if (!n) return cudaSuccess;
    sa::remap_kernel<<<(n + 255) / 256, 256, 0, s>>>(sa_, n, map, quoted_by_text, sa_out, quoted_bits_by_rank);
    return cudaGetLastError();
}

cudaError_t sa_launch_query_file(const uint8_t* file, uint64_t file_size, const uint32_t* sa_file, uint32_t n,
                                 uint32_t max_suffix_length, const uint8_t* qbytes, const uint64_t* qoff, uint32_t nq,
                                 uint32_t* start, uint32_t* end, cudaStream_t s)
{
    if (!nq) return cudaSuccess;
    sa::query_file_kernel<<<(nq + 127) / 128, 128, 0, s>>>(file, file_size, sa_file, n, max_suffix_length, qbytes, qoff, nq,
                                                           start, end);
    return cudaGetLastError();
}
