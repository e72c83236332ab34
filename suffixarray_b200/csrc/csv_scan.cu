// csv_scan.cu -- data-parallel CSV column scanner (sm_100a).
//
// Replaces the byte state machine of the reference's mmap column reader (engine.c:522-605) for
// REGULAR files -- RFC-4180 quoting, no NUL bytes -- on which that state machine is equivalent to
//   in_quotes(p)  = parity of the '"' bytes in front of p
//   column(p)     = (number of ',' / '\n' outside quotes in front of p) mod num_columns
//                   (the reference never resets the column at a row end either, engine.c:547,602)
// so the whole file is scanned with three tile-parallel passes over the device copy of the file
// (quote counts -> delimiter counts -> emitted bytes), each followed by a small prefix scan.
// Emitted for the wanted column, in file order (= the reference's three append_* streams):
//   lower-cased cell bytes, a '\n' per cell at the delimiter's offset, the file offset of every
//   emitted byte with the "cell was quoted" flag in bit 31; `""` inside a quoted cell emits one '"'.
// Anything the equivalence does not cover sets an irregularity flag and sa_build_csv falls back to
// the reference-shaped host scan (sa_csv.cu): a NUL byte, a quote opening in the middle of a cell,
// a closing quote not followed by , \n " or the end of the file, an unterminated quote.
#include "common.cuh"
#include "kernels.h"

namespace sa {

constexpr int CSV_THREADS = 256;
constexpr int CSV_TILE = CSV_THREADS * 16;

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* ws, uint32_t* total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    __syncthreads();                       // ws may still be read from a previous scan
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    uint32_t pre = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < CSV_THREADS / 32; ++w) {
        const uint32_t x = ws[w];
        if (w < warp) pre += x;
        tot += x;
    }
    *total = tot;
    return pre + inc - v;
}

struct Bytes16 {
    uint32_t w[4];
    __device__ __forceinline__ uint32_t at(int j) const { return (w[j >> 2] >> (8 * (j & 3))) & 0xFFu; }
};

// bit j set when byte j of the 16 equals c
__device__ __forceinline__ uint32_t eq_mask(const Bytes16& b, uint32_t c)
{
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t x = b.w[k] ^ (c * 0x01010101u);
        // zero-byte detector, exact per byte
        const uint32_t z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);
        m |= (((z >> 7) & 1u) | ((z >> 14) & 2u) | ((z >> 21) & 4u) | ((z >> 28) & 8u)) << (4 * k);
    }
    return m;
}

__device__ __forceinline__ uint32_t prefix_xor16(uint32_t m)   // bit j = parity of bits 0..j
{
    m ^= m << 1; m ^= m << 2; m ^= m << 4; m ^= m << 8;
    return m & 0xFFFFu;
}

__device__ __forceinline__ Bytes16 load16(const uint8_t* f, uint64_t p)
{
    const uint4 v = *reinterpret_cast<const uint4*>(f + p);
    Bytes16 b;
    b.w[0] = v.x; b.w[1] = v.y; b.w[2] = v.z; b.w[3] = v.w;
    return b;
}

// valid-byte mask of a 16-byte group starting at p (bytes < size)
__device__ __forceinline__ uint32_t valid_mask(uint64_t p, uint64_t size)
{
    if (p + 16 <= size) return 0xFFFFu;
    if (p >= size) return 0u;
    return (1u << (uint32_t)(size - p)) - 1u;
}

// pass 1: '"' bytes per tile; NUL bytes are an irregularity
__global__ void __launch_bounds__(CSV_THREADS)
csv_quotes_kernel(const uint8_t* __restrict__ f, uint64_t size, uint32_t* __restrict__ tile_q, uint32_t* __restrict__ flags)
{
    __shared__ uint32_t ws[CSV_THREADS / 32];
    const uint64_t p = (uint64_t)blockIdx.x * CSV_TILE + threadIdx.x * 16;
    const Bytes16 b = load16(f, p);
    const uint32_t vm = valid_mask(p, size);
    const uint32_t q = eq_mask(b, '"') & vm;
    if (eq_mask(b, 0) & vm) atomicOr(flags, SA_CSV_IRREGULAR_NUL);
    uint32_t tot;
    block_excl_scan(__popc(q), ws, &tot);
    if (threadIdx.x == 0) tile_q[blockIdx.x] = tot;
}

// pass 2: ',' and '\n' outside quotes per tile (tile_q holds the exclusive quote counts now)
__global__ void __launch_bounds__(CSV_THREADS)
csv_delims_kernel(const uint8_t* __restrict__ f, uint64_t size, const uint32_t* __restrict__ tile_q,
                  uint32_t* __restrict__ tile_d)
{
    __shared__ uint32_t ws[CSV_THREADS / 32];
    const uint64_t p = (uint64_t)blockIdx.x * CSV_TILE + threadIdx.x * 16;
    const Bytes16 b = load16(f, p);
    const uint32_t vm = valid_mask(p, size);
    const uint32_t q = eq_mask(b, '"') & vm;
    uint32_t tot;
    const uint32_t qpre = block_excl_scan(__popc(q), ws, &tot) + tile_q[blockIdx.x];
    const uint32_t inq_before = ((prefix_xor16(q) << 1) & 0xFFFFu) ^ ((qpre & 1u) ? 0xFFFFu : 0u);
    const uint32_t d = (eq_mask(b, ',') | eq_mask(b, '\n')) & vm & ~inq_before;
    block_excl_scan(__popc(d), ws, &tot);
    if (threadIdx.x == 0) tile_d[blockIdx.x] = tot;
}

// pass 3: count (WRITE = false) or write (WRITE = true) the wanted column's stream.
// tile_q / tile_d: exclusive per-tile quote / delimiter counts; tile_e: per-tile emit counts out
// (WRITE = false) or exclusive emit offsets in (WRITE = true).  totals[0] = all quotes,
// totals[1] = all delimiters (for the end-of-file cell).
template <bool WRITE>
__global__ void __launch_bounds__(CSV_THREADS)
csv_emit_kernel(const uint8_t* __restrict__ f, uint64_t size, const uint32_t* __restrict__ tile_q,
                const uint32_t* __restrict__ tile_d, uint32_t* __restrict__ tile_e, uint32_t wanted,
                uint32_t num_columns, uint8_t* __restrict__ text, uint32_t* __restrict__ map,
                uint32_t* __restrict__ flags)
{
    __shared__ uint32_t ws[CSV_THREADS / 32];
    const uint64_t p = (uint64_t)blockIdx.x * CSV_TILE + threadIdx.x * 16;
    const Bytes16 b = load16(f, p);
    const uint32_t vm = valid_mask(p, size);
    const uint32_t q = eq_mask(b, '"') & vm;
    uint32_t tot;
    const uint32_t qpre = block_excl_scan(__popc(q), ws, &tot) + tile_q[blockIdx.x];
    const uint32_t inq_before = ((prefix_xor16(q) << 1) & 0xFFFFu) ^ ((qpre & 1u) ? 0xFFFFu : 0u);
    const uint32_t draw = (eq_mask(b, ',') | eq_mask(b, '\n')) & vm;
    const uint32_t d = draw & ~inq_before;
    const uint32_t dpre = block_excl_scan(__popc(d), ws, &tot) + tile_d[blockIdx.x];
    // neighbours of the group: the byte in front and the byte behind (0 outside the file)
    const uint32_t prev = p ? (uint32_t)f[p - 1] : (uint32_t)',';      // the file start behaves like a cell start
    const uint32_t next = (p + 16 < size) ? (uint32_t)f[p + 16] : 0u;

    // bytes of the wanted column: column(p + j) == wanted
    uint32_t colmask = 0;
    {
        uint32_t col = dpre % num_columns;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (col == wanted) colmask |= 1u << j;
            if ((d >> j) & 1u) { ++col; if (col == num_columns) col = 0; }
        }
    }
    const uint32_t qnext = ((q >> 1) | ((next == '"') ? 0x8000u : 0u)) & 0xFFFFu;   // byte j+1 is a quote
    const uint32_t qprev = ((q << 1) | ((prev == '"') ? 1u : 0u)) & 0xFFFFu;        // byte j-1 is a quote
    // regularity: opening quotes start a cell (or re-open after an escape), closing quotes end one
    {
        const uint32_t dprev_raw = ((draw << 1) | ((prev == ',' || prev == '\n') ? 1u : 0u)) & 0xFFFFu;
        const uint32_t opening = q & ~inq_before;
        if (opening & ~(dprev_raw | qprev)) atomicOr(flags, SA_CSV_IRREGULAR_QUOTE);
        const uint32_t closing = q & inq_before;
        uint32_t dnext_raw = (draw >> 1) | ((next == ',' || next == '\n') ? 0x8000u : 0u);
        // the byte behind the last file byte counts as a delimiter
        if (p + 16 >= size && size > p) dnext_raw |= 1u << (uint32_t)(size - p - 1);
        if (closing & ~(dnext_raw | qnext)) atomicOr(flags, SA_CSV_IRREGULAR_QUOTE);
    }
    // emitted bytes
    const uint32_t escaped = q & inq_before & qnext;                  // first quote of a "" pair inside quotes
    uint32_t emit = colmask & vm & ((~q & 0xFFFFu) | escaped);        // cell bytes + sentinels, no bare quotes
    // end of the file inside the wanted cell: one more sentinel at offset `size`
    bool eof_sentinel = false;
    if (size >= p && size < p + 16 && size > 0) {
        const uint32_t j = (uint32_t)(size - p);                      // slot of the virtual byte
        const uint32_t last = j ? b.at(j - 1) : prev;
        const bool last_inq = j ? ((inq_before >> (j - 1)) & 1u) : false;   // state in front of the last byte
        const bool last_is_delim = (last == ',' || last == '\n') && !last_inq && !(j == 0 && p == 0);
        // column and quote state at the end of the file
        uint32_t col = dpre % num_columns;
        for (uint32_t t = 0; t < j; ++t)
            if ((d >> t) & 1u) { ++col; if (col == num_columns) col = 0; }
        const uint32_t qtot = qpre + __popc(q);
        if (qtot & 1u) atomicOr(flags, SA_CSV_IRREGULAR_QUOTE);       // unterminated quote
        eof_sentinel = !last_is_delim && col == wanted;
    }
    const uint32_t cnt = __popc(emit) + (eof_sentinel ? 1u : 0u);
    const uint32_t epre = block_excl_scan(cnt, ws, &tot);
    if (!WRITE) {
        if (threadIdx.x == 0) tile_e[blockIdx.x] = tot;
        return;
    }
    uint64_t o = (uint64_t)tile_e[blockIdx.x] + epre;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        if (!((emit >> j) & 1u)) continue;
        uint32_t c = b.at(j);
        uint32_t qf;
        if ((d >> j) & 1u) {                       // the cell's delimiter -> row sentinel
            c = '\n';
            qf = (qprev >> j) & 1u;                // a quoted cell ends with its closing quote
        } else {
            qf = (inq_before >> j) & 1u;
            if (c >= 'A' && c <= 'Z') c += 32;
        }
        text[o] = (uint8_t)c;
        map[o] = (uint32_t)(p + j) | (qf << 31);     // offsets are < 2^31: the quoted flag rides in bit 31
        ++o;
    }
    if (eof_sentinel) {
        const uint32_t j = (uint32_t)(size - p);
        const uint32_t last = j ? b.at(j - 1) : prev;
        text[o] = '\n';
        map[o] = (uint32_t)size | ((last == '"') ? 0x80000000u : 0u);
    }
}

// Row table of the FILE: bit p of the mask = file byte p is a '\n' outside quotes (a record end), one
// u16 per 16-byte group = one u64 per 64 file bytes, the layout of the text's rank structure
// (newline_write_kernel, query.cu), so hits_kernel turns the file offset of a hit into its row index
// -- header line = row 0 -- with two loads.  Counting rows in the FILE, not '\n' bytes in the column
// text, keeps row ids right for cells that hold a raw newline inside quotes and for ragged rows
// (the reference seeks the record of a hit from its file offset too, engine.c:1217-1323).
// pass A: masks + per-tile counts
__global__ void __launch_bounds__(CSV_THREADS)
csv_rows_kernel(const uint8_t* __restrict__ f, uint64_t size, const uint32_t* __restrict__ tile_q,
                uint16_t* __restrict__ rowbits16, uint32_t* __restrict__ tile_r)
{
    __shared__ uint32_t ws[CSV_THREADS / 32];
    const uint64_t p = (uint64_t)blockIdx.x * CSV_TILE + threadIdx.x * 16;
    const Bytes16 b = load16(f, p);
    const uint32_t vm = valid_mask(p, size);
    const uint32_t q = eq_mask(b, '"') & vm;
    uint32_t tot;
    const uint32_t qpre = block_excl_scan(__popc(q), ws, &tot) + tile_q[blockIdx.x];
    const uint32_t inq_before = ((prefix_xor16(q) << 1) & 0xFFFFu) ^ ((qpre & 1u) ? 0xFFFFu : 0u);
    const uint32_t nl = eq_mask(b, '\n') & vm & ~inq_before;
    rowbits16[p >> 4] = (uint16_t)nl;
    block_excl_scan(__popc(nl), ws, &tot);
    if (threadIdx.x == 0) tile_r[blockIdx.x] = tot;
}

// pass B: rows in front of every 64-byte block (tile_r holds the exclusive per-tile counts now)
__global__ void __launch_bounds__(CSV_THREADS)
csv_rowrank_kernel(const uint16_t* __restrict__ rowbits16, const uint32_t* __restrict__ tile_r,
                   uint32_t* __restrict__ rowrank)
{
    __shared__ uint32_t ws[CSV_THREADS / 32];
    const uint64_t g = (uint64_t)blockIdx.x * CSV_THREADS + threadIdx.x;      // 16-byte group of the file
    uint32_t tot;
    const uint32_t pre = block_excl_scan(__popc((uint32_t)rowbits16[g]), ws, &tot) + tile_r[blockIdx.x];
    if ((threadIdx.x & 3) == 0) rowrank[g >> 2] = pre;
}

}  // namespace sa

uint32_t sa_csv_tiles(uint64_t size) { return (uint32_t)((size + 1 + sa::CSV_TILE - 1) / sa::CSV_TILE); }
uint64_t sa_csv_padded_bytes(uint64_t size) { return (uint64_t)sa_csv_tiles(size) * sa::CSV_TILE + 16; }

cudaError_t sa_launch_csv_quotes(const uint8_t* f, uint64_t size, uint32_t* tile_q, uint32_t* flags, cudaStream_t s)
{
    sa::csv_quotes_kernel<<<sa_csv_tiles(size), sa::CSV_THREADS, 0, s>>>(f, size, tile_q, flags);
    return cudaGetLastError();
}

cudaError_t sa_launch_csv_delims(const uint8_t* f, uint64_t size, const uint32_t* tile_q, uint32_t* tile_d, cudaStream_t s)
{
    sa::csv_delims_kernel<<<sa_csv_tiles(size), sa::CSV_THREADS, 0, s>>>(f, size, tile_q, tile_d);
    return cudaGetLastError();
}

cudaError_t sa_launch_csv_emit(bool write, const uint8_t* f, uint64_t size, const uint32_t* tile_q, const uint32_t* tile_d,
                               uint32_t* tile_e, uint32_t wanted, uint32_t num_columns, uint8_t* text, uint32_t* map,
                               uint32_t* flags, cudaStream_t s)
{
    if (write)
        sa::csv_emit_kernel<true><<<sa_csv_tiles(size), sa::CSV_THREADS, 0, s>>>(f, size, tile_q, tile_d, tile_e, wanted,
                                                                                num_columns, text, map, flags);
    else
        sa::csv_emit_kernel<false><<<sa_csv_tiles(size), sa::CSV_THREADS, 0, s>>>(f, size, tile_q, tile_d, tile_e, wanted,
                                                                                 num_columns, text, map, flags);
    return cudaGetLastError();
}

// rowbits: sa_csv_rowbits_bytes(size) bytes, rowrank: sa_csv_rowrank_bytes(size) bytes
uint64_t sa_csv_rowbits_bytes(uint64_t size) { return (uint64_t)sa_csv_tiles(size) * sa::CSV_THREADS * sizeof(uint16_t); }
uint64_t sa_csv_rowrank_bytes(uint64_t size) { return (uint64_t)sa_csv_tiles(size) * (sa::CSV_THREADS / 4) * sizeof(uint32_t); }

cudaError_t sa_launch_csv_rows(const uint8_t* f, uint64_t size, const uint32_t* tile_q, unsigned long long* rowbits,
                               uint32_t* tile_r, cudaStream_t s)
{
    sa::csv_rows_kernel<<<sa_csv_tiles(size), sa::CSV_THREADS, 0, s>>>(f, size, tile_q, reinterpret_cast<uint16_t*>(rowbits),
                                                                       tile_r);
    return cudaGetLastError();
}

cudaError_t sa_launch_csv_rowrank(const unsigned long long* rowbits, uint64_t size, const uint32_t* tile_r, uint32_t* rowrank,
                                  cudaStream_t s)
{
    sa::csv_rowrank_kernel<<<sa_csv_tiles(size), sa::CSV_THREADS, 0, s>>>(reinterpret_cast<const uint16_t*>(rowbits), tile_r,
                                                                          rowrank);
    return cudaGetLastError();
}
