// sa_csv.cu -- CSV column feeder + device remap behind sa_build_csv.
//
// Replaces construct_truncated_suffix_array_from_csv_partitioned_mmap[_full]
// (engine.c:1454-1482 -> :461-654).  The file is read into a pinned ring and copied to the device
// asynchronously; the column scan (engine.c:522-605) runs there as three tile-parallel passes
// (csv_scan.cu).  Files outside RFC-4180 quoting, where that scan and the reference's byte state
// machine could differ, take the reference-shaped host scan below instead (into pinned buffers,
// async H2D).  The build runs on the device; the two post-build loops of the reference
// (quoted-bit permutation, serial, engine.c:632-645; SA[i] = map[SA[i]], engine.c:647-651) are
// one gather kernel (remap_kernel, query.cu).
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "sa_index.h"

namespace {

struct Mapped {
    const uint8_t* p = nullptr;
    size_t size = 0;
    int fd = -1;
    ~Mapped()
    {
        if (p) munmap(const_cast<uint8_t*>(p), size);
        if (fd >= 0) close(fd);
    }
};

struct Pinned {
    void* p = nullptr;
    ~Pinned() { if (p) cudaFreeHost(p); }
    cudaError_t alloc(size_t bytes) { return cudaMallocHost(&p, bytes ? bytes : 1); }
    template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

inline uint8_t lower_ascii(uint8_t c) { return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c; }

// Header cells of the first record (engine.c:26-96 splits on ',' and strips one pair of quotes).
std::vector<std::string> header_cells(const uint8_t* f, size_t size)
{
    std::vector<std::string> cells;
    std::string cur;
    size_t i = 0;
    bool in_q = false;
    for (; i < size; ++i) {
        const uint8_t c = f[i];
        if (in_q) {
            if (c == '"') {
                if (i + 1 < size && f[i + 1] == '"') { cur.push_back('"'); ++i; }
                else in_q = false;
            } else cur.push_back((char)c);
        } else if (c == '"') in_q = true;
        else if (c == ',') { cells.push_back(cur); cur.clear(); }
        else if (c == '\n') break;
        else if (c == '\r' && i + 1 < size && f[i + 1] == '\n') continue;
        else cur.push_back((char)c);
    }
    cells.push_back(cur);
    return cells;
}

// at(i): file byte, 0 past the end (the reference relies on the zero tail of the last page).
struct FileView {
    const uint8_t* f;
    size_t size;
    inline uint8_t at(size_t i) const { return i < size ? f[i] : 0; }
};

// One pass over the file extracting `column_idx` of every record (the header row included, as
// the reference does: its scan starts at byte 0, engine.c:513).  Same outputs as the reference's
// three append_* streams: lower-cased text with a '\n' per row, file offset per text byte
// (u32, as buffer_u32 truncates), quoted flag per text byte.
// Partitions (engine.c:1469-1481): the scan starts at `start` as if a record began there (col = 0, outside
// quotes) and runs while pos < end_point; a cell that straddles end_point is copied to its end (the
// inner loops of engine.c:553-598 do not look at end_point).  *stop = where the scan stopped =
// global_byte_end_idx (engine.c:611), the start of the next partition.
uint64_t scan_column(FileView v, uint32_t column_idx, uint32_t num_columns, uint8_t* text, uint32_t* map,
                     uint8_t* quoted, size_t start = 0, size_t end_point = (size_t)-1, size_t* stop = nullptr,
                     uint64_t cap = ~0ull)
{
    const size_t end = end_point < v.size ? end_point : v.size;
    size_t pos = start;
    uint64_t out = 0;
    uint32_t col = 0;
    struct StopAt { size_t* p; size_t& pos; ~StopAt() { if (p) *p = pos; } } stop_at{stop, pos};
    auto emit = [&](uint8_t c, size_t where, uint8_t q) {
        if (out < cap) { text[out] = c; map[out] = (uint32_t)where; quoted[out] = q; }
        ++out;
    };
    while (pos < end) {
        // skip cells in front of the wanted column (quote aware), engine.c:523-551
        while (col != column_idx) {
            if (v.at(pos) == '"') {
                ++pos;
                for (;;) {
                    if (v.at(pos) == '"') {
                        if (v.at(pos + 1) == '"') { pos += 2; if (pos >= end) return out; continue; }
                        ++pos; if (pos >= end) return out;
                        break;
                    }
                    ++pos; if (pos >= end) return out;
                }
            }
            const uint8_t c = v.at(pos);
            if (c == ',' || c == '\n') col = (col + 1) % num_columns;
            ++pos; if (pos >= end) return out;
        }
        const uint8_t is_q = v.at(pos) == '"';
        if (is_q) {                                   // engine.c:555-586
            ++pos;
            for (;;) {
                const uint8_t c = v.at(pos), c1 = v.at(pos + 1);
                if (c == '"' && c1 == '"') { emit(lower_ascii(c), pos, 1); pos += 2; continue; }
                if (c == '"' && (c1 == '\n' || c1 == ',' || c1 == 0)) { ++pos; break; }
                emit(lower_ascii(c), pos, 1);
                ++pos;
                if (pos >= v.size) break;             // unterminated quote (the reference runs off the mapping)
            }
        } else {                                      // engine.c:588-598
            for (;;) {
                const uint8_t c = v.at(pos);
                if (c == '\n' || c == 0 || c == ',') break;
                emit(lower_ascii(c), pos, 0);
                ++pos;
            }
        }
        emit('\n', pos, is_q);                        // row sentinel, engine.c:600-604
        ++pos;
        col = (col + 1) % num_columns;
    }
    return out;
}

}  // namespace

// File bytes -> device: pread into a two-slot pinned ring, one async H2D copy per slot; the read
// of slot i+1 overlaps the copy of slot i.
static int stream_file_to_device(sa_index* idx, int fd, size_t size, uint8_t* d_file)
{
    constexpr size_t SLOT = 32u << 20;
    cudaStream_t s = idx->stream;
    if (!idx->h_ring[0]) {
        for (int i = 0; i < 2; ++i) {
            SA_CU(cudaMallocHost(&idx->h_ring[i], SLOT));
            SA_CU(cudaEventCreateWithFlags(&idx->ring_ev[i], cudaEventDisableTiming));
        }
        idx->ring_bytes = SLOT;
    }
    size_t off = 0;
    for (int slot = 0; off < size; slot ^= 1) {
        SA_CU(cudaEventSynchronize(idx->ring_ev[slot]));          // the slot's previous copy has drained
        const size_t len = size - off < SLOT ? size - off : SLOT;
        // the page-cache copy is the slow part of this loop: split the slot over a few reader threads
        constexpr int MAX_READERS = 16;
        static const int READERS = [] {
            // default: one reader per host core, at most 16 (797 MB file on a 16-core box: read + H2D 33 ms
            // with 4 readers, 26 with 8, 19.5 with 16; SA_CSV_READERS overrides)
            const char* e = getenv("SA_CSV_READERS");
            int v = e ? atoi(e) : (int)std::thread::hardware_concurrency();
            return v < 1 ? 1 : (v > MAX_READERS ? MAX_READERS : v);
        }();
        uint8_t* dst = static_cast<uint8_t*>(idx->h_ring[slot]);
        std::atomic<int> failed{0};
        auto read_part = [&](size_t b, size_t e) {
            while (b < e) {
                const ssize_t r = pread(fd, dst + b, e - b, (off_t)(off + b));
                if (r <= 0) { failed = 1; return; }
                b += (size_t)r;
            }
        };
        const size_t part = ((len + READERS - 1) / READERS + 4095) & ~(size_t)4095;
        std::thread th[MAX_READERS - 1];
        int started = 0;
        for (int t = 1; t < READERS && (size_t)t * part < len; ++t, ++started)
            th[t - 1] = std::thread(read_part, (size_t)t * part, std::min(len, (size_t)(t + 1) * part));
        read_part(0, std::min(len, part));
        for (int t = 0; t < started; ++t) th[t].join();
        if (failed) { sa_set_error("read failed near offset %zu", off); return SA_ERR_IO; }
        SA_CU(cudaMemcpyAsync(d_file + off, idx->h_ring[slot], len, cudaMemcpyHostToDevice, s));
        SA_CU(cudaEventRecord(idx->ring_ev[slot], s));
        off += len;
    }
    return SA_OK;
}

// The file on the device (on src): upload, quote parity per tile, the row table of the file, and --
// unless the host scanner is forced -- the delimiter counts per tile.  Shared by every column of a
// multi-column build.
static int csv_load_device(sa_index* src, int fd, size_t size, bool want_delims)
{
    cudaStream_t s = src->stream;
    uint32_t* sm = src->small.as<uint32_t>();
    const uint32_t tiles = sa_csv_tiles(size);
    const uint64_t padded = sa_csv_padded_bytes(size);
    SA_CU(src->csv_file.ensure(padded));
    SA_CU(src->csv_tq.ensure((size_t)tiles * 4));
    SA_CU(src->csv_td.ensure((size_t)tiles * 4));
    SA_CU(src->csv_te.ensure((size_t)tiles * 4));
    SA_CU(src->csv_tr.ensure((size_t)tiles * 4));
    SA_CU(src->rowbits.ensure(sa_csv_rowbits_bytes(size)));
    SA_CU(src->rowrank.ensure(sa_csv_rowrank_bytes(size)));
    uint8_t* d_file = src->csv_file.as<uint8_t>();
    const size_t tail = size & ~(size_t)15;
    SA_CU(cudaMemsetAsync(d_file + tail, 0, padded - tail, s));
    SA_CU(cudaMemsetAsync(sm + SM_CSV, 0, 8 * sizeof(uint32_t), s));
    {
        ScopedPhase ph(src, &src->stats.ms_h2d);
        int rc = stream_file_to_device(src, fd, size, d_file);
        if (rc != SA_OK) return rc;
    }
    uint32_t *tq = src->csv_tq.as<uint32_t>(), *td = src->csv_td.as<uint32_t>(), *tr = src->csv_tr.as<uint32_t>();
    ScopedPhase ph(src, &src->stats.ms_csv_scan);
    SA_CU(sa_launch_csv_quotes(d_file, size, tq, sm + SM_CSV, s));
    SA_CU(sa_launch_scan_u32(tq, tiles, sm + SM_CSV + 1, s));
    SA_CU(sa_launch_csv_rows(d_file, size, tq, src->rowbits.as<unsigned long long>(), tr, s));
    SA_CU(sa_launch_scan_u32(tr, tiles, sm + SM_CSV + 4, s));
    SA_CU(sa_launch_csv_rowrank(src->rowbits.as<unsigned long long>(), size, tr, src->rowrank.as<uint32_t>(), s));
    src->stats.kernel_launches += 5;
    if (want_delims) {
        SA_CU(sa_launch_csv_delims(d_file, size, tq, td, s));
        SA_CU(sa_launch_scan_u32(td, tiles, sm + SM_CSV + 2, s));
        src->stats.kernel_launches += 2;
    }
    return SA_OK;
}

// Device column scan (csv_scan.cu) of column `col` of the file held by src into dst's text / map.
// *regular = false when the file needs the host state machine; otherwise dst->n is the stream length.
static int scan_column_device(sa_index* src, sa_index* dst, size_t size, uint32_t col, uint32_t num_columns, bool* regular)
{
    cudaStream_t s = src->stream;
    uint32_t* sm = src->small.as<uint32_t>();
    const uint32_t tiles = sa_csv_tiles(size);
    const uint8_t* d_file = src->csv_file.as<uint8_t>();
    uint32_t *tq = src->csv_tq.as<uint32_t>(), *td = src->csv_td.as<uint32_t>(), *te = src->csv_te.as<uint32_t>();
    {
        ScopedPhase ph(src, &dst->stats.ms_csv_scan);
        SA_CU(sa_launch_csv_emit(false, d_file, size, tq, td, te, col, num_columns, nullptr, nullptr, sm + SM_CSV, s));
        SA_CU(sa_launch_scan_u32(te, tiles, sm + SM_CSV + 3, s));
        dst->stats.kernel_launches += 2;
    }
    SA_CU(cudaMemcpyAsync(src->h_small, sm + SM_CSV, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SA_CU(cudaStreamSynchronize(s));
    const uint32_t irregular = src->h_small[0];
    const uint64_t n = src->h_small[3];
    *regular = irregular == 0;
    if (irregular) return SA_OK;
    if (n > SA_MAX_INDEX_BYTES) { sa_set_error("column text of %llu bytes exceeds the 2^31 limit", (unsigned long long)n); return SA_ERR_LIMIT; }
    dst->n = n; dst->tail = 0;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(dst->text.ensure(bytes));
    SA_CU(dst->map.ensure((size_t)n * 4 + 4));
    SA_CU(cudaMemsetAsync(dst->text.as<uint8_t>() + ((size_t)n & ~(size_t)15), 0, bytes - ((size_t)n & ~(size_t)15), s));
    {
        ScopedPhase ph(src, &dst->stats.ms_csv_scan);
        SA_CU(sa_launch_csv_emit(true, d_file, size, tq, td, te, col, num_columns, dst->text.as<uint8_t>(),
                                 dst->map.as<uint32_t>(), sm + SM_CSV, s));
        dst->stats.kernel_launches += 1;
    }
    SA_CU(cudaStreamSynchronize(s));          // dst builds on its own stream
    return SA_OK;
}

// The reference-shaped host scan into pinned buffers (irregular files, or sa_debug_set csv_scanner = 1)
static int scan_column_host(sa_index* idx, const Mapped& m, uint32_t col, uint32_t num_columns)
{
    const size_t cap = m.size + 2;
    Pinned h_text, h_map, h_q;
    cudaError_t e = h_text.alloc(cap);
    if (e == cudaSuccess) e = h_map.alloc(cap * 4);
    if (e == cudaSuccess) e = h_q.alloc(cap);
    if (e != cudaSuccess) return sa_cuda_fail(e, "pinned staging buffers");
    const uint64_t n = scan_column(FileView{m.p, m.size}, col, num_columns, h_text.as<uint8_t>(), h_map.as<uint32_t>(),
                                   h_q.as<uint8_t>());
    if (n > SA_MAX_INDEX_BYTES) { sa_set_error("column text of %llu bytes exceeds the 2^31 limit", (unsigned long long)n); return SA_ERR_LIMIT; }
    idx->n = n; idx->tail = 0;
    cudaStream_t s = idx->stream;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    SA_CU(idx->map.ensure((size_t)n * 4 + 4));
    SA_CU(idx->quoted_text.ensure((size_t)n + 1));
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + ((size_t)n & ~(size_t)15), 0, bytes - ((size_t)n & ~(size_t)15), s));
    if (n) {
        ScopedPhase ph(idx, &idx->stats.ms_h2d);
        SA_CU(cudaMemcpyAsync(idx->text.p, h_text.p, n, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->map.p, h_map.p, n * 4, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->quoted_text.p, h_q.p, n, cudaMemcpyHostToDevice, s));
    }
    SA_CU(cudaStreamSynchronize(s));     // the pinned staging buffers are released on return
    return SA_OK;
}

static int open_csv(const char* csv_path, Mapped& m, std::vector<std::string>& cells, bool allow_big = false)
{
    m.fd = open(csv_path, O_RDONLY);
    if (m.fd < 0) { sa_set_error("cannot open %s", csv_path); return SA_ERR_IO; }
    struct stat sb;
    if (fstat(m.fd, &sb) != 0) { sa_set_error("fstat failed on %s", csv_path); return SA_ERR_IO; }
    m.size = (size_t)sb.st_size;
    if (m.size == 0) { sa_set_error("%s is empty", csv_path); return SA_ERR_IO; }
    if (m.size >= (1ull << 31) && !allow_big) {
        sa_set_error("%s: files >= 2 GiB span several reference partitions (engine.c:1437): build them one by one with "
                     "sa_build_csv_partition", csv_path);
        return SA_ERR_LIMIT;
    }
    if (m.size >= (1ull << 32) - 4096) {
        sa_set_error("%s: file offsets do not fit the uint32 suffix array (the reference truncates them, engine.c:560)", csv_path);
        return SA_ERR_LIMIT;
    }
    void* mp = mmap(nullptr, m.size, PROT_READ, MAP_PRIVATE, m.fd, 0);
    if (mp == MAP_FAILED) { sa_set_error("mmap failed on %s", csv_path); return SA_ERR_IO; }
    m.p = static_cast<const uint8_t*>(mp);
    madvise(mp, m.size, MADV_SEQUENTIAL);
    cells = header_cells(m.p, m.size);
    return SA_OK;
}

static int find_column(const std::vector<std::string>& cells, const char* column, const char* csv_path, uint32_t* out)
{
    for (size_t i = 0; i < cells.size(); ++i)
        if (cells[i] == column) { *out = (uint32_t)i; return SA_OK; }     // strcmp match, engine.c:1424-1435
    sa_set_error("column '%s' not found in %s", column, csv_path);
    return SA_ERR_IO;
}

// Column `col` of the file loaded on src -> dst's text / map (/ quoted_text on the host route).
// *file_regular: in/out, false once the file has been found irregular (or the host scanner is forced).
static int scan_one_column(sa_index* src, sa_index* dst, const Mapped& m, uint32_t col, uint32_t num_columns,
                           const char* csv_path, bool* file_regular, bool* regular)
{
    *regular = false;
    if (*file_regular) {
        int rc = scan_column_device(src, dst, m.size, col, num_columns, regular);
        if (rc != SA_OK) return rc;
        if (!*regular && src->dbg_csv_scanner == 2) {
            sa_set_error("%s is not a regular CSV file for the device scanner (flags 0x%x)", csv_path, src->h_small[0]);
            return SA_ERR_IO;
        }
        *file_regular = *regular;                 // irregular quoting is a property of the file, not of the column
    }
    if (!*regular) return scan_column_host(dst, m, col, num_columns);
    return SA_OK;
}

// The two post-build loops of the reference (engine.c:632-651) on the complete ORDER-E array of idx.
static int csv_remap(sa_index* idx, bool regular)
{
    const uint64_t n = idx->n;
    cudaStream_t s = idx->stream;
    SA_CU(idx->sa_file.ensure((size_t)(n ? n : 1) * 4));
    SA_CU(idx->qbits.ensure((((size_t)n + 31) / 32) * 4 + 4));
    ScopedPhase ph(idx, &idx->stats.ms_remap);
    SA_CU(sa_launch_remap(idx->sa_final.as<uint32_t>(), (uint32_t)n, idx->map.as<uint32_t>(),
                          regular ? nullptr : idx->quoted_text.as<uint8_t>(), idx->sa_file.as<uint32_t>(),
                          idx->qbits.as<uint32_t>(), s));
    idx->stats.kernel_launches += 1;
    return SA_OK;
}

// Several columns of ONE file, one index per column (the reference's entry takes
// `char** search_cols, uint32_t num_search_cols` but only ever indexes search_cols[0], engine.c:1476,
// suffix_array.pyx:197 "TODO: Allow multiple columns").  The file is read, uploaded and analysed once
// (quote parity, delimiter counts, row table) on idxs[0]; every column then costs one emit pass, its
// build and its remap.
extern "C" int sa_build_csv_columns(sa_index* const* idxs, const char* const* columns, uint32_t ncols, const char* csv_path,
                                    uint32_t flags)
{
    if (!idxs || !columns || !csv_path || ncols == 0) { sa_set_error("sa_build_csv_columns: NULL argument"); return SA_ERR_INVALID; }
    if (ncols > 64) { sa_set_error("sa_build_csv_columns: at most 64 columns per call"); return SA_ERR_LIMIT; }
    std::vector<sa_index*> order(idxs, idxs + ncols);
    for (uint32_t c = 0; c < ncols; ++c) {
        if (!idxs[c] || !columns[c]) { sa_set_error("sa_build_csv_columns: NULL entry %u", c); return SA_ERR_INVALID; }
        if (idxs[c]->device != idxs[0]->device) { sa_set_error("sa_build_csv_columns: all indexes must live on one device"); return SA_ERR_INVALID; }
    }
    std::sort(order.begin(), order.end());
    if (std::adjacent_find(order.begin(), order.end()) != order.end()) {
        sa_set_error("sa_build_csv_columns: one index per column (a handle was passed twice)");
        return SA_ERR_INVALID;
    }
    std::vector<std::unique_lock<std::shared_mutex>> locks;      // address order: concurrent callers cannot deadlock
    for (sa_index* p : order) locks.emplace_back(p->mu);
    sa_index* src = idxs[0];
    SA_CU(cudaSetDevice(src->device));
    Mapped m;
    std::vector<std::string> cells;
    int rc = open_csv(csv_path, m, cells);
    if (rc != SA_OK) return rc;
    std::vector<uint32_t> col_of(ncols);
    for (uint32_t c = 0; c < ncols; ++c) {
        rc = find_column(cells, columns[c], csv_path, &col_of[c]);
        if (rc != SA_OK) return rc;
    }
    for (uint32_t c = 0; c < ncols; ++c) {
        sa_index* idx = idxs[c];
        idx->timing = (flags & SA_FLAG_TIMING) != 0;
        idx->built = false;
        sa_invalidate_results(idx);
        idx->shard_prepared = false;
        idx->shard_csv = false;
        idx->stats.ms_h2d = 0.f;
        idx->stats.ms_csv_scan = 0.f;
        idx->stats.csv_host_scan = 0;
        idx->stats.kernel_launches = 0;
    }
    rc = csv_load_device(src, m.fd, m.size, src->dbg_csv_scanner != 1);
    if (rc != SA_OK) return rc;
    bool file_regular = src->dbg_csv_scanner != 1;
    for (uint32_t c = 0; c < ncols; ++c) {
        sa_index* idx = idxs[c];
        bool regular = false;
        rc = scan_one_column(src, idx, m, col_of[c], (uint32_t)cells.size(), csv_path, &file_regular, &regular);
        if (rc != SA_OK) return rc;
        cudaStream_t s = idx->stream;
        const uint64_t scan_launches = idx->stats.kernel_launches;   // sa_build_resident clears the stats
        // the scan already lower-cased the column (tolower, engine.c:559,581,593)
        rc = sa_build_resident(idx, flags & ~SA_FLAG_LOWERCASE);
        if (rc != SA_OK) return rc;
        idx->stats.kernel_launches += scan_launches;
        idx->built = false;
        idx->stats.csv_host_scan = regular ? 0u : 1u;
        rc = csv_remap(idx, regular);
        if (rc != SA_OK) return rc;
        if (idx != src) {                          // every index answers hits with row ids: its own copy of the row table
            SA_CU(idx->rowbits.ensure(sa_csv_rowbits_bytes(m.size)));
            SA_CU(idx->rowrank.ensure(sa_csv_rowrank_bytes(m.size)));
            SA_CU(cudaMemcpyAsync(idx->rowbits.p, src->rowbits.p, sa_csv_rowbits_bytes(m.size), cudaMemcpyDeviceToDevice, s));
            SA_CU(cudaMemcpyAsync(idx->rowrank.p, src->rowrank.p, sa_csv_rowrank_bytes(m.size), cudaMemcpyDeviceToDevice, s));
            // ... and of the file (file-semantics probes, record materialisation)
            SA_CU(idx->csv_file.ensure(sa_csv_padded_bytes(m.size)));
            SA_CU(cudaMemcpyAsync(idx->csv_file.p, src->csv_file.p, sa_csv_padded_bytes(m.size), cudaMemcpyDeviceToDevice, s));
        }
        SA_CU(cudaStreamSynchronize(s));
        sa_time_collect(idx);
        if (idx != src) sa_time_collect(src);      // scan phases of this column were recorded on src's stream
        idx->csv = true;
        idx->csv_file_bytes = m.size;
        idx->built = true;
    }
    return SA_OK;
}

extern "C" int sa_build_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags)
{
    if (!idx || !csv_path || !column) { sa_set_error("sa_build_csv: NULL argument"); return SA_ERR_INVALID; }
    return sa_build_csv_columns(&idx, &column, 1, csv_path, flags);
}

// ------------------------------------------------------------------------------ CSV column over several GPUs
// The sharded build (sa_shard.cu) takes its text from a CSV column: every rank reads and scans the
// file itself (the scan is ~1 % of the build; nothing has to be exchanged), sa_shard_begin_csv leaves
// the rank where sa_shard_begin would, and once the complete ORDER-E array is installed
// (sa_shard_mark_built / sa_adopt_sa[_parts]) sa_csv_finalize_shard runs the remap on it.
extern "C" int sa_shard_begin_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags)
{
    if (!idx || !csv_path || !column) { sa_set_error("sa_shard_begin_csv: NULL argument"); return SA_ERR_INVALID; }
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    Mapped m;
    std::vector<std::string> cells;
    int rc = open_csv(csv_path, m, cells);
    if (rc != SA_OK) return rc;
    uint32_t col = 0;
    rc = find_column(cells, column, csv_path, &col);
    if (rc != SA_OK) return rc;
    idx->timing = (flags & SA_FLAG_TIMING) != 0;
    idx->built = false;
    sa_invalidate_results(idx);
    idx->csv = false;
    idx->shard_prepared = false;
    idx->shard_csv = false;
    memset(&idx->stats, 0, sizeof idx->stats);
    rc = csv_load_device(idx, m.fd, m.size, idx->dbg_csv_scanner != 1);
    if (rc != SA_OK) return rc;
    bool file_regular = idx->dbg_csv_scanner != 1, regular = false;
    rc = scan_one_column(idx, idx, m, col, (uint32_t)cells.size(), csv_path, &file_regular, &regular);
    if (rc != SA_OK) return rc;
    const uint64_t n = idx->n;
    if (n == 0) { sa_set_error("sa_shard_begin_csv: the column is empty"); return SA_ERR_INVALID; }
    idx->lprime = idx->L < n ? idx->L : (uint32_t)n;
    idx->stats.n = n; idx->stats.max_suffix_length = idx->L; idx->stats.lprime = idx->lprime;
    idx->stats.csv_host_scan = regular ? 0u : 1u;
    rc = sa_stage_prepare(idx, flags & ~SA_FLAG_LOWERCASE);
    if (rc != SA_OK) return rc;
    if (idx->lprime == 0) { sa_set_error("sa_shard_begin_csv: max_suffix_length 0 needs no sharded build"); return SA_ERR_INVALID; }
    idx->csv_file_bytes = m.size;
    idx->shard_csv = true;
    idx->shard_csv_regular = regular;
    idx->shard_prepared = true;
    return SA_OK;
}

// ------------------------------------------------------------------------------ partitions of a big file
// Files of 2 GiB and more: the reference cuts the FILE every 2^31 bytes and builds one suffix array per
// partition, each partition resuming where the previous one stopped (engine.c:1437-1438, :1469-1481; the
// scan of a partition starts with col_idx = 0 and outside quotes, wherever that is).  One call builds one
// partition into one handle: partition_start = 0 for the first, then the *partition_end of the previous
// call; *partition_end == file size (returned in *file_bytes) means it was the last.  The array holds
// ABSOLUTE file offsets (u32, as engine.c:560 stores them), so hits and records resolve against the one
// row table of the whole file, of which every partition's handle keeps a copy on the device.  The column
// scan runs on the host (the byte state machine of engine.c:522-605 on the mmap'd file): the
// tile-parallel device scanner assumes the scan starts at a record boundary, which a partition cut
// does not guarantee.  Also accepts files below 2 GiB (one partition).
extern "C" int sa_build_csv_partition(sa_index* idx, const char* csv_path, const char* column, uint64_t partition_start,
                                      uint32_t flags, uint64_t* partition_end, uint64_t* file_bytes)
{
    if (!idx || !csv_path || !column || !partition_end) { sa_set_error("sa_build_csv_partition: NULL argument"); return SA_ERR_INVALID; }
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    Mapped m;
    std::vector<std::string> cells;
    int rc = open_csv(csv_path, m, cells, true);
    if (rc != SA_OK) return rc;
    if (file_bytes) *file_bytes = m.size;
    if (partition_start >= m.size) { sa_set_error("sa_build_csv_partition: partition_start beyond the end of the file"); return SA_ERR_INVALID; }
    uint32_t col = 0;
    rc = find_column(cells, column, csv_path, &col);
    if (rc != SA_OK) return rc;
    idx->timing = (flags & SA_FLAG_TIMING) != 0;
    idx->built = false;
    sa_invalidate_results(idx);
    idx->shard_prepared = false;
    idx->shard_csv = false;
    memset(&idx->stats, 0, sizeof idx->stats);
    const uint64_t TWO_GB = 1ull << 31;
    const uint64_t length = std::min<uint64_t>(TWO_GB, m.size - partition_start);     // engine.c:487
    const size_t end_point = (size_t)(partition_start + length);                       // engine.c:513
    // size of the column text of this partition: a first pass that only counts, then the real one
    size_t stop = 0;
    const uint64_t n = scan_column(FileView{m.p, m.size}, col, (uint32_t)cells.size(), nullptr, nullptr, nullptr,
                                   (size_t)partition_start, end_point, &stop, 0);
    if (n > SA_MAX_INDEX_BYTES) { sa_set_error("column text of %llu bytes exceeds the 2^31 limit", (unsigned long long)n); return SA_ERR_LIMIT; }
    *partition_end = stop;
    std::vector<uint8_t> h_text((size_t)n + 1), h_q((size_t)n + 1);
    std::vector<uint32_t> h_map((size_t)n + 1);
    scan_column(FileView{m.p, m.size}, col, (uint32_t)cells.size(), h_text.data(), h_map.data(), h_q.data(), (size_t)partition_start,
                end_point, nullptr, n);
    // the whole file on the device: row table (hits -> rows), file-semantics probes, records
    rc = csv_load_device(idx, m.fd, m.size, false);
    if (rc != SA_OK) return rc;
    cudaStream_t s = idx->stream;
    idx->n = n; idx->tail = 0;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    SA_CU(idx->map.ensure((size_t)n * 4 + 4));
    SA_CU(idx->quoted_text.ensure((size_t)n + 1));
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + ((size_t)n & ~(size_t)15), 0, bytes - ((size_t)n & ~(size_t)15), s));
    if (n) {
        SA_CU(cudaMemcpyAsync(idx->text.p, h_text.data(), n, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->map.p, h_map.data(), n * 4, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->quoted_text.p, h_q.data(), n, cudaMemcpyHostToDevice, s));
    }
    SA_CU(cudaStreamSynchronize(s));
    const float h2d = idx->stats.ms_h2d;
    rc = sa_build_resident(idx, flags & ~SA_FLAG_LOWERCASE);
    if (rc != SA_OK) return rc;
    idx->stats.ms_h2d = h2d;
    idx->built = false;
    idx->stats.csv_host_scan = 1u;
    rc = csv_remap(idx, false);
    if (rc != SA_OK) return rc;
    SA_CU(cudaStreamSynchronize(s));
    sa_time_collect(idx);
    idx->csv = true;
    idx->csv_file_bytes = m.size;
    idx->built = true;
    return SA_OK;
}

// A CSV index restored by sa_load has the arrays and the row table but not the file bytes: put them
// back on the device (file-semantics queries and record materialisation read them).
extern "C" int sa_attach_csv_file(sa_index* idx, const char* csv_path)
{
    if (!idx || !csv_path) { sa_set_error("sa_attach_csv_file: NULL argument"); return SA_ERR_INVALID; }
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    if (!idx->built || !idx->csv) { sa_set_error("sa_attach_csv_file: no CSV index built"); return SA_ERR_INVALID; }
    SA_CU(cudaSetDevice(idx->device));
    Mapped m;
    std::vector<std::string> cells;
    int rc = open_csv(csv_path, m, cells);
    if (rc != SA_OK) return rc;
    if (m.size != idx->csv_file_bytes) {
        sa_set_error("%s has %zu bytes, the indexed file had %llu", csv_path, m.size, (unsigned long long)idx->csv_file_bytes);
        return SA_ERR_IO;
    }
    const uint64_t padded = sa_csv_padded_bytes(m.size);
    SA_CU(idx->csv_file.ensure(padded));
    uint8_t* d_file = idx->csv_file.as<uint8_t>();
    const size_t tail = m.size & ~(size_t)15;
    SA_CU(cudaMemsetAsync(d_file + tail, 0, padded - tail, idx->stream));
    rc = stream_file_to_device(idx, m.fd, m.size, d_file);
    if (rc != SA_OK) return rc;
    SA_CU(cudaStreamSynchronize(idx->stream));
    return SA_OK;
}

int sa_csv_finalize_shard(sa_index* idx)
{
    int rc = csv_remap(idx, idx->shard_csv_regular);
    if (rc != SA_OK) return rc;
    SA_CU(cudaStreamSynchronize(idx->stream));
    sa_time_collect(idx);
    idx->csv = true;
    return SA_OK;
}
