// sa_csv.cu -- CSV column feeder + device remap behind sa_build_csv.
//
// Replaces construct_truncated_suffix_array_from_csv_partitioned_mmap[_full]
// (engine.c:1454-1482 -> :461-654).  The column scan stays on the host (a branchy byte state
// machine, engine.c:522-605) but writes straight into PINNED buffers that are copied to the
// device asynchronously; the build runs on the device; the two post-build loops of the
// reference (quoted-bit permutation, serial, engine.c:632-645; SA[i] = map[SA[i]],
// engine.c:647-651) are one gather kernel (remap_kernel, query.cu).
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <string>
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
uint64_t scan_column(FileView v, uint32_t column_idx, uint32_t num_columns, uint8_t* text, uint32_t* map,
                     uint8_t* quoted)
{
    const size_t end = v.size;
    size_t pos = 0;
    uint64_t out = 0;
    uint32_t col = 0;
    auto emit = [&](uint8_t c, size_t where, uint8_t q) {
        text[out] = c; map[out] = (uint32_t)where; quoted[out] = q; ++out;
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
                if (pos >= end) break;                // unterminated quote (the reference runs off the mapping)
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

extern "C" int sa_build_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags)
{
    if (!idx || !csv_path || !column) { sa_set_error("sa_build_csv: NULL argument"); return SA_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    Mapped m;
    m.fd = open(csv_path, O_RDONLY);
    if (m.fd < 0) { sa_set_error("cannot open %s", csv_path); return SA_ERR_IO; }
    struct stat sb;
    if (fstat(m.fd, &sb) != 0) { sa_set_error("fstat failed on %s", csv_path); return SA_ERR_IO; }
    m.size = (size_t)sb.st_size;
    if (m.size == 0) { sa_set_error("%s is empty", csv_path); return SA_ERR_IO; }
    if (m.size >= (1ull << 31)) {
        sa_set_error("%s: files >= 2 GiB span several reference partitions (engine.c:1437); not supported yet", csv_path);
        return SA_ERR_LIMIT;
    }
    void* mp = mmap(nullptr, m.size, PROT_READ, MAP_PRIVATE, m.fd, 0);
    if (mp == MAP_FAILED) { sa_set_error("mmap failed on %s", csv_path); return SA_ERR_IO; }
    m.p = static_cast<const uint8_t*>(mp);
    madvise(mp, m.size, MADV_SEQUENTIAL);

    const std::vector<std::string> cells = header_cells(m.p, m.size);
    int col = -1;
    for (size_t i = 0; i < cells.size(); ++i)
        if (cells[i] == column) { col = (int)i; break; }     // strcmp match, engine.c:1424-1435
    if (col < 0) { sa_set_error("column '%s' not found in %s", column, csv_path); return SA_ERR_IO; }

    const size_t cap = m.size + 2;
    Pinned h_text, h_map, h_q;
    cudaError_t e = h_text.alloc(cap);
    if (e == cudaSuccess) e = h_map.alloc(cap * 4);
    if (e == cudaSuccess) e = h_q.alloc(cap);
    if (e != cudaSuccess) return sa_cuda_fail(e, "pinned staging buffers");

    const uint64_t n = scan_column(FileView{m.p, m.size}, (uint32_t)col, (uint32_t)cells.size(), h_text.as<uint8_t>(),
                                   h_map.as<uint32_t>(), h_q.as<uint8_t>());
    if (n > SA_MAX_INDEX_BYTES) { sa_set_error("column text of %llu bytes exceeds the 2^31 limit", (unsigned long long)n); return SA_ERR_LIMIT; }

    idx->timing = (flags & SA_FLAG_TIMING) != 0;
    idx->n = n;
    cudaStream_t s = idx->stream;
    const size_t bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    SA_CU(idx->map.ensure((size_t)n * 4));
    SA_CU(idx->quoted_text.ensure((size_t)n));
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + ((size_t)n & ~(size_t)15), 0, bytes - ((size_t)n & ~(size_t)15), s));
    idx->stats.ms_h2d = 0.f;
    if (n) {
        ScopedPhase ph(idx, &idx->stats.ms_h2d);
        SA_CU(cudaMemcpyAsync(idx->text.p, h_text.p, n, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->map.p, h_map.p, n * 4, cudaMemcpyHostToDevice, s));
        SA_CU(cudaMemcpyAsync(idx->quoted_text.p, h_q.p, n, cudaMemcpyHostToDevice, s));
    }
    // the scan already lower-cased the column (tolower, engine.c:559,581,593)
    int rc = sa_build_resident(idx, flags & ~SA_FLAG_LOWERCASE);
    if (rc != SA_OK) return rc;
    idx->built = false;
    SA_CU(idx->sa_file.ensure((size_t)n * 4));
    SA_CU(idx->qbits.ensure((((size_t)n + 31) / 32) * 4 + 4));
    {
        ScopedPhase ph(idx, &idx->stats.ms_remap);
        SA_CU(sa_launch_remap(idx->sa_final.as<uint32_t>(), (uint32_t)n, idx->map.as<uint32_t>(),
                              idx->quoted_text.as<uint8_t>(), idx->sa_file.as<uint32_t>(), idx->qbits.as<uint32_t>(), s));
        idx->stats.kernel_launches += 1;
    }
    SA_CU(cudaStreamSynchronize(s));     // pinned staging buffers are released on return
    sa_time_collect(idx);
    idx->csv = true;
    idx->built = true;
    return SA_OK;
}
