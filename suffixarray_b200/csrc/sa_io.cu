// sa_io.cu -- save / load of a built index.
//
// The reference only sketches persistence (write_suffix_array / read_into_suffix_array_file,
// engine.c:1098-1165, unreachable from Python; TODO.md:1,4).  Two things are provided here:
//   * sa_save / sa_load: one self-describing container with everything a query needs (text,
//     ORDER-E positions, and for CSV indexes the file offsets, quoted bits and the file's row table), so a loaded index
//     answers queries without re-sorting; the record table is rebuilt on the device at load time.
//   * sa_export_reference_files: the reference's own layout, for tools written against it:
//       <sa file>      u64 global_byte_start_idx, u64 global_byte_end_idx, u32 max_suffix_length,
//                      u32 n, u32 suffix_array[n]                              (engine.c:1121-1126)
//       <quoted file>  u32 capacity_bytes, u8 bits[capacity_bytes], LSB first   (engine.c:1098-1101)
#include <stdio.h>
#include <string.h>

#include <vector>

#include <sys/stat.h>

#include "sa_index.h"

namespace {

constexpr size_t IO_CHUNK = 64u << 20;

struct FileHeader {
    char magic[8];
    uint32_t version;
    uint32_t max_suffix_length;
    uint64_t n;
    uint32_t flags;          // bit 0: CSV index (file offsets + quoted bits follow)
    uint32_t reserved;
    uint64_t csv_file_bytes;
};
const char MAGIC[8] = {'S', 'A', 'C', 'U', 'D', 'A', '0', '1'};

struct File {
    FILE* f = nullptr;
    ~File() { if (f) fclose(f); }
};

int write_device(sa_index* idx, FILE* f, const void* dev, size_t bytes, std::vector<uint8_t>& buf)
{
    const uint8_t* p = static_cast<const uint8_t*>(dev);
    for (size_t off = 0; off < bytes; off += IO_CHUNK) {
        const size_t len = bytes - off < IO_CHUNK ? bytes - off : IO_CHUNK;
        SA_CU(cudaMemcpyAsync(buf.data(), p + off, len, cudaMemcpyDeviceToHost, idx->stream));
        SA_CU(cudaStreamSynchronize(idx->stream));
        if (fwrite(buf.data(), 1, len, f) != len) { sa_set_error("write failed"); return SA_ERR_IO; }
    }
    return SA_OK;
}

int read_device(sa_index* idx, FILE* f, void* dev, size_t bytes, std::vector<uint8_t>& buf)
{
    uint8_t* p = static_cast<uint8_t*>(dev);
    for (size_t off = 0; off < bytes; off += IO_CHUNK) {
        const size_t len = bytes - off < IO_CHUNK ? bytes - off : IO_CHUNK;
        if (fread(buf.data(), 1, len, f) != len) { sa_set_error("file is truncated"); return SA_ERR_IO; }
        SA_CU(cudaMemcpyAsync(p + off, buf.data(), len, cudaMemcpyHostToDevice, idx->stream));
        SA_CU(cudaStreamSynchronize(idx->stream));
    }
    return SA_OK;
}

}  // namespace

extern "C" int sa_save(sa_index* idx, const char* path)
{
    if (!idx || !path) { sa_set_error("sa_save: NULL argument"); return SA_ERR_INVALID; }
    if (!idx->built) { sa_set_error("sa_save: index not built"); return SA_ERR_INVALID; }
    if (idx->sharded_sa) { sa_set_error("sa_save: the array of a SA_FLAG_SHARDED_SA index is spread over several GPUs"); return SA_ERR_INVALID; }
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    File out;
    out.f = fopen(path, "wb");
    if (!out.f) { sa_set_error("cannot create %s", path); return SA_ERR_IO; }
    FileHeader h{};
    memcpy(h.magic, MAGIC, 8);
    h.version = 2;
    h.max_suffix_length = idx->L;
    h.n = idx->n;
    h.flags = idx->csv ? 1u : 0u;
    h.csv_file_bytes = idx->csv_file_bytes;
    if (fwrite(&h, sizeof h, 1, out.f) != 1) { sa_set_error("write failed"); return SA_ERR_IO; }
    std::vector<uint8_t> buf(IO_CHUNK);
    const size_t n = (size_t)idx->n;
    int rc = write_device(idx, out.f, idx->text.p, n, buf);
    if (rc == SA_OK) rc = write_device(idx, out.f, idx->sa_final.p, n * 4, buf);
    if (rc == SA_OK && idx->csv) rc = write_device(idx, out.f, idx->sa_file.p, n * 4, buf);
    if (rc == SA_OK && idx->csv) rc = write_device(idx, out.f, idx->qbits.p, ((n + 31) / 32) * 4, buf);
    // the row table of the CSV file (hits -> row index): restored as is, the file itself is not needed to load
    if (rc == SA_OK && idx->csv) rc = write_device(idx, out.f, idx->rowbits.p, sa_csv_rowbits_bytes(idx->csv_file_bytes), buf);
    if (rc == SA_OK && idx->csv) rc = write_device(idx, out.f, idx->rowrank.p, sa_csv_rowrank_bytes(idx->csv_file_bytes), buf);
    return rc;
}

extern "C" int sa_load(sa_index* idx, const char* path)
{
    if (!idx || !path) { sa_set_error("sa_load: NULL argument"); return SA_ERR_INVALID; }
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    File in;
    in.f = fopen(path, "rb");
    if (!in.f) { sa_set_error("cannot open %s", path); return SA_ERR_IO; }
    FileHeader h{};
    if (fread(&h, sizeof h, 1, in.f) != 1 || memcmp(h.magic, MAGIC, 8) != 0 || h.version != 2) {
        sa_set_error("%s is not a libsa_cuda index file", path);
        return SA_ERR_IO;
    }
    if (h.max_suffix_length != idx->L) {
        sa_set_error("%s was built with max_suffix_length %u, this handle has %u", path, h.max_suffix_length, idx->L);
        return SA_ERR_INVALID;
    }
    // sa_save writes any built index, including sharded ones of up to 2^32 - 1025 entries
    if (h.n > 0xFFFFFFFFull - 1024ull) { sa_set_error("%s: n exceeds the uint32 position range", path); return SA_ERR_LIMIT; }
    if ((h.flags & 1u) && (h.csv_file_bytes == 0 || h.csv_file_bytes >= (1ull << 32))) {
        sa_set_error("%s: implausible CSV file size %llu in the header", path, (unsigned long long)h.csv_file_bytes);
        return SA_ERR_IO;
    }
    {   // the file must hold what the header promises
        struct stat sb;
        uint64_t need = sizeof h + (uint64_t)h.n * 5;
        if (h.flags & 1u)
            need += (uint64_t)h.n * 4 + ((h.n + 31) / 32) * 4 + sa_csv_rowbits_bytes(h.csv_file_bytes) +
                    sa_csv_rowrank_bytes(h.csv_file_bytes);
        if (fstat(fileno(in.f), &sb) != 0 || (uint64_t)sb.st_size < need) {
            sa_set_error("%s is truncated (%llu bytes needed)", path, (unsigned long long)need);
            return SA_ERR_IO;
        }
    }
    idx->built = false;
    sa_invalidate_results(idx);
    idx->shard_prepared = false;
    idx->shard_csv = false;
    idx->csv_file.release();            // a restored CSV index holds no copy of the file (sa_attach_csv_file)
    idx->n = h.n; idx->tail = 0;
    const size_t n = (size_t)h.n;
    memset(&idx->stats, 0, sizeof idx->stats);
    idx->lprime = idx->L < (uint32_t)n ? idx->L : (uint32_t)n;
    idx->stats.n = n; idx->stats.max_suffix_length = idx->L; idx->stats.lprime = idx->lprime;
    idx->timing = false;
    cudaStream_t s = idx->stream;
    const size_t bytes = (n + SA_TEXT_PAD + 255) & ~(size_t)255;
    SA_CU(idx->text.ensure(bytes));
    SA_CU(cudaMemsetAsync(idx->text.as<uint8_t>() + (n & ~(size_t)15), 0, bytes - (n & ~(size_t)15), s));
    SA_CU(idx->sa_final.ensure((n ? n : 1) * 4));
    std::vector<uint8_t> buf(IO_CHUNK);
    int rc = read_device(idx, in.f, idx->text.p, n, buf);
    if (rc == SA_OK) rc = read_device(idx, in.f, idx->sa_final.p, n * 4, buf);
    if (rc != SA_OK) return rc;
    idx->csv = false;
    if (h.flags & 1u) {
        SA_CU(idx->sa_file.ensure((n ? n : 1) * 4));
        SA_CU(idx->qbits.ensure(((n + 31) / 32) * 4 + 4));
        rc = read_device(idx, in.f, idx->sa_file.p, n * 4, buf);
        if (rc == SA_OK) rc = read_device(idx, in.f, idx->qbits.p, ((n + 31) / 32) * 4, buf);
        if (rc != SA_OK) return rc;
        idx->csv_file_bytes = h.csv_file_bytes;
        SA_CU(idx->rowbits.ensure(sa_csv_rowbits_bytes(h.csv_file_bytes)));
        SA_CU(idx->rowrank.ensure(sa_csv_rowrank_bytes(h.csv_file_bytes)));
        rc = read_device(idx, in.f, idx->rowbits.p, sa_csv_rowbits_bytes(h.csv_file_bytes), buf);
        if (rc == SA_OK) rc = read_device(idx, in.f, idx->rowrank.p, sa_csv_rowrank_bytes(h.csv_file_bytes), buf);
        if (rc != SA_OK) return rc;
    }
    {   // positions / offsets out of range would become out-of-bounds device reads in the query kernels
        unsigned int* bad = reinterpret_cast<unsigned int*>(idx->small.as<uint32_t>() + SM_ERR);
        SA_CU(cudaMemsetAsync(bad, 0, sizeof(unsigned int), s));
        SA_CU(sa_launch_count_out_of_range(idx->sa_final.as<uint32_t>(), n, n, bad, s));
        if (h.flags & 1u) SA_CU(sa_launch_count_out_of_range(idx->sa_file.as<uint32_t>(), n, h.csv_file_bytes, bad, s));
        SA_CU(cudaMemcpyAsync(idx->h_small, bad, sizeof(unsigned int), cudaMemcpyDeviceToHost, s));
        SA_CU(cudaStreamSynchronize(s));
        if (idx->h_small[0]) {
            sa_set_error("%s is corrupt: %u array entries out of range", path, idx->h_small[0]);
            return SA_ERR_IO;
        }
    }
    if (n) {
        rc = sa_stage_prepare(idx, 0);          // record table (+ alphabet codes, unused by queries)
        if (rc != SA_OK) return rc;
        SA_CU(cudaStreamSynchronize(s));
    } else {
        idx->num_nl = 0;
        idx->stats.num_docs = 1;
    }
    idx->csv = (h.flags & 1u) != 0;
    idx->built = true;
    return SA_OK;
}

extern "C" int sa_export_reference_files(sa_index* idx, const char* sa_path, const char* quoted_path)
{
    if (!idx || !sa_path) { sa_set_error("sa_export_reference_files: NULL argument"); return SA_ERR_INVALID; }
    if (!idx->built) { sa_set_error("sa_export_reference_files: index not built"); return SA_ERR_INVALID; }
    std::unique_lock<std::shared_mutex> lock(idx->mu);
    SA_CU(cudaSetDevice(idx->device));
    const size_t n = (size_t)idx->n;
    if (n >= (1ull << 32)) { sa_set_error("n does not fit the reference's u32 count"); return SA_ERR_LIMIT; }
    std::vector<uint8_t> buf(IO_CHUNK);
    {
        File out;
        out.f = fopen(sa_path, "wb");
        if (!out.f) { sa_set_error("cannot create %s", sa_path); return SA_ERR_IO; }
        const uint64_t start = 0, end = idx->csv ? idx->csv_file_bytes : (uint64_t)n;
        const uint32_t L = idx->L, n32 = (uint32_t)n;
        if (fwrite(&start, 8, 1, out.f) != 1 || fwrite(&end, 8, 1, out.f) != 1 || fwrite(&L, 4, 1, out.f) != 1 ||
            fwrite(&n32, 4, 1, out.f) != 1) { sa_set_error("write failed"); return SA_ERR_IO; }
        int rc = write_device(idx, out.f, idx->csv ? idx->sa_file.p : idx->sa_final.p, n * 4, buf);
        if (rc != SA_OK) return rc;
    }
    if (quoted_path) {
        File out;
        out.f = fopen(quoted_path, "wb");
        if (!out.f) { sa_set_error("cannot create %s", quoted_path); return SA_ERR_IO; }
        const uint32_t cap = (uint32_t)((n + 7) / 8);
        if (fwrite(&cap, 4, 1, out.f) != 1) { sa_set_error("write failed"); return SA_ERR_IO; }
        if (idx->csv) {
            int rc = write_device(idx, out.f, idx->qbits.p, cap, buf);
            if (rc != SA_OK) return rc;
        } else {                                       // documents: no cell is quoted
            std::vector<uint8_t> zeros(cap, 0);
            if (cap && fwrite(zeros.data(), 1, cap, out.f) != cap) { sa_set_error("write failed"); return SA_ERR_IO; }
        }
    }
    return SA_OK;
}
