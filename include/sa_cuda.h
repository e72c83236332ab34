/*
 * sa_cuda.h -- C ABI of libsa_cuda.so: truncated suffix-array build + batched substring query
 * on one B200 (sm_100a).  Plain pointers and sizes only; no torch / C++ types.
 *
 * Each entry point names the reference interface it replaces (paths under
 * /root/reference/suffix_array/).  The reference binds these through
 * `cdef extern from "engine.h"` in suffix_array.pyx:37-101; INTEGRATION.md shows the
 * replacement `cdef extern from "sa_cuda.h"` block.
 *
 * Conventions: every function returns SA_OK (0) or a negative sa_status; the message of the
 * last failure on the calling thread is sa_last_error().  The library never prints and never
 * calls exit() (the reference does both, e.g. engine.c:125,477,865).  There is no CPU
 * fallback: without a CUDA device every call fails with SA_ERR_CUDA.
 * Pointers named d_* are DEVICE pointers; all others are HOST pointers.
 */
#ifndef SA_CUDA_H
#define SA_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SA_CUDA_ABI_VERSION 2

typedef enum sa_status {
    SA_OK = 0,
    SA_ERR_INVALID = -1,   /* bad argument / call order */
    SA_ERR_OOM = -2,       /* host or device allocation failed */
    SA_ERR_CUDA = -3,      /* CUDA runtime / kernel failure */
    SA_ERR_IO = -4,        /* file could not be opened / mapped, column not found */
    SA_ERR_LIMIT = -5,     /* n >= 2^32 - 1024 or max_suffix_length > 4096 */
    SA_ERR_INTERNAL = -6   /* an internal invariant failed (device-side guard tripped) */
} sa_status;

/* build flags */
#define SA_FLAG_LOWERCASE 1u   /* ASCII A-Z -> a-z on the device copy of the text: replaces
                                  lowercase_string, suffix_array.pyx:103-107 */
#define SA_FLAG_TIMING    2u   /* record per-phase CUDA-event times into sa_stats */
#define SA_FLAG_DEVICE_TEXT 4u /* sa_shard_begin / sa_shard_build: `text` is a device pointer (copied device to device) */
#define SA_FLAG_SHARDED_SA 8u  /* sa_shard_build / sa_build_text_multi: do NOT replicate the finished array: every rank keeps
                                  its key range ("bucket-ranged SA shards") and queries read the other ranges from the
                                  peers' memory over NVLink (CUDA IPC / peer access).  Saves (N-1)/N * 4n bytes of
                                  NVLink traffic per GPU per build; the peers' indexes must stay alive while this one is
                                  queried, and sa_save / sa_device_sa are not available on such an index. */

/* Opaque index: owns the device text, the SA, the doc-offset table and all scratch.
 * Replaces SuffixArray_struct (engine.h:123-130) / SuffixArrayIndex (engine.h:163-172). */
typedef struct sa_index sa_index;

typedef struct sa_stats {
    uint64_t n;                 /* text bytes == SA entries */
    uint32_t max_suffix_length; /* L as given */
    uint32_t lprime;            /* min(L, n), engine.c:841 */
    uint32_t rounds;            /* refinement rounds after round 0 */
    uint32_t radix_passes;      /* onesweep scatter launches of the last build */
    uint64_t kernel_launches;   /* kernels launched by the last build / query call */
    uint64_t refined_elems;     /* sum over rounds of elements re-sorted */
    uint64_t finish_runs;       /* <=32 buckets re-sorted by the 17-byte window key */
    uint64_t num_docs;          /* '\n' count + 1 */
    float ms_total;             /* device time of the last build (SA_FLAG_TIMING) */
    float ms_h2d, ms_pack, ms_hist, ms_sort0, ms_refine, ms_finish, ms_docs, ms_remap;
    float ms_onesweep_sum;      /* sum of all onesweep scatter launches */
    float ms_query, ms_hits;    /* last query batch (device time) */
    uint64_t onesweep_elems;    /* total elements moved by those launches */
    uint32_t code_bits;         /* bits per suffix byte in the sort keys (8 = raw bytes) */
    uint32_t csv_host_scan;     /* sa_build_csv: 1 when the file took the host column scan (irregular quoting / NUL bytes) */
    float ms_csv_scan;          /* sa_build_csv: device column scan (three tile passes + scans) */
    uint32_t reserved1;
} sa_stats;

const char* sa_last_error(void);
int sa_abi_version(void);

/* Lifecycle.  Replaces init_suffix_array / init_suffix_array_byte_idxs (engine.c:309-344) and
 * free_suffix_array (engine.c:346).  device: CUDA ordinal. */
int sa_create(sa_index** out, uint32_t max_suffix_length, int device);
void sa_destroy(sa_index* idx);

/* Build from a host text of n bytes ('\n'-joined documents, suffix_array.pyx:143).
 * Replaces construct_truncated_suffix_array(const char*, SuffixArray_struct*), engine.h:217 /
 * engine.c:837-866, producing the identical uint32 order.  H2D copy included. */
int sa_build_text(sa_index* idx, const uint8_t* text, uint64_t n, uint32_t flags);

/* One partition of a longer text.  The reference passes `text + idx * 2 GiB` of ONE contiguous buffer
 * (suffix_array.pyx:176-180), so suffixes near the end of a partition are sorted and compared on the
 * first bytes of the next partition.  text holds n + tail bytes; the n first are indexed, up to 960
 * tail bytes are kept as look-ahead (enough for max_suffix_length <= 900; a longer tail is cut). */
int sa_build_text_tail(sa_index* idx, const uint8_t* text, uint64_t n, uint64_t tail, uint32_t flags);

/* Same, text already resident in HBM (copied device-to-device into the index). */
int sa_build_text_device(sa_index* idx, const uint8_t* d_text, uint64_t n, uint32_t flags);

/* CSV path.  Replaces construct_truncated_suffix_array_from_csv_partitioned_mmap_full
 * (engine.h:209-215, engine.c:1454-1482 -> :461-654): mmap the file, scan `column` into
 * pinned text / file-offset / quoted buffers (engine.c:522-605), build, then gather
 * SA[i] = map[SA[i]] and the per-rank quoted bits on the device (engine.c:632-651).
 * Single partition (file < 2 GiB), as the reference handles one partition per 2 GiB. */
int sa_build_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags);

/* Files of 2 GiB and more (also accepts smaller ones): the reference cuts the FILE every 2^31 bytes into
 * partitions with one suffix array each, every partition resuming where the previous one stopped
 * (engine.c:1437-1438, :1469-1481).  One call = one partition into one handle: partition_start = 0, then
 * the *partition_end of the previous call, until *partition_end == *file_bytes.  The array holds absolute
 * u32 file offsets (engine.c:560), so the file must stay below 4 GiB.  Host column scan (the reference's
 * state machine; a partition need not start at a record boundary). */
int sa_build_csv_partition(sa_index* idx, const char* csv_path, const char* column, uint64_t partition_start,
                           uint32_t flags, uint64_t* partition_end, uint64_t* file_bytes);

/* Several columns of one file, one index per column (idxs[c] indexes columns[c]; distinct handles on
 * one device).  The reference's entry already takes `char** search_cols, uint32_t num_search_cols`
 * (engine.h:209-215) but indexes search_cols[0] only (engine.c:1476; suffix_array.pyx:197 "TODO:
 * Allow multiple columns").  The file is read, uploaded and analysed once (quote parity, delimiter
 * counts, row table); each column then costs its emit pass, its build and its remap.
 * sa_build_csv(idx, path, column, flags) == sa_build_csv_columns(&idx, &column, 1, path, flags).
 * Every column name is resolved before anything is touched; if a later stage fails, the handles of
 * the columns already finished stay built and the others are left unbuilt (sa_size() == 0). */
int sa_build_csv_columns(sa_index* const* idxs, const char* const* columns, uint32_t ncols, const char* csv_path,
                         uint32_t flags);

/* Accessors.  sa_copy_sa_u32 returns what SuffixArray_struct.suffix_array holds after the
 * build (text positions; FILE OFFSETS after sa_build_csv, engine.c:647-651). */
uint64_t sa_size(const sa_index* idx);
int sa_copy_sa_u32(sa_index* idx, uint32_t* out, uint64_t cap);
int sa_copy_sa_i64(sa_index* idx, int64_t* out, uint64_t cap);
int sa_copy_text_positions_u32(sa_index* idx, uint32_t* out, uint64_t cap); /* pre-remap SA (CSV) */
int sa_copy_quoted_bits(sa_index* idx, uint8_t* out_bits, uint64_t cap_bytes); /* LSB-first, engine.c:264-286 */
int sa_copy_text(sa_index* idx, uint8_t* out, uint64_t cap);   /* the (lower-cased) text that was indexed */
const uint32_t* sa_device_sa(const sa_index* idx);              /* device pointer, n entries */
int sa_get_stats(const sa_index* idx, sa_stats* out);

/* Batched query.  Query i is qbytes[qoff[i] .. qoff[i+1]) (cut at its first NUL, as strlen
 * does, engine.c:874).  Replaces get_substring_positions (engine.h:229-233, engine.c:869-918):
 * start[i], end[i] are the inclusive rank range with the reference's UINT32_MAX sentinels.
 * hit_off[nq+1] is the exclusive prefix sum of min(k, end-start+1) (engine.c:1181-1185);
 * hit_off[nq] = total hits.  Any output pointer may be NULL.  Results stay on the device for
 * sa_query_hits. */
int sa_query_ranges(sa_index* idx, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq,
                    uint32_t k, uint32_t* start, uint32_t* end, uint64_t* hit_off);

/* Hits of the last sa_query_ranges call, in rank order, duplicates kept (engine.c:1187).
 * Text index: hit_pos = SA[rank] (text position), hit_doc = record index = number of '\n' before it
 * (the record engine.c:1189-1191 scans back to).  CSV index: hit_pos = SA[rank] = offset in the FILE
 * (engine.c:647-651), hit_doc = row of the file that holds that offset, counting record ends ('\n'
 * outside quotes) in front of it -- header line = row 0 -- i.e. the record the reference seeks from
 * the offset (engine.c:1217-1323); cells with raw newlines inside quotes and ragged rows keep their
 * file row.  Replaces the hit loop of get_matching_records / get_matching_records_file
 * (engine.c:1185-1211, :1356-1388). */
int sa_query_hits(sa_index* idx, uint32_t* hit_pos, uint32_t* hit_doc, uint64_t cap);

/* ---- Per-call query results (ABI 2).  An sa_result owns the device scratch of ONE query batch
 * (queries, ranges, hit offsets, hits, records), a stream and two event pairs.  Any number of
 * threads may query one built index at the same time, each through its own sa_result (the index
 * is only read; the reference's query functions are re-entrant in the same way, engine.c:869-999).
 * sa_query_ranges / sa_query_hits above are the same calls on a result object the index keeps
 * for callers that do not care (one batch at a time per index). */
typedef struct sa_result sa_result;
int sa_result_create(sa_index* idx, sa_result** out);
void sa_result_destroy(sa_result* res);

/* query flags */
#define SA_QUERY_FILE_SEMANTICS 1u /* CSV index only: the probes of get_substring_positions_file
                                      (engine.c:920-999) instead of get_substring_positions: each probe
                                      compares the LOWER-CASED RAW FILE BYTES at offset SA[mid] (a probe that
                                      runs past the end of a cell sees `,next cell` / `""`, not '\n' + the next
                                      row's cell), `start` / `end` are only set when cmp == 0 (:949-958,
                                      :983-991), the second bisection starts at first = start (:969), and a
                                      read that crosses the end of the file keeps the bytes of the previous
                                      probe in the tail of the compare buffer (fread into the same malloc'd
                                      line, :943-944).  This is the only query the reference's Python class
                                      reaches (suffix_array.pyx:221-247 -> get_matching_records_file). */
#define SA_QUERY_DEDUPE 2u         /* sa_query_hits_r / sa_query_records_size: per query, keep the first hit (in
                                      rank order) of every record id -- the "segmented dedupe" of the
                                      north star.  The reference keeps duplicates (engine.c:1187), so this is
                                      off unless asked for. */
#define SA_RECORDS_REFERENCE_COMPAT 4u /* sa_query_records_size: cut CSV records the way
                                      get_matching_records_file does (engine.c:1217-1323, :1369-1385): the last
                                      byte in front of the record's '\n' is dropped (:1314 returns
                                      `num_chars_ahead + char_idx - 1`), and for hits at file offsets >= 4096 so
                                      is the record's first byte (:1262 starts the backward scan at
                                      line[4095] = file[idx - 1] but :1277 subtracts one more).  Default: the
                                      whole line without its '\n'. */

int sa_query_ranges_r(sa_index* idx, sa_result* res, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq,
                      uint32_t k, uint32_t flags, uint32_t* start, uint32_t* end, uint64_t* hit_off);
/* Hits of the ranges held by res.  flags: SA_QUERY_DEDUPE or 0.  With dedupe the lists are compacted:
 * hit_off_out[nq + 1] (required then, optional otherwise) receives the offsets of the lists actually
 * written; cap must hold the un-deduplicated total (hit_off[nq] of sa_query_ranges_r). */
int sa_query_hits_r(sa_index* idx, sa_result* res, uint32_t flags, uint32_t* hit_pos, uint32_t* hit_doc, uint64_t cap,
                    uint64_t* hit_off_out);
/* Record materialisation in the library (replaces the record copies of get_matching_records[_file],
 * engine.c:1207-1210, :1382-1387, and the per-hit Python loop above them): every hit's record -- the
 * document of a text index, the line of the FILE of a CSV index -- gathered on the device into one
 * packed buffer.  sa_query_records_size expands the hits (flags: SA_QUERY_DEDUPE,
 * SA_RECORDS_REFERENCE_COMPAT) and returns the number of records and of bytes;
 * sa_query_records copies them out: record j is out[rec_off[j] .. rec_off[j + 1]) (no terminator),
 * rec_off has num_records + 1 entries, hit_off_out[nq + 1] (optional) the per-query record offsets
 * and hit_doc (optional) the record ids. */
int sa_query_records_size(sa_index* idx, sa_result* res, uint32_t flags, uint64_t* num_records, uint64_t* num_bytes);
int sa_query_records(sa_index* idx, sa_result* res, uint8_t* out, uint64_t cap_bytes, uint64_t* rec_off,
                     uint64_t cap_records, uint64_t* hit_off_out, uint32_t* hit_doc);
/* device time of the last ranges / hits+records calls on res (CUDA events), kernels launched */
int sa_result_stats(const sa_result* res, float* ms_query, float* ms_hits, uint64_t* kernel_launches);

/* CSV index: the row table of the file.  sa_num_rows = record ends ('\n' outside quotes) + 1;
 * row_start[r] = offset of row r's first byte (row 0 = the header line), row_start[num_rows] = file size.
 * Replaces the host-side newline scans of engine.c:1217-1323 for callers that cut rows themselves. */
uint64_t sa_num_rows(const sa_index* idx);
int sa_copy_row_starts(sa_index* idx, uint64_t* out, uint64_t cap);
/* A CSV index restored by sa_load holds no copy of the file: attach it again (read + upload) before
 * SA_QUERY_FILE_SEMANTICS queries or sa_query_records.  The file must be the one that was indexed. */
int sa_attach_csv_file(sa_index* idx, const char* csv_path);

/* Device-resident variants for callers that keep batches in HBM (bench `value` path). */
int sa_query_ranges_device(sa_index* idx, const uint8_t* d_qbytes, const uint64_t* d_qoff, uint64_t nq,
                           uint64_t qbytes_len, uint32_t k, uint32_t* d_start, uint32_t* d_end);

/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The hit-offset scan of a batch whose ranges are already on the device (sa_query_ranges_device):
 * d_hit_off[q] = sum over q' < q of min(k, end[q'] - start[q'] + 1) (0 for a miss, start == UINT32_MAX;
 * engine.c:1181-1185), d_hit_off[nq] = the total -- the numbers sa_query_ranges_r returns in hit_off.
 * Calls no cudaMalloc / cudaFree / allocating routine and does not synchronise: three kernels are queued
 * on `stream` (a cudaStream_t; NULL = the default stream) and the call returns.  Its scratch is a fixed
 * __device__ array inside the library (128 KiB + 8 B on every GPU that loads it, held until exit):
 * at most SA_NOALLOC_MAX_QUERIES queries per call, more -> SA_ERR_LIMIT.  All calls share that array:
 * calls that overlap on the GPU (different streams without an event between them) race; calls on one
 * stream are ordered.  The first call on a device looks up the array's device address
 * (cudaGetSymbolAddress): make it outside graph capture. */
#define SA_NOALLOC_MAX_QUERIES (1ull << 24)
int sa_hit_offsets_noalloc(sa_index* idx, const uint32_t* d_start, const uint32_t* d_end, uint64_t nq, uint32_t k,
                           uint64_t* d_hit_off, void* stream);

/* Document table: doc_start[num_docs] = text offset of each record's first byte. */
uint64_t sa_num_docs(const sa_index* idx);
int sa_copy_doc_starts(sa_index* idx, uint32_t* out, uint64_t cap);

/* ---- One suffix array over several GPUs (sample sort; one sa_index per GPU / process, the text
 * replicated).  No reference counterpart: the reference only cuts the text into independent 2 GiB
 * partitions (suffix_array.pyx:148-180).  Call order on every rank:
 *   sa_shard_begin -> sa_shard_pack (+ all-reduce of the 65536-bin histogram, splitter choice on
 *   the host) -> sa_shard_split -> all-to-all of the (key, position) records over NCCL ->
 *   sa_shard_sort -> sa_shard_edges (+ all-gather of the edges) -> sa_shard_finish ->
 *   all-gather of the slices -> sa_adopt_sa (queries).  suffixarray_b200/sharded.py drives it. */
int sa_shard_begin(sa_index* idx, const uint8_t* text, uint64_t n, uint32_t flags);
/* Same starting point with the text taken from a CSV column (BASELINE config 3 at 2/4/8 GPUs): every
 * rank reads and scans the file itself (sa_build_csv's feeder), the stages that follow are unchanged,
 * and when the complete array is installed (sa_shard_mark_built / sa_adopt_sa[_parts]) the rank runs
 * the reference's two post-build loops on it (SA[i] = map[SA[i]], quoted bits; engine.c:632-651).
 * sa_shard_text_size / sa_shard_copy_text (any byte range) give the host driver the few windows of
 * the column text it needs for the boundary LCPs. */
int sa_shard_begin_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags);
uint64_t sa_shard_text_size(const sa_index* idx);
int sa_shard_copy_text(sa_index* idx, uint64_t offset, uint64_t len, uint8_t* out);
int sa_shard_pack(sa_index* idx, uint64_t pos_begin, uint64_t pos_end, uint64_t* d_keys, uint32_t* d_pos,
                  uint32_t* d_hist16);
int sa_shard_split(sa_index* idx, const uint64_t* d_keys, const uint32_t* d_pos, uint64_t m,
                   const uint8_t* bin2rank, int nranks, uint64_t* d_keys_out, uint32_t* d_pos_out,
                   uint64_t* counts_out);
int sa_shard_sort(sa_index* idx, const uint64_t* d_keys, const uint32_t* d_pos, uint64_t m);
int sa_shard_edges(sa_index* idx, uint32_t edge, uint32_t* head_pos, uint16_t* head_lcp, uint32_t* tail_pos,
                   uint16_t* tail_lcp);
int sa_shard_finish(sa_index* idx, const uint32_t* left_pos, const uint16_t* left_lcp, uint32_t nleft,
                    uint32_t lcp_first, const uint32_t* right_pos, const uint16_t* right_lcp, uint32_t nright,
                    uint32_t* d_sa_out);
int sa_adopt_sa(sa_index* idx, const uint32_t* d_sa_full, uint64_t n);
int sa_adopt_sa_parts(sa_index* idx, const uint32_t* d_parts, uint64_t stride, const uint64_t* counts, int nparts);
/* Fused split + exchange for ranks on one node: instead of sa_shard_split + NCCL all-to-all, the
 * split kernel stores every record straight into its destination rank's sort buffers, mapped
 * through CUDA IPC (peer stores over NVLink).  sa_shard_count -> all-gather of the counts ->
 * sa_shard_recv_prepare (exports two 64-byte handles) -> all-gather of the handles ->
 * sa_shard_scatter_p2p -> barrier -> sa_shard_sort_recv. */
int sa_shard_count(sa_index* idx, const uint64_t* d_keys, uint64_t m, const uint8_t* bin2rank, int nranks,
                   uint64_t* counts_out);
int sa_shard_recv_prepare(sa_index* idx, uint64_t m_recv, uint8_t* handle_keys, uint8_t* handle_pos);
int sa_shard_scatter_p2p(sa_index* idx, const uint64_t* d_keys, const uint32_t* d_pos, uint64_t m, int nranks,
                         int my_rank, const uint8_t* handles, const uint64_t* dst_offset);
int sa_shard_sort_recv(sa_index* idx, uint64_t m);

/* Finish + replicate, overlapped (no all-gather, no adopt copy): the shard is finished in chunks
 * straight into this rank's final array while the previous chunk is pushed into every peer's final
 * array over NVLink (a copy kernel on a second stream storing into CUDA IPC mappings).
 * sa_shard_final_prepare exports this rank's array (handle_final: 64 bytes);
 * sa_shard_finish_bcast takes all ranks' handles (64 bytes each) and the global rank of this
 * shard's first record; after a barrier over the ranks sa_shard_mark_built makes the index
 * queryable.  Same halo arguments as sa_shard_finish. */
int sa_shard_final_prepare(sa_index* idx, uint8_t* handle_final);
int sa_shard_finish_bcast(sa_index* idx, const uint32_t* left_pos, const uint16_t* left_lcp, uint32_t nleft,
                          uint32_t lcp_first, const uint32_t* right_pos, const uint16_t* right_lcp, uint32_t nright,
                          int nranks, int my_rank, const uint8_t* handles, uint64_t global_offset);
int sa_shard_mark_built(sa_index* idx);

/* ---- The native driver of the multi-GPU build (sa_multi.cu): the whole pipeline above in ONE call
 * per rank.  The small exchanges (histogram sum, counts, buffer handles, shard edges) and the barriers
 * go through a node-local communicator; text slices, records and array slices move GPU to GPU as peer
 * stores over NVLink.  Each rank uploads only ITS slice of the text.
 *   sa_comm_create_local: the ranks are threads of one process (out: nranks handles)
 *   sa_comm_create_shm:   one process per GPU on one node (torchrun); `name` ("/..." unique per job)
 *                         names a POSIX shared-memory object rank 0 creates
 * sa_shard_build: text_slice = text[slice_off, slice_off + slice_len) (host pointer, or a device
 * pointer with SA_FLAG_DEVICE_TEXT); the slices of the ranks tile [0, n) in rank order.  phase_ms
 * (optional, 8 floats): text, prepare, pack+hist, split, exchange, sort+refine, edges, finish+replicate.
 * Afterwards every rank's index holds the text and the complete array (a replica for queries).
 * sa_build_text_multi: the same over the GPUs (handles) of ONE process, a thread per handle. */
typedef struct sa_comm sa_comm;
int sa_comm_create_local(int nranks, sa_comm** out);
int sa_comm_create_shm(const char* name, int rank, int nranks, sa_comm** out);
void sa_comm_destroy(sa_comm* c);
int sa_comm_rank(const sa_comm* c);
int sa_comm_size(const sa_comm* c);
int sa_comm_barrier(sa_comm* c);
int sa_shard_build(sa_index* idx, sa_comm* c, const uint8_t* text_slice, uint64_t slice_off, uint64_t slice_len,
                   uint64_t n, uint32_t flags, float* phase_ms);
int sa_shard_info(const sa_index* idx, uint64_t* shard_records, uint64_t* exchange_bytes_sent);
int sa_build_text_multi(sa_index* const* idxs, int ndev, const uint8_t* text, uint64_t n, uint32_t flags);

/* Persistence (the reference only sketches it: write_suffix_array / read_into_suffix_array_file,
 * engine.c:1098-1165, TODO.md).  sa_save writes one container file (text, ORDER-E positions, and for
 * CSV indexes the file offsets + quoted bits); sa_load restores it into a handle created with the
 * same max_suffix_length -- the index then answers queries without sorting anything.
 * sa_export_reference_files writes the reference's own two-file layout (u64 start, u64 end, u32 L,
 * u32 n, u32 SA[n]; u32 capacity + quoted bits); quoted_path may be NULL. */
int sa_save(sa_index* idx, const char* path);
int sa_load(sa_index* idx, const char* path);
int sa_export_reference_files(sa_index* idx, const char* sa_path, const char* quoted_path);

/* Debug / parity hooks (used by tests to localise a mismatch; not part of the drop-in). */
int sa_debug_copy_lcp(sa_index* idx, uint16_t* out, uint64_t cap);
int sa_debug_set(sa_index* idx, const char* key, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* SA_CUDA_H */
