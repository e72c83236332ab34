// kernels.h -- host-side launch wrappers of the sm_100a kernels (internal to libsa_cuda.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

// ---- radix.cu
uint32_t sa_onesweep_tile_items();
cudaError_t sa_launch_hist(const uint64_t* keys, uint32_t n, int shift0, int npasses, uint32_t* hist,
                           int num_sms, cudaStream_t s);
cudaError_t sa_launch_scan_hist(const uint32_t* hist, uint32_t n, int npasses, uint32_t* base, uint32_t* skip,
                                cudaStream_t s);
cudaError_t sa_launch_onesweep(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                               uint32_t* vals_out, uint32_t n, int shift, const uint32_t* digit_base,
                               uint32_t* status, uint32_t* tile_counter, uint32_t* err, bool identity_vals,
                               cudaStream_t s);

// ---- refine.cu
constexpr uint32_t SA_SCAN_TILE = 4096;   // ranks per block in the group analysis / compaction kernels
cudaError_t sa_launch_iota(uint32_t* out, uint32_t n, cudaStream_t s);
cudaError_t sa_launch_lowercase(uint8_t* text, uint64_t n, cudaStream_t s);
// presence (0/1) of every byte value in text[0, n)
cudaError_t sa_launch_byte_hist(const uint8_t* text, uint64_t n, uint32_t* hist256, int num_sms, cudaStream_t s);
// codes = table[text] over nbytes_padded (multiple of 16) bytes
cudaError_t sa_launch_encode(const uint8_t* text, uint64_t nbytes_padded, const uint8_t* table, uint8_t* codes,
                             cudaStream_t s);
// src = text (ci.bits == 8) or the code array; cb = suffix bytes per key
cudaError_t sa_launch_pack0(const uint8_t* src, uint32_t n, int cb, sa::CodeInfo ci, uint64_t* keys, cudaStream_t s);
// also writes the tie mask of the first refinement round (bit r = lcp[r] == cb)
cudaError_t sa_launch_lcp0(const uint64_t* sorted_keys, uint32_t n, int cb, int bits, uint32_t lprime, uint16_t* lcp,
                           uint32_t* tie_words, uint32_t num_tie_words, cudaStream_t s);
// tie bit r = (lcp[r] == depth); the mask has one zero word in front (bit r lives at bit r+32)
cudaError_t sa_launch_tie_mask(const uint16_t* lcp, uint32_t n, uint32_t depth, uint32_t* tie_words,
                               uint32_t num_words, cudaStream_t s);
// amask: ranks in tied groups of more than 32 members; hmask: first rank of each such group;
// counts[2*tile], counts[2*tile+1]: per-tile popcounts (tile = SA_SCAN_TILE ranks)
cudaError_t sa_launch_group_flags(const uint32_t* tie_words, uint32_t nwords_data, uint32_t* amask,
                                  uint32_t* hmask, uint32_t* counts, cudaStream_t s);
// exclusive scan of `num` (a,b) pairs in place; totals[0..1] = sums
cudaError_t sa_launch_scan_pairs(uint32_t* counts, uint32_t num, uint32_t* totals, cudaStream_t s);
cudaError_t sa_launch_compact_gather(const uint32_t* amask, const uint32_t* hmask, uint32_t nwords_data,
                                     const uint32_t* tile_offsets, const uint32_t* sa, const uint8_t* text,
                                     uint32_t depth, int cb, int gbits, sa::CodeInfo ci, uint32_t* slots,
                                     uint64_t* ckeys, uint32_t* cvals, cudaStream_t s);
cudaError_t sa_launch_writeback(const uint32_t* slots, const uint64_t* ckeys, const uint32_t* cvals, uint32_t m,
                                uint32_t depth, int cb, int gbits, int bits, uint32_t lprime, uint32_t* sa,
                                uint16_t* lcp, uint32_t* tie_next, uint32_t num_tie_words, cudaStream_t s);

// ---- finish.cu
cudaError_t sa_launch_finish(const uint32_t* sa_in, const uint16_t* lcp, const uint8_t* text, uint32_t n,
                             uint32_t lprime, uint32_t r_begin, uint32_t r_end, const sa::FinishOut& outs,
                             unsigned long long* run_counter, uint32_t* err, cudaStream_t s, uint32_t block0 = 0,
                             uint32_t nblocks = 0xFFFFFFFFu);
uint32_t sa_finish_blocks(uint32_t n);       // blocks of sa_finish_block_ranks() consecutive ranks
uint32_t sa_finish_block_ranks();

// ---- query.cu
uint32_t sa_newline_tiles(uint32_t n);
cudaError_t sa_launch_newline_count(const uint8_t* text, uint32_t n, uint32_t* counts, cudaStream_t s);
cudaError_t sa_launch_scan_u32(uint32_t* counts, uint32_t num, uint32_t* total, cudaStream_t s);
cudaError_t sa_launch_newline_write(const uint8_t* text, uint32_t n, const uint32_t* tile_offsets,
                                    uint32_t* nlpos, unsigned long long* nlbits, uint32_t* nlrank, cudaStream_t s,
                                    uint32_t count_base = 0, uint32_t pos_base = 0);
cudaError_t sa_launch_query(const uint8_t* text, const uint32_t* sa, uint32_t n, uint32_t max_suffix_length,
                            const uint8_t* qbytes, const uint64_t* qoff, uint32_t nq, uint32_t* start,
                            uint32_t* end, int mode /* 0 auto, 1 thread per query, 2 eight lanes per query */,
                            cudaStream_t s, const void* qtop = nullptr, int qtop_levels = 0);
// probe cache of the top `levels` levels of the bisection tree (16 bytes per node, 2^levels entries)
cudaError_t sa_launch_qtop_build(const uint8_t* text, const uint32_t* sa, uint32_t n, int levels, void* qtop, cudaStream_t s);
cudaError_t sa_launch_hit_counts(const uint32_t* start, const uint32_t* end, uint32_t nq, uint32_t k, uint32_t n,
                                 uint32_t* counts, cudaStream_t s);
// exclusive scan of u32 counts into u64 offsets (off[num] = total)
cudaError_t sa_launch_scan_u32_to_u64(const uint32_t* counts, uint32_t num, uint64_t* off, uint32_t* block_sums,
                                      cudaStream_t s);
cudaError_t sa_launch_hits(const uint32_t* sa, const uint32_t* start, const uint64_t* hit_off, uint32_t nq,
                           const unsigned long long* nlbits, const uint32_t* nlrank, uint32_t* hit_pos,
                           uint32_t* hit_doc, cudaStream_t s);
cudaError_t sa_launch_remap(const uint32_t* sa, uint32_t n, const uint32_t* map, const uint8_t* quoted_by_text,
                            uint32_t* sa_out, uint32_t* quoted_bits_by_rank, cudaStream_t s);

// ---- csv_scan.cu (device copy of the file, zero padded to sa_csv_padded_bytes)
uint32_t sa_csv_tiles(uint64_t size);
uint64_t sa_csv_padded_bytes(uint64_t size);
cudaError_t sa_launch_csv_quotes(const uint8_t* f, uint64_t size, uint32_t* tile_q, uint32_t* flags, cudaStream_t s);
cudaError_t sa_launch_csv_delims(const uint8_t* f, uint64_t size, const uint32_t* tile_q, uint32_t* tile_d, cudaStream_t s);
cudaError_t sa_launch_csv_emit(bool write, const uint8_t* f, uint64_t size, const uint32_t* tile_q, const uint32_t* tile_d,
                               uint32_t* tile_e, uint32_t wanted, uint32_t num_columns, uint8_t* text, uint32_t* map,
                               uint32_t* flags, cudaStream_t s);
// row table of the file (record ends = '\n' outside quotes): bit mask + rows in front of every 64-byte block
uint64_t sa_csv_rowbits_bytes(uint64_t size);
uint64_t sa_csv_rowrank_bytes(uint64_t size);
cudaError_t sa_launch_csv_rows(const uint8_t* f, uint64_t size, const uint32_t* tile_q, unsigned long long* rowbits,
                               uint32_t* tile_r, cudaStream_t s);
cudaError_t sa_launch_csv_rowrank(const unsigned long long* rowbits, uint64_t size, const uint32_t* tile_r, uint32_t* rowrank,
                                  cudaStream_t s);

// ---- query.cu: file-semantics bisection (get_substring_positions_file, engine.c:920-999)
cudaError_t sa_launch_query_file(const uint8_t* file, uint64_t file_size, const uint32_t* sa_file, uint32_t n,
                                 uint32_t max_suffix_length, const uint8_t* qbytes, const uint64_t* qoff, uint32_t nq,
                                 uint32_t* start, uint32_t* end, cudaStream_t s);

// ---- records.cu: segmented dedupe of hit lists, record table of a CSV file, record materialisation
// in-place: per query the first ucount[q] entries of its segment of hit_pos / hit_doc become the hits
// whose record id occurs for the first time (rank order kept).  table: scratch of 4 u32 per hit, only
// touched by queries with more than 1024 hits.
cudaError_t sa_launch_dedupe_mark(const uint64_t* hit_off, uint32_t nq, uint32_t* hit_pos, uint32_t* hit_doc,
                                  uint32_t* ucount, uint32_t* table, cudaStream_t s);
cudaError_t sa_launch_segment_compact(const uint64_t* hit_off, const uint64_t* uoff, uint32_t nq, const uint32_t* hit_pos,
                                      const uint32_t* hit_doc, uint32_t* upos, uint32_t* udoc, cudaStream_t s);
// positions of the set bits of a rank structure (bits: one per byte, 64 per word; rank: set bits in
// front of every word) -> pos[rank + i]
cudaError_t sa_launch_rank_positions(const unsigned long long* bits, const uint32_t* rank, uint64_t nwords, uint32_t* pos,
                                     cudaStream_t s);
// per hit: source offset + length of its record.  ends: ascending positions of the record-ending '\n'
// (num_ends of them), total: bytes of the source (text or file); compat: see SA_RECORDS_REFERENCE_COMPAT
cudaError_t sa_launch_record_bounds(const uint32_t* hit_pos, const uint32_t* hit_doc, uint64_t num, const uint32_t* ends,
                                    uint64_t num_ends, uint64_t total, bool compat, unsigned long long* src_off,
                                    uint32_t* len, cudaStream_t s);
cudaError_t sa_launch_record_gather(const uint8_t* src, const unsigned long long* src_off, const uint64_t* rec_off,
                                    uint64_t num, uint8_t* out, cudaStream_t s);
cudaError_t sa_launch_count_out_of_range(const uint32_t* a, uint64_t n, uint64_t limit, unsigned int* bad, cudaStream_t s);
// ---- query.cu: the array as key ranges on several GPUs (sa::SaView, common.cuh)
cudaError_t sa_launch_query_sharded(const uint8_t* text, const sa::SaView& view, uint32_t n, uint32_t max_suffix_length,
                                    const uint8_t* qbytes, const uint64_t* qoff, uint32_t nq, uint32_t* start, uint32_t* end,
                                    cudaStream_t s, const void* qtop, int qtop_levels);
cudaError_t sa_launch_qtop_build_sharded(const uint8_t* text, const sa::SaView& view, uint32_t n, int levels, void* qtop,
                                         cudaStream_t s);
cudaError_t sa_launch_hits_sharded(const sa::SaView& view, const uint32_t* start, const uint64_t* hit_off, uint32_t nq,
                                   const unsigned long long* nlbits, const uint32_t* nlrank, uint32_t* hit_pos,
                                   uint32_t* hit_doc, cudaStream_t s);
