// sa_index.h -- the opaque handle behind sa_cuda.h and small host utilities.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

#include "../../include/sa_cuda.h"
#include "common.cuh"
#include "kernels.h"

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap && p) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        if (bytes == 0) bytes = 256;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// layout of the small device scratch block (u32 words)
enum SmallSlot : uint32_t {
    SM_HIST = 0,            // 8*256
    SM_BASE = 2048,         // 8*256
    SM_SKIP = 4096,         // 8
    SM_TOTALS = 4104,       // 2
    SM_TILECTR = 4106,      // 1
    SM_ERR = 4107,          // 1
    SM_RUNCTR = 4108,       // 2 (u64, 8-byte aligned)
    SM_NLTOTAL = 4110,      // 1
    SM_BYTEHIST = 4112,     // 256 presence flags
    SM_TABLE = 4368,        // 256 bytes = 64 words: byte -> code
    SM_CSV = 4432,          // 8: irregularity flags, total quotes, total delimiters, total emitted bytes
    SM_WORDS = 4440
};

constexpr uint64_t SA_MAX_INDEX_BYTES = 1ull << 31;   // records per index: look-back prefixes are 31-bit
constexpr size_t H_SMALL_WORDS = 1024;   // [0,64) misc, [256,512) byte presence, [512,576) code table

struct PhaseTimer {
    cudaEvent_t a = nullptr, b = nullptr;
    float* target = nullptr;
};

struct sa_index;

// Per-call query state (sa_cuda.h "Per-call query results"): device scratch of one batch, its own
// stream and events.  Created by sa_result_create, bound to one index.
struct sa_result {
    sa_index* idx = nullptr;
    int device = 0;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    DevBuf qbytes, qoff, qstart, qend, qcounts, hitoff, blocksums, hitpos, hitdoc;
    DevBuf ucount, uoff, upos, udoc, table;     // dedupe: unique counts / compacted offsets / lists, hash scratch
    DevBuf reclen, recoff, recbytes, recsrc;    // records: lengths, offsets, packed bytes, source offsets
    uint64_t nq = 0, total_hits = 0, generation = 0;
    uint32_t k = 0, range_flags = 0;
    bool have_ranges = false;
    // after an expansion (sa_query_hits_r / sa_query_records_size)
    bool have_hits = false, deduped = false, have_records = false;
    uint64_t out_hits = 0, rec_bytes = 0;
    cudaEvent_t ev[4] = {};          // query begin/end, hits begin/end
    float ms_query = 0.f, ms_hits = 0.f;
    uint64_t launches = 0;
    uint64_t* h_pin = nullptr;       // pinned, 8 words
};

struct sa_index {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    uint32_t L = 0;
    uint64_t n = 0;
    uint32_t tail = 0;          // bytes of look-ahead text kept behind the n indexed positions (sa_build_text_tail)
    uint32_t lprime = 0;
    bool built = false;
    bool csv = false;
    std::shared_mutex mu;       // builds / loads: exclusive; queries (one sa_result each): shared

    // persistent after a build
    DevBuf text;        // n + SA_TEXT_PAD bytes, zero padded
    DevBuf sa_final;    // u32[n] text positions in ORDER-E
    DevBuf nlbits, nlrank;   // rank structure over the '\n' bytes: u64 bit mask + u32 count in front, per 64 text bytes
    DevBuf nlpos;       // u32[num_nl] positions of '\n'
    uint32_t num_nl = 0;
    DevBuf sa_file;     // CSV: u32[n] file offsets (engine.c:647-651)
    DevBuf qbits;       // CSV: quoted bit per rank
    DevBuf map, quoted_text;   // CSV feeder outputs on the device
    DevBuf csv_file, csv_tq, csv_td, csv_te, csv_tr;   // device copy of the file + per-tile counts (csv_scan.cu)
    DevBuf rowbits, rowrank;   // CSV: rank structure over the record ends of the FILE (hits -> row index)
    void* h_ring[2] = {nullptr, nullptr};      // pinned staging ring of the file reader
    cudaEvent_t ring_ev[2] = {nullptr, nullptr};
    size_t ring_bytes = 0;
    uint64_t csv_file_bytes = 0;               // size of the CSV file of the last sa_build_csv

    // build scratch (kept between builds so repeated builds do not re-allocate)
    DevBuf keys[2], vals[2], aux[2], lcp, tie, amask, hmask, counts, status, small, nl_counts;
    DevBuf codes;       // compressed-alphabet copy of the text (code_bits < 8), build scratch
    sa::CodeInfo code{8, 0x100u, 0x100u};
    int cb0 = 8, max_chars = 8;   // suffix bytes in the round-0 key / per key word

    // native multi-GPU driver (sa_multi.cu): peers' buffers this rank stores into, mapped through CUDA IPC
    struct PeerMapping {
        void* ptr = nullptr;
        cudaIpcMemHandle_t h{};
        bool mapped = false;
    };
    PeerMapping peer_text[16], peer_nlbits[16], peer_nlrank[16];
    DevBuf pack_keys, pack_pos, hist16;       // round-0 records of this rank's slice of positions, splitter histogram
    uint8_t* h_big = nullptr;                 // pinned, one mailbox (256 KiB + 8 KiB)
    cudaStream_t push_stream = nullptr;       // high priority: peer pushes that must not queue behind the sort kernels
    uint64_t shard_m_last = 0, shard_exchange_bytes = 0;
    // SA_FLAG_SHARDED_SA: the array stays distributed; `view` routes a global rank to its owner's memory
    bool shard_no_replicate = false;          // set for the duration of a build
    bool sharded_sa = false;                  // the built index answers queries through `view`
    sa::SaView view{};
    uint32_t* shard_peer_base[16] = {};       // every rank's final array as this rank addresses it (finish_bcast)

    // sharded (multi-GPU) build state, see sa_shard.cu
    bool shard_prepared = false;
    bool shard_csv = false;          // the sharded build's text is a CSV column (sa_shard_begin_csv): remap when the array is complete
    bool shard_csv_regular = false;  // ... scanned on the device (quoted flags ride in the offset map)
    uint32_t shard_m = 0;          // records this rank sorts
    uint32_t* shard_sa = nullptr;  // their positions in prefix order (points into vals[])
    DevBuf ext_sa, ext_lcp, ext_out, split_table, split_offsets;
    // peers' receive buffers (their keys[0] / vals[0]) mapped through CUDA IPC for the fused split + exchange
    void* peer_keys[16] = {};
    void* peer_pos[16] = {};
    cudaIpcMemHandle_t peer_hk[16] = {}, peer_hp[16] = {};
    bool peer_mapped[16] = {};
    void* peer_final[16] = {};                 // peers' sa_final arrays (fused finish + replicate)
    cudaIpcMemHandle_t peer_hf[16] = {};
    bool peer_final_mapped[16] = {};
    cudaStream_t copy_streams[16] = {};        // replication copies (one stream per peer) run beside the finishing kernels
    cudaEvent_t chunk_ev[16] = {};
    cudaEvent_t ext_ev[17] = {};               // sharded finish: piece c of the shard staged into the ext arrays (+ "inputs ready")

    // queries: per-call state lives in sa_result; `generation` changes with every build / load so a
    // result of an older array cannot be expanded against a newer one
    uint64_t generation = 1;
    sa_result* default_res = nullptr;   // the result object behind sa_query_ranges / sa_query_hits
    std::mutex res_mu;                  // creates default_res
    float ms_query_last = 0.f, ms_hits_last = 0.f;   // mirrored into sa_stats for the legacy calls
    // probe cache of the batched bisection (query.cu "qtop"), built on the first big batch after a build
    DevBuf qtop;
    int qtop_levels = 0;
    bool qtop_ready = false;
    std::mutex qtop_mu;
    int64_t dbg_qtop_levels = -1;   // -1: automatic (<= 20); 0: off
    // CSV: positions of the record ends of the file (built on first use from rowbits / rowrank)
    DevBuf rowpos;
    uint64_t num_row_ends = 0;
    bool rowpos_ready = false;
    std::mutex rowpos_mu;

    uint32_t* h_small = nullptr;   // pinned mirror for tiny read-backs (H_SMALL_WORDS words)

    sa_stats stats{};
    std::vector<PhaseTimer> timers;
    std::vector<cudaEvent_t> event_pool;
    size_t events_used = 0;
    bool timing = false;

    // debug knobs (sa_debug_set)
    int64_t dbg_max_rounds = -1;    // stop refining after this many rounds (-1: no limit)
    int64_t dbg_skip_finish = 0;    // leave the phase-1 order (no <=32 bucket step)
    int64_t dbg_force_gbits = 0;    // use at least this many group-id bits (exercises small cb)
    int64_t dbg_query_kernel = 0;   // 0: by batch size; 1: one thread per query; 2: eight lanes per query
    int64_t dbg_same_process = 0;   // sharded build: peers live in this process (raw pointers instead of IPC)
    int64_t dbg_replicate_memcpy = 0;   // sharded finish: replicate with cudaMemcpyAsync instead of replicate_kernel
    int64_t dbg_timeline = 0;       // sharded finish: print when each chunk was finished / pushed (stderr, CUDA events)
    int64_t dbg_replicate_ctas = 0; // sharded finish: k > 0 = 512-thread push kernel with k blocks (0: 64); k < 0 = the 128-thread co-resident kernel
    int64_t dbg_text_push = 0;      // sharded build: 0 = text slices travel on the copy engines beside pack / exchange; 1 = blocking push kernel
    int64_t dbg_split_gamma_x100 = -1;   // sharded build: extra cost of a record whose first-sort key repeats, in 1/100 records (-1: 170; 0: equal counts)
    int64_t dbg_replicate_fused = 1;    // sharded finish: 1 = finish_kernel stores into every replica itself; 0 = replicate_kernel pushes finished chunks
    int64_t dbg_ext_fill = 0;       // sharded finish: 1 = one kernel copies the whole shard into the ext arrays first (round-1 behaviour)
    int64_t dbg_finish_chunks = 0;  // sharded finish: chunks the shard is finished / pushed in (0: 8; at most 16)
    int64_t dbg_csv_scanner = 0;    // 0: device scan, host scan for irregular files; 1: host scan; 2: device scan or fail
    int64_t dbg_sort0_chars = -1;   // -1: automatic; else at most this many characters in the first sort's keys
    int64_t dbg_code_bits = -1;     // -1: choose from the alphabet; 8: never compress; 1..7: at least this width
};

void sa_set_error(const char* fmt, ...);
int sa_cuda_fail(cudaError_t e, const char* what);

#define SA_CU(call)                                             \
    do {                                                        \
        cudaError_t e__ = (call);                               \
        if (e__ != cudaSuccess) return sa_cuda_fail(e__, #call); \
    } while (0)

// phase timing (no-ops unless SA_FLAG_TIMING): a ScopedPhase records two events on the index
// stream; sa_time_collect() synchronises and adds the elapsed times to their targets.
cudaEvent_t sa_take_event(sa_index* idx);
void sa_time_collect(sa_index* idx);

struct ScopedPhase {
    sa_index* idx;
    cudaEvent_t a = nullptr, b = nullptr;
    float* target;
    ScopedPhase(sa_index* i, float* tgt) : idx(i), target(tgt)
    {
        if (!idx->timing) return;
        a = sa_take_event(idx);
        b = sa_take_event(idx);
        cudaEventRecord(a, idx->stream);
    }
    ~ScopedPhase()
    {
        if (!idx->timing) return;
        cudaEventRecord(b, idx->stream);
        idx->timers.push_back(PhaseTimer{a, b, target});
    }
};

int sa_build_resident(sa_index* idx, uint32_t flags);
int sa_stage_prepare(sa_index* idx, uint32_t flags);
int sa_stage_alphabet(sa_index* idx, const uint32_t* present, uint64_t enc_lo = 0, uint64_t enc_hi = ~0ull);
int sa_stage_encode_range(sa_index* idx, uint64_t lo, uint64_t hi);
const uint8_t* sa_key_source(const sa_index* idx);
int sa_ensure_scratch(sa_index* idx, uint32_t m);
int sa_stage_sort_refine(sa_index* idx, uint32_t m, bool identity, uint32_t** sa_sorted);
int sa_stage_finish(sa_index* idx, const uint32_t* sa, const uint16_t* lcp, uint32_t m, uint32_t r_begin,
                    uint32_t r_end, uint32_t* out);
int sa_stage_finish_to(sa_index* idx, const uint32_t* sa, const uint16_t* lcp, uint32_t m, uint32_t r_begin,
                       uint32_t r_end, const sa::FinishOut& outs, uint32_t block0 = 0, uint32_t nblocks = 0xFFFFFFFFu);
int sa_check_device_error(sa_index* idx);
int sa_csv_finalize_shard(sa_index* idx);
inline void sa_invalidate_results(sa_index* idx) { idx->generation += 1; idx->rowpos_ready = false; idx->qtop_ready = false; idx->sharded_sa = false; }   // sa_csv.cu: SA[i] = map[SA[i]] + quoted bits on the complete array of a sharded CSV build
