// sa_multi.cu -- the native driver of the multi-GPU build: ONE call per rank (sa_shard_build) runs the
// whole sample-sort pipeline of sa_shard.cu, with the small host-side exchanges (histogram sum, counts,
// buffer handles, shard edges) and the barriers done through a node-local communicator instead of a
// Python loop over torch.distributed.  Bulk data never touches it: text slices, (key, position) records
// and finished array slices move GPU to GPU as peer stores over NVLink.
//
//   text      every rank uploads ITS slice; after one barrier it pulls the few KB of its own position range that
//             another rank uploaded and hands its slice to the copy engines (one copy per peer): the all-gather
//             of the text over NVLink runs beside the next four stages (the whole text is never sent over PCIe
//             more than once per node)
//   prepare   newline table + byte presence on the rank's slice of positions, exchanged / pushed;
//             the alphabet codes are computed locally (cheaper than NVLink): own range now, the rest before the sort
//   pack      round-0 keys of the slice + two 65536-bin histograms of a sample (keys, keys with a twin in the
//             sample); summed over ranks -> splitters by equal share of the expected work
//   exchange  split_scatter_p2p_kernel: records stored straight into the owner's sort buffers
//   sort      onesweep LSD sort + tie refinement of the received key range
//   edges     first / last 128 records of every shard; junction LCPs computed on the device
//   finish    <= 32 bucket finish; every finished block leaves finish_kernel as 16-byte stores into this rank's
//             AND every peer's final array (the replication is fused into the kernel)
//
// Communicators: sa_comm_create_local (ranks are threads of one process: sa_build_text_multi,
// SuffixArray(devices=[...])) and sa_comm_create_shm (one process per GPU on one node -- torchrun: a
// POSIX shared-memory block; CUDA IPC handles travel through it).  Every wait has a timeout and an
// abort flag, so a failing rank cannot hang its peers.
//
// The reference has no counterpart: its only split is the sequential 2 GiB partitioning
// (suffix_array.pyx:148-180).
#include <fcntl.h>
#include <sched.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "sa_index.h"

#define MG_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            sa_set_error(__VA_ARGS__);         \
            return (code);                     \
        }                                      \
    } while (0)

// ------------------------------------------------------------------------------ communicator
namespace {

constexpr size_t MAILBOX_BYTES = 2 * 65536 * sizeof(uint32_t) + 8192;   // two histograms, or a few KB of small data
constexpr double COMM_TIMEOUT_S = 120.0;

struct CommShared {
    std::atomic<uint32_t> ready;
    std::atomic<uint32_t> arrive;
    std::atomic<uint32_t> gen;
    std::atomic<uint32_t> abort_flag;
    std::atomic<uint32_t> attached;
    uint32_t nranks;
    uint32_t pad[10];
    // 2 slots x nranks mailboxes follow
};

double now_s()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

}  // namespace

struct sa_comm {
    CommShared* sh = nullptr;
    int rank = 0, nranks = 1;
    bool local = false;          // ranks are threads of this process (raw device pointers instead of IPC handles)
    size_t bytes = 0;
    std::string shm_name;
    uint32_t op = 0;             // collective counter: mailbox slot = op & 1
    uint8_t* box(int slot, int r) const
    {
        return reinterpret_cast<uint8_t*>(sh) + sizeof(CommShared) + ((size_t)slot * nranks + r) * MAILBOX_BYTES;
    }
};

static size_t comm_bytes(int nranks) { return sizeof(CommShared) + 2 * (size_t)nranks * MAILBOX_BYTES; }

extern "C" int sa_comm_create_local(int nranks, sa_comm** out)
{
    MG_REQUIRE(out && nranks >= 1 && nranks <= 16, SA_ERR_INVALID, "sa_comm_create_local: 1..16 ranks");
    const size_t bytes = comm_bytes(nranks);
    void* mem = calloc(1, bytes);
    MG_REQUIRE(mem, SA_ERR_OOM, "host allocation failed");
    CommShared* sh = new (mem) CommShared();
    sh->nranks = (uint32_t)nranks;
    sh->attached.store((uint32_t)nranks);
    sh->ready.store(1);
    for (int r = 0; r < nranks; ++r) {
        sa_comm* c = new (std::nothrow) sa_comm();
        if (!c) { sa_set_error("host allocation failed"); return SA_ERR_OOM; }
        c->sh = sh; c->rank = r; c->nranks = nranks; c->local = true; c->bytes = bytes;
        out[r] = c;
    }
    return SA_OK;
}

// One process per GPU on one node.  `name` ("/something", unique per job) names a POSIX shared-memory
// object: rank 0 creates it, the others attach (waiting for it to appear).
extern "C" int sa_comm_create_shm(const char* name, int rank, int nranks, sa_comm** out)
{
    MG_REQUIRE(out && name && name[0] == '/', SA_ERR_INVALID, "sa_comm_create_shm: name must start with '/'");
    MG_REQUIRE(nranks >= 1 && nranks <= 16 && rank >= 0 && rank < nranks, SA_ERR_INVALID, "sa_comm_create_shm: bad rank");
    *out = nullptr;
    const size_t bytes = comm_bytes(nranks);
    int fd = -1;
    if (rank == 0) {
        shm_unlink(name);
        fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
        MG_REQUIRE(fd >= 0, SA_ERR_IO, "shm_open(%s) failed", name);
        if (ftruncate(fd, (off_t)bytes) != 0) { close(fd); shm_unlink(name); sa_set_error("ftruncate failed"); return SA_ERR_IO; }
    } else {
        const double t0 = now_s();
        for (;;) {
            fd = shm_open(name, O_RDWR, 0600);
            if (fd >= 0) {
                struct stat sb;
                if (fstat(fd, &sb) == 0 && (size_t)sb.st_size >= bytes) break;
                close(fd);
                fd = -1;
            }
            MG_REQUIRE(now_s() - t0 < COMM_TIMEOUT_S, SA_ERR_IO, "sa_comm_create_shm: rank 0 never created %s", name);
            usleep(1000);
        }
    }
    void* mem = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    MG_REQUIRE(mem != MAP_FAILED, SA_ERR_IO, "mmap of %s failed", name);
    CommShared* sh = static_cast<CommShared*>(mem);
    if (rank == 0) {
        new (sh) CommShared();            // the object is zero-filled by ftruncate
        sh->nranks = (uint32_t)nranks;
        sh->ready.store(1, std::memory_order_release);
    } else {
        const double t0 = now_s();
        while (sh->ready.load(std::memory_order_acquire) != 1) {
            if (now_s() - t0 > COMM_TIMEOUT_S) { munmap(mem, bytes); sa_set_error("sa_comm_create_shm: rank 0 never initialised %s", name); return SA_ERR_IO; }
            usleep(200);
        }
        if (sh->nranks != (uint32_t)nranks) { munmap(mem, bytes); sa_set_error("sa_comm_create_shm: %s was created for %u ranks", name, sh->nranks); return SA_ERR_INVALID; }
    }
    sh->attached.fetch_add(1);
    sa_comm* c = new (std::nothrow) sa_comm();
    if (!c) { munmap(mem, bytes); sa_set_error("host allocation failed"); return SA_ERR_OOM; }
    c->sh = sh; c->rank = rank; c->nranks = nranks; c->local = false; c->bytes = bytes; c->shm_name = name;
    *out = c;
    return SA_OK;
}

extern "C" void sa_comm_destroy(sa_comm* c)
{
    if (!c) return;
    if (c->local) {
        if (c->sh->attached.fetch_sub(1) == 1) free(c->sh);
    } else {
        const bool last = c->sh->attached.fetch_sub(1) == 1;
        munmap(c->sh, c->bytes);
        if (last || c->rank == 0) shm_unlink(c->shm_name.c_str());
    }
    delete c;
}

extern "C" int sa_comm_rank(const sa_comm* c) { return c ? c->rank : -1; }
extern "C" int sa_comm_size(const sa_comm* c) { return c ? c->nranks : 0; }

static void comm_abort(sa_comm* c) { c->sh->abort_flag.store(1, std::memory_order_release); }

static int comm_barrier(sa_comm* c)
{
    CommShared* sh = c->sh;
    if (sh->abort_flag.load(std::memory_order_acquire)) { sa_set_error("a peer rank failed (communicator aborted)"); return SA_ERR_INTERNAL; }
    if (c->nranks == 1) return SA_OK;
    const uint32_t g = sh->gen.load(std::memory_order_acquire);
    if (sh->arrive.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)c->nranks) {
        sh->arrive.store(0, std::memory_order_relaxed);
        sh->gen.store(g + 1, std::memory_order_release);
        return SA_OK;
    }
    const double t0 = now_s();
    uint32_t spins = 0;
    while (sh->gen.load(std::memory_order_acquire) == g) {
        if (sh->abort_flag.load(std::memory_order_acquire)) { sa_set_error("a peer rank failed (communicator aborted)"); return SA_ERR_INTERNAL; }
        if ((++spins & 1023u) == 0) {
            if (now_s() - t0 > COMM_TIMEOUT_S) { comm_abort(c); sa_set_error("barrier timed out after %.0f s", COMM_TIMEOUT_S); return SA_ERR_INTERNAL; }
            sched_yield();
        }
    }
    return SA_OK;
}

extern "C" int sa_comm_barrier(sa_comm* c)
{
    MG_REQUIRE(c, SA_ERR_INVALID, "NULL communicator");
    return comm_barrier(c);
}

// Every rank contributes `bytes` (<= MAILBOX_BYTES); *all receives a pointer to nranks consecutive
// mailboxes (stride MAILBOX_BYTES), valid until the rank's next-but-one collective.
static int comm_allgather(sa_comm* c, const void* mine, size_t bytes, const uint8_t** all)
{
    if (bytes > MAILBOX_BYTES) { sa_set_error("allgather payload too large"); return SA_ERR_INTERNAL; }
    const int slot = (int)(c->op++ & 1u);
    memcpy(c->box(slot, c->rank), mine, bytes);
    int rc = comm_barrier(c);
    if (rc != SA_OK) return rc;
    *all = c->box(slot, 0);
    return SA_OK;
}

// ------------------------------------------------------------------------------ kernels
namespace sa {

// ASCII lower-casing of text[off, off + len): whole 16-byte vectors inside the range by vector, the
// ragged head / tail bytewise (the neighbouring bytes belong to slices other ranks are writing).
__global__ void __launch_bounds__(256) lowercase_range_kernel(uint8_t* __restrict__ text, uint64_t off, uint64_t len)
{
    const uint64_t a = (off + 15) & ~15ull, b = (off + len) & ~15ull;       // vector body [a, b)
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    if (a < b) {
        uint4* v16 = reinterpret_cast<uint4*>(text + a);
        const uint64_t nvec = (b - a) >> 4;
        for (uint64_t i = tid; i < nvec; i += nthreads) {
            uint4 v = v16[i];
            uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t x = w[j], hi = x & 0x80808080u, lo = x & 0x7F7F7F7Fu;
                w[j] = x | (((lo + 0x3F3F3F3Fu) & ~(lo + 0x25252525u) & ~hi & 0x80808080u) >> 2);
            }
            v16[i] = v;
        }
    }
    if (tid < 32) {
        const uint64_t head_end = a < off + len ? a : off + len;
        for (uint64_t p = off + tid; p < head_end; p += 32) { const uint8_t ch = text[p]; if (ch >= 'A' && ch <= 'Z') text[p] = ch + 32; }
        const uint64_t tail_begin = b > head_end ? b : head_end;
        for (uint64_t p = tail_begin + tid; p < off + len; p += 32) { const uint8_t ch = text[p]; if (ch >= 'A' && ch <= 'Z') text[p] = ch + 32; }
    }
}

struct PushTable {
    uint8_t* dst[15];
    int n;
};

// src[0, count) -> dst[g][0, count) for every peer: one read of local HBM, 16-byte peer stores (all
// buffers share the alignment of the offset), ragged head / tail bytewise.
__global__ void __launch_bounds__(512) push_bytes_kernel(const uint8_t* __restrict__ src, uint64_t count, PushTable t)
{
    const uint64_t head = min(count, (uint64_t)((16u - ((uintptr_t)src & 15u)) & 15u));
    const uint64_t nvec = (count - head) >> 4;
    const uint64_t tail0 = head + (nvec << 4);
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint4* sv = reinterpret_cast<const uint4*>(src + head);
    for (uint64_t i = tid; i < nvec; i += nthreads) {
        const uint4 v = __ldcs(sv + i);
#pragma unroll 1
        for (int g = 0; g < t.n; ++g) reinterpret_cast<uint4*>(t.dst[g] + head)[i] = v;
    }
    for (uint64_t i = tid; i < head + (count - tail0); i += nthreads) {
        const uint64_t j = i < head ? i : tail0 + (i - head);
        const uint8_t v = src[j];
        for (int g = 0; g < t.n; ++g) t.dst[g][j] = v;
    }
}

// LCP (bytes, capped at lprime) of the masked prefixes of suffixes p and q as adjacent records have it:
// equal through a terminator ('\n' / 0xFF) = tied for good (lprime).  Bytes past n are zero.
__global__ void junction_lcp_kernel(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ pairs,
                                    uint32_t npairs, uint32_t lprime, uint32_t* __restrict__ out)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= npairs) return;
    const uint64_t p = pairs[2 * t], q = pairs[2 * t + 1];
    uint32_t r = lprime;
    for (uint32_t i = 0; i < lprime; ++i) {
        const uint8_t a = p + i < n ? text[p + i] : 0, b = q + i < n ? text[q + i] : 0;
        if (a != b) { r = i; break; }
        if (a == 10 || a == 255) break;
    }
    out[t] = r;
}

}  // namespace sa

// Cost-weighted splitters: how many of the slice's sampled keys (the sample hist16_kernel takes) have a twin --
// another sampled key with the same 64 bits -- per top-16-bit bin.  Records whose whole first-sort key repeats
// are the ones the refinement rounds work on, and they cost ~2.7x a record that is placed by the first sort
// (profiles/r02_multi_gpu_phases.md), so key ranges with many of them get fewer records.  Two passes over the
// sample with a table of 2^24 counters addressed by a hash of the key (9 % of the slots taken at 1.5 M samples:
// a false twin now and then only adds a near-uniform offset).
namespace sa {
constexpr int TWIN_TABLE_BITS = 24;
__device__ __forceinline__ uint32_t twin_slot(uint64_t key) { return (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> (64 - TWIN_TABLE_BITS)); }

__global__ void twin_count_kernel(const uint64_t* __restrict__ keys, uint32_t m, uint32_t stride_keys, uint32_t* __restrict__ table)
{
    const uint32_t ms = (m + stride_keys - 1) / stride_keys;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ms; i += gridDim.x * blockDim.x)
        atomicAdd(table + twin_slot(__ldg(keys + (size_t)i * stride_keys)), 1u);
}

__global__ void twin_hist_kernel(const uint64_t* __restrict__ keys, uint32_t m, uint32_t stride_keys,
                                 const uint32_t* __restrict__ table, uint32_t* __restrict__ hist)
{
    const uint32_t ms = (m + stride_keys - 1) / stride_keys;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ms; i += gridDim.x * blockDim.x) {
        const uint64_t k = __ldg(keys + (size_t)i * stride_keys);
        if (table[twin_slot(k)] >= 2u) atomicAdd(hist + (uint32_t)(k >> 48), 1u);
    }
}
}  // namespace sa

// ------------------------------------------------------------------------------ peer buffers
namespace {

enum PeerBuf { PB_TEXT = 0, PB_FINAL, PB_NLBITS, PB_NLRANK, PB_KEYS, PB_VALS, PB_COUNT };

struct Export {                    // what a rank publishes about one buffer
    uint8_t handle[64];
};

void export_buffer(sa_index* idx, bool local, void* p, uint8_t out[64])
{
    memset(out, 0, 64);
    if (local || idx->dbg_same_process) {
        memcpy(out, "SAMEPROC", 8);
        memcpy(out + 8, &p, sizeof p);
    } else {
        cudaIpcMemHandle_t h;
        if (p && cudaIpcGetMemHandle(&h, p) == cudaSuccess) memcpy(out, &h, 64);
    }
}

int map_peer(sa_index::PeerMapping& pm, const uint8_t h[64], void** out)
{
    if (memcmp(h, "SAMEPROC", 8) == 0) {
        memcpy(out, h + 8, sizeof(void*));
        return SA_OK;
    }
    if (!pm.mapped || memcmp(&pm.h, h, 64) != 0) {
        if (pm.mapped) { cudaIpcCloseMemHandle(pm.ptr); pm.mapped = false; }
        memcpy(&pm.h, h, 64);
        SA_CU(cudaIpcOpenMemHandle(&pm.ptr, pm.h, cudaIpcMemLazyEnablePeerAccess));
        pm.mapped = true;
    }
    *out = pm.ptr;
    return SA_OK;
}

void unmap_peer(sa_index::PeerMapping& pm)
{
    if (pm.mapped) { cudaIpcCloseMemHandle(pm.ptr); pm.mapped = false; pm.ptr = nullptr; }
}

// Before any rank re-allocates a buffer its peers have mapped, they close their mappings (freeing
// exported memory that importers still hold open is undefined).  grow[k] = this rank is about to.
int announce_growth(sa_index* idx, sa_comm* c, const uint8_t grow[PB_COUNT])
{
    const uint8_t* all = nullptr;
    int rc = comm_allgather(c, grow, PB_COUNT, &all);
    if (rc != SA_OK) return rc;
    bool any = false;
    for (int g = 0; g < c->nranks; ++g) {
        const uint8_t* f = all + (size_t)g * MAILBOX_BYTES;
        for (int k = 0; k < PB_COUNT; ++k) {
            if (!f[k]) continue;
            any = true;
            if (g == c->rank) continue;
            if (k == PB_TEXT) unmap_peer(idx->peer_text[g]);
            else if (k == PB_NLBITS) unmap_peer(idx->peer_nlbits[g]);
            else if (k == PB_NLRANK) unmap_peer(idx->peer_nlrank[g]);
            else if (k == PB_FINAL) { if (idx->peer_final_mapped[g]) { cudaIpcCloseMemHandle(idx->peer_final[g]); idx->peer_final_mapped[g] = false; } }
            else if (idx->peer_mapped[g]) {       // keys + positions are mapped as a pair
                cudaIpcCloseMemHandle(idx->peer_keys[g]);
                cudaIpcCloseMemHandle(idx->peer_pos[g]);
                idx->peer_mapped[g] = false;
            }
        }
    }
    return any ? comm_barrier(c) : SA_OK;
}

bool must_grow(const DevBuf& b, size_t bytes) { return !b.p || b.cap < bytes; }

struct Lap {
    sa_index* idx;
    float* out;
    double t0;
    int k = 0;
    Lap(sa_index* i, float* o) : idx(i), out(o), t0(now_s()) {}
    void mark()
    {
        const double t = now_s();
        if (out && k < 16) out[k] = (float)((t - t0) * 1e3);
        ++k;
        t0 = t;
    }
};

constexpr uint32_t EDGE = 128;

struct EdgeBlock {                 // what each rank publishes about its sorted shard
    uint64_t m;
    uint32_t k;
    uint32_t pad;
    uint32_t head_pos[EDGE], tail_pos[EDGE];
    uint16_t head_lcp[EDGE], tail_lcp[EDGE];
};

}  // namespace

// ------------------------------------------------------------------------------ the build, one rank
// text_slice: slice_len bytes = text[slice_off, slice_off + slice_len) (host pointer, or a device pointer
// with SA_FLAG_DEVICE_TEXT); the slices of all ranks tile [0, n) in rank order.  phase_ms (optional,
// 8 floats): text, prepare, pack+hist, split, exchange, sort+refine, edges, finish+replicate (host
// wall time of this rank, barriers included).
extern "C" int sa_shard_build(sa_index* idx, sa_comm* c, const uint8_t* text_slice, uint64_t slice_off, uint64_t slice_len,
                              uint64_t n, uint32_t flags, float* phase_ms)
{
    MG_REQUIRE(idx && c, SA_ERR_INVALID, "sa_shard_build: NULL argument");
    MG_REQUIRE(n > 0 && n < 0xFFFFFFFFull - 1024ull, SA_ERR_LIMIT, "sa_shard_build: n out of range");
    MG_REQUIRE(slice_off <= n && slice_len <= n - slice_off && (text_slice || slice_len == 0), SA_ERR_INVALID,
               "sa_shard_build: bad slice");
    const int R = c->nranks, me = c->rank;
    int rc = SA_OK;
    auto fail = [&](int code) { comm_abort(c); return code; };
#define MG_TRY(expr)                                 \
    do {                                             \
        rc = (expr);                                 \
        if (rc != SA_OK) return fail(rc);            \
    } while (0)
#define MG_CU(call)                                                          \
    do {                                                                     \
        cudaError_t e__ = (call);                                            \
        if (e__ != cudaSuccess) return fail(sa_cuda_fail(e__, #call));       \
    } while (0)

    if (cudaSetDevice(idx->device) != cudaSuccess) return fail(sa_cuda_fail(cudaGetLastError(), "cudaSetDevice"));
    cudaStream_t s = idx->stream;
    Lap lap(idx, phase_ms);
    if (c->local) idx->dbg_same_process = 1;
    idx->shard_no_replicate = (flags & SA_FLAG_SHARDED_SA) != 0 && R > 1;
    idx->sharded_sa = false;

    // ---- state of a fresh sharded build (what sa_shard_begin sets)
    {
        std::unique_lock<std::shared_mutex> lock(idx->mu);
        idx->timing = (flags & SA_FLAG_TIMING) != 0;
        idx->n = n; idx->tail = 0;
        idx->built = false;
        sa_invalidate_results(idx);
        idx->csv = false;
        idx->shard_prepared = false;
        idx->shard_csv = false;
        memset(&idx->stats, 0, sizeof idx->stats);
        idx->lprime = idx->L < n ? idx->L : (uint32_t)n;
        idx->stats.n = n; idx->stats.max_suffix_length = idx->L; idx->stats.lprime = idx->lprime;
    }
    if (idx->lprime == 0) { sa_set_error("sa_shard_build: max_suffix_length 0 needs no sharded build"); return fail(SA_ERR_INVALID); }
    if (!idx->h_big) MG_CU(cudaMallocHost(reinterpret_cast<void**>(&idx->h_big), MAILBOX_BYTES));

    // ---- buffers every peer stores into: text, final array, newline rank structure
    const size_t text_bytes = ((size_t)n + SA_TEXT_PAD + 255) & ~(size_t)255;
    const size_t nlwords = ((size_t)n >> 6) + 1;
    {
        uint8_t grow[PB_COUNT] = {};
        grow[PB_TEXT] = must_grow(idx->text, text_bytes);
        grow[PB_FINAL] = must_grow(idx->sa_final, (size_t)n * 4);
        grow[PB_NLBITS] = must_grow(idx->nlbits, nlwords * 8);
        grow[PB_NLRANK] = must_grow(idx->nlrank, nlwords * 4);
        MG_TRY(announce_growth(idx, c, grow));
        MG_CU(idx->text.ensure(text_bytes));
        MG_CU(idx->sa_final.ensure((size_t)n * 4));
        MG_CU(idx->nlbits.ensure(nlwords * 8));
        MG_CU(idx->nlrank.ensure(nlwords * 4));
    }
    uint8_t* peer_text[16] = {};
    unsigned long long* peer_nlbits[16] = {};
    uint32_t* peer_nlrank[16] = {};
    uint8_t final_handles[16 * 64];
    uint64_t slice_lo[16] = {}, slice_hi[16] = {};       // every rank's slice of the text
    {
        Export ex[5];
        export_buffer(idx, c->local, idx->text.p, ex[0].handle);
        export_buffer(idx, c->local, idx->sa_final.p, ex[1].handle);
        export_buffer(idx, c->local, idx->nlbits.p, ex[2].handle);
        export_buffer(idx, c->local, idx->nlrank.p, ex[3].handle);
        memset(ex[4].handle, 0, 64);
        memcpy(ex[4].handle, &slice_off, 8);
        memcpy(ex[4].handle + 8, &slice_len, 8);
        const uint8_t* all = nullptr;
        MG_TRY(comm_allgather(c, ex, sizeof ex, &all));
        for (int g = 0; g < R; ++g) {
            const Export* e = reinterpret_cast<const Export*>(all + (size_t)g * MAILBOX_BYTES);
            memcpy(final_handles + (size_t)g * 64, e[1].handle, 64);
            uint64_t len = 0;
            memcpy(&slice_lo[g], e[4].handle, 8);
            memcpy(&len, e[4].handle + 8, 8);
            slice_hi[g] = slice_lo[g] + len;
            if (g == me) {
                peer_text[g] = idx->text.as<uint8_t>();
                peer_nlbits[g] = idx->nlbits.as<unsigned long long>();
                peer_nlrank[g] = idx->nlrank.as<uint32_t>();
                continue;
            }
            void* p = nullptr;
            MG_TRY(map_peer(idx->peer_text[g], e[0].handle, &p));
            peer_text[g] = static_cast<uint8_t*>(p);
            MG_TRY(map_peer(idx->peer_nlbits[g], e[2].handle, &p));
            peer_nlbits[g] = static_cast<unsigned long long*>(p);
            MG_TRY(map_peer(idx->peer_nlrank[g], e[3].handle, &p));
            peer_nlrank[g] = static_cast<uint32_t*>(p);
        }
    }
    if (!idx->push_stream) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        MG_CU(cudaStreamCreateWithPriority(&idx->push_stream, cudaStreamNonBlocking, hi));
    }

    // ---- text: own slice up, lower-cased, pushed to every peer
    uint8_t* d_text = idx->text.as<uint8_t>();
    {
        const size_t pad_from = (size_t)n & ~(size_t)15;
        // the bytes [pad_from, n) belong to the last slice: zero first, then the slice lands on top
        MG_CU(cudaMemsetAsync(d_text + pad_from, 0, text_bytes - pad_from, s));
        if (slice_len) {
            ScopedPhase ph(idx, &idx->stats.ms_h2d);
            MG_CU(cudaMemcpyAsync(d_text + slice_off, text_slice, slice_len,
                                  (flags & SA_FLAG_DEVICE_TEXT) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
        }
        if ((flags & SA_FLAG_LOWERCASE) && slice_len) {
            sa::lowercase_range_kernel<<<idx->num_sms * 4, 256, 0, s>>>(d_text, slice_off, slice_len);
            MG_CU(cudaGetLastError());
        }
        // peers zero their own padding before anything of mine can land in it: the tail bytes [pad_from, n)
        // of the LAST slice are the only overlap, and they are pushed after the barrier below
        MG_CU(cudaStreamSynchronize(s));
        MG_TRY(comm_barrier(c));          // every rank's own slice sits in its own buffer
    }
    // This rank prepares and packs the positions [b_lo, b_hi) (4096-aligned bounds; they need not coincide with
    // the slice it uploaded) and reads up to 64 bytes beyond them.
    auto bound = [&](int g) -> uint64_t { return g >= R ? n : (g <= 0 ? 0 : (((uint64_t)g * n / R) & ~4095ull)); };
    const uint64_t b_lo = bound(me), b_hi = bound(me + 1);
    const uint64_t own_hi = (std::min<uint64_t>(text_bytes, b_hi + 64) + 15) & ~15ull;    // end of the early encode
    const bool async_text = R > 1 && idx->dbg_text_push == 0;
    if (R > 1) {
        if (async_text) {
            // 1. pull the part of [b_lo, b_hi + 64) other ranks uploaded straight from their buffers (a few KB when
            //    the slices follow the bounds); 2. send this slice to every peer with the copy engines, beside the
            //    prepare / pack / split / exchange kernels: the whole text is only needed by the refinement rounds.
            const uint64_t need_hi = std::min<uint64_t>(n, b_hi + 64);
            for (int g = 0; g < R; ++g) {
                if (g == me) continue;
                const uint64_t lo = std::max(b_lo, slice_lo[g]), hi = std::min(need_hi, slice_hi[g]);
                if (lo < hi) MG_CU(cudaMemcpyAsync(d_text + lo, peer_text[g] + lo, hi - lo, cudaMemcpyDeviceToDevice, s));
            }
            if (slice_len) {
                int lo = 0, hi = 0;
                cudaDeviceGetStreamPriorityRange(&lo, &hi);
                for (int k = 1; k < R; ++k) {
                    const int g = (me + k) % R;
                    if (!idx->copy_streams[g]) MG_CU(cudaStreamCreateWithPriority(&idx->copy_streams[g], cudaStreamNonBlocking, hi));
                    MG_CU(cudaMemcpyAsync(peer_text[g] + slice_off, d_text + slice_off, slice_len, cudaMemcpyDeviceToDevice,
                                          idx->copy_streams[g]));
                }
            }
        } else {
            if (slice_len) {
                sa::PushTable pt{};
                for (int k = 1; k < R; ++k) pt.dst[pt.n++] = peer_text[(me + k) % R] + slice_off;
                sa::push_bytes_kernel<<<idx->num_sms * 2, 512, 0, s>>>(d_text + slice_off, slice_len, pt);
                MG_CU(cudaGetLastError());
                MG_CU(cudaStreamSynchronize(s));
            }
            MG_TRY(comm_barrier(c));          // the whole text is on every GPU
        }
    }
    lap.mark();

    // ---- prepare: newline table + byte presence on this rank's slice of POSITIONS (4096-aligned)
    const uint32_t m_local = (uint32_t)(b_hi - b_lo);
    uint32_t* sm = idx->small.as<uint32_t>();
    MG_CU(cudaMemsetAsync(sm + SM_ERR, 0, 4 * sizeof(uint32_t), s));
    uint32_t nl_before = 0, nl_total = 0;
    {
        const uint32_t tiles = sa_newline_tiles(m_local);
        MG_CU(idx->nl_counts.ensure((size_t)(tiles ? tiles : 1) * sizeof(uint32_t)));
        struct Small { uint32_t nl; uint32_t present[256]; } mine{};
        if (m_local) {
            MG_CU(sa_launch_newline_count(d_text + b_lo, m_local, idx->nl_counts.as<uint32_t>(), s));
            MG_CU(sa_launch_scan_u32(idx->nl_counts.as<uint32_t>(), tiles, sm + SM_NLTOTAL, s));
            MG_CU(cudaMemcpyAsync(idx->h_small + 8, sm + SM_NLTOTAL, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
            MG_CU(sa_launch_byte_hist(d_text + b_lo, m_local, sm + SM_BYTEHIST, idx->num_sms, s));
            MG_CU(cudaMemcpyAsync(idx->h_small + 256, sm + SM_BYTEHIST, 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
            MG_CU(cudaStreamSynchronize(s));
            mine.nl = idx->h_small[8];
            memcpy(mine.present, idx->h_small + 256, sizeof mine.present);
        }
        const uint8_t* all = nullptr;
        MG_TRY(comm_allgather(c, &mine, sizeof mine, &all));
        uint32_t present[256] = {};
        for (int g = 0; g < R; ++g) {
            const Small* o = reinterpret_cast<const Small*>(all + (size_t)g * MAILBOX_BYTES);
            if (g < me) nl_before += o->nl;
            nl_total += o->nl;
            for (int b = 0; b < 256; ++b) present[b] |= o->present[b];
        }
        if (m_local) {
            // rank structure of the slice, written into the local arrays at the slice's place ...
            MG_CU(sa_launch_newline_write(d_text + b_lo, m_local, idx->nl_counts.as<uint32_t>(), nullptr,
                                          idx->nlbits.as<unsigned long long>() + (b_lo >> 6), idx->nlrank.as<uint32_t>() + (b_lo >> 6),
                                          s, nl_before, (uint32_t)b_lo));
            // ... and pushed to the peers (n/8 + n/16 bytes in total over all ranks)
            if (R > 1) {
                const uint64_t w0 = b_lo >> 6, w1 = me == R - 1 ? (uint64_t)nlwords : (b_hi >> 6);
                sa::PushTable pb{}, pr{};
                for (int k = 1; k < R; ++k) {
                    const int g = (me + k) % R;
                    pb.dst[pb.n++] = reinterpret_cast<uint8_t*>(peer_nlbits[g] + w0);
                    pr.dst[pr.n++] = reinterpret_cast<uint8_t*>(peer_nlrank[g] + w0);
                }
                sa::push_bytes_kernel<<<idx->num_sms, 512, 0, s>>>(reinterpret_cast<const uint8_t*>(idx->nlbits.as<unsigned long long>() + w0),
                                                                   (w1 - w0) * 8, pb);
                sa::push_bytes_kernel<<<idx->num_sms, 512, 0, s>>>(reinterpret_cast<const uint8_t*>(idx->nlrank.as<uint32_t>() + w0),
                                                                   (w1 - w0) * 4, pr);
                MG_CU(cudaGetLastError());
            }
        }
        {
            std::unique_lock<std::shared_mutex> lock(idx->mu);
            idx->num_nl = nl_total;
            idx->stats.num_docs = (uint64_t)nl_total + 1;
            // alphabet codes, locally (one pass over HBM; cheaper than receiving them): the whole text, or -- while
            // the peers' slices are still on their way -- this rank's range now and the rest before the sort
            rc = async_text ? sa_stage_alphabet(idx, present, b_lo, own_hi) : sa_stage_alphabet(idx, present);
        }
        if (rc != SA_OK) return fail(rc);
        MG_CU(cudaStreamSynchronize(s));
        MG_TRY(comm_barrier(c));          // every slice of the rank structure has landed
        MG_CU(idx->nlpos.ensure((size_t)(nl_total ? nl_total : 1) * sizeof(uint32_t)));
        MG_CU(sa_launch_rank_positions(idx->nlbits.as<unsigned long long>(), idx->nlrank.as<uint32_t>(), nlwords,
                                       idx->nlpos.as<uint32_t>(), s));
        idx->shard_prepared = true;
    }
    lap.mark();

    // ---- pack: round-0 keys of the slice + top-16-bit histogram; sum over the ranks
    MG_CU(idx->pack_keys.ensure((size_t)(m_local ? m_local : 1) * 8));
    MG_CU(idx->pack_pos.ensure((size_t)(m_local ? m_local : 1) * 4));
    MG_CU(idx->hist16.ensure(2 * 65536 * 4));                 // sampled keys per bin | sampled keys with a twin per bin
    MG_CU(cudaMemsetAsync(idx->hist16.p, 0, 2 * 65536 * 4, s));
    MG_TRY(sa_shard_pack(idx, b_lo, b_hi, idx->pack_keys.as<uint64_t>(), idx->pack_pos.as<uint32_t>(), idx->hist16.as<uint32_t>()));
    // sa_debug_set("split_gamma_x100", g): a record whose key repeats counts 1 + g / 100 (default 1.7, profiles/r02_multi_gpu_phases.md); 0 = equal record counts
    const uint64_t gamma100 = idx->dbg_split_gamma_x100 >= 0 ? (uint64_t)idx->dbg_split_gamma_x100 : 170u;
    if (gamma100 && m_local && R > 1) {
        const size_t tbytes = ((size_t)4 << sa::TWIN_TABLE_BITS);
        MG_CU(idx->aux[1].ensure(tbytes));
        MG_CU(cudaMemsetAsync(idx->aux[1].p, 0, tbytes, s));
        const uint32_t stride_keys = m_local >= (16u << 20) ? 64u : 16u;       // the sample of hist16_kernel (sa_shard_pack)
        sa::twin_count_kernel<<<idx->num_sms * 8, 256, 0, s>>>(idx->pack_keys.as<uint64_t>(), m_local, stride_keys, idx->aux[1].as<uint32_t>());
        sa::twin_hist_kernel<<<idx->num_sms * 8, 256, 0, s>>>(idx->pack_keys.as<uint64_t>(), m_local, stride_keys, idx->aux[1].as<uint32_t>(),
                                                             idx->hist16.as<uint32_t>() + 65536);
        MG_CU(cudaGetLastError());
        idx->stats.kernel_launches += 2;
    }
    std::vector<uint8_t> bin2rank(65536);
    {
        MG_CU(cudaMemcpyAsync(idx->h_big, idx->hist16.p, 2 * 65536 * 4, cudaMemcpyDeviceToHost, s));
        MG_CU(cudaStreamSynchronize(s));
        const uint8_t* all = nullptr;
        MG_TRY(comm_allgather(c, idx->h_big, 2 * 65536 * 4, &all));
        // per bin: records and records with a twin (sample counts x the rank's sample stride; both stay below n < 2^32,
        // so the sums over the ranks run in 32 bits -- the compiler vectorises these loops, the 64-bit form cost 0.4 ms)
        std::vector<uint32_t> acc(2 * 65536, 0);
        for (int g = 0; g < R; ++g) {
            const uint32_t* __restrict__ h = reinterpret_cast<const uint32_t*>(all + (size_t)g * MAILBOX_BYTES);
            const uint32_t w = (bound(g + 1) - bound(g)) >= (16u << 20) ? 64u : 16u;      // the rank's sample stride (sa_shard_pack)
            uint32_t* __restrict__ a = acc.data();
            for (int b = 0; b < 2 * 65536; ++b) a[b] += w * h[b];
        }
        // bins in order; rank g gets the bins whose first record falls into the g-th equal share of the cost
        // (in 1/100 records: 100 per record + gamma100 per record with a twin)
        const uint32_t* cntv = acc.data();
        const uint32_t* twv = acc.data() + 65536;
        uint64_t total = 0, before = 0, r = 0;
        for (int b = 0; b < 65536; ++b) total += 100ull * cntv[b] + gamma100 * twv[b];
        for (int b = 0; b < 65536; ++b) {
            // r = min(R - 1, floor(before * R / total)), stepped instead of divided (before only grows)
            while (r + 1 < (uint64_t)R && (r + 1) * total <= before * (uint64_t)R) ++r;
            bin2rank[b] = (uint8_t)r;
            before += 100ull * cntv[b] + gamma100 * twv[b];
        }
    }
    lap.mark();

    // ---- split: counts per destination, receive buffers sized and exported
    uint64_t counts[16] = {}, dst_off[16] = {}, m_recv = 0;
    std::vector<uint8_t> handles((size_t)R * 128, 0);
    {
        MG_TRY(sa_shard_count(idx, idx->pack_keys.as<uint64_t>(), m_local, bin2rank.data(), R, counts));
        const uint8_t* all = nullptr;
        MG_TRY(comm_allgather(c, counts, sizeof(uint64_t) * 16, &all));
        for (int src = 0; src < R; ++src) {
            const uint64_t* cs = reinterpret_cast<const uint64_t*>(all + (size_t)src * MAILBOX_BYTES);
            m_recv += cs[me];
            if (src < me)
                for (int d = 0; d < R; ++d) dst_off[d] += cs[d];
        }
        if (m_recv > SA_MAX_INDEX_BYTES) { sa_set_error("sa_shard_build: %llu records for one GPU exceed the 2^31 limit", (unsigned long long)m_recv); return fail(SA_ERR_LIMIT); }
        uint8_t grow[PB_COUNT] = {};
        const size_t mr = (size_t)(m_recv ? m_recv : 1);
        grow[PB_KEYS] = must_grow(idx->keys[0], mr * 8);
        grow[PB_VALS] = must_grow(idx->vals[0], mr * 4);
        MG_TRY(announce_growth(idx, c, grow));
        uint8_t mine[128];
        MG_TRY(sa_shard_recv_prepare(idx, m_recv, mine, mine + 64));
        MG_TRY(comm_allgather(c, mine, 128, &all));
        for (int g = 0; g < R; ++g) memcpy(handles.data() + (size_t)g * 128, all + (size_t)g * MAILBOX_BYTES, 128);
    }
    lap.mark();

    // ---- exchange: every record straight into its owner's sort buffers
    MG_TRY(sa_shard_scatter_p2p(idx, idx->pack_keys.as<uint64_t>(), idx->pack_pos.as<uint32_t>(), m_local, R, me, handles.data(),
                                dst_off));
    if (async_text)
        for (int g = 0; g < R; ++g)
            if (g != me && idx->copy_streams[g]) MG_CU(cudaStreamSynchronize(idx->copy_streams[g]));   // my slice has left
    MG_TRY(comm_barrier(c));                  // all peer stores have landed (records, and every slice of the text)
    if (async_text) {
        MG_TRY(sa_stage_encode_range(idx, 0, b_lo));
        MG_TRY(sa_stage_encode_range(idx, own_hi, text_bytes));
    }
    lap.mark();

    // ---- sort + refine the received key range
    MG_TRY(sa_shard_sort_recv(idx, m_recv));
    lap.mark();

    // ---- edges: first / last records of every shard; halos for the bucket finish
    std::vector<uint32_t> left_pos, right_pos;
    std::vector<uint16_t> left_lcp, right_lcp;
    uint32_t lcp_first = 0;
    uint64_t global_offset = 0;
    uint64_t shard_ms[16] = {};
    {
        EdgeBlock eb{};
        eb.m = m_recv;
        eb.k = (uint32_t)(m_recv < EDGE ? m_recv : EDGE);
        if (eb.k) MG_TRY(sa_shard_edges(idx, EDGE, eb.head_pos, eb.head_lcp, eb.tail_pos, eb.tail_lcp));
        const uint8_t* all = nullptr;
        MG_TRY(comm_allgather(c, &eb, sizeof eb, &all));
        auto E = [&](int g) { return reinterpret_cast<const EdgeBlock*>(all + (size_t)g * MAILBOX_BYTES); };
        for (int g = 0; g < me; ++g) global_offset += E(g)->m;
        for (int g = 0; g < R; ++g) shard_ms[g] = E(g)->m;
        if (m_recv) {
            // which records form the halos (index logic of sharded.py assemble_halos), and the pairs whose
            // LCP across a shard boundary has to be computed from the text
            std::vector<uint32_t> pairs;            // (p, q) ...
            std::vector<int> fix_left, fix_right;   // where each pair's LCP goes: index into left_lcp / right_lcp, -1 = lcp_first
            const EdgeBlock* mine = E(me);
            uint32_t need = EDGE, nxt_first = mine->head_pos[0];
            std::vector<std::vector<uint32_t>> lp;  // chunks, nearest first
            std::vector<std::vector<uint16_t>> ll;
            std::vector<int> chunk_pair;            // pair index whose LCP belongs to the junction right of chunk i
            for (int g = me - 1; g >= 0 && need > 0; --g) {
                const EdgeBlock* e = E(g);
                const uint32_t k = need < e->k ? need : e->k;
                if (k == 0) continue;
                lp.emplace_back(e->tail_pos + (e->k - k), e->tail_pos + e->k);
                ll.emplace_back(e->tail_lcp + (e->k - k), e->tail_lcp + e->k);
                pairs.push_back(e->tail_pos[e->k - 1]);
                pairs.push_back(nxt_first);
                chunk_pair.push_back((int)pairs.size() / 2 - 1);
                nxt_first = e->tail_pos[e->k - k];
                need -= k;
                if (k < e->m) break;              // this rank still has records in front: the halo is full
            }
            need = EDGE;
            uint32_t prev_last = mine->tail_pos[mine->k - 1];
            std::vector<int> right_pair_at;       // index in right_lcp that receives pair i's LCP
            std::vector<int> right_pair_id;
            for (int g = me + 1; g < R && need > 0; ++g) {
                const EdgeBlock* e = E(g);
                const uint32_t k = need < e->k ? need : e->k;
                if (k == 0) continue;
                right_pair_at.push_back((int)right_pos.size());
                pairs.push_back(prev_last);
                pairs.push_back(e->head_pos[0]);
                right_pair_id.push_back((int)pairs.size() / 2 - 1);
                right_pos.insert(right_pos.end(), e->head_pos, e->head_pos + k);
                right_lcp.insert(right_lcp.end(), e->head_lcp, e->head_lcp + k);
                prev_last = e->head_pos[k - 1];
                need -= k;
                if (k < e->m) break;
            }
            // junction LCPs on the device (the host holds only this rank's slice of the text)
            std::vector<uint32_t> jl(pairs.size() / 2, 0);
            if (!jl.empty()) {
                const uint32_t np = (uint32_t)jl.size();
                MG_CU(idx->aux[0].ensure((size_t)np * 12 + 64));
                uint32_t* d_pairs = idx->aux[0].as<uint32_t>();
                uint32_t* d_out = d_pairs + 2 * np;
                MG_CU(cudaMemcpyAsync(d_pairs, pairs.data(), (size_t)np * 8, cudaMemcpyHostToDevice, s));
                sa::junction_lcp_kernel<<<(np + 63) / 64, 64, 0, s>>>(d_text, n, d_pairs, np, idx->lprime, d_out);
                MG_CU(cudaGetLastError());
                MG_CU(cudaMemcpyAsync(jl.data(), d_out, (size_t)np * 4, cudaMemcpyDeviceToHost, s));
                MG_CU(cudaStreamSynchronize(s));
            }
            // left halo in global order: farthest chunk first; the junction right of chunk i is lcp_first for the
            // nearest chunk, else the first LCP of the chunk that follows it
            for (int i = (int)lp.size() - 1; i >= 0; --i) {
                left_pos.insert(left_pos.end(), lp[i].begin(), lp[i].end());
                left_lcp.insert(left_lcp.end(), ll[i].begin(), ll[i].end());
            }
            {
                size_t at = left_pos.size();
                for (size_t i = 0; i < lp.size(); ++i) {      // nearest first
                    const uint32_t v = jl[(size_t)chunk_pair[i]];
                    if (i == 0) lcp_first = v;
                    else left_lcp[at] = (uint16_t)v;          // first record of the chunk collected before (nearer)
                    at -= lp[i].size();
                }
            }
            for (size_t i = 0; i < right_pair_at.size(); ++i) right_lcp[(size_t)right_pair_at[i]] = (uint16_t)jl[(size_t)right_pair_id[i]];
        }
    }
    lap.mark();

    // ---- finish + replicate, then everyone's stores have landed
    MG_TRY(sa_shard_finish_bcast(idx, left_pos.data(), left_lcp.data(), (uint32_t)left_pos.size(), lcp_first, right_pos.data(),
                                 right_lcp.data(), (uint32_t)right_pos.size(), R, me, final_handles, global_offset));
    MG_TRY(comm_barrier(c));
    if (idx->shard_no_replicate) {
        // the array stays distributed: rank g's key range lives in rank g's final array
        sa::SaView v{};
        v.n = R;
        uint64_t at = 0;
        for (int g = 0; g < R; ++g) {
            v.base[g] = idx->shard_peer_base[g];
            v.bound[g] = (uint32_t)at;
            at += shard_ms[g];
        }
        for (int g = R; g <= 16; ++g) v.bound[g] = (uint32_t)n;
        idx->view = v;
        idx->sharded_sa = true;
        idx->shard_no_replicate = false;
    }
    MG_TRY(sa_shard_mark_built(idx));
    idx->shard_m_last = m_recv;
    idx->shard_exchange_bytes = 0;
    for (int d = 0; d < R; ++d)
        if (d != me) idx->shard_exchange_bytes += counts[d] * 12;
    lap.mark();
#undef MG_TRY
#undef MG_CU
    return SA_OK;
}

extern "C" int sa_shard_info(const sa_index* idx, uint64_t* shard_records, uint64_t* exchange_bytes_sent)
{
    MG_REQUIRE(idx, SA_ERR_INVALID, "NULL argument");
    if (shard_records) *shard_records = idx->shard_m_last;
    if (exchange_bytes_sent) *exchange_bytes_sent = idx->shard_exchange_bytes;
    return SA_OK;
}

// ------------------------------------------------------------------------------ all GPUs of one process
// One index over `ndev` GPUs driven from ONE process: a thread per handle, peer access between the
// devices, raw device pointers instead of IPC handles.  Every handle ends up with the text and the
// complete array (a replica that answers queries).  Handles may also share a device (the test tier
// runs the sharded pipeline on one GPU this way).
extern "C" int sa_build_text_multi(sa_index* const* idxs, int ndev, const uint8_t* text, uint64_t n, uint32_t flags)
{
    MG_REQUIRE(idxs && ndev >= 1 && ndev <= 16 && (text || n == 0), SA_ERR_INVALID, "sa_build_text_multi: bad argument");
    for (int g = 0; g < ndev; ++g) MG_REQUIRE(idxs[g], SA_ERR_INVALID, "sa_build_text_multi: NULL handle %d", g);
    if (ndev == 1 || n == 0) {                // nothing to shard: every handle builds the text itself
        for (int g = 0; g < ndev; ++g) {
            int rc = sa_build_text(idxs[g], text, n, flags);
            if (rc != SA_OK) return rc;
        }
        return SA_OK;
    }
    for (int a = 0; a < ndev; ++a) {
        for (int b = 0; b < ndev; ++b) {
            if (idxs[a]->device == idxs[b]->device) continue;
            int can = 0;
            SA_CU(cudaDeviceCanAccessPeer(&can, idxs[a]->device, idxs[b]->device));
            MG_REQUIRE(can, SA_ERR_CUDA, "GPU %d cannot access GPU %d (no peer access)", idxs[a]->device, idxs[b]->device);
            SA_CU(cudaSetDevice(idxs[a]->device));
            cudaError_t e = cudaDeviceEnablePeerAccess(idxs[b]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return sa_cuda_fail(e, "cudaDeviceEnablePeerAccess");
            cudaGetLastError();
        }
    }
    std::vector<sa_comm*> comms((size_t)ndev, nullptr);
    int rc = sa_comm_create_local(ndev, comms.data());
    if (rc != SA_OK) return rc;
    std::vector<int> rcs((size_t)ndev, SA_OK);
    std::vector<std::string> errs((size_t)ndev);
    std::vector<std::thread> th;
    for (int g = 0; g < ndev; ++g) {
        th.emplace_back([&, g] {
            const uint64_t a = (uint64_t)g * n / ndev, b = (uint64_t)(g + 1) * n / ndev;
            rcs[(size_t)g] = sa_shard_build(idxs[g], comms[(size_t)g], text + a, a, b - a, n, flags, nullptr);
            if (rcs[(size_t)g] != SA_OK) errs[(size_t)g] = sa_last_error();
        });
    }
    for (auto& t : th) t.join();
    for (sa_comm* c : comms) sa_comm_destroy(c);
    for (int g = 0; g < ndev; ++g)
        if (rcs[(size_t)g] != SA_OK && errs[(size_t)g].find("communicator aborted") == std::string::npos) {
            sa_set_error("rank %d: %s", g, errs[(size_t)g].c_str());
            return rcs[(size_t)g];
        }
    for (int g = 0; g < ndev; ++g)
        if (rcs[(size_t)g] != SA_OK) { sa_set_error("rank %d: %s", g, errs[(size_t)g].c_str()); return rcs[(size_t)g]; }
    return SA_OK;
}
