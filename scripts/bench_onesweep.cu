// bench_onesweep.cu -- standalone tuning harness for the onesweep scatter pass (sm_100a).
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
//              -o scripts/bench_onesweep scripts/bench_onesweep.cu
// Run  :  scripts/bench_onesweep [n_records] [reps]
// For each compile-time configuration: full 8-pass LSD sort of (u64 key, u32 index) records on
// (a) uniformly random keys and (b) text-like keys (8 bytes of a Zipf-ish word stream),
// verified for order + stability, reporting the average pass time and 24 B/record bandwidth.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "../suffixarray_b200/csrc/radix.cu"
#include "onesweep_variants.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

typedef cudaError_t (*launch_fn)(const uint64_t*, const uint32_t*, uint64_t*, uint32_t*, uint32_t, int,
                                 const uint32_t*, uint32_t*, uint32_t*, uint32_t*, bool, cudaStream_t);
struct Variant { const char* name; launch_fn fn; int tile; };

#define V2(T, I, M, R, W) Variant{"v2 " #T "x" #I " minb" #M " rank" #R " lbw" #W, sa::launch_onesweep2_cfg<T, I, M, R, W>, T * I}
#define V3(T, I, M, R, W) Variant{"v3pf " #T "x" #I " minb" #M " rank" #R " lbw" #W, sa::launch_onesweep2_cfg<T, I, M, R, W, true>, T * I}
#define V4(I, M, W, R) Variant{"v4ws 256x" #I " minb" #M " lbw" #W " lbregs" #R, sa::launch_onesweep4_cfg<I, M, W, R>, 256 * I}
#define V(T, I, M, R) Variant{#T "x" #I " minb" #M " rank" #R, sa::launch_onesweep_cfg<T, I, M, R>, T * I}

__global__ void pack_text_keys(const uint8_t* text, uint32_t n, uint64_t* keys)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) keys[p] = sa::load_be64(text, p);
}

__global__ void check_sorted(const uint64_t* k, const uint32_t* v, uint32_t n, unsigned long long* bad)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 || i >= n) return;
    if (k[i - 1] > k[i] || (k[i - 1] == k[i] && v[i - 1] >= v[i])) atomicAdd(bad, 1ull);
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static inline uint64_t rnd()
{
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return rng_state;
}

int main(int argc, char** argv)
{
    const uint32_t n = argc > 1 ? (uint32_t)atoll(argv[1]) : 64u * 1000 * 1000;
    const int reps = argc > 2 ? atoi(argv[2]) : 3;
    const char* only = argc > 3 ? argv[3] : nullptr;      // run only variants whose name contains this
    const int first_input = argc > 4 ? atoi(argv[4]) : 0;
    std::vector<Variant> variants = {
        V(256, 16, 2, 0), V(384, 8, 3, 0), V(512, 8, 2, 1),
        V3(256, 16, 2, 0, 8), V3(256, 16, 2, 1, 8), V3(256, 16, 2, 0, 4), V3(512, 8, 2, 0, 8), V3(512, 8, 2, 1, 8),
        V3(256, 16, 2, 0, 16), V3(256, 16, 2, 0, 32), V3(256, 16, 2, 1, 16), V3(256, 16, 2, 1, 32),
        V3(512, 16, 1, 0, 8), V3(512, 16, 1, 0, 4), V3(1024, 8, 1, 0, 8), V3(256, 10, 3, 0, 8), V3(512, 12, 1, 0, 8),
        V4(12, 2, 8, 32), V4(12, 2, 4, 24), V4(14, 2, 8, 32), V4(14, 2, 4, 24), V4(16, 2, 8, 32), V4(16, 2, 4, 24), V4(10, 2, 8, 32),
        V3(256, 12, 2, 0, 8), V3(384, 8, 2, 0, 8), V3(256, 8, 4, 0, 8), V3(256, 8, 3, 0, 8), V3(384, 10, 2, 0, 8),
        V2(256, 16, 2, 0, 1), V2(256, 16, 2, 0, 4), V2(256, 16, 2, 0, 8), V2(256, 16, 2, 1, 4),
        V2(384, 8, 3, 0, 4), V2(384, 8, 3, 0, 8), V2(512, 8, 2, 0, 4), V2(512, 8, 2, 1, 4), V2(512, 8, 2, 1, 8),
        V2(256, 8, 4, 0, 4), V2(256, 8, 4, 1, 4), V2(256, 12, 3, 0, 4), V2(384, 10, 2, 0, 4), V2(512, 6, 3, 0, 4),
        V2(1024, 4, 1, 1, 4),
    };
    // ---- inputs
    std::vector<uint64_t> h_rand(n);
    for (uint32_t i = 0; i < n; ++i) h_rand[i] = rnd();
    std::vector<uint8_t> h_text((size_t)n + 64, 0);
    {
        std::vector<std::string> vocab(8192);
        for (auto& w : vocab) { int len = 2 + (int)(rnd() % 8); for (int j = 0; j < len; ++j) w.push_back('a' + (char)(rnd() % 26)); }
        size_t p = 0;
        while (p < n) {
            double u = (double)(rnd() >> 11) / 9007199254740992.0;
            size_t wi = (size_t)(u * u * u * vocab.size());            // skewed pick
            const std::string& w = vocab[wi];
            for (char c : w) { if (p < n) h_text[p++] = (uint8_t)c; }
            if (p < n) h_text[p++] = (rnd() % 12 == 0) ? '\n' : ' ';
        }
    }
    uint64_t *d_src, *d_k[2];
    uint32_t *d_v[2], *d_status, *d_small;
    uint8_t* d_text;
    unsigned long long* d_bad;
    CK(cudaMalloc(&d_src, (size_t)n * 8));
    for (int i = 0; i < 2; ++i) { CK(cudaMalloc(&d_k[i], (size_t)n * 8)); CK(cudaMalloc(&d_v[i], (size_t)n * 4)); }
    CK(cudaMalloc(&d_status, ((size_t)n / 1024 + 2) * 256 * 4));
    CK(cudaMalloc(&d_small, 8192 * 4));
    CK(cudaMalloc(&d_text, (size_t)n + 64));
    CK(cudaMalloc(&d_bad, 8));
    CK(cudaMemcpy(d_text, h_text.data(), (size_t)n + 64, cudaMemcpyHostToDevice));
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    cudaEvent_t ev[18];
    for (auto& e : ev) CK(cudaEventCreate(&e));
    uint32_t* hist = d_small; uint32_t* base = d_small + 2048; uint32_t* skip = d_small + 4096;
    uint32_t* tilectr = d_small + 4200; uint32_t* err = d_small + 4201;

    for (int input = first_input; input < 2; ++input) {
        if (input == 0) CK(cudaMemcpy(d_src, h_rand.data(), (size_t)n * 8, cudaMemcpyHostToDevice));
        else { pack_text_keys<<<(n + 255) / 256, 256, 0, s>>>(d_text, n, d_src); CK(cudaStreamSynchronize(s)); }
        printf("== input: %s, n = %u\n", input == 0 ? "random u64" : "text-like (8-byte big-endian windows)", n);
        for (const Variant& v : variants) {
            if (only && !strstr(v.name, only)) continue;
            float best = 1e30f, first_pass = 0;
            unsigned long long bad = 0;
            uint32_t herr = 0;
            for (int r = 0; r < reps; ++r) {
                CK(cudaMemcpyAsync(d_k[0], d_src, (size_t)n * 8, cudaMemcpyDeviceToDevice, s));
                CK(cudaMemsetAsync(err, 0, 4, s));
                CK(sa_launch_hist(d_k[0], n, 0, 8, hist, 148, s));
                CK(sa_launch_scan_hist(hist, n, 8, base, skip, s));
                int cur = 0;
                for (int p = 0; p < 8; ++p) {
                    CK(cudaEventRecord(ev[2 * p], s));
                    CK(v.fn(d_k[cur], d_v[cur], d_k[cur ^ 1], d_v[cur ^ 1], n, 8 * p, base + 256 * p, d_status, tilectr,
                            err, p == 0, s));
                    CK(cudaEventRecord(ev[2 * p + 1], s));
                    cur ^= 1;
                }
                CK(cudaMemsetAsync(d_bad, 0, 8, s));
                check_sorted<<<(n + 255) / 256, 256, 0, s>>>(d_k[cur], d_v[cur], n, d_bad);
                CK(cudaStreamSynchronize(s));
                CK(cudaMemcpy(&bad, d_bad, 8, cudaMemcpyDeviceToHost));
                CK(cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost));
                float tot = 0, ms;
                for (int p = 1; p < 8; ++p) { CK(cudaEventElapsedTime(&ms, ev[2 * p], ev[2 * p + 1])); tot += ms; }
                CK(cudaEventElapsedTime(&first_pass, ev[0], ev[1]));
                if (tot / 7 < best) best = tot / 7;
            }
#ifdef OS_PHASE_CLOCKS
            {
                unsigned long long ph[8], z[8] = {0};
                CK(cudaMemcpyFromSymbol(ph, sa::os_phase_cycles, sizeof ph));
                CK(cudaMemcpyToSymbol(sa::os_phase_cycles, z, sizeof z));
                double tot = 0; for (int i = 0; i < 8; ++i) tot += (double)ph[i];
                printf("   thread-0 cycle shares: zero+sync %.1f%% | load %.1f%% | rank+sync %.1f%% | scan %.1f%% | reorder %.1f%% | lookback(t0)/v4:issue next loads %.1f%% | sync(after lookback)/v4:wait offsets %.1f%% | store %.1f%%   (cycles/tile %.0f)\n",
                       100 * ph[0] / tot, 100 * ph[1] / tot, 100 * ph[2] / tot, 100 * ph[3] / tot, 100 * ph[4] / tot, 100 * ph[5] / tot,
                       100 * ph[6] / tot, 100 * ph[7] / tot, tot / (8.0 * reps * ((n + v.tile - 1) / v.tile)));
            }
#endif
            printf("%-22s tile %5d  pass %.3f ms  %7.1f GB/s  (identity pass %.3f ms)  %s%s\n", v.name, v.tile, best,
                   24.0 * n / (best * 1e-3) / 1e9, first_pass, bad ? "NOT SORTED " : "ok", herr ? " DEVERR" : "");
            fflush(stdout);
        }
    }
    return 0;
}
