// bench_onesweep_wide.cu -- EXPERIMENT: does a wider digit pay on this path's keys?  (sm_100a)
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
//              -o scripts/bench_onesweep_wide scripts/bench_onesweep_wide.cu
// Run  :  scripts/bench_onesweep_wide [n_records] [reps] [name filter]
// Sorts (60-bit key, u32 index) records -- the first sort of a build: 12 characters of 5-bit codes,
// key bits [4, 64) -- with the product's pass (8 passes of 8 bits, onesweep4_kernel) and with the
// RB-bit generalisation in onesweep_wide.cuh under several digit plans, on (a) text-like keys and
// (b) uniformly random 60-bit keys.  Every run is checked for order and stability; the figure of
// merit is the time of the WHOLE sort (all passes), since the plans differ in pass count.
// The CPU model that motivates it: scripts/digit_model.py, profiles/r01_digit_model.md.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "../suffixarray_b200/csrc/radix.cu"
#include "onesweep_wide.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int MAXP = 8, MAXBINS = 1024, LOW_BIT = 4;

typedef cudaError_t (*pass_fn)(const uint64_t*, const uint32_t*, uint64_t*, uint32_t*, uint32_t, int, int, const uint32_t*,
                               uint32_t*, uint32_t*, uint32_t*, bool, cudaStream_t);
struct Variant { const char* name; pass_fn fn; int tile; std::vector<int> widths; };

// the product's kernel behind the same signature (8-bit digits only)
static cudaError_t product_pass(const uint64_t* ki, const uint32_t* vi, uint64_t* ko, uint32_t* vo, uint32_t n, int shift,
                                int nbits, const uint32_t* base, uint32_t* status, uint32_t* ctr, uint32_t* err, bool ident,
                                cudaStream_t s)
{
    if (nbits != 8) return cudaErrorInvalidValue;
    return sa::launch_onesweep4_cfg<12, 2, 4, 24>(ki, vi, ko, vo, n, shift, base, status, ctr, err, ident, s);
}

#define W(RB, I, M, L, R, ...) Variant{"wide rb" #RB " 256x" #I " minb" #M " lbw" #L " lbregs" #R " plan " #__VA_ARGS__, \
                                       sa::launch_onesweepw_cfg<RB, I, M, L, R>, 256 * I, {__VA_ARGS__}}

struct PlanArgs { int shift[MAXP]; uint32_t mask[MAXP]; int np; };

__global__ void hist_plan_kernel(const uint64_t* __restrict__ keys, uint32_t n, PlanArgs pa, uint32_t* __restrict__ hist)
{
    __shared__ uint32_t h[MAXP * MAXBINS];
    for (int i = threadIdx.x; i < pa.np * MAXBINS; i += blockDim.x) h[i] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint64_t k = keys[i];
        for (int p = 0; p < pa.np; ++p) atomicAdd(&h[p * MAXBINS + ((uint32_t)(k >> pa.shift[p]) & pa.mask[p])], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < pa.np * MAXBINS; i += blockDim.x)
        if (h[i]) atomicAdd(&hist[i], h[i]);
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
    const uint32_t n = argc > 1 ? (uint32_t)atoll(argv[1]) : 100u * 1000 * 1000;
    const int reps = argc > 2 ? atoi(argv[2]) : 3;
    const char* only = argc > 3 ? argv[3] : nullptr;
    std::vector<Variant> variants = {
        Variant{"product v4ws 256x12 (8 passes of 8)", product_pass, 3072, {8, 8, 8, 8, 8, 8, 8, 8}},
        W(8, 12, 2, 4, 24, 8, 8, 8, 8, 8, 8, 8, 8),            // control: the generalised kernel at 8 bits
        W(9, 12, 2, 4, 24, 8, 8, 8, 9, 9, 9, 9),
        W(9, 12, 2, 2, 24, 8, 8, 8, 9, 9, 9, 9),
        W(9, 12, 2, 4, 32, 8, 8, 8, 9, 9, 9, 9),
        W(9, 12, 2, 4, 24, 9, 9, 9, 9, 8, 8, 8),
        W(10, 12, 2, 2, 24, 10, 10, 10, 10, 10, 10),
        W(10, 12, 2, 2, 32, 10, 10, 10, 10, 10, 10),
        W(10, 12, 2, 1, 24, 10, 10, 10, 10, 10, 10),
        W(10, 10, 2, 2, 24, 10, 10, 10, 10, 10, 10),
        W(10, 24, 1, 2, 32, 10, 10, 10, 10, 10, 10),           // one CTA per SM, 6144-record tiles
        W(10, 20, 1, 2, 32, 10, 10, 10, 10, 10, 10),
    };
    // ---- inputs: 60-bit keys in bits [4, 64)
    std::vector<uint64_t> h_text(n), h_rand(n);
    {
        std::vector<uint8_t> codes((size_t)n + 16, 0);
        std::vector<std::string> vocab(8192);
        for (auto& w : vocab) { int len = 2 + (int)(rnd() % 8); for (int j = 0; j < len; ++j) w.push_back((char)(3 + rnd() % 26)); }
        size_t p = 0;
        while (p < n) {
            double u = (double)(rnd() >> 11) / 9007199254740992.0;
            const std::string& w = vocab[(size_t)(u * u * u * vocab.size())];            // skewed pick
            for (char c : w) { if (p < n) codes[p++] = (uint8_t)c; }
            if (p < n) codes[p++] = (rnd() % 12 == 0) ? 1 : 2;                            // '\n' : ' '
        }
        uint64_t k = 0;
        for (int j = 0; j < 11; ++j) k = (k << 5) | codes[j];
        for (uint32_t i = 0; i < n; ++i) {
            k = ((k << 5) | codes[(size_t)i + 11]) & ((1ull << 60) - 1);
            h_text[i] = k << LOW_BIT;
        }
        for (uint32_t i = 0; i < n; ++i) h_rand[i] = rnd() << LOW_BIT;
    }
    uint64_t *d_src, *d_k[2];
    uint32_t *d_v[2], *d_status, *d_hist, *d_base, *d_misc;
    unsigned long long* d_bad;
    CK(cudaMalloc(&d_src, (size_t)n * 8));
    for (int i = 0; i < 2; ++i) { CK(cudaMalloc(&d_k[i], (size_t)n * 8)); CK(cudaMalloc(&d_v[i], (size_t)n * 4)); }
    CK(cudaMalloc(&d_status, ((size_t)n / 2048 + 2) * MAXBINS * 4));     // tiles are >= 2560 records
    CK(cudaMalloc(&d_hist, MAXP * MAXBINS * 4));
    CK(cudaMalloc(&d_base, MAXP * MAXBINS * 4));
    CK(cudaMalloc(&d_misc, 64));
    CK(cudaMalloc(&d_bad, 8));
    uint32_t* tilectr = d_misc; uint32_t* err = d_misc + 1;
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    cudaEvent_t ev[2 * MAXP + 2];
    for (auto& e : ev) CK(cudaEventCreate(&e));
    std::vector<uint32_t> h_hist(MAXP * MAXBINS), h_base(MAXP * MAXBINS);

    for (int input = 0; input < 2; ++input) {
        CK(cudaMemcpy(d_src, (input == 0 ? h_text : h_rand).data(), (size_t)n * 8, cudaMemcpyHostToDevice));
        printf("== input: %s, n = %u\n", input == 0 ? "text-like (12 characters x 5-bit codes)" : "random 60-bit", n);
        for (const Variant& v : variants) {
            if (only && !strstr(v.name, only)) continue;
            const int np = (int)v.widths.size();
            PlanArgs pa{};
            pa.np = np;
            int sh = LOW_BIT, covered = 0;
            for (int p = 0; p < np; ++p) { pa.shift[p] = sh; pa.mask[p] = (1u << v.widths[p]) - 1u; sh += v.widths[p]; covered += v.widths[p]; }
            if (covered < 60 || np > MAXP) { printf("%s: plan does not cover 60 bits\n", v.name); continue; }
            float best = 1e30f, best_pass[MAXP] = {0}, hist_ms = 0;
            unsigned long long bad = 0;
            uint32_t herr = 0;
            bool launch_failed = false;
            for (int r = 0; r < reps && !launch_failed; ++r) {
                CK(cudaMemcpyAsync(d_k[0], d_src, (size_t)n * 8, cudaMemcpyDeviceToDevice, s));
                CK(cudaMemsetAsync(err, 0, 4, s));
                CK(cudaMemsetAsync(d_hist, 0, MAXP * MAXBINS * 4, s));
                CK(cudaEventRecord(ev[2 * MAXP], s));
                hist_plan_kernel<<<148 * 4, 512, 0, s>>>(d_k[0], n, pa, d_hist);
                CK(cudaEventRecord(ev[2 * MAXP + 1], s));
                CK(cudaMemcpyAsync(h_hist.data(), d_hist, MAXP * MAXBINS * 4, cudaMemcpyDeviceToHost, s));
                CK(cudaStreamSynchronize(s));
                for (int p = 0; p < np; ++p) {
                    uint32_t run = 0;
                    for (int d = 0; d < MAXBINS; ++d) { h_base[p * MAXBINS + d] = run; run += h_hist[p * MAXBINS + d]; }
                }
                CK(cudaMemcpyAsync(d_base, h_base.data(), MAXP * MAXBINS * 4, cudaMemcpyHostToDevice, s));
                int cur = 0;
                for (int p = 0; p < np; ++p) {
                    CK(cudaEventRecord(ev[2 * p], s));
                    cudaError_t le = v.fn(d_k[cur], d_v[cur], d_k[cur ^ 1], d_v[cur ^ 1], n, pa.shift[p], v.widths[p],
                                          d_base + p * MAXBINS, d_status, tilectr, err, p == 0, s);
                    if (le != cudaSuccess) { printf("%s: launch failed: %s\n", v.name, cudaGetErrorString(le)); launch_failed = true; break; }
                    CK(cudaEventRecord(ev[2 * p + 1], s));
                    cur ^= 1;
                }
                if (launch_failed) { cudaGetLastError(); break; }
                CK(cudaMemsetAsync(d_bad, 0, 8, s));
                check_sorted<<<(n + 255) / 256, 256, 0, s>>>(d_k[cur], d_v[cur], n, d_bad);
                CK(cudaStreamSynchronize(s));
                CK(cudaMemcpy(&bad, d_bad, 8, cudaMemcpyDeviceToHost));
                CK(cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost));
                float tot = 0, ms[MAXP];
                for (int p = 0; p < np; ++p) { CK(cudaEventElapsedTime(&ms[p], ev[2 * p], ev[2 * p + 1])); tot += ms[p]; }
                CK(cudaEventElapsedTime(&hist_ms, ev[2 * MAXP], ev[2 * MAXP + 1]));
                if (tot < best) { best = tot; memcpy(best_pass, ms, sizeof ms); }
            }
            if (launch_failed) continue;
            printf("%-72s tile %5d  sort %6.3f ms (%d passes:", v.name, v.tile, best, np);
            for (int p = 0; p < np; ++p) printf(" %.3f", best_pass[p]);
            printf(")  hist %.3f ms  %s%s\n", hist_ms, bad ? "NOT SORTED " : "ok", herr ? " DEVERR" : "");
            fflush(stdout);
        }
    }
    return 0;
}
