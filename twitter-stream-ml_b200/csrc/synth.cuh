// synth.cuh -- synthetic micro-batches generated straight into HBM (K6; measurement tool for the
// SURVEY.md 8d configs, no reference counterpart).  Counter-based, so the bits depend only on
// (seed, global row, column) and match oracle/twtml_oracle.c's generators exactly; products and
// sums are left unfused on purpose (the host generator is built with -ffp-contract=off).
#pragma once
#include "common.cuh"

namespace twtml {

// one CTA per row: features written coalesced; the pad / label slots are cleared
static __global__ void __launch_bounds__(256) k_synth_dense_x(double* __restrict__ X, long long ldx, int d, uint64_t seed,
                                                       long long row0, double inv_sqrt_d, long long n)
{
    for (long long i = blockIdx.x; i < n; i += gridDim.x) {
        const uint64_t r = (uint64_t)(row0 + i);
        double* row = X + i * ldx;
        for (int j = threadIdx.x; j < (int)ldx; j += blockDim.x) {
            double v = 0.0;
            if (j < d) v = __dmul_rn(__dsub_rn(__dmul_rn(2.0, u01(seed, r * (uint64_t)d + (uint64_t)j)), 1.0), inv_sqrt_d);
            row[j] = v;
        }
    }
}

// one thread per row: y = sum_j X[r][j] * w*[j] in index order + noise, regenerated from the
// counters (no read of X), written to the label column
static __global__ void __launch_bounds__(256) k_synth_dense_y(double* __restrict__ X, long long ldx, int d, uint64_t seed,
                                                       long long row0, double inv_sqrt_d, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t r = (uint64_t)(row0 + i);
    double acc = 0.0;
    for (int j = 0; j < d; ++j) {
        const double x = __dmul_rn(__dsub_rn(__dmul_rn(2.0, u01(seed, r * (uint64_t)d + (uint64_t)j)), 1.0), inv_sqrt_d);
        const double ws = __dmul_rn(__dsub_rn(__dmul_rn(2.0, u01(seed + 1, (uint64_t)j)), 1.0), inv_sqrt_d);
        acc = __dadd_rn(acc, __dmul_rn(x, ws));
    }
    const double noise = __dmul_rn(0.01, __dsub_rn(__dmul_rn(2.0, u01(seed + 2, r)), 1.0));
    X[i * ldx + d] = __dadd_rn(acc, noise);
}

// ---- synthetic CSR (config 5, uniform columns): the row recipe of twtml_oracle_synth_csr.
// Pass 1 counts distinct columns per row, pass 2 (after an exclusive scan) fills.
__device__ __forceinline__ int synth_csr_row(uint64_t seed, uint64_t r, int d_text, int* cols /*<=40*/)
{
    const uint64_t h = splitmix64(seed ^ (r * 0x100000001b3ull));
    const int k = 1 + (int)(h % 40);
    for (int t = 0; t < k; ++t) cols[t] = (int)(splitmix64(h + (uint64_t)t) % (uint64_t)d_text);
    for (int a = 1; a < k; ++a) { // insertion sort
        const int v = cols[a];
        int b = a - 1;
        while (b >= 0 && cols[b] > v) { cols[b + 1] = cols[b]; --b; }
        cols[b + 1] = v;
    }
    int u = 0;
    for (int t = 0; t < k; ++t)
        if (t == 0 || cols[t] != cols[t - 1]) cols[u++] = cols[t];
    return u;
}

static __global__ void __launch_bounds__(128) k_synth_csr_count(long long* __restrict__ rownnz, long long n, int d_text,
                                                         uint64_t seed, long long row0)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int cols[40];
    rownnz[i] = synth_csr_row(seed, (uint64_t)(row0 + i), d_text, cols) + 4;
}

static __global__ void __launch_bounds__(128) k_synth_csr_fill(const long long* __restrict__ rowptr, int* __restrict__ colidx,
                                                        double* __restrict__ vals, double* __restrict__ labels,
                                                        long long n, int d_text, uint64_t seed, long long row0)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t r = (uint64_t)(row0 + i);
    int cols[40];
    const int u = synth_csr_row(seed, r, d_text, cols);
    const uint64_t h = splitmix64(seed ^ (r * 0x100000001b3ull));
    long long o = rowptr[i];
    for (int t = 0; t < u; ++t) {
        colidx[o] = cols[t];
        vals[o] = (double)(1 + (int)(splitmix64(h ^ (uint64_t)cols[t]) % 3));
        ++o;
    }
    for (int t = 0; t < 4; ++t) {
        colidx[o] = d_text + t;
        vals[o] = __dmul_rn(u01(seed + 7 + (uint64_t)t, r), 1e-6);
        ++o;
    }
    labels[i] = 100.0 + floor(__dmul_rn(u01(seed + 3, r), 901.0));
}

} // namespace twtml
