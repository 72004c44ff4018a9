// csr.cuh -- the sparse (CSR) form of the per-iteration pass (K3).
//
// MLlib's sparse branch of LeastSquaresGradient.compute [U] is, per row,
//     diff = sum_k values(k) * w(indices(k)) - y      (BLAS.dot sparse.dense, stored order)
//     gradSum(indices(k)) += diff * values(k)         (BLAS.axpy)
// The first line is a row-wise pass, the second a scatter.  A scatter with floating-point atomics
// would make the sums depend on the order the hardware happened to serve them, so instead the
// batch keeps its transpose (CSC, built once at ingest, entries of a column in row order) and every
// iteration runs
//   k_csr_rows   rows -> diff[r] (+ lossSum, count, statistics per row block, fixed order)
//   k_csr_cols   column tasks (<= kSeg entries of one column): sum diff[row] * value, fixed order
//   k_csr_fold   per column: add its tasks' sums in task order -> gradSum[j]
// Cost: the values are read twice per iteration (12 B/nnz in each form); see DESIGN.md.
#pragma once
#include "common.cuh"
#include "dense.cuh" // IterMode

namespace twtml {

constexpr int kCsrThreads = 256;
constexpr int kCsrGroups = kCsrThreads / 8; // 8 lanes per row
constexpr int kSeg = 1024;                  // entries per column task

struct CsrIterArgs {
    const long long* rowptr;
    const int* colidx;
    const double* vals;
    const double* labels;
    const double* w;
    const TrainState* state;
    double* diff;          // n
    double* partials;      // nblocks x 2 (lossSum, count)
    double* stat_partials; // nblocks x kStatStride
    double* out_pred;
    double* out_real;
    Plan plan;
    double frac;
    unsigned long long row0_global;
    int iter_override;
};

template <int MODE>
__global__ void __launch_bounds__(kCsrThreads) k_csr_rows(const CsrIterArgs a)
{
    constexpr bool kTrain = (MODE == MODE_TRAIN || MODE == MODE_TRAIN_STATS);
    constexpr bool kStats = (MODE == MODE_TRAIN_STATS || MODE == MODE_PREDICT_STATS);
    if (kTrain && a.state->done) return;
    __shared__ double sred[kCsrGroups * kStatStride];
    const int t = threadIdx.x, grp = t >> 3, l8 = t & 7;
    const unsigned long long iter_seed =
        splitmix64((unsigned long long)(42 + (a.iter_override > 0 ? a.iter_override : (kTrain ? a.state->iter : 0))));
    const bool sampling = kTrain && a.frac < 1.0;

    for (long long blk = blockIdx.x; blk < a.plan.nblocks; blk += gridDim.x) {
        long long row0;
        int nrows;
        block_geom(a.plan, blk, row0, nrows);
        double my_loss = 0.0, my_cnt = 0.0;
        Stat7 my_stat;
        stat_zero(my_stat);
        for (int r0 = 0; r0 < nrows; r0 += kCsrGroups) {
            const int r = r0 + grp;
            const bool valid = r < nrows;
            const long long row = row0 + r;
            double acc = 0.0;
            if (valid) {
                const long long b = a.rowptr[row], e = a.rowptr[row + 1];
                for (long long k = b + l8; k < e; k += 8) acc = fma(a.vals[k], a.w[a.colidx[k]], acc);
            }
            acc += shfl_xor_f64(acc, 4);
            acc += shfl_xor_f64(acc, 2);
            acc += shfl_xor_f64(acc, 1);
            if (valid && l8 == 0) {
                const double y = a.labels[row];
                if (kTrain) {
                    double diff = 0.0;
                    const bool keep = !sampling || row_sampled(iter_seed, a.row0_global + (unsigned long long)row, a.frac);
                    if (keep) {
                        diff = acc - y;
                        my_loss += diff * diff / 2.0;
                        my_cnt += 1.0;
                    }
                    a.diff[row] = diff;
                }
                if (kStats) {
                    const double pr = round_half_up(acc + 0.0);
                    stat_add(my_stat, y, pr);
                    if (a.out_pred) a.out_pred[row] = pr;
                    if (a.out_real) a.out_real[row] = y;
                } else if (MODE == MODE_PREDICT) {
                    a.out_pred[row] = acc + 0.0;
                }
            }
        }
        if (kTrain) {
            if (l8 == 0) { sred[2 * grp] = my_loss; sred[2 * grp + 1] = my_cnt; }
            __syncthreads();
            if (t == 0) {
                double ls = 0.0, cn = 0.0;
                for (int q = 0; q < kCsrGroups; ++q) { ls += sred[2 * q]; cn += sred[2 * q + 1]; }
                a.partials[blk * 2] = ls;
                a.partials[blk * 2 + 1] = cn;
            }
            __syncthreads();
        }
        if (kStats) {
            if (l8 == 0) stat_store(sred + grp * kStatStride, my_stat);
            __syncthreads();
            if (t == 0) {
                Stat7 tot;
                stat_zero(tot);
                for (int q = 0; q < kCsrGroups; ++q) {
                    Stat7 s;
                    stat_load(s, sred + q * kStatStride);
                    stat_merge(tot, s);
                }
                stat_store(a.stat_partials + blk * kStatStride, tot);
            }
            __syncthreads();
        }
    }
}

// one warp per column task
static __global__ void __launch_bounds__(256) k_csr_cols(const int* __restrict__ rowidx, const double* __restrict__ cvals,
                                                  const long long* __restrict__ colptr, const int* __restrict__ task_col,
                                                  const long long* __restrict__ task_beg, long long ntasks,
                                                  const double* __restrict__ diff, double* __restrict__ task_part,
                                                  const TrainState* gate)
{
    if (gate && gate->done) return;
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long tsk = warp0; tsk < ntasks; tsk += nwarps) {
        const long long beg = task_beg[tsk];
        const long long cend = colptr[task_col[tsk] + 1];
        const long long end = beg + kSeg < cend ? beg + kSeg : cend;
        double acc = 0.0;
        for (long long e = beg + lane; e < end; e += 32) acc = fma(diff[rowidx[e]], cvals[e], acc);
        for (int m = 16; m >= 1; m >>= 1) acc += shfl_xor_f64(acc, m);
        if (lane == 0) task_part[tsk] = acc;
    }
}

// one warp per column: gradSum[j] = sum of its tasks in task order (0 for an empty column)
static __global__ void __launch_bounds__(256) k_csr_fold(const long long* __restrict__ taskptr, const double* __restrict__ task_part,
                                                  int d, double* __restrict__ g, const TrainState* gate)
{
    if (gate && gate->done) return;
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long j = warp0; j < d; j += nwarps) {
        const long long b = taskptr[j], e = taskptr[j + 1];
        double acc = 0.0;
        for (long long k = b + lane; k < e; k += 32) acc += task_part[k];
        for (int m = 16; m >= 1; m >>= 1) acc += shfl_xor_f64(acc, m);
        if (lane == 0) g[j] = acc;
    }
}

// ---- transpose build helpers (ingest time) ---------------------------------------------------
static __global__ void k_iota(int* __restrict__ p, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = (int)i;
}

// sorted column keys -> colptr[0..d]
static __global__ void k_colptr(const int* __restrict__ keys, long long nnz, int d, long long* __restrict__ colptr)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e > nnz) return;
    const int prev = e == 0 ? -1 : keys[e - 1];
    const int cur = e == nnz ? d : keys[e];
    for (int j = prev + 1; j <= cur; ++j) colptr[j] = e;
}

// perm -> (row of the entry by binary search in rowptr, value)
static __global__ void k_csc_gather(const int* __restrict__ perm, const long long* __restrict__ rowptr, long long n,
                             const double* __restrict__ vals, long long nnz, int* __restrict__ rowidx, double* __restrict__ cvals)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nnz) return;
    const long long src = perm[e];
    long long lo = 0, hi = n; // largest r with rowptr[r] <= src
    while (hi - lo > 1) {
        const long long mid = (lo + hi) >> 1;
        if (rowptr[mid] <= src) lo = mid; else hi = mid;
    }
    rowidx[e] = (int)lo;
    cvals[e] = vals[src];
}

static __global__ void k_task_count(const long long* __restrict__ colptr, int d, long long* __restrict__ ntasks)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < d) ntasks[j] = (colptr[j + 1] - colptr[j] + kSeg - 1) / kSeg;
    if (j == d) ntasks[j] = 0;
}

static __global__ void k_task_fill(const long long* __restrict__ colptr, const long long* __restrict__ taskptr, int d,
                            int* __restrict__ task_col, long long* __restrict__ task_beg)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d) return;
    long long t = taskptr[j];
    for (long long b = colptr[j]; b < colptr[j + 1]; b += kSeg, ++t) { task_col[t] = j; task_beg[t] = b; }
}

} // namespace twtml
