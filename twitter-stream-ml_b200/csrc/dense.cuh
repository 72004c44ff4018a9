// dense.cuh -- the fused per-iteration kernel of the dense path (K2) and its predict-only forms (K5).
//
// One launch = one pass over this rank's rows = one execution of the treeAggregate in
// GradientDescent.runMiniBatchSGD [U] with seqOp = LeastSquaresGradient.compute [U]:
//     diff = x.w - y;  gradSum += diff * x;  lossSum += diff^2 / 2;  count += 1
// and, in the *_STATS modes, the predict + statistics block of LinearRegression.scala:56-65 from
// the same loaded rows (those use the pre-training weights, which is what iteration 1 holds).
//
// Layout in HBM: row-major, row stride ldx = round_up(d + 1, 2) doubles; columns [0, d) are the
// features, column d is the label, the rest is a zero pad.  A run of consecutive rows is therefore
// one contiguous, 16-byte-aligned byte range, which is what lets a whole stage be fetched by ONE 1-D
// bulk TMA copy (cp.async.bulk, no tensor map) that carries the labels with it.
//
// CTA = 512 threads.  Thread t owns a pair of adjacent columns (col2) of one "row lane" rl; a stage
// holds R = RPT * RL rows.  Per stage:
//   phase 1  each thread reads its column pair of its RPT rows from shared memory ONCE (LDS.128)
//            into registers, forms partial dot products with its two weights (kept in registers for
//            the whole kernel), and the warp folds RPT partials with a reduce-scatter butterfly;
//   (bar A)  R threads add the warp partials of their row in warp order, subtract the label
//            (read from the stage), and publish diff[r]; loss / count / statistics accumulate there;
//   (bar B)  thread 0 re-arms the stage and issues the TMA for the stage S steps ahead;
//   phase 2  g[col] += diff[r] * x[r][col] from the registers loaded in phase 1.
// So every byte of X crosses HBM -> SMEM once and SMEM -> registers once per iteration.
// At the end of a row block the column owners write the block's partial vector
// [g_0 .. g_{ldx-1}, lossSum, count] to scratch; nothing is accumulated with atomics, the later
// reduction order is fixed (reduce.cuh), hence bit-reproducible results.
#pragma once
#include "common.cuh"

namespace twtml {

enum IterMode { MODE_TRAIN = 0, MODE_TRAIN_STATS = 1, MODE_PREDICT = 2, MODE_PREDICT_STATS = 3 };

struct DenseIterArgs {
    const double* X;       // n x ldx
    const double* w;       // current weights, >= d doubles
    const TrainState* state;
    double* partials;      // nblocks x pv      (train modes)
    double* stat_partials; // nblocks x kStatStride (stats modes)
    double* out_pred;      // n (PREDICT: raw; *_STATS: rounded) or null
    double* out_real;      // n or null
    Plan plan;
    long long ldx;
    int d, c2, c2p, rl, nstages;
    unsigned stage_stride;  // bytes between stages in shared memory
    double frac;            // miniBatchFraction
    unsigned long long row0_global; // global index of local row 0 (sampler key)
    int iter_override;      // >0: use this iteration number for the sampler instead of state->iter
};

template <int RPT>
__device__ __forceinline__ void warp_fold(double (&v)[RPT], int lane, double& out, int& which)
{
    // reduce-scatter butterfly: after it, every lane of an 8-lane (RPT=4) / 16-lane (RPT=2) group
    // holds the warp total of row `which`.
    if constexpr (RPT == 4) {
        const bool hi = lane & 16;
        double s0 = hi ? v[0] : v[2], s1 = hi ? v[1] : v[3];
        double k0 = hi ? v[2] : v[0], k1 = hi ? v[3] : v[1];
        k0 += shfl_xor_f64(s0, 16);
        k1 += shfl_xor_f64(s1, 16);
        const bool hi2 = lane & 8;
        double s = hi2 ? k0 : k1, k = hi2 ? k1 : k0;
        k += shfl_xor_f64(s, 8);
        k += shfl_xor_f64(k, 4);
        k += shfl_xor_f64(k, 2);
        k += shfl_xor_f64(k, 1);
        out = k;
        which = (hi ? 2 : 0) + (hi2 ? 1 : 0);
    } else if constexpr (RPT == 2) {
        const bool hi = lane & 16;
        double s = hi ? v[0] : v[1], k = hi ? v[1] : v[0];
        k += shfl_xor_f64(s, 16);
        k += shfl_xor_f64(k, 8);
        k += shfl_xor_f64(k, 4);
        k += shfl_xor_f64(k, 2);
        k += shfl_xor_f64(k, 1);
        out = k;
        which = hi ? 1 : 0;
    } else {
        double k = v[0];
        k += shfl_xor_f64(k, 16);
        k += shfl_xor_f64(k, 8);
        k += shfl_xor_f64(k, 4);
        k += shfl_xor_f64(k, 2);
        k += shfl_xor_f64(k, 1);
        out = k;
        which = 0;
    }
}

// shared-memory carve-up (offsets in bytes from the dynamic smem base)
struct IterSmem {
    static constexpr unsigned kBars = 0;                        // 8 x u64
    static constexpr unsigned kWpart = 64;                      // kIterWarps x 4 doubles
    static constexpr unsigned kDiff = kWpart + kIterWarps * 4 * 8; // 64 doubles
    static constexpr unsigned kRed = kDiff + 64 * 8;            // 64 x (lossSum, count) of the per-row threads at a block's end
    static constexpr unsigned kGbuf = kRed + 64 * 2 * 8;        // kIterThreads x 2 doubles (row-lane fold)
    static constexpr unsigned kStat = kGbuf + kIterThreads * 16;  // 64 x kStatStride doubles: the per-row threads' StatCounter tuples
    static constexpr unsigned kStages = ((kStat + 64 * kStatStride * 8) + 127) & ~127u;
};

// RL1: one row lane (c2p == 512 or CP > 1): the common wide-row case, with compile-time offsets.
template <int CP, int MODE, bool RL1>
__global__ void __launch_bounds__(kIterThreads, (CP == 1) ? 2 : 1) k_dense_iter(const DenseIterArgs a)
{
    constexpr bool kTrain = (MODE == MODE_TRAIN || MODE == MODE_TRAIN_STATS);
    constexpr bool kStats = (MODE == MODE_TRAIN_STATS || MODE == MODE_PREDICT_STATS);
    constexpr int RPT = 4 / CP;
    static_assert(CP == 1 || RL1, "several column passes imply one row lane");
    if (kTrain && a.state->done) return;

    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + IterSmem::kBars);
    double* wpart = reinterpret_cast<double*>(smem + IterSmem::kWpart);
    double* sdiff = reinterpret_cast<double*>(smem + IterSmem::kDiff);
    double* sred = reinterpret_cast<double*>(smem + IterSmem::kRed);
    double* gbuf = reinterpret_cast<double*>(smem + IterSmem::kGbuf);
    double* sstat = reinterpret_cast<double*>(smem + IterSmem::kStat); // kept in shared memory: only R threads use them, registers are per thread
    unsigned char* stages = smem + IterSmem::kStages;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int d = a.d, C2 = a.c2, S = a.nstages;
    const int RL = RL1 ? 1 : a.rl;
    const int R = RPT * RL;
    const long long ldx = a.ldx;
    const int pv = (int)ldx + 2;
    const int rl = RL1 ? 0 : t / a.c2p;
    const int wpr = RL1 ? kIterWarps : a.c2p / 32; // warps per row lane

    // Column ownership.  A thread without a column (past the row's end, or in an unused row lane)
    // shadows column pair 0 of row lane 0 with zero weights: it loads valid finite data, adds 0 to
    // the dot products and never stores its gradient, so the hot loop carries no predicates.
    int col2[CP];
    double w0[CP], w1[CP];
    bool live[CP];
    unsigned xoff[CP]; // byte offset of this thread's first row inside a stage
#pragma unroll
    for (int p = 0; p < CP; ++p) {
        col2[p] = RL1 ? (t + p * kIterThreads) : (t - rl * a.c2p);
        live[p] = rl < RL && col2[p] < C2;
        const int c0 = 2 * col2[p];
        // the label / pad columns carry zero weights: they add exactly 0 to x.w; their gradient slots are never read
        w0[p] = (live[p] && c0 < d) ? a.w[c0] : 0.0;
        w1[p] = (live[p] && c0 + 1 < d) ? a.w[c0 + 1] : 0.0;
        xoff[p] = live[p] ? (unsigned)(((long long)rl * ldx + c0) * 8) : 0u;
    }
    const unsigned rstride = (unsigned)((long long)RL * ldx * 8); // between a thread's consecutive rows
    const int drow = (RL1 || rl < RL) ? rl : 0;

    const unsigned long long iter_seed =
        splitmix64((unsigned long long)(42 + (a.iter_override > 0 ? a.iter_override : (kTrain ? a.state->iter : 0))));
    const bool sampling = kTrain && a.frac < 1.0;

    // rows beyond a ragged stage's end are read as whatever the slot held before: make that finite
    for (unsigned o = (unsigned)t * 16; o < (unsigned)S * a.stage_stride; o += kIterThreads * 16)
        *reinterpret_cast<double2*>(stages + o) = make_double2(0.0, 0.0);
    if (kStats && t < 64)
        for (int q = 0; q < kStatStride; ++q) sstat[t * kStatStride + q] = 0.0;
    if (t == 0) {
        for (int s = 0; s < S; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic writes above, async-proxy (TMA) writes next
    __syncthreads();

    // ---- producer cursor (thread 0 only), S stages ahead of the consumer cursor
    const uint64_t policy = l2_evict_first_policy();
    long long pblk = blockIdx.x;
    int pit = 0;
    auto produce = [&](int slot) {
        if (pblk >= a.plan.nblocks) return;
        long long row0;
        int nrows;
        block_geom(a.plan, pblk, row0, nrows);
        const int r0 = pit * R;
        const int nr = min(R, nrows - r0);
        const unsigned bytes = (unsigned)((long long)nr * ldx * 8);
        mbar_expect_tx(&full[slot], bytes);
        tma_load_1d(stages + (size_t)slot * a.stage_stride, a.X + (row0 + r0) * ldx, bytes, &full[slot], policy);
        if (r0 + R >= nrows) { pblk += gridDim.x; pit = 0; } else { ++pit; }
    };
    if (t == 0)
        for (int s = 0; s < S; ++s) produce(s);

    // ---- per-thread accumulators
    double g0[CP], g1[CP];
#pragma unroll
    for (int p = 0; p < CP; ++p) g0[p] = g1[p] = 0.0;
    double my_loss = 0.0, my_cnt = 0.0; // threads t < R: their row of every stage
    // thread t < R finishes row t of each stage: that is row lane t % RL, k-th row t / RL
    const double* my_wpart = wpart + ((RL1 ? 0 : (t % RL) * wpr) * 4 + (RL1 ? t : t / RL));
    const unsigned yoff = (unsigned)(((long long)t * ldx + d) * 8);

    int slot = 0;
    unsigned parity = 0;
    const unsigned char* st = stages;
    for (long long blk = blockIdx.x; blk < a.plan.nblocks; blk += gridDim.x) {
        long long row0;
        int nrows;
        block_geom(a.plan, blk, row0, nrows);
        for (int r0 = 0; r0 < nrows; r0 += R) {
            mbar_wait(&full[slot], parity);

            // ---- phase 1: registers <- smem, partial dots
            double2 x[RPT][CP];
            double part[RPT];
#pragma unroll
            for (int k = 0; k < RPT; ++k) {
                part[k] = 0.0;
#pragma unroll
                for (int p = 0; p < CP; ++p) {
                    const double2 v = *reinterpret_cast<const double2*>(st + xoff[p] + k * rstride);
                    x[k][p] = v;
                    part[k] = fma(v.x, w0[p], part[k]);
                    part[k] = fma(v.y, w1[p], part[k]);
                }
            }
            double tot;
            int which;
            warp_fold<RPT>(part, lane, tot, which);
            constexpr int grp = 32 / RPT;
            if ((lane & (grp - 1)) == 0) wpart[warp * 4 + which] = tot;
            __syncthreads(); // (A)

            // ---- one thread per row: finish the dot product, residual, loss, statistics
            if (t < R) {
                double diff = 0.0;
                if (t < nrows - r0) {
                    double dot = 0.0;
                    if (RL1) {
#pragma unroll
                        for (int q = 0; q < kIterWarps; ++q) dot += my_wpart[q * 4];
                    } else {
                        for (int q = 0; q < wpr; ++q) dot += my_wpart[q * 4];
                    }
                    const double y = *reinterpret_cast<const double*>(st + yoff);
                    const long long lrow = row0 + r0 + t;
                    if (kTrain) {
                        const bool keep = !sampling || row_sampled(iter_seed, a.row0_global + (unsigned long long)lrow, a.frac);
                        if (keep) {
                            diff = dot - y;
                            my_loss += diff * diff / 2.0;
                            my_cnt += 1.0;
                        }
                    }
                    if (kStats) {
                        const double pr = round_half_up(dot + 0.0);
                        Stat7 ms;
                        stat_load(ms, sstat + t * kStatStride);
                        stat_add(ms, y, pr);
                        stat_store(sstat + t * kStatStride, ms);
                        if (a.out_pred) a.out_pred[lrow] = pr;
                        if (a.out_real) a.out_real[lrow] = y;
                    } else if (MODE == MODE_PREDICT) {
                        a.out_pred[lrow] = dot + 0.0;
                    }
                }
                if (kTrain) sdiff[t] = diff; // 0 for rows past the end: they add +-0 to the gradient
            }
            __syncthreads(); // (B) the stage is no longer read from here on

            if (t == 0) produce(slot);

            // ---- phase 2: gradient from the registers of phase 1
            if (kTrain) {
                double df[RPT];
                if (RL1 && RPT == 4) {
                    const double2 d01 = *reinterpret_cast<const double2*>(sdiff), d23 = *reinterpret_cast<const double2*>(sdiff + 2);
                    df[0] = d01.x; df[1] = d01.y; df[2 % RPT] = d23.x; df[3 % RPT] = d23.y;
                } else {
#pragma unroll
                    for (int k = 0; k < RPT; ++k) df[k] = sdiff[k * RL + drow];
                }
#pragma unroll
                for (int k = 0; k < RPT; ++k) {
#pragma unroll
                    for (int p = 0; p < CP; ++p) {
                        g0[p] = fma(df[k], x[k][p].x, g0[p]);
                        g1[p] = fma(df[k], x[k][p].y, g1[p]);
                    }
                }
            }
            st += a.stage_stride;
            if (++slot == S) { slot = 0; parity ^= 1u; st = stages; }
        }

        // ---- end of the row block: write its partial vector
        if (kTrain) {
            double* out = a.partials + blk * (long long)pv;
            if (RL1) {
#pragma unroll
                for (int p = 0; p < CP; ++p)
                    if (live[p]) *reinterpret_cast<double2*>(out + 2 * col2[p]) = make_double2(g0[p], g1[p]);
            } else { // fold the row lanes in lane order
                if (live[0]) *reinterpret_cast<double2*>(gbuf + 2 * t) = make_double2(g0[0], g1[0]);
            }
            if (t < R) { sred[2 * t] = my_loss; sred[2 * t + 1] = my_cnt; }
            __syncthreads();
            if (!RL1 && rl == 0 && live[0]) {
                double2 acc = make_double2(0.0, 0.0);
                for (int q = 0; q < RL; ++q) {
                    const double2 v = *reinterpret_cast<const double2*>(gbuf + 2 * (q * a.c2p + col2[0]));
                    acc.x += v.x;
                    acc.y += v.y;
                }
                *reinterpret_cast<double2*>(out + 2 * col2[0]) = acc;
            }
            if (t == 0) {
                double ls = 0.0, cn = 0.0;
                for (int q = 0; q < R; ++q) { ls += sred[2 * q]; cn += sred[2 * q + 1]; }
                out[ldx] = ls;
                out[ldx + 1] = cn;
            }
#pragma unroll
            for (int p = 0; p < CP; ++p) g0[p] = g1[p] = 0.0;
            my_loss = my_cnt = 0.0;
            if (kStats || !RL1) __syncthreads(); // sred / gbuf are reused below / by the next block
        }
        if (kStats) {
            __syncthreads();
            if (t == 0) {
                Stat7 acc;
                stat_zero(acc);
                for (int q = 0; q < R; ++q) {
                    Stat7 s;
                    stat_load(s, sstat + q * kStatStride);
                    stat_merge(acc, s);
                }
                stat_store(a.stat_partials + blk * kStatStride, acc);
            }
            __syncthreads();
            if (t < R)
                for (int q = 0; q < kStatStride; ++q) sstat[t * kStatStride + q] = 0.0;
        }
    }
}

} // namespace twtml
