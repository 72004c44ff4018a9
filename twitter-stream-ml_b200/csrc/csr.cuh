// csr.cuh -- the sparse (CSR) form of the per-iteration pass (K3).
//
// MLlib's sparse branch of LeastSquaresGradient.compute [U] is, per row,
//     diff = sum_k values(k) * w(indices(k)) - y      (BLAS.dot sparse.dense, stored order)
//     gradSum(indices(k)) += diff * values(k)         (BLAS.axpy)
// The first line is a row-wise pass, the second a scatter.  A scatter with floating-point atomics
// would make the sums depend on the order the hardware happened to serve them, so instead the
// batch keeps its transpose (CSC, built once at ingest, entries of a column in row order) and every
// iteration runs
//   k_csr_rows      rows -> diff[r] (+ lossSum, count, statistics per row block, fixed order)
//   k_csc_windows   one warp per window of 256 consecutive CSC entries: products diff[row] * value,
//                   segmented inclusive scan keyed by the column-head bit mask; columns that lie
//                   inside the window are written to gradSum directly, the pieces of a column that
//                   crosses a window edge go to two per-window slots
//   k_csc_edges     one warp per window that ends such a column: adds its pieces in window order
// Work is balanced by entries, not by columns, so heavily skewed columns (char-bigram hashes) cost
// the same as uniform ones.  Cost: the values are read twice per iteration (12 B/nnz in each
// form); see DESIGN.md.
#pragma once
#include "common.cuh"
#include "dense.cuh" // IterMode

namespace twtml {

constexpr int kCsrThreads = 256; // LPR lanes per row (4 / 8 / 16 by average row length), kCsrThreads / LPR rows at a time
constexpr int kCsrUnroll = 4;                // entries per lane and row whose loads and gathers are in flight together
constexpr int kWin = 256;                   // CSC entries per window (8 head-mask words)

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

template <int MODE, int LPR>
__global__ void __launch_bounds__(kCsrThreads) k_csr_rows(const CsrIterArgs a)
{
    constexpr int kCsrGroups = kCsrThreads / LPR;
    constexpr bool kTrain = (MODE == MODE_TRAIN || MODE == MODE_TRAIN_STATS);
    constexpr bool kStats = (MODE == MODE_TRAIN_STATS || MODE == MODE_PREDICT_STATS);
    if (kTrain && a.state->done) return;
    __shared__ double sred[kCsrGroups * kStatStride];
    const int t = threadIdx.x, grp = t / LPR, l8 = t % LPR;
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
        // two rows per 8-lane group and two entries per lane in flight: the loads of a row's entries
        // are issued before the dependent gathers of w, four gathers outstanding per lane
        for (int r0 = 0; r0 < nrows; r0 += 2 * kCsrGroups) {
          double accs[2];
          bool valids[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) { valids[h] = r0 + h * kCsrGroups + grp < nrows; accs[h] = 0.0; }
          {
            long long k[2], e[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const long long row = row0 + r0 + h * kCsrGroups + grp;
                k[h] = valids[h] ? a.rowptr[row] + l8 : 0;
                e[h] = valids[h] ? a.rowptr[row + 1] : 0;
            }
            constexpr int U = kCsrUnroll; // entries per lane and row in flight
            while (k[0] < e[0] || k[1] < e[1]) {
                int ci[2][U];
                double vv[2][U];
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const long long kk = k[h] + LPR * u;
                        const bool p = kk < e[h];
                        ci[h][u] = p ? a.colidx[kk] : 0;
                        vv[h][u] = p ? a.vals[kk] : 0.0;
                    }
                double ww[2][U];
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int u = 0; u < U; ++u) ww[h][u] = a.w[ci[h][u]];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
#pragma unroll
                    for (int u = 0; u < U; ++u) accs[h] = fma(vv[h][u], ww[h][u], accs[h]); // entries k, k+LPR, k+2 LPR, ... in this order per lane
                    k[h] += U * LPR;
                }
            }
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int r = r0 + h * kCsrGroups + grp;
            const bool valid = valids[h];
            const long long row = row0 + r;
            double acc = accs[h];
#pragma unroll
            for (int mk = LPR / 2; mk >= 1; mk >>= 1) acc += shfl_xor_f64(acc, mk);
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

// The transpose and its window index (built once per batch, csc_build in twtml_sparse.cu):
//   rowidx / cvals   entries sorted by (column, row)
//   headbits         bit e = entry e is the first of its column; bit nnz is set too (a virtual head
//                    that closes the last column); zero words pad the array to (nwin + 1) * 8
//   nzcol            the non-empty columns in order; win_first_nz[w] = heads before window w
//   wkey[w]          column of the window's first entry; colptr as usual
struct CscView {
    const int* rowidx;
    const double* cvals;
    const unsigned* headbits;
    const int* nzcol;
    const long long* win_first_nz;
    const int* wkey;
    const long long* colptr;
    long long nnz, nwin;
};

static __global__ void __launch_bounds__(256) k_csc_windows(const CscView c, const double* __restrict__ diff, double* __restrict__ g,
                                                            double* __restrict__ first_part, double* __restrict__ last_part,
                                                            const TrainState* gate)
{
    if (gate && gate->done) return;
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const unsigned le = lane == 31 ? 0xffffffffu : ((2u << lane) - 1u); // bits 0..lane
    for (long long w = warp0; w < c.nwin; w += nwarps) {
        const long long base = w * kWin;
        const unsigned myword = lane < 9 ? c.headbits[w * 8 + lane] : 0u; // this window's 8 words + the next one's first
        const long long colbase = c.win_first_nz[w];
        int heads_before = 0;
        double carry = 0.0;      // running sum of the segment left open by the previous round
        bool carry_head = false; // ... and whether that segment's head lies in this window
        unsigned hnext = 0;
        // all of the window's loads first (8 streaming pairs, then 8 gathers of diff in flight per lane)
        double prod[8];
        int ri[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const long long e = base + i * 32 + lane;
            const bool in = e < c.nnz;
            ri[i] = in ? c.rowidx[e] : -1;
            prod[i] = in ? c.cvals[e] : 0.0;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) prod[i] = ri[i] >= 0 ? diff[ri[i]] * prod[i] : 0.0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const unsigned hm = __shfl_sync(0xffffffffu, myword, i);
            hnext = __shfl_sync(0xffffffffu, myword, i + 1);
            const long long e = base + i * 32 + lane;
            double v = prod[i];
            const unsigned hb = hm & le;
            const int seg_start = hb ? 31 - __clz(hb) : -1;
#pragma unroll
            for (int dl = 1; dl < 32; dl <<= 1) { // segmented inclusive scan: a fixed summation tree
                const double up = __shfl_up_sync(0xffffffffu, v, dl);
                if (lane >= dl && lane - dl >= seg_start) v += up;
            }
            if (seg_start < 0) v += carry;
            const bool next_is_head = lane < 31 ? ((hm >> (lane + 1)) & 1u) : (hnext & 1u);
            if (e < c.nnz && next_is_head) { // this entry closes its column's run
                if (seg_start >= 0 || carry_head) g[c.nzcol[colbase + heads_before + __popc(hb) - 1]] = v;
                else first_part[w] = v; // the run came in from an earlier window and ends here
            }
            const double v31 = __shfl_sync(0xffffffffu, v, 31);
            if (hnext & 1u) { carry = 0.0; carry_head = false; }
            else { carry = v31; carry_head = carry_head || hm != 0u; }
            heads_before += __popc(hm);
        }
        if (lane == 0 && !(hnext & 1u)) { // a run leaves through the window's end
            if (carry_head) last_part[w] = carry;
            else first_part[w] = carry; // the whole window is the middle of one long column
        }
    }
}

// One warp per window: if a column that came in from an earlier window ends here, add its pieces
// (the tail piece of its first window, then the through / head pieces in window order).
static __global__ void __launch_bounds__(256) k_csc_edges(const CscView c, const double* __restrict__ first_part,
                                                          const double* __restrict__ last_part, double* __restrict__ g,
                                                          const TrainState* gate)
{
    if (gate && gate->done) return;
    // One thread looks at one window; a column that spans just two windows (the common case) is
    // finished by that thread, longer spans are then folded by the whole warp, lane-strided with a
    // fixed butterfly, one after the other.
    const int lane = threadIdx.x & 31;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nt = (long long)gridDim.x * blockDim.x;
    for (long long wbase = t0 - lane; wbase + 1 < c.nwin; wbase += nt) {
        const long long w = wbase + lane + 1;
        long long wa = -1; // >= 0: a column that came in from window wa ends in window w
        int col = 0;
        if (w < c.nwin && !(c.headbits[w * 8] & 1u)) {
            unsigned any = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) any |= c.headbits[w * 8 + i];
            if (any || (c.headbits[w * 8 + 8] & 1u)) {
                col = c.wkey[w];
                wa = c.colptr[col] / kWin;
                if (wa == w - 1) { g[col] = last_part[wa] + first_part[w]; wa = -1; }
            }
        }
        unsigned todo = __ballot_sync(0xffffffffu, wa >= 0);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const long long wa_s = __shfl_sync(0xffffffffu, wa, src), w_s = wbase + src + 1;
            const int col_s = __shfl_sync(0xffffffffu, col, src);
            double acc = 0.0;
            for (long long u = wa_s + 1 + lane; u <= w_s; u += 32) acc += first_part[u];
            for (int m = 16; m >= 1; m >>= 1) acc += shfl_xor_f64(acc, m);
            if (lane == 0) g[col_s] = last_part[wa_s] + acc;
        }
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

// head bit of every entry (and the virtual head at e == nnz), one 32-bit word per warp
static __global__ void k_csc_headbits(const int* __restrict__ keys, long long nnz, unsigned* __restrict__ headbits)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool head = e <= nnz && (e == 0 || e == nnz || keys[e] != keys[e - 1]);
    const unsigned m = __ballot_sync(0xffffffffu, head);
    if ((threadIdx.x & 31) == 0 && e <= nnz) headbits[e >> 5] = m;
}

// per window: number of real heads (the virtual one at nnz excluded), and the key of its first entry
static __global__ void k_csc_winmeta(const unsigned* __restrict__ headbits, const int* __restrict__ keys, long long nnz, long long nwin,
                                     long long* __restrict__ winheads, int* __restrict__ wkey)
{
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (w > nwin) return;
    if (w == nwin) { winheads[w] = 0; return; }
    int cnt = 0;
    for (int i = 0; i < 8; ++i) cnt += __popc(headbits[w * 8 + i]);
    const long long lo = w * kWin, hi = lo + kWin;
    if (nnz >= lo && nnz < hi) cnt -= 1; // the virtual head sits in this window
    winheads[w] = cnt;
    wkey[w] = lo < nnz ? keys[lo] : 0;
}

static __global__ void k_csc_nzcol(const int* __restrict__ keys, const unsigned* __restrict__ headbits,
                                   const long long* __restrict__ win_first_nz, long long nnz, int* __restrict__ nzcol)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nnz) return;
    const unsigned word = headbits[e >> 5];
    const int bit = (int)(e & 31);
    if (!((word >> bit) & 1u)) return;
    const long long w = e / kWin;
    int before = __popc(word & ((1u << bit) - 1u));
    for (long long i = w * 8; i < (e >> 5); ++i) before += __popc(headbits[i]);
    nzcol[win_first_nz[w] + before] = keys[e];
}

} // namespace twtml
