// csr_narrow.cuh -- CSR batches over a narrow feature space (d <= 4096; the reference's own 1004).
//
// With few columns a whole gradient vector fits in shared memory several times over.  One kernel then does
// BOTH halves of LeastSquaresGradient.compute [U] from a single read of the CSR entries, with no transpose:
//   * every warp owns a contiguous range of rows (fixed at ingest: k_narrow_bounds balances entries + rows)
//     and a PRIVATE gradient accumulator of d doubles in shared memory;
//   * a row is handled by the whole warp: its entries go to registers (4 per lane, prefetched two rows
//     ahead), the dot product with the weights (shared memory) is folded by an xor butterfly so that all
//     lanes hold the same residual, and diff * value is added to the warp's accumulator straight from those
//     registers.  Indices inside a row are strictly increasing (checked at ingest; the featuriser emits
//     them that way), so the lanes of one row never meet on a column, and rows follow each other in order:
//     the sum per column is a fixed sequence, no atomics;
//   * at the end the CTA adds its warps' accumulators in warp order and writes one partial vector
//     [g, lossSum, count]; k_shard_reduce adds the CTAs' vectors in CTA order.
// Per iteration and entry that is the 12 B of CSR from HBM and nothing else (the two-pass form reads 24 B
// per entry and gathers residuals through L2).  The grid and the row ranges are functions of the device and
// the batch only, so results are bit-reproducible run to run.
#pragma once
#include "common.cuh"
#include "dense.cuh" // IterMode

namespace twtml {

constexpr int kNarrowMaxD = 4096;
constexpr int kNarrowMaxWarps = 16;
constexpr int kNarrowK = 4;       // entries per lane held in registers (rows up to 128 entries are read once)
constexpr int kNarrowRedDoubles = kNarrowMaxWarps * (2 + kStatStride);

struct NarrowArgs {
    const long long* rowptr; // n + 1
    const long long* bounds; // grid * warps + 1: first row of every warp
    const int* colidx;
    const double* vals;
    const double* labels;
    const double* w;
    const TrainState* state;
    double* partials;      // gridDim.x x (d + 2)
    double* stat_partials; // gridDim.x x kStatStride
    double* out_pred;
    double* out_real;
    long long n;
    int d, warps;
    double frac;
    unsigned long long row0_global;
    int iter_override;
};

// warps per CTA: as many private accumulators as shared memory holds beside the weights
__host__ __device__ inline int narrow_warps(int d)
{
    const long long room = 227ll * 1024 - 8ll * d - 8ll * kNarrowRedDoubles - 256;
    const long long w = room / (8ll * (d > 0 ? d : 1));
    return (int)(w > kNarrowMaxWarps ? kNarrowMaxWarps : w);
}
__host__ __device__ inline size_t narrow_smem_bytes(int d, int warps)
{
    return (size_t)8 * d + 8 * kNarrowRedDoubles + (size_t)8 * d * warps;
}

struct NarrowRow {
    int c[kNarrowK];
    double v[kNarrowK];
    int kb, ne; // first entry and number of entries (all of them, also those beyond the registers)
    double y;
};

__device__ __forceinline__ void narrow_fetch(const NarrowArgs& a, long long r, long long r1, int kb, int ke, int lane, NarrowRow& o)
{
    o.kb = kb;
    o.ne = ke - kb;
    o.y = 0.0;
#pragma unroll
    for (int i = 0; i < kNarrowK; ++i) { o.c[i] = 0; o.v[i] = 0.0; }
    if (r < r1) {
        o.y = a.labels[r];
#pragma unroll
        for (int i = 0; i < kNarrowK; ++i) {
            const int k = kb + lane + 32 * i;
            if (k < ke) { o.c[i] = __ldcs(a.colidx + k); o.v[i] = __ldcs(a.vals + k); }
        }
    }
}

// Stat7 tuples of a warp -> lane 0, along a fixed tree
__device__ __forceinline__ void narrow_warp_stat(Stat7& s)
{
#pragma unroll
    for (int m = 1; m < 32; m <<= 1) {
        Stat7 o;
        o.n = __shfl_down_sync(0xffffffffu, s.n, m);
        o.mu_y = __shfl_down_sync(0xffffffffu, s.mu_y, m);
        o.m2_y = __shfl_down_sync(0xffffffffu, s.m2_y, m);
        o.mu_p = __shfl_down_sync(0xffffffffu, s.mu_p, m);
        o.m2_p = __shfl_down_sync(0xffffffffu, s.m2_p, m);
        o.mu_e = __shfl_down_sync(0xffffffffu, s.mu_e, m);
        o.bad = __shfl_down_sync(0xffffffffu, s.bad, m);
        if (((threadIdx.x & 31) & (2 * m - 1)) == 0) stat_merge(s, o);
    }
}

template <int MODE>
__global__ void __launch_bounds__(kNarrowMaxWarps * 32, 1) k_csr_narrow(const NarrowArgs a)
{
    constexpr bool kTrain = (MODE == MODE_TRAIN || MODE == MODE_TRAIN_STATS);
    constexpr bool kStats = (MODE == MODE_TRAIN_STATS || MODE == MODE_PREDICT_STATS);
    constexpr bool kPerRow = kStats || MODE == MODE_PREDICT; // something is written per row
    if (kTrain && a.state->done) return;

    extern __shared__ __align__(16) unsigned char smem[];
    double* w_s = reinterpret_cast<double*>(smem);
    double* sred = w_s + a.d;               // per warp: (lossSum, count), then the stats tuples
    double* gacc = sred + kNarrowRedDoubles; // warps x d
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, W = a.warps, d = a.d;
    double* my_g = gacc + (size_t)warp * d;
    const unsigned long long iter_seed =
        splitmix64((unsigned long long)(42 + (a.iter_override > 0 ? a.iter_override : (kTrain ? a.state->iter : 0))));
    const bool sampling = kTrain && a.frac < 1.0;

    for (int j = t; j < d; j += blockDim.x) w_s[j] = a.w[j];
    if (kTrain)
        for (int j = t; j < W * d; j += blockDim.x) gacc[j] = 0.0;
    __syncthreads();

    const long long gw = (long long)blockIdx.x * W + warp;
    const long long R0 = a.bounds[gw], R1 = a.bounds[gw + 1];
    double my_loss = 0.0, my_cnt = 0.0, keep_acc = 0.0, keep_y = 0.0;
    Stat7 my_stat;
    stat_zero(my_stat);
    if (R0 < R1) {
        // Row pointers: lane i holds those of row base + i for two chunks of 32 rows (entry indices fit 32 bits, checked
        // at ingest); a row's bounds come out of them by shuffle, so nothing on the path to a row's loads waits on memory.
        long long base = R0;
        int lo1, hi1, lo2, hi2;
        auto load_chunk = [&](long long b0, int& lo, int& hi) {
            const long long i0 = b0 + lane;
            lo = (int)a.rowptr[i0 < R1 ? i0 : R1];
            hi = (int)a.rowptr[i0 + 1 < R1 ? i0 + 1 : R1];
        };
        load_chunk(base, lo1, hi1);
        load_chunk(base + 32, lo2, hi2);
        auto fetch = [&](long long rr, NarrowRow& o) {
            const int sel = (int)(rr - base);
            const int kb = __shfl_sync(0xffffffffu, sel < 32 ? lo1 : lo2, sel & 31);
            const int ke = __shfl_sync(0xffffffffu, sel < 32 ? hi1 : hi2, sel & 31);
            narrow_fetch(a, rr, R1, kb, ke, lane, o);
        };
        auto process = [&](const NarrowRow& cur, long long r) {
            const int ne = cur.ne;
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < kNarrowK; ++i)
                if (lane + 32 * i < ne) acc = fma(cur.v[i], w_s[cur.c[i]], acc);
            if (ne > 32 * kNarrowK)
                for (int k = cur.kb + 32 * kNarrowK + lane; k < cur.kb + ne; k += 32) acc = fma(a.vals[k], w_s[a.colidx[k]], acc);
#pragma unroll
            for (int mk = 16; mk >= 1; mk >>= 1) acc += shfl_xor_f64(acc, mk); // the same bits in every lane
            if (kTrain) {
                const bool keep = !sampling || row_sampled(iter_seed, a.row0_global + (unsigned long long)r, a.frac);
                if (keep) {
                    const double diff = acc - cur.y;
                    my_loss += diff * diff / 2.0;
                    my_cnt += 1.0;
#pragma unroll
                    for (int i = 0; i < kNarrowK; ++i)
                        if (lane + 32 * i < ne) my_g[cur.c[i]] += diff * cur.v[i];
                    if (ne > 32 * kNarrowK)
                        for (int k = cur.kb + 32 * kNarrowK + lane; k < cur.kb + ne; k += 32) my_g[a.colidx[k]] += diff * a.vals[k];
                    __syncwarp(); // the next row's lanes may meet these columns
                }
            }
            if (kPerRow) { // lane i keeps row i of every 32 and writes them out together
                const int slot = (int)((r - R0) & 31);
                if (lane == slot) { keep_acc = acc; keep_y = cur.y; }
                if (slot == 31 || r == R1 - 1) {
                    const long long row = r - slot + lane;
                    if (row <= r) {
                        if (kStats) {
                            const double pr = round_half_up(keep_acc + 0.0);
                            stat_add(my_stat, keep_y, pr);
                            if (a.out_pred) a.out_pred[row] = pr;
                            if (a.out_real) a.out_real[row] = keep_y;
                        } else {
                            a.out_pred[row] = keep_acc + 0.0;
                        }
                    }
                }
            }
        };
        // Four row buffers in rotation, written out by hand (a register copy would wait for the loads it copies): the
        // entries of a row are requested three rows before its arithmetic.
        NarrowRow A, B, C, D;
        fetch(R0, A);
        fetch(R0 + 1, B);
        fetch(R0 + 2, C);
        for (long long r = R0; r < R1; r += 4) {
            if (r - base == 32) { // the next chunk of row pointers was requested 32 rows ago
                base += 32;
                lo1 = lo2;
                hi1 = hi2;
                load_chunk(base + 32, lo2, hi2);
            }
            fetch(r + 3, D);
            process(A, r);
            if (r + 1 >= R1) break;
            fetch(r + 4, A);
            process(B, r + 1);
            if (r + 2 >= R1) break;
            fetch(r + 5, B);
            process(C, r + 2);
            if (r + 3 >= R1) break;
            fetch(r + 6, C);
            process(D, r + 3);
        }
    }
    __syncthreads();

    // ---- this CTA's partial vector: the warps' accumulators in warp order
    if (kTrain) {
        double* out = a.partials + (long long)blockIdx.x * (d + 2);
        for (int j = t; j < d; j += blockDim.x) {
            double s = gacc[j];
            for (int q = 1; q < W; ++q) s += gacc[(size_t)q * d + j];
            out[j] = s;
        }
        if (lane == 0) { sred[2 * warp] = my_loss; sred[2 * warp + 1] = my_cnt; }
        __syncthreads();
        if (t == 0) {
            double ls = 0.0, cn = 0.0;
            for (int q = 0; q < W; ++q) { ls += sred[2 * q]; cn += sred[2 * q + 1]; }
            out[d] = ls;
            out[d + 1] = cn;
        }
    }
    if (kStats) {
        double* sst = sred + 2 * kNarrowMaxWarps;
        narrow_warp_stat(my_stat);
        if (lane == 0) stat_store(sst + warp * kStatStride, my_stat);
        __syncthreads();
        if (t == 0) {
            Stat7 tot;
            stat_zero(tot);
            for (int q = 0; q < W; ++q) {
                Stat7 s;
                stat_load(s, sst + q * kStatStride);
                stat_merge(tot, s);
            }
            stat_store(a.stat_partials + (long long)blockIdx.x * kStatStride, tot);
        }
    }
}

// ---- ingest: first row of every warp, balancing entries + 4 per row (a row costs about that beside its entries)
static __global__ void k_narrow_bounds(const long long* __restrict__ rowptr, long long n, long long G, long long* __restrict__ bounds)
{
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g > G) return;
    if (g == 0) { bounds[0] = 0; return; }
    if (g == G) { bounds[G] = n; return; }
    const long long total = rowptr[n] + 4 * n;
    const long long tgt = (long long)(((__int128)total * g) / G);
    long long lo = 0, hi = n; // first r with rowptr[r] + 4 r >= tgt
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (rowptr[mid] + 4 * mid >= tgt) hi = mid; else lo = mid + 1;
    }
    bounds[g] = lo;
}

} // namespace twtml
