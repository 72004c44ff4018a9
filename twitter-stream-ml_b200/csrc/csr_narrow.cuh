// csr_narrow.cuh -- CSR batches over a narrow feature space (d <= 4096; the reference's own 1004).
//
// With few columns a whole gradient vector fits in shared memory several times over.  One kernel then does
// BOTH halves of LeastSquaresGradient.compute [U] from a single read of the CSR entries, with no transpose:
//   * every warp owns a contiguous range of rows (fixed at ingest: k_narrow_bounds balances entries + rows)
//     and a PRIVATE gradient accumulator of d doubles in shared memory;
//   * a row is handled by the whole warp: its entries go to registers (3 per lane, requested a pair of
//     rows ahead), the dot product with the weights (shared memory) is folded by an xor butterfly so that all
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
constexpr int kNarrowK = 3;       // entries per lane held in registers (rows up to 96 entries are read once)
constexpr long long kNarrowRowCost = 96;
constexpr int kNarrowRedDoubles = kNarrowMaxWarps * (2 + kStatStride);

struct NarrowArgs {
    const long long* rowptr; // n + 1
    const long long* bounds; // grid + 1: first row of every CTA
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

constexpr int kNarrowSlack = 32 * kNarrowK; // colidx / vals carry this many zeroed entries past the last one

struct NarrowRow {
    int c[kNarrowK];
    double v[kNarrowK];
    int kb, ne; // first entry and number of entries (all of them, also those beyond the registers)
};

__device__ __forceinline__ double lds_f64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory"); }

// Request a row's first 32 * kNarrowK entries: no guards at all.  Lanes past the row's end read the entries that follow it
// in the stream (the arrays end in kNarrowSlack zeroed entries) and never use them.
__device__ __forceinline__ void narrow_fetch(const NarrowArgs& a, int kb, int ke, int lane, NarrowRow& o)
{
    o.kb = kb;
    o.ne = ke - kb;
    const int* pc = a.colidx + (kb + lane);
    const double* pv = a.vals + (kb + lane);
#pragma unroll
    for (int i = 0; i < kNarrowK; ++i) {
        o.c[i] = __ldcs(pc + 32 * i);
        o.v[i] = __ldcs(pv + 32 * i);
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

    // The CTA owns rows [C0, C1); its warps take them round robin (warp w: C0 + w, C0 + w + W, ...), so that the CTA as a
    // whole walks one contiguous stretch of the entry arrays: neighbouring warps read neighbouring lines.  Below, rows are
    // counted per warp: local row j is row C0 + warp + W j, and R0 = 0, R1 = the warp's number of rows.
    const long long C0 = a.bounds[blockIdx.x], C1 = a.bounds[blockIdx.x + 1];
    const long long R0 = 0, R1 = C1 - C0 > warp ? (C1 - C0 - warp + W - 1) / W : 0;
    auto rowof = [&](long long j) { return C0 + warp + (long long)W * j; };
    double my_loss = 0.0, my_cnt = 0.0, keep_acc = 0.0, keep_y = 0.0;
    Stat7 my_stat;
    stat_zero(my_stat);
    if (R0 < R1) {
        // Row pointers and labels: lane i holds those of local row base + i for two chunks of 32 rows (entry indices fit 32
        // bits, checked at ingest); a row's come out of them by shuffle, so nothing on the path to a row's loads waits on memory.
        const uint32_t ws = smem_u32(w_s), gs = smem_u32(my_g);
        long long base = R0;
        int lo1, hi1, lo2, hi2;
        double y1, y2;
        auto load_chunk = [&](long long b0, int& lo, int& hi, double& y) {
            const long long j0 = b0 + lane;
            const bool in = j0 < R1;
            const long long row = rowof(in ? j0 : R1 - 1);
            lo = (int)a.rowptr[in ? row : C1];
            hi = (int)a.rowptr[in ? row + 1 : C1];
            y = a.labels[row];
        };
        load_chunk(base, lo1, hi1, y1);
        load_chunk(base + 32, lo2, hi2, y2);
        auto fetch = [&](long long rr, NarrowRow& o) {
            const int sel = (int)(rr - base);
            const int kb = __shfl_sync(0xffffffffu, sel < 32 ? lo1 : lo2, sel & 31);
            const int ke = __shfl_sync(0xffffffffu, sel < 32 ? hi1 : hi2, sel & 31);
            narrow_fetch(a, kb, ke, lane, o);
        };
        auto dot = [&](const NarrowRow& cur) { // this lane's share of w . x
            const int ne = cur.ne;
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < kNarrowK; ++i)
                if (lane + 32 * i < ne) acc = fma(cur.v[i], lds_f64(ws + 8u * (unsigned)cur.c[i]), acc);
            if (ne > 32 * kNarrowK)
                for (int k = cur.kb + 32 * kNarrowK + lane; k < cur.kb + ne; k += 32) acc = fma(a.vals[k], w_s[a.colidx[k]], acc);
            return acc;
        };
        auto finish = [&](const NarrowRow& cur, const double acc, long long r) {
            const int ne = cur.ne;
            const int sel = (int)(r - base);
            const double y = __shfl_sync(0xffffffffu, sel < 32 ? y1 : y2, sel & 31);
            if (kTrain) {
                const bool keep = !sampling || row_sampled(iter_seed, a.row0_global + (unsigned long long)rowof(r), a.frac);
                if (keep) {
                    const double diff = acc - y;
                    my_loss += diff * diff / 2.0;
                    my_cnt += 1.0;
#pragma unroll
                    for (int i = 0; i < kNarrowK; ++i)
                        if (lane + 32 * i < ne) {
                            const uint32_t ga = gs + 8u * (unsigned)cur.c[i];
                            sts_f64(ga, lds_f64(ga) + diff * cur.v[i]);
                        }
                    if (ne > 32 * kNarrowK)
                        for (int k = cur.kb + 32 * kNarrowK + lane; k < cur.kb + ne; k += 32) my_g[a.colidx[k]] += diff * a.vals[k];
                    __syncwarp(); // the next row's lanes may meet these columns
                }
            }
            if (kPerRow) { // lane i keeps row i of every 32 and writes them out together
                const int slot = (int)((r - R0) & 31);
                if (lane == slot) { keep_acc = acc; keep_y = y; }
                if (slot == 31 || r == R1 - 1) {
                    const long long row = rowof(r - slot + lane);
                    if (lane <= slot) {
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
        // Rows go two at a time: their dot products and butterflies are independent chains, interleaved; the additions
        // to the accumulator follow in row order.  Four row buffers, rotated by hand (a register copy would wait for the
        // loads it copies): a pair's entries are requested while the pair before it is worked on.
        auto pair = [&](const NarrowRow& X, const NarrowRow& Y, long long r) {
            double ax = dot(X), ay = dot(Y);
#pragma unroll
            for (int mk = 16; mk >= 1; mk >>= 1) { // the same bits in every lane
                const double ux = shfl_xor_f64(ax, mk), uy = shfl_xor_f64(ay, mk);
                ax += ux;
                ay += uy;
            }
            finish(X, ax, r);
            if (r + 1 < R1) finish(Y, ay, r + 1);
        };
        NarrowRow A, B, C, D;
        fetch(R0, A);
        fetch(R0 + 1, B);
        fetch(R0 + 2, C);
        fetch(R0 + 3, D);
        for (long long r = R0; r < R1; r += 4) {
            if (r - base == 32) { // the next chunk of row pointers and labels was requested 32 rows ago
                base += 32;
                lo1 = lo2;
                hi1 = hi2;
                y1 = y2;
                load_chunk(base + 32, lo2, hi2, y2);
            }
            pair(A, B, r);
            fetch(r + 4, A);
            fetch(r + 5, B);
            if (r + 2 >= R1) break;
            pair(C, D, r + 2);
            fetch(r + 6, C);
            fetch(r + 7, D);
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

// ---- ingest: first row of every warp, balancing entries + kNarrowRowCost per row (a row's fixed work -- bounds, butterfly,
// bookkeeping -- costs about as much as that many entries)
static __global__ void k_narrow_bounds(const long long* __restrict__ rowptr, long long n, long long G, long long* __restrict__ bounds)
{
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g > G) return;
    if (g == 0) { bounds[0] = 0; return; }
    if (g == G) { bounds[G] = n; return; }
    const long long total = rowptr[n] + kNarrowRowCost * n;
    const long long tgt = (long long)(((__int128)total * g) / G);
    long long lo = 0, hi = n; // first r with rowptr[r] + cost * r >= tgt
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (rowptr[mid] + kNarrowRowCost * mid >= tgt) hi = mid; else lo = mid + 1;
    }
    bounds[g] = lo;
}

} // namespace twtml
