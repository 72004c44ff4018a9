// csr_narrow.cuh -- CSR batches over a narrow feature space (d <= 2048; the reference's own 1004).
//
// With few columns a whole gradient vector fits in shared memory several times over.  One kernel then does
// BOTH halves of LeastSquaresGradient.compute [U] from a single read of the CSR entries, with no transpose:
//   * a CTA owns a contiguous range of rows (fixed at ingest: k_narrow_bounds balances entries + rows), hence
//     one contiguous stretch of the entry arrays.  A producer warp streams that stretch through a ring of
//     kNarrowStages tiles of kNarrowTile entries in shared memory, two 1-D bulk TMA copies per tile (values,
//     column indices), each tile with a "full" mbarrier (the copies' bytes) and an "empty" one (one arrival
//     per consumer warp);
//   * the consumer warps take the CTA's rows round robin, each with a PRIVATE gradient accumulator of d
//     doubles in shared memory.  A row is handled by a whole warp: the dot product with the weights (shared
//     memory) is folded by an xor butterfly so that all lanes hold the same residual, and diff * value is
//     added to the warp's accumulator.  Indices inside a row are strictly increasing (checked at ingest; the
//     featuriser emits them that way), so the lanes of one row never meet on a column, and a warp's rows
//     follow each other in order: the sum per column is a fixed sequence, no atomics;
//   * at the end the CTA adds its warps' accumulators in warp order and writes one partial vector
//     [g, lossSum, count]; k_shard_reduce adds the CTAs' vectors in CTA order.
// Per iteration and entry that is the 12 B of CSR from HBM and nothing else (the two-pass form reads 24 B
// per entry and gathers residuals through L2).  The grid and the row ranges are functions of the device and
// the batch only, so results are bit-reproducible run to run.
//
// Ring protocol (per consumer warp: `waited` tiles observed full, `released` tiles given back, both in tile
// order).  A tile is given back only after it was observed full: the producer fills tile t after every warp
// gave back t - kNarrowStages, so an arrival for t can never land in the phase of t - kNarrowStages.  A row of
// at most d <= kNarrowMaxD entries spans at most kNarrowStages - 1 tiles, so a warp waiting for a row's last
// tile never holds the tile the producer needs to get there.
#pragma once
#include "common.cuh"
#include "dense.cuh" // IterMode

namespace twtml {

constexpr int kNarrowMaxD = 2048;
constexpr int kNarrowMaxWarps = 16;                   // consumer warps (one more warp produces: 544 threads)
constexpr int kNarrowThreadsMax = (kNarrowMaxWarps + 1) * 32;
constexpr int kNarrowK = 3;                           // entries per lane kept in registers between the two halves
constexpr int kNarrowTileLog = 10;
constexpr int kNarrowTile = 1 << kNarrowTileLog;      // entries per tile
constexpr int kNarrowStages = 4;                      // (kNarrowStages - 1) * kNarrowTile >= kNarrowMaxD + kNarrowTile - 1
constexpr int kNarrowRing = kNarrowStages * kNarrowTile;
constexpr int kNarrowSlack = 8;                       // colidx / vals carry a few entries past the last one (copies are 16-byte wide)
constexpr long long kNarrowRowCost = 96;              // a row's fixed work costs about as much as this many entries
constexpr int kNarrowRedDoubles = kNarrowMaxWarps * (2 + kStatStride);
constexpr size_t kNarrowRingBytes = 128 + (size_t)kNarrowRing * 12; // barriers | values | column indices

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

// consumer warps per CTA: as many private accumulators as shared memory holds beside the ring and the weights
__host__ __device__ inline int narrow_warps(int d)
{
    const long long room = 227ll * 1024 - (long long)kNarrowRingBytes - 8ll * d - 8ll * kNarrowRedDoubles - 256;
    const long long w = room / (8ll * (d > 0 ? d : 1));
    return (int)(w > kNarrowMaxWarps ? kNarrowMaxWarps : w);
}
__host__ __device__ inline size_t narrow_smem_bytes(int d, int warps)
{
    return kNarrowRingBytes + (size_t)8 * d + 8 * kNarrowRedDoubles + (size_t)8 * d * warps;
}

__device__ __forceinline__ uint32_t keep_u32(uint32_t x)
{
    asm volatile("mov.u32 %0, %0;" : "+r"(x));
    return x;
}
__device__ __forceinline__ double lds_f64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ int lds_s32(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory"); }
__device__ __forceinline__ void mbar_arrive_addr(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
// One 32-entry slot of a row, branch-free (predicated): if (active) { c = col[..]; v = val[..]; acc += v * w[c]; }
__device__ __forceinline__ void narrow_slot_dot(int active, uint32_t caddr, uint32_t vaddr, uint32_t ws, int& c, double& v, double& acc)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .u32 wa;\n"
        ".reg .f64 wv;\n"
        "setp.ne.s32 p, %3, 0;\n"
        "@p ld.shared.s32 %0, [%4];\n"
        "@p ld.shared.f64 %1, [%5];\n"
        "@p mad.lo.u32 wa, %0, 8, %6;\n"
        "@p ld.shared.f64 wv, [wa];\n"
        "@p fma.rn.f64 %2, %1, wv, %2;\n"
        "}\n"
        : "+r"(c), "+d"(v), "+d"(acc)
        : "r"(active), "r"(caddr), "r"(vaddr), "r"(ws));
}
// ... and of the second half: if (active) g[c] += diff * v
__device__ __forceinline__ void narrow_slot_add(int active, int c, double v, double diff, uint32_t gs)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .u32 ga;\n"
        ".reg .f64 g;\n"
        "setp.ne.s32 p, %0, 0;\n"
        "@p mad.lo.u32 ga, %1, 8, %2;\n"
        "@p ld.shared.f64 g, [ga];\n"
        "@p fma.rn.f64 g, %3, %4, g;\n"
        "@p st.shared.f64 [ga], g;\n"
        "}\n" ::"r"(active),
        "r"(c), "r"(gs), "d"(diff), "d"(v)
        : "memory");
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
__global__ void __launch_bounds__(kNarrowThreadsMax, 1) k_csr_narrow(const NarrowArgs a)
{
    constexpr bool kTrain = (MODE == MODE_TRAIN || MODE == MODE_TRAIN_STATS);
    constexpr bool kStats = (MODE == MODE_TRAIN_STATS || MODE == MODE_PREDICT_STATS);
    constexpr bool kPerRow = kStats || MODE == MODE_PREDICT; // something is written per row
    constexpr unsigned kMask = kNarrowRing - 1;
    if (kTrain && a.state->done) return;

    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + kNarrowStages;
    double* val_ring = reinterpret_cast<double*>(smem + 128);
    int* col_ring = reinterpret_cast<int*>(smem + 128 + (size_t)kNarrowRing * 8);
    double* w_s = reinterpret_cast<double*>(smem + kNarrowRingBytes);
    double* sred = w_s + a.d;                // per warp: (lossSum, count), then the stats tuples
    double* gacc = sred + kNarrowRedDoubles; // warps x d
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, W = a.warps, d = a.d;
    const unsigned long long iter_seed =
        splitmix64((unsigned long long)(42 + (a.iter_override > 0 ? a.iter_override : (kTrain ? a.state->iter : 0))));
    const bool sampling = kTrain && a.frac < 1.0;

    // the CTA's rows [C0, C1) and its stretch [E0, E1) of the entry arrays (entry indices fit 32 bits, checked at ingest);
    // tile tl holds entries [Ea + tl * kNarrowTile, + kNarrowTile), Ea = E0 rounded down to a 16-byte boundary of colidx
    const long long C0 = a.bounds[blockIdx.x], C1 = a.bounds[blockIdx.x + 1];
    const int E0 = (int)a.rowptr[C0], E1 = (int)a.rowptr[C1], Ea = E0 & ~3;
    const int ntiles = E1 > E0 ? (E1 - Ea + kNarrowTile - 1) >> kNarrowTileLog : 0;
    const uint64_t policy = l2_evict_first_policy();
    auto fill = [&](int tl) { // producer lane only
        const int s = tl & (kNarrowStages - 1);
        const int start = Ea + tl * kNarrowTile;
        const int cnt = min(kNarrowTile, (E1 - start + 3) & ~3);
        mbar_expect_tx(&full[s], (uint32_t)cnt * 12u);
        tma_load_1d(val_ring + s * kNarrowTile, a.vals + start, (uint32_t)cnt * 8u, &full[s], policy);
        tma_load_1d(col_ring + s * kNarrowTile, a.colidx + start, (uint32_t)cnt * 4u, &full[s], policy);
    };
    if (t == W * 32) { // the producer lane: barriers, then the first tiles are on their way while the CTA sets up
        for (int s = 0; s < kNarrowStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], (uint32_t)W); }
        mbar_fence_init();
        for (int tl = 0; tl < min(ntiles, kNarrowStages); ++tl) fill(tl);
    }
    for (int j = t; j < d; j += blockDim.x) w_s[j] = a.w[j];
    if (kTrain)
        for (int j = t; j < W * d; j += blockDim.x) gacc[j] = 0.0;
    __syncthreads();

    double my_loss = 0.0, my_cnt = 0.0, keep_acc = 0.0, keep_y = 0.0;
    Stat7 my_stat;
    stat_zero(my_stat);

    if (warp == W) {
        // ---- producer: one lane keeps the ring full
        if (lane == 0) {
            for (int tl = kNarrowStages; tl < ntiles; ++tl) {
                mbar_wait(&empty[tl & (kNarrowStages - 1)], (uint32_t)((tl / kNarrowStages - 1) & 1));
                fill(tl);
            }
        }
    } else {
        // ---- consumers: warp w takes rows C0 + w, C0 + w + W, ...; local row j is row C0 + w + W j (row counts fit 32 bits,
        // checked at ingest)
        const int span = (int)(C1 - C0);
        const int R1 = span > warp ? (span - warp + W - 1) / W : 0;
        auto rowof = [&](int j) { return C0 + (warp + W * j); };
        // shared-memory addresses, kept in registers (opaque to the compiler, which would otherwise rebuild them at every use)
        const uint32_t ws = keep_u32(smem_u32(w_s)), gs = keep_u32(smem_u32(gacc + (size_t)warp * d)), vs = keep_u32(smem_u32(val_ring)),
                       cs = keep_u32(smem_u32(col_ring)), fb = keep_u32(smem_u32(full)), eb = keep_u32(smem_u32(empty));
        unsigned waited = 0, released = 0;
        auto wait_tile = [&]() {
            mbar_wait_addr(fb + 8u * (waited & (kNarrowStages - 1)), (waited / kNarrowStages) & 1u);
            ++waited;
        };
        auto release_tile = [&]() { // only a tile that was observed full
            if (waited <= released) wait_tile();
            if (lane == 0) mbar_arrive_addr(eb + 8u * (released & (kNarrowStages - 1)));
            ++released;
        };
        // Row pointers and labels: lane i holds those of local row 32 c + i for the current chunk c of 32 rows and for the
        // next one; a row's come out of them by shuffle, so a row never waits on a load of its own bounds.
        int lo1 = 0, hi1 = 0, lo2, hi2;
        double y1 = 0.0, y2;
        auto load_chunk = [&](int b0, int& lo, int& hi, double& y) {
            const int j0 = b0 + lane;
            const bool in = j0 < R1;
            const long long row = in ? rowof(j0) : C0; // C0 < C1 <= n here: a valid label index
            lo = (int)a.rowptr[in ? row : C1];
            hi = (int)a.rowptr[in ? row + 1 : C1];
            y = a.labels[row];
        };
        if (R1 > 0) load_chunk(0, lo2, hi2, y2);
        for (int j = 0; j < R1; ++j) {
            const int sel = j & 31;
            if (sel == 0) { // the chunk requested 32 rows ago becomes the current one
                lo1 = lo2;
                hi1 = hi2;
                y1 = y2;
                if (j + 32 < R1) load_chunk(j + 32, lo2, hi2, y2);
            }
            const int kb = __shfl_sync(0xffffffffu, lo1, sel), ke = __shfl_sync(0xffffffffu, hi1, sel);
            const double y = __shfl_sync(0xffffffffu, y1, sel);
            const int ne = ke - kb;
            if (ne > 0) { // give back the tiles before the row's first entry, see those up to its last entry full
                const unsigned tb = (unsigned)(kb - Ea) >> kNarrowTileLog, te = (unsigned)(ke - 1 - Ea) >> kNarrowTileLog;
#pragma unroll 1
                while (released < tb) release_tile();
#pragma unroll 1
                while (waited <= te) wait_tile();
            }
            // w . x: the first kNarrowK entries of every lane stay in registers for the second half
            const unsigned o0 = (unsigned)(kb - Ea + lane);
            int c[kNarrowK];
            double v[kNarrowK];
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < kNarrowK; ++i) {
                const unsigned idx = (o0 + 32u * i) & kMask;
                c[i] = 0;
                v[i] = 0.0;
                narrow_slot_dot(lane + 32 * i < ne, cs + 4u * idx, vs + 8u * idx, ws, c[i], v[i], acc);
            }
#pragma unroll 1
            for (int k = lane + 32 * kNarrowK; k < ne; k += 32) {
                const unsigned idx = (unsigned)(kb - Ea + k) & kMask;
                acc = fma(lds_f64(vs + 8u * idx), lds_f64(ws + 8u * (unsigned)lds_s32(cs + 4u * idx)), acc);
            }
#pragma unroll
            for (int mk = 16; mk >= 1; mk >>= 1) acc += shfl_xor_f64(acc, mk); // the same bits in every lane
            if (kTrain) {
                const bool keep = !sampling || row_sampled(iter_seed, a.row0_global + (unsigned long long)rowof(j), a.frac);
                if (keep) {
                    const double diff = acc - y;
                    my_loss += diff * diff / 2.0;
                    my_cnt += 1.0;
#pragma unroll
                    for (int i = 0; i < kNarrowK; ++i) narrow_slot_add(lane + 32 * i < ne, c[i], v[i], diff, gs);
#pragma unroll 1
                    for (int k = lane + 32 * kNarrowK; k < ne; k += 32) {
                        const unsigned idx = (unsigned)(kb - Ea + k) & kMask;
                        const uint32_t ga = gs + 8u * (unsigned)lds_s32(cs + 4u * idx);
                        sts_f64(ga, fma(diff, lds_f64(vs + 8u * idx), lds_f64(ga)));
                    }
                    __syncwarp(); // the next row's lanes may meet these columns
                }
            }
            if (kPerRow) { // lane i keeps row i of every 32 and writes them out together
                if (lane == sel) { keep_acc = acc; keep_y = y; }
                if (sel == 31 || j == R1 - 1) {
                    if (lane <= sel) {
                        const long long row = rowof(j - sel + lane);
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
        }
#pragma unroll 1
        while (released < (unsigned)ntiles) release_tile(); // let the producer (and the slower warps) finish the stream
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
        if (lane == 0 && warp < W) { sred[2 * warp] = my_loss; sred[2 * warp + 1] = my_cnt; }
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
        if (lane == 0 && warp < W) stat_store(sst + warp * kStatStride, my_stat);
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

// ---- ingest: the order of a row's entries inside the copy the single-pass kernel streams -----------------------------
// The kernel's shared-memory traffic, not HBM, is its busiest unit (profiles/r02_ncu_csr_narrow_*.csv: the LSU data pipe is
// 63-66 % busy and 40 % of its wavefronts are bank conflicts).  The conflicts come from the gathers of w[col] and the
// read-modify-writes of g[col]: 8-byte accesses of a half-warp are conflict-free only if its 16 columns fall into 16
// different 8-byte banks, i.e. differ mod 16, and the columns of 16 consecutive entries of a row are as good as random.
// But a row's entries may be visited in ANY order (no column occurs twice in a row), so ingest lays them out for the
// kernel: sort the row by (col mod 16) -- keeping column order inside a class -- and deal that sequence round robin over
// the row's 16-entry half-slots.  Entries of one class then sit in different half-slots (as long as the class has no
// more entries than the row has half-slots), and each half-warp access touches (nearly) 16 different banks.
// The batch keeps the caller's order as well: export, the two-pass kernels and the checks read that one.
static __global__ void __launch_bounds__(256) k_narrow_permute(const long long* __restrict__ rowptr, long long n, const int* __restrict__ colidx,
                                                                const double* __restrict__ vals, int* __restrict__ pcol, double* __restrict__ pval)
{
    __shared__ int cnt[8][16];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const long long nw = (long long)gridDim.x * 8;
    for (long long r = (long long)blockIdx.x * 8 + wib; r < n; r += nw) {
        const long long kb = rowptr[r];
        const int m = (int)(rowptr[r + 1] - kb);
        if (m <= 16) { // a single half-slot: nothing to arrange
            if (lane < m) { pcol[kb + lane] = colidx[kb + lane]; pval[kb + lane] = vals[kb + lane]; }
            continue;
        }
        if (lane < 16) cnt[wib][lane] = 0;
        __syncwarp();
        for (int b = 0; b < m; b += 32) { // entries per class
            const int e = b + lane;
            const bool in = e < m;
            const int cls = in ? (colidx[kb + e] & 15) : 16 + lane; // lanes past the row's end: keys of their own
            const unsigned peers = __match_any_sync(0xffffffffu, cls);
            if (in && (peers & lt) == 0u) cnt[wib][cls] += __popc(peers); // the group's lowest lane counts for it
            __syncwarp();
        }
        const int tot = lane < 16 ? cnt[wib][lane] : 0;
        int incl = tot;
#pragma unroll
        for (int dl = 1; dl < 16; dl <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, incl, dl);
            if (lane >= dl) incl += up;
        }
        __syncwarp();
        if (lane < 16) cnt[wib][lane] = incl - tot; // where the class's next entry goes in the class-sorted sequence
        __syncwarp();
        const int H = (m + 15) >> 4, rr = m - 16 * (H - 1); // half-slots; entries of the last one
        for (int b = 0; b < m; b += 32) {
            const int e = b + lane;
            const bool in = e < m;
            const int c = in ? colidx[kb + e] : 0;
            const int cls = in ? (c & 15) : 16 + lane;
            const unsigned peers = __match_any_sync(0xffffffffu, cls);
            const int within = __popc(peers & lt);
            const int k = in ? cnt[wib][cls] + within : 0;
            __syncwarp();
            if (in && within == 0) cnt[wib][cls] += __popc(peers);
            __syncwarp();
            if (in) { // position k of the sorted sequence -> (half-slot, place): places below rr exist in all H half-slots
                int hs, tpl;
                if (k < rr * H) {
                    tpl = k / H;
                    hs = k - tpl * H;
                } else {
                    const int k2 = k - rr * H, q2 = k2 / (H - 1);
                    tpl = rr + q2;
                    hs = k2 - q2 * (H - 1);
                }
                const long long dst = kb + hs * 16 + tpl;
                pcol[dst] = c;
                pval[dst] = vals[kb + e];
            }
        }
        __syncwarp();
    }
}

// ---- ingest: first row of every CTA, balancing entries + kNarrowRowCost per row (a row's fixed work -- bounds, butterfly,
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
