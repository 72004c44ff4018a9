// reduce.cuh -- fixed-order reductions behind the iteration kernels (K4) and the statistics block.
//
//   k_shard_reduce   block partials -> one partial vector per shard   (combOp of treeAggregate [U])
//   [NCCL all-gather of the shard vectors when the job has more than one rank]
//   k_update         sum the shard vectors in shard order, then SimpleUpdater.compute [U] and
//                    GradientDescent.isConverged [U]; one CTA, so the order is fixed
//   k_stats_shard / k_stats_final   Chan merges of the StatCounter tuples in block / shard order,
//                    then the roundings of LinearRegression.scala:63-65
#pragma once
#include "common.cuh"

namespace twtml {

// grid (ceil(pv/32), spr), block (32, 32).  Thread (x, y): column bx*32+x, slice y of the shard's
// blocks, summed in block order; then the 32 slice sums are added in slice order.
static __global__ void __launch_bounds__(1024) k_shard_reduce(const double* __restrict__ partials, int pv, Plan plan,
                                                       double* __restrict__ shard_out /* spr x pv */,
                                                       const TrainState* state)
{
    if (state && state->done) return;
    __shared__ double sm[32][33];
    const int s = blockIdx.y;
    const int col = blockIdx.x * 32 + threadIdx.x;
    const long long b0 = plan.shard_blk0[s], B = plan.shard_blk0[s + 1] - b0;
    const long long lo = b0 + B * threadIdx.y / 32, hi = b0 + B * (threadIdx.y + 1) / 32;
    double acc = 0.0;
    if (col < pv) {
        long long b = lo;
        for (; b + 4 <= hi; b += 4) {
            const double v0 = partials[(b + 0) * pv + col], v1 = partials[(b + 1) * pv + col];
            const double v2 = partials[(b + 2) * pv + col], v3 = partials[(b + 3) * pv + col];
            acc += v0;
            acc += v1;
            acc += v2;
            acc += v3;
        }
        for (; b < hi; ++b) acc += partials[b * pv + col];
    }
    sm[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && col < pv) {
        double tot = 0.0;
#pragma unroll
        for (int q = 0; q < 32; ++q) tot += sm[q][threadIdx.x];
        shard_out[(long long)s * pv + col] = tot;
    }
}

// ---- exchange of the shard vectors over peer memory (NVLink) ------------------------------------
// Every rank owns one exchange buffer: lines[2 parities][kMaxShards * kPvCap] of 16 bytes, then the number of the
// last exchange this rank completed.  All ranks map all buffers (CUDA IPC between processes, direct peer access
// between threads of one process).  A line carries one double together with the number of the exchange it belongs
// to: {low word, seq} {high word, seq}, two 8-byte halves, each stored and loaded whole.  In exchange number seq a
// rank's k_shard_reduce_p2p stores every element of its shard vectors as such a line straight into EVERY rank's
// buffer (parity seq & 1); the update kernel of a rank reads a line until both halves show seq, and then holds the
// value.  Data and "it has arrived" travel in the same store, so the exchange costs one trip over NVLink and needs
// no fence, no counter and no separate flag (the first form of this exchange -- vectors, a system-scope fence, a
// flag per writer, another fence on the reading side -- took two round trips and measured slower than NCCL's
// all-gather on the 8 KB vectors of the 1004-column model: profiles/r02_SUMMARY.md).
// The sequence number lives ON THE DEVICE (one word of the local buffer, advanced by the update kernel
// after it has read the vectors): it counts the exchanges that really ran, not the launches the host
// enqueued -- after an early convergence the remaining launches of a call are no-ops and leave it alone.
// Every rank takes the same decisions from the same bits, so the words stay equal across ranks.  Two
// parities then suffice: a rank starts exchange seq + 2 only after its update kernel has read every line of
// exchange seq + 1, which a peer sends only after its own update kernel of exchange seq has finished reading
// parity seq & 1.  The buffer starts zeroed and seq at 1, so a line never written never matches.
constexpr int kPvCap = 4104;
constexpr size_t kXbufLines = (size_t)2 * kMaxShards * kPvCap;
constexpr size_t kXbufBytes = kXbufLines * 16 + 64;

struct PeerExchange {
    double* peer[kMaxShards]; // base of rank r's exchange buffer, as mapped here
    int world, rank;
};

__host__ __device__ __forceinline__ ulonglong2* xbuf_lines(double* base, int parity)
{
    return reinterpret_cast<ulonglong2*>(base) + (size_t)parity * kMaxShards * kPvCap;
}
// local word behind the lines: the number of the last exchange this rank completed
__host__ __device__ __forceinline__ unsigned long long* xbuf_seq(double* base) { return reinterpret_cast<unsigned long long*>(base) + 2 * kXbufLines; }

__device__ __forceinline__ void ll_store(ulonglong2* line, double v, unsigned tag)
{
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v), t = (unsigned long long)tag << 32;
    const unsigned long long lo = (bits & 0xffffffffull) | t, hi = (bits >> 32) | t;
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(line), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ ulonglong2 ll_load(const ulonglong2* line)
{
    ulonglong2 v;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(line) : "memory");
    return v;
}
__device__ __forceinline__ bool ll_ready(const ulonglong2& v, unsigned tag) { return (unsigned)(v.x >> 32) == tag && (unsigned)(v.y >> 32) == tag; }
__device__ __forceinline__ double ll_value(const ulonglong2& v)
{
    return __longlong_as_double((long long)((v.x & 0xffffffffull) | (v.y << 32)));
}

static __global__ void __launch_bounds__(1024) k_shard_reduce_p2p(const double* __restrict__ partials, int pv, Plan plan, PeerExchange px,
                                                           const TrainState* state)
{
    if (state && state->done) return;
    __shared__ double sm[32][33];
    const int s = blockIdx.y;
    const int col = blockIdx.x * 32 + threadIdx.x;
    const long long b0 = plan.shard_blk0[s], B = plan.shard_blk0[s + 1] - b0;
    const long long lo = b0 + B * threadIdx.y / 32, hi = b0 + B * (threadIdx.y + 1) / 32;
    double acc = 0.0;
    if (col < pv) {
        long long b = lo;
        for (; b + 4 <= hi; b += 4) {
            const double v0 = partials[(b + 0) * pv + col], v1 = partials[(b + 1) * pv + col];
            const double v2 = partials[(b + 2) * pv + col], v3 = partials[(b + 3) * pv + col];
            acc += v0;
            acc += v1;
            acc += v2;
            acc += v3;
        }
        for (; b < hi; ++b) acc += partials[b * pv + col];
    }
    sm[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    const unsigned long long seq = *xbuf_seq(px.peer[px.rank]) + 1ull; // advanced only by the update kernel that follows on this stream
    const int parity = (int)(seq & 1ull);
    // the 32 columns of this CTA go to the world's buffers, one warp row per destination rank
    if ((int)threadIdx.y < px.world && col < pv) {
        double tot = 0.0;
#pragma unroll
        for (int q = 0; q < 32; ++q) tot += sm[q][threadIdx.x];
        const size_t slot = ((size_t)px.rank * plan.spr + s) * pv + col; // packed nshards x pv like the NCCL path
        for (int r = threadIdx.y; r < px.world; r += 32) ll_store(xbuf_lines(px.peer[r], parity) + slot, tot, (unsigned)seq);
    }
}

// (lossSum, count) pairs of the CSR row blocks -> out[0..1]; one CTA: 1024 slices in block order, then a fixed tree
__device__ __forceinline__ double block_sum_1024(double v, double* sm32);
static __global__ void __launch_bounds__(1024) k_pair_reduce(const double* __restrict__ pairs, long long nb, double* __restrict__ out,
                                                      const TrainState* gate)
{
    __shared__ double sm32[32];
    if (gate && gate->done) return;
    const long long lo = nb * threadIdx.x / 1024, hi = nb * (threadIdx.x + 1) / 1024;
    double a = 0.0, c = 0.0;
    for (long long b = lo; b < hi; ++b) { a += pairs[2 * b]; c += pairs[2 * b + 1]; }
    a = block_sum_1024(a, sm32);
    c = block_sum_1024(c, sm32);
    if (threadIdx.x == 0) { out[0] = a; out[1] = c; }
}

struct UpdateArgs {
    const double* shards; // nshards x pv   (all ranks' shard vectors, shard order)
    double* w;            // d
    TrainState* state;
    double* loss_hist;    // max_iter
    int nshards, pv, d, ldx;
    int max_iter, updater;
    double step, reg, tol;
    // peer exchange (xbuf != null): the nshards x pv lines of exchange seq = *xbuf_seq + 1 are read from this rank's
    // exchange buffer as they arrive and left in `shards` (then this rank's staging copy)
    double* xbuf;
    int* host_done; // mapped host word set when the loop is over, so a single-GPU host can stop enqueuing no-op iterations
};

constexpr int kPeerTimeoutMark = 0x7E57; // TrainState::pad after a peer never showed up

// Peer exchange: collects this exchange's lines into a.shards as their writers (this GPU among them) deliver them.
// Returns false when a line did not arrive within ~10 s (the kernel then stops the training loop).  A thread asks for
// four lines before it looks at the first, so the waits of one thread overlap.
__device__ __forceinline__ bool wait_for_peers(UpdateArgs& a, unsigned long long& seq)
{
    __shared__ int timed_out;
    seq = 0ull;
    if (!a.xbuf) return true;
    seq = *xbuf_seq(a.xbuf) + 1ull;
    const unsigned tag = (unsigned)seq;
    const ulonglong2* lines = xbuf_lines(a.xbuf, (int)(seq & 1ull));
    double* stage = const_cast<double*>(a.shards);
    const int total = a.nshards * a.pv;
    if (threadIdx.x == 0) timed_out = 0;
    __syncthreads();
    const long long t0 = clock64();
    constexpr int U = 4;
    for (int base = 0; base < total; base += U * (int)blockDim.x) {
        ulonglong2 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int idx = base + u * (int)blockDim.x + (int)threadIdx.x;
            if (idx < total) v[u] = ll_load(lines + idx);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int idx = base + u * (int)blockDim.x + (int)threadIdx.x;
            if (idx >= total) continue;
            while (!ll_ready(v[u], tag)) {
                if (clock64() - t0 > 20000000000ll) { timed_out = 1; break; }
                v[u] = ll_load(lines + idx);
            }
            stage[idx] = ll_value(v[u]);
        }
    }
    __syncthreads(); // the staging copy is complete for every thread of this CTA
    if (timed_out) {
        if (threadIdx.x == 0) { a.state->done = 1; a.state->pad = kPeerTimeoutMark; }
        return false;
    }
    return true;
}

__device__ __forceinline__ double block_sum_1024(double v, double* sm32)
{
    // fixed tree: butterfly inside each warp, then warp 0 folds the 32 warp sums the same way
    for (int m = 16; m >= 1; m >>= 1) v += shfl_xor_f64(v, m);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sm32[warp] = v;
    __syncthreads();
    double r = sm32[lane];
    for (int m = 16; m >= 1; m >>= 1) r += shfl_xor_f64(r, m);
    return r; // identical in every thread
}

// Updaters [U] optimization/Updater.scala: 0 Simple (the reference path), 1 SquaredL2, 2 L1, on one
// column.  a*b+c is deliberately NOT fused here (the JVM does not fuse).
__device__ __forceinline__ void update_column(const UpdateArgs& a, int j, double cnt, double this_step, double& dn2, double& wn2,
                                              double& l1)
{
    double g = 0.0;
    for (int s = 0; s < a.nshards; ++s) g += __ldcg(a.shards + (long long)s * a.pv + j);
    g = g / cnt;
    const double wo = a.w[j];
    double wn;
    if (a.updater == 1) {
        wn = __dmul_rn(wo, 1.0 - this_step * a.reg);
        wn = __dadd_rn(wn, __dmul_rn(-this_step, g));
    } else {
        wn = __dadd_rn(wo, __dmul_rn(-this_step, g));
        if (a.updater == 2) {
            const double shr = a.reg * this_step;
            const double m = fabs(wn) - shr;
            const double sg = (double)((wn > 0.0) - (wn < 0.0));
            wn = sg * (m > 0.0 ? m : 0.0);
        }
    }
    a.w[j] = wn;
    const double df = wo - wn;
    dn2 += df * df;
    wn2 += wn * wn;
    l1 += fabs(wn);
}

__device__ __forceinline__ void shard_totals(const UpdateArgs& a, double& loss, double& cnt)
{
    loss = cnt = 0.0;
    for (int s = 0; s < a.nshards; ++s) {
        loss += __ldcg(a.shards + (long long)s * a.pv + a.ldx);
        cnt += __ldcg(a.shards + (long long)s * a.pv + a.ldx + 1);
    }
}

// loss history, regVal, GradientDescent.isConverged [U], iteration counter
__device__ __forceinline__ void finish_iteration(const UpdateArgs& a, double loss, double cnt, double dn2, double wn2, double l1)
{
    TrainState* st = a.state;
    const int i = st->iter;
    bool converged = false;
    if (cnt > 0.0) {
        const bool had_previous = st->nhist >= 1; // previousWeights != None [U]: an earlier iteration of this batch had m > 0
        a.loss_hist[st->nhist] = loss / cnt + st->reg_val;
        st->nhist += 1;
        const double wn = sqrt(wn2);
        st->reg_val = a.updater == 1 ? 0.5 * a.reg * wn * wn : (a.updater == 2 ? l1 * a.reg : 0.0);
        if (had_previous) converged = sqrt(dn2) < a.tol * fmax(wn, 1.0);
    }
    st->last_count = cnt;
    st->iter = i + 1;
    if (converged || i + 1 > a.max_iter) {
        st->done = 1;
        if (a.host_done) { *reinterpret_cast<volatile int*>(a.host_done) = 1; __threadfence_system(); }
    }
}

// Narrow models (the dense path): one CTA of 1024 threads does the whole update.
static __global__ void __launch_bounds__(1024) k_update(const UpdateArgs a_in)
{
    __shared__ double sm32[32];
    UpdateArgs a = a_in;
    if (a.state->done) return;
    unsigned long long seq;
    if (!wait_for_peers(a, seq)) return;
    double cnt, loss;
    shard_totals(a, loss, cnt);
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    if (cnt > 0.0) {
        const double this_step = a.step / sqrt((double)a.state->iter);
        for (int j = threadIdx.x; j < a.d; j += blockDim.x) update_column(a, j, cnt, this_step, dn2, wn2, l1);
    }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32); // (its barriers: every thread has read the vectors by now)
    if (threadIdx.x == 0) {
        finish_iteration(a, loss, cnt, dn2, wn2, l1);
        if (a.xbuf) *xbuf_seq(a.xbuf) = seq; // this exchange is complete on this rank
    }
}

// One GPU, narrow model: the two launches above in one.  grid (ceil(pv/32), spr) of 1024 threads as k_shard_reduce; every
// CTA leaves its 32 columns of its shard's vector, and the CTA that finishes last (a counter; every other CTA's vector is
// then complete and visible) goes on to be k_update.  The update reads spr vectors of pv doubles, not the block partials,
// so the extra work of that one CTA is small; one launch and one dependent-launch gap per SGD iteration are saved, which
// is what a 13-row batch interval or a 70 us tweet iteration notices.  Same arithmetic in the same order as the pair.
// P2P (several ranks): the vectors go out as lines of the peer exchange instead (k_shard_reduce_p2p), and the last CTA
// collects every rank's lines before it updates (k_update with a.xbuf set).
template <bool P2P>
static __global__ void __launch_bounds__(1024) k_shard_reduce_update(const double* __restrict__ partials, int pv, Plan plan,
                                                              double* __restrict__ shard_out, unsigned* __restrict__ counter,
                                                              const UpdateArgs a_in, const PeerExchange px)
{
    UpdateArgs a = a_in;
    __shared__ double sm[32][33];
    __shared__ double sm32[32];
    __shared__ int is_last;
    if (a.state->done) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int s = blockIdx.y;
    const int col = blockIdx.x * 32 + tx;
    const long long b0 = plan.shard_blk0[s], B = plan.shard_blk0[s + 1] - b0;
    const long long lo = b0 + B * ty / 32, hi = b0 + B * (ty + 1) / 32;
    double acc = 0.0;
    if (col < pv) {
        long long b = lo;
        for (; b + 4 <= hi; b += 4) {
            const double v0 = partials[(b + 0) * pv + col], v1 = partials[(b + 1) * pv + col];
            const double v2 = partials[(b + 2) * pv + col], v3 = partials[(b + 3) * pv + col];
            acc += v0;
            acc += v1;
            acc += v2;
            acc += v3;
        }
        for (; b < hi; ++b) acc += partials[b * pv + col];
    }
    sm[ty][tx] = acc;
    __syncthreads();
    if (P2P) {
        const unsigned long long seq = *xbuf_seq(px.peer[px.rank]) + 1ull; // read before this CTA's ticket: the update that advances it starts after the last ticket
        if (ty < px.world && col < pv) {
            double tot = 0.0;
#pragma unroll
            for (int q = 0; q < 32; ++q) tot += sm[q][tx];
            const size_t slot = ((size_t)px.rank * plan.spr + s) * pv + col;
            for (int r = ty; r < px.world; r += 32) ll_store(xbuf_lines(px.peer[r], (int)(seq & 1ull)) + slot, tot, (unsigned)seq);
        }
    } else if (ty == 0 && col < pv) {
        double tot = 0.0;
#pragma unroll
        for (int q = 0; q < 32; ++q) tot += sm[q][tx];
        shard_out[(long long)s * pv + col] = tot;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned total = gridDim.x * gridDim.y;
        is_last = atomicAdd(counter, 1u) == total - 1u;
        if (is_last) *counter = 0u;
    }
    __syncthreads();
    if (!is_last) return;
    unsigned long long seq = 0ull;
    if (P2P) {
        if (!wait_for_peers(a, seq)) return; // (lines announce themselves: no fence)
    } else {
        __threadfence();
    }
    double cnt, loss;
    shard_totals(a, loss, cnt);
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    if (cnt > 0.0) {
        const double this_step = a.step / sqrt((double)a.state->iter);
        for (int j = threadIdx.x; j < a.d; j += blockDim.x) update_column(a, j, cnt, this_step, dn2, wn2, l1);
    }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) {
        finish_iteration(a, loss, cnt, dn2, wn2, l1);
        if (P2P) *xbuf_seq(a.xbuf) = seq; // this exchange is complete on this rank
    }
}

// Wide models (hashed text, d ~ 2^20): the columns are spread over many CTAs, each leaving its three
// norm partials; k_update_finish adds them in CTA order and advances the state.  kWideCols columns
// per CTA, fixed, so the partial sums do not depend on the grid the launch happens to get.
constexpr int kWideCols = 4096;
static __global__ void __launch_bounds__(1024) k_update_wide(const UpdateArgs a, double* __restrict__ part /* gridDim.x x 3 */)
{
    __shared__ double sm32[32];
    if (a.state->done) return;
    double cnt, loss;
    shard_totals(a, loss, cnt);
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    if (cnt > 0.0) {
        const double this_step = a.step / sqrt((double)a.state->iter);
        const int lo = blockIdx.x * kWideCols, hi = min(a.d, lo + kWideCols);
        for (int j = lo + threadIdx.x; j < hi; j += blockDim.x) update_column(a, j, cnt, this_step, dn2, wn2, l1);
    }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) { part[blockIdx.x * 3] = dn2; part[blockIdx.x * 3 + 1] = wn2; part[blockIdx.x * 3 + 2] = l1; }
}

// ---- wide models on several ranks: every rank owns a slice of the columns ------------------------
// After an all-to-all of gradient slices (NCCL send/recv; pure data movement), rank r holds, for its own
// columns [lo, hi), the W partial sums in rank order: it adds them in that order, updates its slice of w,
// and an all-gather hands every rank the whole new w.  Each rank then computes the two norms of
// isConverged [U] over all columns itself (same data, same kernel: same bits everywhere).
struct SliceArgs {
    const double* slices; // W x slice_len: rank r's sums for my columns at slices + r * slice_len
    const double* pairs;  // W x 2: every rank's (lossSum, count)
    double* w;
    const TrainState* state;
    int W, slice_len, lo, hi;
    int hi_all; // model width d
    int updater;
    double step, reg;
};

static __global__ void __launch_bounds__(1024) k_update_slice(const SliceArgs a)
{
    if (a.state->done) return;
    double cnt = 0.0;
    for (int r = 0; r < a.W; ++r) cnt += a.pairs[2 * r + 1];
    if (!(cnt > 0.0)) return;
    const double this_step = a.step / sqrt((double)a.state->iter);
    const int j0 = a.lo + blockIdx.x * kWideCols, j1 = min(a.hi, j0 + kWideCols);
    for (int j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
        double g = 0.0;
        for (int r = 0; r < a.W; ++r) g += a.slices[(long long)r * a.slice_len + (j - a.lo)];
        g = g / cnt;
        const double wo = a.w[j];
        double wn;
        if (a.updater == 1) {
            wn = __dmul_rn(wo, 1.0 - this_step * a.reg);
            wn = __dadd_rn(wn, __dmul_rn(-this_step, g));
        } else {
            wn = __dadd_rn(wo, __dmul_rn(-this_step, g));
            if (a.updater == 2) {
                const double shr = a.reg * this_step;
                const double m = fabs(wn) - shr;
                const double sg = (double)((wn > 0.0) - (wn < 0.0));
                wn = sg * (m > 0.0 ? m : 0.0);
            }
        }
        a.w[j] = wn;
    }
}

// ---- the same over peer memory: all-to-all + sliced update + all-gather in ONE kernel -------------------------------
// Every rank owns a wide exchange buffer, mapped by all ranks:  G[dpad] (its gradient vector, lossSum and count at
// G[d], G[d + 1]) | WX[dpad] (the new weights, written by the ranks that own the columns) | flag_g[8] | flag_w[8] |
// CTA counter | sequence number.  Per iteration, after the rank's column pass has left its gradient in G:
//   k_wide_signal        tells every peer "my G of exchange seq is complete" (flag_g[me] on the peer := seq);
//   k_update_slice_p2p   waits for all flag_g, then for each of ITS columns loads the W partial sums straight from the
//                        peers' G over NVLink (plain peer loads, added in rank order), applies the updater and stores
//                        the new weight into EVERY rank's WX (peer stores); its last CTA raises flag_w[me] everywhere;
//   k_norms_wide_p2p     waits for all flag_w, moves WX into the model's weights and takes the convergence norms on the
//                        way (old weight still in place: no copy of the previous weights is kept);
//   k_slice_finish       as before; it also advances the buffer's sequence number.
// No NCCL call, no staging copy; transfers are the 2 x (W-1)/W x 8 d bytes per rank that the all-to-all + all-gather
// moved, but issued by the SMs that consume / produce them.  Reuse is safe without a second parity: a rank rewrites G
// only after its k_norms_wide_p2p saw every flag_w (so every peer has finished reading G), and peers write WX(seq + 1)
// only after this rank signalled G(seq + 1), which follows its own reading of WX(seq) in stream order.
struct WideExchange {
    double* peer[kMaxShards]; // base of rank r's wide exchange buffer, as mapped here
    long long dpad;           // doubles per region
    int world, rank;
};
__host__ __device__ __forceinline__ double* wx_G(double* base) { return base; }
__host__ __device__ __forceinline__ double* wx_W(double* base, long long dpad) { return base + dpad; }
__host__ __device__ __forceinline__ unsigned long long* wx_flag_g(double* base, long long dpad) { return reinterpret_cast<unsigned long long*>(base + 2 * dpad); }
__host__ __device__ __forceinline__ unsigned long long* wx_flag_w(double* base, long long dpad) { return wx_flag_g(base, dpad) + kMaxShards; }
__host__ __device__ __forceinline__ unsigned* wx_counter(double* base, long long dpad) { return reinterpret_cast<unsigned*>(wx_flag_g(base, dpad) + 2 * kMaxShards); }
__host__ __device__ __forceinline__ unsigned long long* wx_seq(double* base, long long dpad) { return wx_flag_g(base, dpad) + 2 * kMaxShards + 1; }
constexpr size_t wide_exchange_bytes(long long dpad) { return (size_t)(2 * dpad) * 8 + (2 * kMaxShards + 2) * 8 + 64; }

__device__ __forceinline__ double ld_sys_f64(const double* p)
{
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

// all `n` flags (written by other GPUs) have reached seq; false after ~10 s
__device__ __forceinline__ bool wait_flags(const unsigned long long* flags, int n, unsigned long long seq)
{
    __shared__ int timed_out;
    if (threadIdx.x == 0) timed_out = 0;
    __syncthreads();
    if (threadIdx.x < n) {
        const volatile unsigned long long* f = flags + threadIdx.x;
        const long long t0 = clock64();
        while (*f < seq)
            if (clock64() - t0 > 20000000000ll) { timed_out = 1; break; }
        __threadfence_system();
    }
    __syncthreads();
    __threadfence_system();
    return timed_out == 0;
}

static __global__ void k_wide_signal(WideExchange wx, const TrainState* state)
{
    if (state->done || threadIdx.x != 0) return;
    double* mine = wx.peer[wx.rank];
    const unsigned long long seq = *wx_seq(mine, wx.dpad) + 1ull;
    __threadfence_system(); // the column pass before this launch wrote G
    for (int r = 0; r < wx.world; ++r) *reinterpret_cast<volatile unsigned long long*>(wx_flag_g(wx.peer[r], wx.dpad) + wx.rank) = seq;
}

static __global__ void __launch_bounds__(1024) k_update_slice_p2p(const SliceArgs a, WideExchange wx, TrainState* state, double* __restrict__ pairs_out)
{
    if (state->done) return;
    double* mine = wx.peer[wx.rank];
    const unsigned long long seq = *wx_seq(mine, wx.dpad) + 1ull;
    if (!wait_flags(wx_flag_g(mine, wx.dpad), wx.world, seq)) {
        if (threadIdx.x == 0) { state->done = 1; state->pad = kPeerTimeoutMark; }
        return;
    }
    const int d = a.hi_all; // model width: (lossSum, count) sit behind it
    double cnt = 0.0;
    for (int r = 0; r < wx.world; ++r) cnt += ld_sys_f64(wx_G(wx.peer[r]) + d + 1);
    if (blockIdx.x == 0 && threadIdx.x < 2 * wx.world) // the W (lossSum, count) pairs for k_slice_finish
        pairs_out[threadIdx.x] = ld_sys_f64(wx_G(wx.peer[threadIdx.x >> 1]) + d + (threadIdx.x & 1));
    if (cnt > 0.0) {
        const double this_step = a.step / sqrt((double)state->iter);
        const int j0 = a.lo + blockIdx.x * kWideCols, j1 = min(a.hi, j0 + kWideCols);
        for (int j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
            double g = 0.0;
            for (int r = 0; r < wx.world; ++r) g += ld_sys_f64(wx_G(wx.peer[r]) + j); // rank order
            g = g / cnt;
            const double wo = a.w[j];
            double wn;
            if (a.updater == 1) {
                wn = __dmul_rn(wo, 1.0 - this_step * a.reg);
                wn = __dadd_rn(wn, __dmul_rn(-this_step, g));
            } else {
                wn = __dadd_rn(wo, __dmul_rn(-this_step, g));
                if (a.updater == 2) {
                    const double shr = a.reg * this_step;
                    const double m = fabs(wn) - shr;
                    const double sg = (double)((wn > 0.0) - (wn < 0.0));
                    wn = sg * (m > 0.0 ? m : 0.0);
                }
            }
            for (int r = 0; r < wx.world; ++r) wx_W(wx.peer[r], wx.dpad)[j] = wn;
        }
    } else { // an empty sample: the weights stand
        const int j0 = a.lo + blockIdx.x * kWideCols, j1 = min(a.hi, j0 + kWideCols);
        for (int j = j0 + threadIdx.x; j < j1; j += blockDim.x) {
            const double wo = a.w[j];
            for (int r = 0; r < wx.world; ++r) wx_W(wx.peer[r], wx.dpad)[j] = wo;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned* counter = wx_counter(mine, wx.dpad);
        if (atomicAdd(counter, 1u) == gridDim.x - 1u) { // last CTA of this rank: its whole slice is on its way everywhere
            *counter = 0u;
            __threadfence_system();
            for (int r = 0; r < wx.world; ++r) *reinterpret_cast<volatile unsigned long long*>(wx_flag_w(wx.peer[r], wx.dpad) + wx.rank) = seq;
        }
    }
}

// WX -> w with the three norms of the step on the way (per-CTA partials over kWideCols columns each, as k_norms_wide)
static __global__ void __launch_bounds__(1024) k_norms_wide_p2p(double* __restrict__ w, WideExchange wx, int d, TrainState* state,
                                                                 double* __restrict__ part)
{
    __shared__ double sm32[32];
    if (state->done) return;
    double* mine = wx.peer[wx.rank];
    const unsigned long long seq = *wx_seq(mine, wx.dpad) + 1ull;
    if (!wait_flags(wx_flag_w(mine, wx.dpad), wx.world, seq)) {
        if (threadIdx.x == 0) { state->done = 1; state->pad = kPeerTimeoutMark; }
        return;
    }
    const double* wxw = wx_W(mine, wx.dpad);
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    const int lo = blockIdx.x * kWideCols, hi = min(d, lo + kWideCols);
    for (int j = lo + threadIdx.x; j < hi; j += blockDim.x) {
        const double wn = __ldcg(wxw + j), df = w[j] - wn; // written by other GPUs: read past L1
        w[j] = wn;
        dn2 += df * df;
        wn2 += wn * wn;
        l1 += fabs(wn);
    }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) { part[blockIdx.x * 3] = dn2; part[blockIdx.x * 3 + 1] = wn2; part[blockIdx.x * 3 + 2] = l1; }
}

// per-CTA partials of ||w_prev - w||^2, ||w||^2, ||w||_1 over kWideCols columns each
static __global__ void __launch_bounds__(1024) k_norms_wide(const double* __restrict__ w, const double* __restrict__ wprev, int d,
                                                     const TrainState* state, double* __restrict__ part)
{
    __shared__ double sm32[32];
    if (state->done) return;
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    const int lo = blockIdx.x * kWideCols, hi = min(d, lo + kWideCols);
    for (int j = lo + threadIdx.x; j < hi; j += blockDim.x) {
        const double wn = w[j], df = wprev[j] - wn;
        dn2 += df * df;
        wn2 += wn * wn;
        l1 += fabs(wn);
    }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) { part[blockIdx.x * 3] = dn2; part[blockIdx.x * 3 + 1] = wn2; part[blockIdx.x * 3 + 2] = l1; }
}

static __global__ void __launch_bounds__(1024) k_slice_finish(const UpdateArgs a, const double* __restrict__ pairs, int W,
                                                       const double* __restrict__ part, int nparts, unsigned long long* seq_word = nullptr)
{
    __shared__ double sm32[32];
    if (a.state->done) return;
    if (seq_word && threadIdx.x == 0) *seq_word += 1ull; // the peer-memory exchange of this iteration is complete on this rank
    double loss = 0.0, cnt = 0.0;
    for (int r = 0; r < W; ++r) { loss += pairs[2 * r]; cnt += pairs[2 * r + 1]; }
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    for (int p = threadIdx.x; p < nparts; p += blockDim.x) { dn2 += part[p * 3]; wn2 += part[p * 3 + 1]; l1 += part[p * 3 + 2]; }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) finish_iteration(a, loss, cnt, dn2, wn2, l1);
}

static __global__ void __launch_bounds__(1024) k_update_finish(const UpdateArgs a, const double* __restrict__ part, int nparts)
{
    __shared__ double sm32[32];
    if (a.state->done) return;
    double cnt, loss;
    shard_totals(a, loss, cnt);
    double dn2 = 0.0, wn2 = 0.0, l1 = 0.0;
    for (int p = threadIdx.x; p < nparts; p += blockDim.x) { dn2 += part[p * 3]; wn2 += part[p * 3 + 1]; l1 += part[p * 3 + 2]; }
    dn2 = block_sum_1024(dn2, sm32);
    wn2 = block_sum_1024(wn2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) finish_iteration(a, loss, cnt, dn2, wn2, l1);
}

// regVal before the first iteration = updater.compute(w, 0, 0, 1, regParam)._2 [U]
static __global__ void __launch_bounds__(1024) k_train_init(TrainState* st, const double* w, int d, int updater, double reg,
                                                     int max_iter)
{
    __shared__ double sm32[32];
    double n2 = 0.0, l1 = 0.0;
    if (updater != 0 && reg != 0.0)
        for (int j = threadIdx.x; j < d; j += blockDim.x) { n2 += w[j] * w[j]; l1 += fabs(w[j]); }
    n2 = block_sum_1024(n2, sm32);
    l1 = block_sum_1024(l1, sm32);
    if (threadIdx.x == 0) {
        st->iter = 1;
        st->done = max_iter < 1 ? 1 : 0;
        st->nhist = 0;
        st->pad = 0;
        const double nrm = sqrt(n2);
        st->reg_val = updater == 1 ? 0.5 * reg * nrm * nrm : (updater == 2 ? l1 * reg : 0.0);
        st->last_count = 0.0;
    }
}

// grid spr, block 256: thread t merges slice t of the shard's block tuples in block order, then
// thread 0 merges the 256 slice tuples in slice order.
static __global__ void __launch_bounds__(256) k_stats_shard(const double* __restrict__ stat_partials, Plan plan,
                                                     double* __restrict__ shard_out /* spr x kStatStride */)
{
    __shared__ double sm[256 * kStatStride];
    const int s = blockIdx.x, t = threadIdx.x;
    const long long b0 = plan.shard_blk0[s], B = plan.shard_blk0[s + 1] - b0;
    const long long lo = b0 + B * t / 256, hi = b0 + B * (t + 1) / 256;
    Stat7 acc;
    stat_zero(acc);
    for (long long b = lo; b < hi; ++b) {
        Stat7 v;
        stat_load(v, stat_partials + b * kStatStride);
        stat_merge(acc, v);
    }
    stat_store(sm + t * kStatStride, acc);
    __syncthreads();
    if (t == 0) {
        Stat7 tot;
        stat_zero(tot);
        for (int q = 0; q < 256; ++q) {
            Stat7 v;
            stat_load(v, sm + q * kStatStride);
            stat_merge(tot, v);
        }
        stat_store(shard_out + (long long)s * kStatStride, tot);
    }
}

// out10: batch, mse, real_stdev, pred_stdev (Utils.round applied), mse_raw, real_stdev_raw,
// pred_stdev_raw, real_mean, pred_mean, nonfinite
static __global__ void k_stats_final(const double* __restrict__ shards, int nshards, double* __restrict__ out10)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    Stat7 tot;
    stat_zero(tot);
    for (int s = 0; s < nshards; ++s) {
        Stat7 v;
        stat_load(v, shards + (long long)s * kStatStride);
        stat_merge(tot, v);
    }
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const double rs = tot.n > 0.0 ? sqrt(tot.m2_y / tot.n) : nan;
    const double ps = tot.n > 0.0 ? sqrt(tot.m2_p / tot.n) : nan;
    const double mse = tot.n > 0.0 ? tot.mu_e : nan;
    out10[0] = tot.n;
    out10[1] = round_half_up(mse);
    out10[2] = round_half_up(rs);
    out10[3] = round_half_up(ps);
    out10[4] = mse;
    out10[5] = rs;
    out10[6] = ps;
    out10[7] = tot.mu_y;
    out10[8] = tot.mu_p;
    out10[9] = tot.bad;
}

} // namespace twtml
