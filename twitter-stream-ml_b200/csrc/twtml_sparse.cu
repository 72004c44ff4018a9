// twtml_sparse.cu -- CSR micro-batches: ingest (from caller CSR, from raw tweets through the device
// featuriser, or synthetic), the transpose they carry for the column pass, and the per-iteration
// launches of csr.cuh.  CUB is used for two pieces of ingest-time plumbing only (a stable radix
// sort by column to build the transpose, exclusive prefix sums); nothing of it runs per iteration.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <mutex>
#include <thread>
#include <type_traits>
#include <vector>

#include "csr.cuh"
#include "csr_narrow.cuh"
#include "featurize.cuh"
#include "host.h"
#include <chrono>
#include "lower_table.h"
#include "reduce.cuh"
#include "synth.cuh"

namespace twtml {

template <class T>
static int dev_alloc(twtml_ctx* ctx, IngestLane& lane, T** p, size_t count)
{
    return pool_alloc(ctx, reinterpret_cast<void**>(p), sizeof(T) * std::max<size_t>(count, 1), lane.stream);
}

// an array that belongs to the batch (given back by batch_free_locked): a block of the context, free of stream ordering
template <class T>
static int batch_array(twtml_ctx* ctx, T** p, size_t count)
{
    return block_alloc(ctx, reinterpret_cast<void**>(p), sizeof(T) * std::max<size_t>(count, 1));
}

static int exclusive_sum(IngestLane& lane, const long long* in, long long* out, long long n)
{
    size_t tmp = 0;
    cudaGetLastError(); // CUB reports any error still pending in the runtime as its own
    TW_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, in, out, n, lane.stream));
    TW_TRY(lane.scan_tmp.reserve(tmp));
    TW_CUDA(cub::DeviceScan::ExclusiveSum(lane.scan_tmp.p, tmp, in, out, n, lane.stream));
    return 0;
}

static unsigned blocks_for(long long n, int threads) { return (unsigned)std::max<long long>(1, (n + threads - 1) / threads); }

// Narrow models: no transpose at all, only the first row of every CTA of the single-pass kernel.
// Returns 1 when the batch does not qualify (the caller then builds the global transpose).
static int csr_finalize_narrow(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b)
{
    const long long n = b->n;
    const int d = (int)b->d;
    if (d > kNarrowMaxD || getenv("TWTML_NO_NARROW")) return 1; // a function of the model width only: every rank decides alike
    const int warps = narrow_warps(d);
    if (warps < 4) return 1;
    // one partial vector per CTA of a grid fixed here: the plan's "blocks" are those CTAs
    const long long grid = std::max<long long>(1, std::min<long long>((n + warps - 1) / warps, (long long)ctx->sm_count));
    TW_TRY(batch_array(ctx, &b->bounds, (size_t)grid + 1));
    k_narrow_bounds<<<blocks_for(grid + 1, 256), 256, 0, lane.stream>>>(b->rowptr, n, grid, b->bounds);
    TW_CUDA(cudaGetLastError());
    static const bool no_permute = getenv("TWTML_NO_NARROW_PERMUTE") != nullptr; // A/B switch
    if (!no_permute && b->nnz > 0) { // the kernel's own copy of the entries, laid out against shared-memory bank conflicts
        TW_TRY(batch_array(ctx, &b->pcolidx, (size_t)b->nnz + kNarrowSlack));
        TW_TRY(batch_array(ctx, &b->pvals, (size_t)b->nnz + kNarrowSlack));
        const unsigned pg = (unsigned)std::min<long long>((n + 7) / 8, (long long)ctx->sm_count * 16);
        k_narrow_permute<<<std::max(1u, pg), 256, 0, lane.stream>>>(b->rowptr, n, b->colidx, b->vals, b->pcolidx, b->pvals);
        TW_CUDA(cudaGetLastError());
    }
    b->narrow = true;
    b->nwarps = warps;
    memset(&b->plan, 0, sizeof(b->plan));
    b->plan.spr = 1;
    b->plan.rb = 1;
    b->plan.nblocks = grid;
    for (int s2 = 1; s2 <= kMaxShards; ++s2) { b->plan.shard_row0[s2] = n; b->plan.shard_blk0[s2] = grid; }
    return 0;
}

// rowptr / colidx / vals / labels are on the device (or on their way, in the lane's stream order): build the plan,
// the transpose and the column tasks.  Everything is enqueued on the lane's stream; nothing here waits for the device.
static int csr_finalize(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b, bool defer_global)
{
    // one shard per rank (the column sums are per rank, not per virtual shard); row blocks small enough to give every SM
    // several CTAs of the row pass even for a modest micro-batch (but at least 32 rows: one step of the widest row form)
    make_plan(b->n, 1, 32, b->plan, 4096);
    if (defer_global && ctx->world > 1) b->global_deferred = true;
    else TW_TRY(ctx_global_rows(ctx, b->n, &b->n_global, &b->row0_global));
    if (b->nnz >= ((long long)1 << 31) || b->n >= ((long long)1 << 31))
        return fail(TWTML_E_INVALID_ARG, "CSR batch too large for 32-bit entry indices (%lld rows, %lld nnz per GPU)", (long long)b->n, (long long)b->nnz);
    {
        const int nrc = csr_finalize_narrow(ctx, lane, b);
        if (nrc <= 0) return nrc; // built (0) or failed (< 0); 1 = not a narrow batch
    }
    cudaStream_t st = lane.stream;
    const int d = (int)b->d;
    const long long nnz = b->nnz;
    const long long nwin = (nnz + kWin - 1) / kWin;
    b->nwin = nwin;
    TW_TRY(batch_array(ctx, &b->colptr, (size_t)d + 1));
    TW_TRY(batch_array(ctx, &b->rowidx, (size_t)nnz));
    TW_TRY(batch_array(ctx, &b->cvals, (size_t)nnz));
    TW_TRY(batch_array(ctx, &b->diff, (size_t)b->n));
    TW_TRY(batch_array(ctx, &b->headbits, (size_t)(nwin + 1) * 8));
    TW_TRY(batch_array(ctx, &b->win_first_nz, (size_t)nwin + 1));
    TW_TRY(batch_array(ctx, &b->wkey, (size_t)nwin + 1));
    TW_TRY(batch_array(ctx, &b->win_first_part, (size_t)nwin + 1));
    TW_TRY(batch_array(ctx, &b->win_last_part, (size_t)nwin + 1));
    TW_TRY(batch_array(ctx, &b->nzcol, (size_t)std::min<long long>(nnz, d) + 1));
    TW_CUDA(cudaMemsetAsync(b->headbits, 0, sizeof(unsigned) * (size_t)(nwin + 1) * 8, st));
    if (nnz > 0) {
        int *keys_out = nullptr, *iota = nullptr, *perm = nullptr;
        long long* winheads = nullptr;
        int rc = dev_alloc(ctx, lane, &keys_out, (size_t)nnz);
        if (!rc) rc = dev_alloc(ctx, lane, &iota, (size_t)nnz);
        if (!rc) rc = dev_alloc(ctx, lane, &perm, (size_t)nnz);
        if (!rc) rc = dev_alloc(ctx, lane, &winheads, (size_t)nwin + 1);
        do {
            if (rc) break;
            k_iota<<<blocks_for(nnz, 256), 256, 0, st>>>(iota, nnz);
            int bits = 1;
            while (((long long)1 << bits) < d) ++bits;
            size_t tmp = 0;
            cudaGetLastError();
            cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, tmp, b->colidx, keys_out, iota, perm, (int)nnz, 0, bits, st);
            if (e != cudaSuccess) { rc = fail(TWTML_E_CUDA, "radix sort sizing: %s", cudaGetErrorString(e)); break; }
            if ((rc = lane.sort_tmp.reserve(tmp))) break;
            // LSD radix sort is stable: entries of a column keep their CSR (row) order
            e = cub::DeviceRadixSort::SortPairs(lane.sort_tmp.p, tmp, b->colidx, keys_out, iota, perm, (int)nnz, 0, bits, st);
            if (e != cudaSuccess) { rc = fail(TWTML_E_CUDA, "radix sort: %s", cudaGetErrorString(e)); break; }
            k_colptr<<<blocks_for(nnz + 1, 256), 256, 0, st>>>(keys_out, nnz, d, b->colptr);
            k_csc_gather<<<blocks_for(nnz, 256), 256, 0, st>>>(perm, b->rowptr, b->n, b->vals, nnz, b->rowidx, b->cvals);
            // window index of the transpose: head bits, heads per window -> first non-empty column per window
            k_csc_headbits<<<blocks_for(nnz + 1, 256), 256, 0, st>>>(keys_out, nnz, b->headbits);
            k_csc_winmeta<<<blocks_for(nwin + 1, 256), 256, 0, st>>>(b->headbits, keys_out, nnz, nwin, winheads, b->wkey);
            rc = exclusive_sum(lane, winheads, b->win_first_nz, nwin + 1);
            if (!rc) k_csc_nzcol<<<blocks_for(nnz, 256), 256, 0, st>>>(keys_out, b->headbits, b->win_first_nz, nnz, b->nzcol);
            if (!rc && cudaGetLastError() != cudaSuccess) rc = fail(TWTML_E_CUDA, "transpose kernels failed to launch");
        } while (0);
        // stream-ordered frees: the temporaries go back to the pool once the kernels above are done with them
        pool_free(ctx, winheads, st);
        pool_free(ctx, keys_out, st);
        pool_free(ctx, iota, st);
        pool_free(ctx, perm, st);
        if (rc) return rc;
    } else {
        TW_CUDA(cudaMemsetAsync(b->colptr, 0, sizeof(long long) * ((size_t)d + 1), st));
    }
    return 0;
}

// the handle alone: what an asynchronous ingest hands back at once
static twtml_batch* csr_shell(twtml_ctx* ctx, int64_t n, int64_t d, int64_t nnz)
{
    twtml_batch* b = new twtml_batch();
    b->ctx = ctx;
    b->csr = true;
    b->n = n;
    b->d = d;
    b->nnz = nnz;
    ctx->live_handles += 1; // released by twtml_batch_free, also on the failure paths of the callers
    return b;
}

static int csr_alloc(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b)
{
    // colidx / vals end in kNarrowSlack spare entries: the narrow kernel's bulk copies are 16 bytes wide
    TW_TRY(batch_array(ctx, &b->rowptr, (size_t)b->n + 1));
    TW_TRY(batch_array(ctx, &b->colidx, (size_t)b->nnz + kNarrowSlack));
    TW_TRY(batch_array(ctx, &b->vals, (size_t)b->nnz + kNarrowSlack));
    TW_TRY(batch_array(ctx, &b->labels, (size_t)b->n));
    return 0;
}

static int new_csr(twtml_ctx* ctx, IngestLane& lane, int64_t n, int64_t d, int64_t nnz, twtml_batch** out)
{
    twtml_batch* b = csr_shell(ctx, n, d, nnz);
    int rc = csr_alloc(ctx, lane, b);
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

// ---- per-iteration launches -----------------------------------------------------------------
int csr_partial_width(const twtml_batch* b) { return (int)b->d + 2; }

static int launch_narrow(twtml_ctx* ctx, const NarrowArgs& a, int mode, unsigned grid)
{
    const size_t smem = narrow_smem_bytes(a.d, a.warps);
#define TW_NARROW(MODE_)                                                                                         \
    do {                                                                                                         \
        auto kern = k_csr_narrow<MODE_>;                                                                         \
        if (!(ctx->narrow_attr_done & (1u << MODE_))) { /* once per context (= per device) and kernel form */    \
            TW_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));        \
            ctx->narrow_attr_done |= 1u << MODE_;                                                                \
        }                                                                                                        \
        kern<<<grid, (a.warps + 1) * 32, smem, ctx->stream>>>(a);                                                \
    } while (0)
    switch (mode) {
    case MODE_TRAIN: TW_NARROW(MODE_TRAIN); break;
    case MODE_TRAIN_STATS: TW_NARROW(MODE_TRAIN_STATS); break;
    case MODE_PREDICT: TW_NARROW(MODE_PREDICT); break;
    default: TW_NARROW(MODE_PREDICT_STATS); break;
    }
#undef TW_NARROW
    TW_CUDA(cudaGetLastError());
    return 0;
}

int csr_launch_iter(twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b, int mode, double* out_pred, double* out_real,
                    int iter_override)
{
    if (b->narrow) {
        NarrowArgs na{};
        na.rowptr = b->rowptr;
        na.colidx = b->pcolidx ? b->pcolidx : b->colidx;
        na.vals = b->pvals ? b->pvals : b->vals;
        na.labels = b->labels;
        na.w = m->d_w;
        na.state = ctx->state.as<TrainState>();
        na.bounds = b->bounds;
        na.partials = ctx->partials.as<double>();
        na.stat_partials = ctx->stat_partials.as<double>();
        na.out_pred = out_pred;
        na.out_real = out_real;
        na.n = b->n;
        na.d = (int)b->d;
        na.warps = b->nwarps;
        na.frac = m->frac;
        na.row0_global = (unsigned long long)b->row0_global;
        na.iter_override = iter_override;
        return launch_narrow(ctx, na, mode, (unsigned)b->plan.nblocks);
    }
    CsrIterArgs a{};
    a.rowptr = b->rowptr;
    a.colidx = b->colidx;
    a.vals = b->vals;
    a.labels = b->labels;
    a.w = m->d_w;
    a.state = ctx->state.as<TrainState>();
    a.diff = b->diff;
    a.partials = ctx->partials.as<double>();
    a.stat_partials = ctx->stat_partials.as<double>();
    a.out_pred = out_pred;
    a.out_real = out_real;
    a.plan = b->plan;
    a.frac = m->frac;
    a.row0_global = (unsigned long long)b->row0_global;
    a.iter_override = iter_override;
    const unsigned grid = (unsigned)std::min<long long>(b->plan.nblocks, (long long)ctx->sm_count * 16);
    if (grid == 0) return 0;
    const double avg = b->n > 0 ? (double)b->nnz / (double)b->n : 0.0; // lanes per row follow the average row length
#define TW_ROWS(MODE_)                                                                        \
    do {                                                                                      \
        if (avg <= 12.0) k_csr_rows<MODE_, 4><<<grid, kCsrThreads, 0, ctx->stream>>>(a);      \
        else if (avg <= 48.0) k_csr_rows<MODE_, 8><<<grid, kCsrThreads, 0, ctx->stream>>>(a); \
        else k_csr_rows<MODE_, 16><<<grid, kCsrThreads, 0, ctx->stream>>>(a);                 \
    } while (0)
    switch (mode) {
    case MODE_TRAIN: TW_ROWS(MODE_TRAIN); break;
    case MODE_TRAIN_STATS: TW_ROWS(MODE_TRAIN_STATS); break;
    case MODE_PREDICT: TW_ROWS(MODE_PREDICT); break;
    default: TW_ROWS(MODE_PREDICT_STATS); break;
    }
#undef TW_ROWS
    TW_CUDA(cudaGetLastError());
    return 0;
}

// column pass + (lossSum, count) fold -> this rank's shard vector [g_0..g_{d-1}, lossSum, count]
int csr_reduce(twtml_ctx* ctx, const twtml_batch* b, double* shard_out, const TrainState* gate)
{
    const int d = (int)b->d;
    if (b->narrow) { // the fused kernel left one [g, lossSum, count] vector per CTA: add them in CTA order
        dim3 grid((d + 2 + 31) / 32, 1), block(32, 32);
        k_shard_reduce<<<grid, block, 0, ctx->stream>>>(ctx->partials.as<double>(), d + 2, b->plan, shard_out, gate);
        TW_CUDA(cudaGetLastError());
        ctx->prof.other_launches += 1;
        return 0;
    }
    // columns without entries keep gradSum = 0; every other column is written exactly once below
    TW_CUDA(cudaMemsetAsync(shard_out, 0, sizeof(double) * (size_t)d, ctx->stream));
    if (b->nwin > 0) {
        CscView c{b->rowidx, b->cvals, b->headbits, b->nzcol, b->win_first_nz, b->wkey, b->colptr, b->nnz, b->nwin};
        const unsigned g1 = (unsigned)std::min<long long>((b->nwin + 7) / 8, (long long)ctx->sm_count * 32);
        k_csc_windows<<<g1, 256, 0, ctx->stream>>>(c, b->diff, shard_out, b->win_first_part, b->win_last_part, gate);
        const unsigned g2 = (unsigned)std::min<long long>((b->nwin + 255) / 256, (long long)ctx->sm_count * 32);
        k_csc_edges<<<g2, 256, 0, ctx->stream>>>(c, b->win_first_part, b->win_last_part, shard_out, gate);
    }
    k_pair_reduce<<<1, 1024, 0, ctx->stream>>>(ctx->partials.as<double>(), b->plan.nblocks, shard_out + d, gate);
    TW_CUDA(cudaGetLastError());
    ctx->prof.other_launches += 3;
    return 0;
}

static int lower_table(twtml_ctx* ctx, const uint16_t** out)
{
    static std::mutex mu; // first use may come from the caller's thread or from the ingest worker
    std::lock_guard<std::mutex> lk(mu);
    if (!ctx->lower.p) {
        std::vector<uint16_t> t(65536);
        for (int c = 0; c < 65536; ++c) t[c] = (uint16_t)c;
        for (int i = 0; i < kLowerPairCount; ++i) t[kLowerPairs[i][0]] = kLowerPairs[i][1];
        DevBuf buf;
        TW_TRY(buf.reserve(sizeof(uint16_t) * 65536));
        cudaError_t e = cudaMemcpy(buf.p, t.data(), sizeof(uint16_t) * 65536, cudaMemcpyHostToDevice); // blocking: the table is complete before anyone sees the pointer
        if (e != cudaSuccess) { buf.release(); return fail(TWTML_E_CUDA, "lower-case table: %s", cudaGetErrorString(e)); }
        ctx->lower = buf;
    }
    *out = ctx->lower.as<uint16_t>();
    return 0;
}

// ---- ingest bodies: the same code serves the blocking calls (lane 0, caller's thread) and the asynchronous ones
// (lane 1, the context's worker thread) ----------------------------------------------------------------------

// Vectors.sparse require()s strictly increasing, in-range indices [U]
static int validate_csr(int64_t n, int64_t d, int64_t nnz, const int64_t* rowptr, const int32_t* colidx)
{
    // the row pointers alone first (a malformed one must be refused before any index is read through it)
    for (int64_t r = 0; r < n; ++r)
        if (rowptr[r] < 0 || rowptr[r] > rowptr[r + 1] || rowptr[r + 1] > nnz)
            return fail(TWTML_E_INVALID_ARG, "row %lld: rowptr must be non-decreasing within [0, rowptr[n]]", (long long)r);
    std::atomic<int64_t> bad_row(-1);
    const unsigned hw = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    const int nt = (int)std::min<int64_t>(hw, std::max<int64_t>(1, n / 4096));
    auto scan = [&](int64_t lo, int64_t hi) { // on all host threads
        for (int64_t r = lo; r < hi && bad_row.load(std::memory_order_relaxed) < 0; ++r) {
            bool ok = true;
            for (int64_t k = rowptr[r]; ok && k < rowptr[r + 1]; ++k)
                ok = colidx[k] >= 0 && colidx[k] < d && (k == rowptr[r] || colidx[k] > colidx[k - 1]);
            if (!ok) bad_row.store(r);
        }
    };
    if (nt <= 1) scan(0, n);
    else {
        std::vector<std::thread> th;
        for (int i = 0; i < nt; ++i) th.emplace_back(scan, n * i / nt, n * (i + 1) / nt);
        for (auto& t : th) t.join();
    }
    if (bad_row.load() >= 0)
        return fail(TWTML_E_INVALID_ARG, "row %lld: indices must be strictly increasing and < %lld", (long long)bad_row.load(), (long long)d);
    return 0;
}

static int csr_ingest(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b, const int64_t* rowptr, const int32_t* colidx, const double* vals,
                      const double* labels, bool defer_global)
{
    std::lock_guard<std::mutex> lane_lock(lane.mu);
    const int64_t n = b->n, nnz = b->nnz;
    TW_TRY(validate_csr(n, b->d, nnz, rowptr, colidx));
    TW_TRY(csr_alloc(ctx, lane, b));
    TW_TRY(upload_plain(ctx, lane, b->rowptr, rowptr, sizeof(int64_t) * (size_t)(n + 1)));
    if (nnz) TW_TRY(upload_plain(ctx, lane, b->colidx, colidx, sizeof(int32_t) * (size_t)nnz));
    if (nnz) TW_TRY(upload_plain(ctx, lane, b->vals, vals, sizeof(double) * (size_t)nnz));
    if (n) TW_TRY(upload_plain(ctx, lane, b->labels, labels, sizeof(double) * (size_t)n));
    TW_TRY(csr_finalize(ctx, lane, b, defer_global));
    if (b->ready) TW_CUDA(cudaEventRecord(b->ready, lane.stream));
    return 0;
}

struct TweetArgs {
    const uint16_t* text_utf16;
    const int64_t* text_off;
    const int32_t *retweet_cnt, *followers, *favourites, *friends;
    const int64_t* created_ms;
    const uint8_t* is_retweet;
    int64_t n, now_ms;
    int32_t rt_begin, rt_end, F;
    uint32_t flags;
};

// rows that pass MllibHelper.filtrate (:89-94): the host can tell without the GPU, so an asynchronously ingested batch
// knows its row count (rdd.count, LinearRegression.scala:59) at once; also refuses texts the featuriser cannot hold
static int tweets_precheck(const TweetArgs& t, int64_t* n_kept)
{
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(std::min(16u, std::max(1u, std::thread::hardware_concurrency())), t.n / 65536));
    std::vector<int64_t> kept((size_t)nt, 0), bad((size_t)nt, -1);
    auto scan = [&](int k) { // a slice of the tweets on its own host thread
        const int64_t lo = t.n * k / nt, hi = t.n * (k + 1) / nt;
        int64_t c = 0;
        for (int64_t i = lo; i < hi; ++i) {
            const int64_t len = t.text_off[i + 1] - t.text_off[i];
            if ((len < 0 || len > kMaxTerms + 1) && bad[(size_t)k] < 0) bad[(size_t)k] = i;
            c += t.is_retweet[i] != 0 && t.retweet_cnt[i] >= t.rt_begin && t.retweet_cnt[i] <= t.rt_end;
        }
        kept[(size_t)k] = c;
    };
    if (nt == 1) scan(0);
    else {
        std::vector<std::thread> th;
        for (int k = 0; k < nt; ++k) th.emplace_back(scan, k);
        for (auto& x : th) x.join();
    }
    int64_t total = 0;
    for (int k = 0; k < nt; ++k) {
        if (bad[(size_t)k] >= 0) {
            const int64_t i = bad[(size_t)k], len = t.text_off[i + 1] - t.text_off[i];
            if (len < 0) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: text offsets decrease at tweet %lld", (long long)i);
            return fail(TWTML_E_INVALID_ARG, "a tweet has %lld UTF-16 units; the featuriser takes at most %d", (long long)len, kMaxTerms + 1);
        }
        total += kept[(size_t)k];
    }
    *n_kept = total;
    return 0;
}

// The address under which the current device can read [p, p + bytes) of the caller's host memory, or null when that memory
// is not page-locked / not mapped for this device.
static const void* device_alias(const void* p, size_t bytes)
{
    cudaPointerAttributes first{}, last{};
    void* dp = nullptr;
    const bool ok = bytes > 0 && cudaPointerGetAttributes(&first, p) == cudaSuccess && first.type == cudaMemoryTypeHost &&
                    cudaPointerGetAttributes(&last, static_cast<const char*>(p) + bytes - 1) == cudaSuccess && last.type == cudaMemoryTypeHost &&
                    cudaHostGetDevicePointer(&dp, const_cast<void*>(p), 0) == cudaSuccess && dp != nullptr;
    cudaGetLastError();
    return ok ? dp : nullptr;
}

static int tweets_ingest(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b, const TweetArgs& t, bool defer_global)
{
    std::lock_guard<std::mutex> lane_lock(lane.mu);
    cudaStream_t st = lane.stream;
    const int64_t n = t.n;
    const int F = t.F;
    if (n == 0) {
        TW_TRY(csr_alloc(ctx, lane, b));
        TW_CUDA(cudaMemsetAsync(b->rowptr, 0, sizeof(long long), st));
        TW_TRY(csr_finalize(ctx, lane, b, defer_global));
        if (b->ready) TW_CUDA(cudaEventRecord(b->ready, st));
        return 0;
    }
    const int64_t units = t.text_off[n];
    // device copies of the tweet columns
    uint16_t* d_text = nullptr;
    long long *d_off = nullptr, *d_created = nullptr, *keep = nullptr, *rownnz = nullptr, *krow = nullptr, *eoff = nullptr;
    int *d_rc = nullptr, *d_fo = nullptr, *d_fa = nullptr, *d_fr = nullptr, *pair_idx = nullptr, *pair_cnt = nullptr;
    uint8_t* d_rt = nullptr;
#define TW_DO(call) do { int rc_ = (call); if (rc_) return rc_; } while (0)
    // TWTML_TRACE_INGEST=1: where one ingest spends its time, on the device (events on the lane's stream) and on this thread
    static const bool trace = getenv("TWTML_TRACE_INGEST") != nullptr;
    cudaEvent_t tev[6] = {};
    const auto h0 = std::chrono::steady_clock::now();
    double hms[16] = {};
    auto mark = [&](int k) { if (trace) { if (!tev[k]) cudaEventCreate(&tev[k]); cudaEventRecord(tev[k], st); } };
    auto hmark = [&](int k) { hms[k] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - h0).count(); };
    // The text of a tweet is needed only if the tweet passes the filter, and the filter does not look at it.  When
    // few tweets pass and the caller's text is page-locked memory this device can address, the featuriser reads the
    // kept tweets' text in place over the host link and the text column is not copied at all (10^6 tweets, 14 % kept:
    // ~20 MB cross the link instead of 140 MB).  b->n is the host filter's count of kept tweets.
    static const bool no_in_place = getenv("TWTML_NO_TEXT_IN_PLACE") != nullptr; // A/B switch
    const uint16_t* text_in_place = nullptr;
    if (units && !no_in_place && b->n * 2 <= n) text_in_place = static_cast<const uint16_t*>(device_alias(t.text_utf16, sizeof(uint16_t) * (size_t)units));
    // The scratch arrays live in one buffer of the lane that is kept from interval to interval (it grows to the largest
    // interval seen): no allocator call per interval.
    auto al = [](size_t bytes) { return (std::max<size_t>(bytes, 1) + 255) & ~(size_t)255; };
    const size_t un = (size_t)units, nn = (size_t)n;
    const size_t total = (text_in_place ? 0 : al(2 * un)) + 5 * al(8 * (nn + 1)) + al(8 * nn) + 4 * al(4 * nn) + al(nn) + 2 * al(4 * un);
    TW_TRY(lane.feat_arena.reserve(total));
    size_t arena_off = 0;
    auto carve = [&](auto** p, size_t count) -> int {
        using T = std::remove_pointer_t<std::remove_pointer_t<decltype(p)>>;
        *p = reinterpret_cast<T*>(lane.feat_arena.as<char>() + arena_off);
        arena_off += al(sizeof(T) * count);
        return arena_off <= lane.feat_arena.cap ? 0 : fail(TWTML_E_CUDA, "featuriser scratch overrun");
    };
    if (!text_in_place) TW_DO(carve(&d_text, un));
    TW_DO(carve(&d_off, nn + 1));
    TW_DO(carve(&d_created, nn));
    TW_DO(carve(&d_rc, nn));
    TW_DO(carve(&d_fo, nn));
    TW_DO(carve(&d_fa, nn));
    TW_DO(carve(&d_fr, nn));
    TW_DO(carve(&d_rt, nn));
    TW_DO(carve(&pair_idx, un));
    TW_DO(carve(&pair_cnt, un));
    TW_DO(carve(&keep, nn + 1));
    TW_DO(carve(&rownnz, nn + 1));
    TW_DO(carve(&krow, nn + 1));
    TW_DO(carve(&eoff, nn + 1));
    hmark(3);
    mark(0);
    hmark(8);
    if (units && !text_in_place) TW_DO(upload_plain(ctx, lane, d_text, t.text_utf16, sizeof(uint16_t) * (size_t)units));
    TW_DO(upload_plain(ctx, lane, d_off, t.text_off, sizeof(int64_t) * (size_t)(n + 1)));
    hmark(6);
    TW_DO(upload_plain(ctx, lane, d_created, t.created_ms, sizeof(int64_t) * (size_t)n));
    hmark(9);
    TW_DO(upload_plain(ctx, lane, d_rc, t.retweet_cnt, sizeof(int32_t) * (size_t)n));
    hmark(10);
    TW_DO(upload_plain(ctx, lane, d_fo, t.followers, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, lane, d_fa, t.favourites, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, lane, d_fr, t.friends, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, lane, d_rt, t.is_retweet, (size_t)n));

    hmark(4);
    const uint16_t* lower = nullptr;
    if (t.flags & TWTML_FLAG_LOWERCASE) TW_DO(lower_table(ctx, &lower));
    TweetCols tc{text_in_place ? text_in_place : d_text, d_off, d_rc, d_fo, d_fa, d_fr, d_created, d_rt};
    cudaMemsetAsync(keep + n, 0, sizeof(long long), st);
    cudaMemsetAsync(rownnz + n, 0, sizeof(long long), st);
    mark(1);
    const long long cta_cap = (long long)ctx->sm_count * 16;
    const unsigned gridA = (unsigned)std::min<long long>((n + kFeatWarps - 1) / kFeatWarps, cta_cap);
    k_feat_terms<<<gridA, kFeatWarps * 32, 0, st>>>(tc, n, t.rt_begin, t.rt_end, F, lower, pair_idx, pair_cnt, keep, rownnz);
    hmark(5);
    mark(2);
    TW_DO(exclusive_sum(lane, keep, krow, n + 1));
    TW_DO(exclusive_sum(lane, rownnz, eoff, n + 1));
    mark(3);
    hmark(0);
    // the one size only the device knows: how many entries the kept rows have (the row count was settled on the host)
    long long tot[2] = {0, 0};
    cudaMemcpyAsync(&tot[0], krow + n, sizeof(long long), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&tot[1], eoff + n, sizeof(long long), cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) return fail(TWTML_E_CUDA, "featuriser failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (tot[0] != b->n) { return fail(TWTML_E_CUDA, "featuriser kept %lld rows, the host filter %lld", tot[0], (long long)b->n); }
    b->nnz = tot[1];
    hmark(1);
    TW_DO(csr_alloc(ctx, lane, b));
    const unsigned gridB = (unsigned)std::min<long long>((n + 3) / 4, cta_cap);
    k_feat_pack<<<gridB, 128, 0, st>>>(tc, n, t.now_ms, F, pair_idx, pair_cnt, keep, krow, rownnz, eoff, b->rowptr, b->colidx, b->vals, b->labels);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(b->rowptr + b->n, eoff + n, sizeof(long long), cudaMemcpyDeviceToDevice, st);
#undef TW_DO
    if (e != cudaSuccess) return fail(TWTML_E_CUDA, "featuriser pack failed: %s", cudaGetErrorString(e));
    mark(4);
    TW_TRY(csr_finalize(ctx, lane, b, defer_global));
    mark(5);
    if (b->ready) TW_CUDA(cudaEventRecord(b->ready, st));
    if (trace) {
        hmark(2);
        cudaStreamSynchronize(st);
        float d[5] = {};
        for (int k = 0; k < 5; ++k) cudaEventElapsedTime(&d[k], tev[k], tev[k + 1]);
        const double hend = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - h0).count();
        fprintf(stderr, "[twtml ingest] device ms: copies %.3f  terms %.3f  scans %.3f  sync+pack %.3f  finalize %.3f | host ms: allocs %.3f  first copy %.3f  copies %.3f  terms %.3f  scans %.3f  sizes back %.3f  all enqueued %.3f  done %.3f\n",
                d[0], d[1], d[2], d[3], d[4], hms[3], hms[6], hms[4], hms[5], hms[0], hms[1], hms[2], hend);
        fprintf(stderr, "[twtml ingest]   host ms: allocs %.3f  event %.3f  copy1 %.3f  copy2 %.3f  copy3 %.3f  all copies %.3f\n", hms[3], hms[8], hms[6], hms[9], hms[10], hms[4]);
        for (int k = 0; k < 6; ++k) cudaEventDestroy(tev[k]);
    }
    return 0;
}

static int check_tweet_args(twtml_ctx* ctx, twtml_batch** out, const TweetArgs& t)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (t.n < 0 || t.F < 1 || t.F > (1 << 30) - 4) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: bad sizes");
    if (t.n > 0 && (!t.text_off || !t.retweet_cnt || !t.followers || !t.favourites || !t.friends || !t.created_ms || !t.is_retweet))
        return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: null column");
    const int64_t units = t.n > 0 ? t.text_off[t.n] : 0;
    if (t.n > 0 && (t.text_off[0] != 0 || units < 0 || (units > 0 && !t.text_utf16))) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: bad text offsets");
    return 0;
}

// an event for a batch that is filled on the copy stream; batch memory is pool memory, so no ordering against the
// compute stream is needed before the first touch (nothing of it existed before)
static int make_ready_event(twtml_batch* b)
{
    TW_CUDA(cudaSetDevice(b->ctx->device)); // an event belongs to the device that is current when it is made
    TW_CUDA(cudaEventCreateWithFlags(&b->ready, cudaEventDisableTiming));
    return 0;
}

} // namespace twtml

using namespace twtml;

extern "C" {

static int32_t check_csr_args(twtml_ctx* ctx, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals,
                              const double* labels, twtml_batch** out, int64_t* nnz_out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d < 1 || d > ((int64_t)1 << 30) || !rowptr || (n > 0 && !labels)) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_csr: bad shape or null data");
    const int64_t nnz = rowptr[n] - rowptr[0];
    if (rowptr[0] != 0 || nnz < 0 || (nnz > 0 && (!colidx || !vals))) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_csr: rowptr must start at 0 and be non-decreasing");
    *nnz_out = nnz;
    return 0;
}

int32_t twtml_batch_from_csr(twtml_ctx* ctx, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals,
                             const double* labels, twtml_batch** out)
{
    int64_t nnz = 0;
    TW_TRY(check_csr_args(ctx, n, d, rowptr, colidx, vals, labels, out, &nnz));
    if (!ctx->members.empty()) return grp::from_csr(ctx, n, d, rowptr, colidx, vals, labels, false, out);
    Guard g(ctx);
    twtml_batch* b = csr_shell(ctx, n, d, nnz);
    int rc = csr_ingest(ctx, ctx->lane[0], b, rowptr, colidx, vals, labels, false);
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(TWTML_E_CUDA, "CSR ingest failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

int32_t twtml_batch_from_csr_async(twtml_ctx* ctx, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals,
                                   const double* labels, twtml_batch** out)
{
    int64_t nnz = 0;
    TW_TRY(check_csr_args(ctx, n, d, rowptr, colidx, vals, labels, out, &nnz));
    if (!ctx->members.empty()) return grp::from_csr(ctx, n, d, rowptr, colidx, vals, labels, true, out);
    twtml_batch* b = csr_shell(ctx, n, d, nnz);
    if (int rc = make_ready_event(b)) { Guard g(ctx); batch_free_locked(b); return rc; }
    submit_ingest(ctx, b, [=]() { return csr_ingest(ctx, ctx->lane[1], b, rowptr, colidx, vals, labels, true); });
    *out = b;
    return 0;
}

int32_t twtml_batch_synth_csr(twtml_ctx* ctx, int64_t n, int32_t d_text, uint64_t seed, int64_t row0, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d_text < 1) return fail(TWTML_E_INVALID_ARG, "twtml_batch_synth_csr: bad shape");
    if (!ctx->members.empty()) return grp::synth_csr(ctx, n, d_text, seed, row0, out);
    Guard g(ctx);
    IngestLane& lane = ctx->lane[0];
    std::lock_guard<std::mutex> lane_lock(lane.mu);
    long long *rownnz = nullptr, *rowptr = nullptr;
    TW_TRY(dev_alloc(ctx, lane, &rownnz, (size_t)n + 1));
    TW_TRY(dev_alloc(ctx, lane, &rowptr, (size_t)n + 1));
    TW_CUDA(cudaMemsetAsync(rownnz, 0, sizeof(long long) * ((size_t)n + 1), ctx->stream));
    if (n > 0) k_synth_csr_count<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(rownnz, n, d_text, seed, row0);
    int rc = exclusive_sum(lane, rownnz, rowptr, n + 1);
    long long nnz = 0;
    if (!rc && cudaMemcpyAsync(&nnz, rowptr + n, sizeof(nnz), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) rc = fail(TWTML_E_CUDA, "nnz copy failed");
    cudaStreamSynchronize(ctx->stream);
    pool_free(ctx, rownnz);
    twtml_batch* b = nullptr;
    if (!rc) rc = new_csr(ctx, lane, n, (int64_t)d_text + 4, nnz, &b);
    if (!rc) {
        cudaMemcpyAsync(b->rowptr, rowptr, sizeof(long long) * ((size_t)n + 1), cudaMemcpyDeviceToDevice, ctx->stream);
        if (n > 0) k_synth_csr_fill<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(b->rowptr, b->colidx, b->vals, b->labels, n, d_text, seed, row0);
        if (cudaGetLastError() != cudaSuccess) rc = fail(TWTML_E_CUDA, "synthetic CSR generator failed to launch");
    }
    pool_free(ctx, rowptr);
    if (!rc) rc = csr_finalize(ctx, lane, b, false);
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(TWTML_E_CUDA, "synthetic CSR batch: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc) { if (b) batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

int32_t twtml_batch_from_tweets(twtml_ctx* ctx, const uint16_t* text_utf16, const int64_t* text_off, const int32_t* retweet_cnt,
                                const int32_t* followers, const int32_t* favourites, const int32_t* friends, const int64_t* created_ms,
                                const uint8_t* is_retweet, int64_t n, int64_t now_ms, int32_t rt_begin, int32_t rt_end,
                                int32_t num_text_features, uint32_t flags, twtml_batch** out)
{
    const TweetArgs t{text_utf16, text_off, retweet_cnt, followers, favourites, friends, created_ms, is_retweet, n, now_ms, rt_begin, rt_end, num_text_features, flags};
    TW_TRY(check_tweet_args(ctx, out, t));
    if (!ctx->members.empty())
        return grp::from_tweets(ctx, grp::TweetCols{text_utf16, text_off, retweet_cnt, followers, favourites, friends, created_ms, is_retweet}, n, now_ms,
                                rt_begin, rt_end, num_text_features, flags, false, out);
    int64_t kept = 0;
    TW_TRY(tweets_precheck(t, &kept));
    Guard g(ctx);
    twtml_batch* b = csr_shell(ctx, kept, (int64_t)num_text_features + 4, 0);
    int rc = tweets_ingest(ctx, ctx->lane[0], b, t, false);
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(TWTML_E_CUDA, "tweet ingest failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

int32_t twtml_batch_from_tweets_async(twtml_ctx* ctx, const uint16_t* text_utf16, const int64_t* text_off, const int32_t* retweet_cnt,
                                      const int32_t* followers, const int32_t* favourites, const int32_t* friends, const int64_t* created_ms,
                                      const uint8_t* is_retweet, int64_t n, int64_t now_ms, int32_t rt_begin, int32_t rt_end,
                                      int32_t num_text_features, uint32_t flags, twtml_batch** out)
{
    const TweetArgs t{text_utf16, text_off, retweet_cnt, followers, favourites, friends, created_ms, is_retweet, n, now_ms, rt_begin, rt_end, num_text_features, flags};
    TW_TRY(check_tweet_args(ctx, out, t));
    if (!ctx->members.empty())
        return grp::from_tweets(ctx, grp::TweetCols{text_utf16, text_off, retweet_cnt, followers, favourites, friends, created_ms, is_retweet}, n, now_ms,
                                rt_begin, rt_end, num_text_features, flags, true, out);
    int64_t kept = 0;
    TW_TRY(tweets_precheck(t, &kept));
    twtml_batch* b = csr_shell(ctx, kept, (int64_t)num_text_features + 4, 0);
    if (int rc = make_ready_event(b)) { Guard g(ctx); batch_free_locked(b); return rc; }
    submit_ingest(ctx, b, [=]() { return tweets_ingest(ctx, ctx->lane[1], b, t, true); });
    *out = b;
    return 0;
}

int32_t twtml_batch_export_csr(const twtml_batch* b, int64_t* rowptr, int32_t* colidx, double* vals, double* labels)
{
    if (!b || !rowptr) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (!b->csr) return fail(TWTML_E_INVALID_ARG, "twtml_batch_export_csr on a dense batch");
    if (!b->members.empty()) return grp::export_csr(b, rowptr, colidx, vals, labels);
    twtml_ctx* ctx = b->ctx;
    Guard g(ctx);
    TW_TRY(settle_batch(ctx, b));
    if (b->ready) TW_CUDA(cudaStreamWaitEvent(ctx->stream, b->ready, 0));
    TW_CUDA(cudaMemcpyAsync(rowptr, b->rowptr, sizeof(int64_t) * (size_t)(b->n + 1), cudaMemcpyDeviceToHost, ctx->stream));
    if (b->nnz && colidx) TW_CUDA(cudaMemcpyAsync(colidx, b->colidx, sizeof(int32_t) * (size_t)b->nnz, cudaMemcpyDeviceToHost, ctx->stream));
    if (b->nnz && vals) TW_CUDA(cudaMemcpyAsync(vals, b->vals, sizeof(double) * (size_t)b->nnz, cudaMemcpyDeviceToHost, ctx->stream));
    if (b->n && labels) TW_CUDA(cudaMemcpyAsync(labels, b->labels, sizeof(double) * (size_t)b->n, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

} // extern "C"
