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
#include <thread>
#include <vector>

#include "csr.cuh"
#include "csr_narrow.cuh"
#include "featurize.cuh"
#include "host.h"
#include "lower_table.h"
#include "reduce.cuh"
#include "synth.cuh"

namespace twtml {

template <class T>
static int dev_alloc(twtml_ctx* ctx, T** p, size_t count)
{
    return pool_alloc(ctx, reinterpret_cast<void**>(p), sizeof(T) * std::max<size_t>(count, 1));
}

static int exclusive_sum(twtml_ctx* ctx, const long long* in, long long* out, long long n)
{
    size_t tmp = 0;
    cudaGetLastError(); // CUB reports any error still pending in the runtime as its own
    TW_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, in, out, n, ctx->stream));
    TW_TRY(ctx->gather.reserve(tmp));
    TW_CUDA(cub::DeviceScan::ExclusiveSum(ctx->gather.p, tmp, in, out, n, ctx->stream));
    return 0;
}

static unsigned blocks_for(long long n, int threads) { return (unsigned)std::max<long long>(1, (n + threads - 1) / threads); }

// Narrow models: no transpose at all, only the first row of every CTA of the single-pass kernel.
// Returns 1 when the batch does not qualify (the caller then builds the global transpose).
static int csr_finalize_narrow(twtml_ctx* ctx, twtml_batch* b)
{
    const long long n = b->n;
    const int d = (int)b->d;
    if (d > kNarrowMaxD || getenv("TWTML_NO_NARROW")) return 1; // a function of the model width only: every rank decides alike
    const int warps = narrow_warps(d);
    if (warps < 4) return 1;
    // one partial vector per CTA of a grid fixed here: the plan's "blocks" are those CTAs
    const long long grid = std::max<long long>(1, std::min<long long>((n + warps - 1) / warps, (long long)ctx->sm_count));
    TW_TRY(dev_alloc(ctx, &b->bounds, (size_t)grid + 1));
    k_narrow_bounds<<<blocks_for(grid + 1, 256), 256, 0, ctx->stream>>>(b->rowptr, n, grid, b->bounds);
    TW_CUDA(cudaGetLastError());
    b->narrow = true;
    b->nwarps = warps;
    memset(&b->plan, 0, sizeof(b->plan));
    b->plan.spr = 1;
    b->plan.rb = 1;
    b->plan.nblocks = grid;
    for (int s2 = 1; s2 <= kMaxShards; ++s2) { b->plan.shard_row0[s2] = n; b->plan.shard_blk0[s2] = grid; }
    return 0;
}

// rowptr / colidx / vals / labels are on the device: build the plan, the transpose and the column tasks
static int csr_finalize(twtml_ctx* ctx, twtml_batch* b)
{
    // one shard per rank (the column sums are per rank, not per virtual shard); row blocks small enough to give every SM
    // several CTAs of the row pass even for a modest micro-batch (but at least 32 rows: one step of the widest row form)
    make_plan(b->n, 1, 32, b->plan, 4096);
    TW_TRY(ctx_global_rows(ctx, b->n, &b->n_global, &b->row0_global));
    if (b->nnz >= ((long long)1 << 31) || b->n >= ((long long)1 << 31))
        return fail(TWTML_E_INVALID_ARG, "CSR batch too large for 32-bit entry indices (%lld rows, %lld nnz per GPU)", (long long)b->n, (long long)b->nnz);
    {
        const int nrc = csr_finalize_narrow(ctx, b);
        if (nrc <= 0) return nrc; // built (0) or failed (< 0); 1 = not a narrow batch
    }
    const int d = (int)b->d;
    const long long nnz = b->nnz;
    const long long nwin = (nnz + kWin - 1) / kWin;
    b->nwin = nwin;
    TW_TRY(dev_alloc(ctx, &b->colptr, (size_t)d + 1));
    TW_TRY(dev_alloc(ctx, &b->rowidx, (size_t)nnz));
    TW_TRY(dev_alloc(ctx, &b->cvals, (size_t)nnz));
    TW_TRY(dev_alloc(ctx, &b->diff, (size_t)b->n));
    TW_TRY(dev_alloc(ctx, &b->headbits, (size_t)(nwin + 1) * 8));
    TW_TRY(dev_alloc(ctx, &b->win_first_nz, (size_t)nwin + 1));
    TW_TRY(dev_alloc(ctx, &b->wkey, (size_t)nwin + 1));
    TW_TRY(dev_alloc(ctx, &b->win_first_part, (size_t)nwin + 1));
    TW_TRY(dev_alloc(ctx, &b->win_last_part, (size_t)nwin + 1));
    TW_TRY(dev_alloc(ctx, &b->nzcol, (size_t)std::min<long long>(nnz, d) + 1));
    TW_CUDA(cudaMemsetAsync(b->headbits, 0, sizeof(unsigned) * (size_t)(nwin + 1) * 8, ctx->stream));
    if (nnz > 0) {
        int *keys_out = nullptr, *iota = nullptr, *perm = nullptr;
        TW_TRY(dev_alloc(ctx, &keys_out, (size_t)nnz));
        TW_TRY(dev_alloc(ctx, &iota, (size_t)nnz));
        TW_TRY(dev_alloc(ctx, &perm, (size_t)nnz));
        int rc = 0;
        do {
            k_iota<<<blocks_for(nnz, 256), 256, 0, ctx->stream>>>(iota, nnz);
            int bits = 1;
            while (((long long)1 << bits) < d) ++bits;
            size_t tmp = 0;
            cudaGetLastError();
            cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, tmp, b->colidx, keys_out, iota, perm, (int)nnz, 0, bits, ctx->stream);
            if (e != cudaSuccess) { rc = fail(TWTML_E_CUDA, "radix sort sizing: %s", cudaGetErrorString(e)); break; }
            if ((rc = ctx->tmp_a.reserve(tmp))) break;
            // LSD radix sort is stable: entries of a column keep their CSR (row) order
            e = cub::DeviceRadixSort::SortPairs(ctx->tmp_a.p, tmp, b->colidx, keys_out, iota, perm, (int)nnz, 0, bits, ctx->stream);
            if (e != cudaSuccess) { rc = fail(TWTML_E_CUDA, "radix sort: %s", cudaGetErrorString(e)); break; }
            k_colptr<<<blocks_for(nnz + 1, 256), 256, 0, ctx->stream>>>(keys_out, nnz, d, b->colptr);
            k_csc_gather<<<blocks_for(nnz, 256), 256, 0, ctx->stream>>>(perm, b->rowptr, b->n, b->vals, nnz, b->rowidx, b->cvals);
            // window index of the transpose: head bits, heads per window -> first non-empty column per window
            k_csc_headbits<<<blocks_for(nnz + 1, 256), 256, 0, ctx->stream>>>(keys_out, nnz, b->headbits);
            long long* winheads = reinterpret_cast<long long*>(iota); // iota is dead from here on; nnz ints >= nwin+1 int64 when nnz >= 2
            long long* winheads_own = nullptr;
            if ((size_t)nnz * sizeof(int) < (size_t)(nwin + 1) * sizeof(long long)) {
                if ((rc = dev_alloc(ctx, &winheads_own, (size_t)nwin + 1))) break;
                winheads = winheads_own;
            }
            k_csc_winmeta<<<blocks_for(nwin + 1, 256), 256, 0, ctx->stream>>>(b->headbits, keys_out, nnz, nwin, winheads, b->wkey);
            rc = exclusive_sum(ctx, winheads, b->win_first_nz, nwin + 1);
            if (!rc) k_csc_nzcol<<<blocks_for(nnz, 256), 256, 0, ctx->stream>>>(keys_out, b->headbits, b->win_first_nz, nnz, b->nzcol);
            if (!rc && cudaGetLastError() != cudaSuccess) rc = fail(TWTML_E_CUDA, "transpose kernels failed to launch");
            cudaStreamSynchronize(ctx->stream);
            pool_free(ctx, winheads_own);
        } while (0);
        cudaStreamSynchronize(ctx->stream);
        pool_free(ctx, keys_out);
        pool_free(ctx, iota);
        pool_free(ctx, perm);
        if (rc) return rc;
    } else {
        TW_CUDA(cudaMemsetAsync(b->colptr, 0, sizeof(long long) * ((size_t)d + 1), ctx->stream));
        TW_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return 0;
}

static int new_csr(twtml_ctx* ctx, int64_t n, int64_t d, int64_t nnz, twtml_batch** out)
{
    twtml_batch* b = new twtml_batch();
    b->ctx = ctx;
    b->csr = true;
    b->n = n;
    b->d = d;
    b->nnz = nnz;
    ctx->live_handles += 1; // released by twtml_batch_free, also on the failure paths below and in the callers
    // colidx / vals end in kNarrowSlack spare entries: the narrow kernel's bulk copies are 16 bytes wide
    int rc = dev_alloc(ctx, &b->rowptr, (size_t)n + 1);
    if (!rc) rc = dev_alloc(ctx, &b->colidx, (size_t)nnz + kNarrowSlack);
    if (!rc) rc = dev_alloc(ctx, &b->vals, (size_t)nnz + kNarrowSlack);
    if (!rc) rc = dev_alloc(ctx, &b->labels, (size_t)n);
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

// ---- per-iteration launches -----------------------------------------------------------------
int csr_partial_width(const twtml_batch* b) { return (int)b->d + 2; }

static int launch_narrow(twtml_ctx* ctx, const NarrowArgs& a, int mode, unsigned grid)
{
    const size_t smem = narrow_smem_bytes(a.d, a.warps);
#define TW_NARROW(MODE_)                                                                                  \
    do {                                                                                                  \
        auto kern = k_csr_narrow<MODE_>;                                                                  \
        TW_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));     \
        kern<<<grid, (a.warps + 1) * 32, smem, ctx->stream>>>(a);                                               \
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
        na.colidx = b->colidx;
        na.vals = b->vals;
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
    if (!ctx->lower.p) {
        std::vector<uint16_t> t(65536);
        for (int c = 0; c < 65536; ++c) t[c] = (uint16_t)c;
        for (int i = 0; i < kLowerPairCount; ++i) t[kLowerPairs[i][0]] = kLowerPairs[i][1];
        TW_TRY(ctx->lower.reserve(sizeof(uint16_t) * 65536));
        TW_CUDA(cudaMemcpyAsync(ctx->lower.p, t.data(), sizeof(uint16_t) * 65536, cudaMemcpyHostToDevice, ctx->stream));
        TW_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    *out = ctx->lower.as<uint16_t>();
    return 0;
}

} // namespace twtml

using namespace twtml;

extern "C" {

int32_t twtml_batch_from_csr(twtml_ctx* ctx, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals,
                             const double* labels, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d < 1 || d > ((int64_t)1 << 30) || !rowptr || (n > 0 && !labels)) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_csr: bad shape or null data");
    const int64_t nnz = rowptr[n] - rowptr[0];
    if (rowptr[0] != 0 || nnz < 0 || (nnz > 0 && (!colidx || !vals))) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_csr: rowptr must start at 0 and be non-decreasing");
    // the row pointers alone first (a malformed one must be refused before any index is read through it)
    for (int64_t r = 0; r < n; ++r)
        if (rowptr[r] < 0 || rowptr[r] > rowptr[r + 1] || rowptr[r + 1] > nnz)
            return fail(TWTML_E_INVALID_ARG, "row %lld: rowptr must be non-decreasing within [0, rowptr[n]]", (long long)r);
    { // Vectors.sparse require()s strictly increasing, in-range indices [U]; checked on all host threads
        std::atomic<int64_t> bad_row(-1);
        const unsigned hw = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
        const int nt = (int)std::min<int64_t>(hw, std::max<int64_t>(1, n / 4096));
        auto scan = [&](int64_t lo, int64_t hi) {
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
    }
    Guard g(ctx);
    twtml_batch* b = nullptr;
    TW_TRY(new_csr(ctx, n, d, nnz, &b));
    int rc = upload_plain(ctx, b->rowptr, rowptr, sizeof(int64_t) * (size_t)(n + 1));
    if (!rc && nnz) rc = upload_plain(ctx, b->colidx, colidx, sizeof(int32_t) * (size_t)nnz);
    if (!rc && nnz) rc = upload_plain(ctx, b->vals, vals, sizeof(double) * (size_t)nnz);
    if (!rc && n) rc = upload_plain(ctx, b->labels, labels, sizeof(double) * (size_t)n);
    if (!rc) rc = csr_finalize(ctx, b);
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

int32_t twtml_batch_synth_csr(twtml_ctx* ctx, int64_t n, int32_t d_text, uint64_t seed, int64_t row0, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d_text < 1) return fail(TWTML_E_INVALID_ARG, "twtml_batch_synth_csr: bad shape");
    Guard g(ctx);
    long long *rownnz = nullptr, *rowptr = nullptr;
    TW_TRY(dev_alloc(ctx, &rownnz, (size_t)n + 1));
    TW_TRY(dev_alloc(ctx, &rowptr, (size_t)n + 1));
    TW_CUDA(cudaMemsetAsync(rownnz, 0, sizeof(long long) * ((size_t)n + 1), ctx->stream));
    if (n > 0) k_synth_csr_count<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(rownnz, n, d_text, seed, row0);
    int rc = exclusive_sum(ctx, rownnz, rowptr, n + 1);
    long long nnz = 0;
    if (!rc && cudaMemcpyAsync(&nnz, rowptr + n, sizeof(nnz), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) rc = fail(TWTML_E_CUDA, "nnz copy failed");
    cudaStreamSynchronize(ctx->stream);
    pool_free(ctx, rownnz);
    twtml_batch* b = nullptr;
    if (!rc) rc = new_csr(ctx, n, (int64_t)d_text + 4, nnz, &b);
    if (!rc) {
        cudaMemcpyAsync(b->rowptr, rowptr, sizeof(long long) * ((size_t)n + 1), cudaMemcpyDeviceToDevice, ctx->stream);
        if (n > 0) k_synth_csr_fill<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(b->rowptr, b->colidx, b->vals, b->labels, n, d_text, seed, row0);
        if (cudaGetLastError() != cudaSuccess) rc = fail(TWTML_E_CUDA, "synthetic CSR generator failed to launch");
    }
    cudaStreamSynchronize(ctx->stream);
    pool_free(ctx, rowptr);
    if (!rc) rc = csr_finalize(ctx, b);
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

int32_t twtml_batch_from_tweets(twtml_ctx* ctx, const uint16_t* text_utf16, const int64_t* text_off, const int32_t* retweet_cnt,
                                const int32_t* followers, const int32_t* favourites, const int32_t* friends, const int64_t* created_ms,
                                const uint8_t* is_retweet, int64_t n, int64_t now_ms, int32_t rt_begin, int32_t rt_end,
                                int32_t num_text_features, uint32_t flags, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || num_text_features < 1 || num_text_features > (1 << 30) - 4) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: bad sizes");
    if (n > 0 && (!text_off || !retweet_cnt || !followers || !favourites || !friends || !created_ms || !is_retweet))
        return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: null column");
    const int64_t units = n > 0 ? text_off[n] : 0;
    if (n > 0 && (text_off[0] != 0 || units < 0 || (units > 0 && !text_utf16))) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_tweets: bad text offsets");
    Guard g(ctx);
    const int F = num_text_features;
    twtml_batch* b = nullptr;
    if (n == 0) {
        TW_TRY(new_csr(ctx, 0, (int64_t)F + 4, 0, &b));
        TW_CUDA(cudaMemsetAsync(b->rowptr, 0, sizeof(long long), ctx->stream));
        int rc = csr_finalize(ctx, b);
        if (rc) { batch_free_locked(b); return rc; }
        *out = b;
        return 0;
    }
    // device copies of the tweet columns
    uint16_t* d_text = nullptr;
    long long *d_off = nullptr, *d_created = nullptr, *keep = nullptr, *rownnz = nullptr, *krow = nullptr, *eoff = nullptr;
    int *d_rc = nullptr, *d_fo = nullptr, *d_fa = nullptr, *d_fr = nullptr, *pair_idx = nullptr, *pair_cnt = nullptr, *d_maxlen = nullptr;
    uint8_t* d_rt = nullptr;
    std::vector<void*> owned;
    auto cleanup = [&]() { for (void* p : owned) pool_free(ctx, p); };
#define TW_OWN(call, ptrvar) do { int rc_ = (call); if (rc_) { cleanup(); return rc_; } owned.push_back(ptrvar); } while (0)
#define TW_DO(call) do { int rc_ = (call); if (rc_) { cleanup(); return rc_; } } while (0)
    TW_OWN(dev_alloc(ctx, &d_text, (size_t)units), d_text);
    TW_OWN(dev_alloc(ctx, &d_off, (size_t)n + 1), d_off);
    TW_OWN(dev_alloc(ctx, &d_created, (size_t)n), d_created);
    TW_OWN(dev_alloc(ctx, &d_rc, (size_t)n), d_rc);
    TW_OWN(dev_alloc(ctx, &d_fo, (size_t)n), d_fo);
    TW_OWN(dev_alloc(ctx, &d_fa, (size_t)n), d_fa);
    TW_OWN(dev_alloc(ctx, &d_fr, (size_t)n), d_fr);
    TW_OWN(dev_alloc(ctx, &d_rt, (size_t)n), d_rt);
    TW_OWN(dev_alloc(ctx, &pair_idx, (size_t)units), pair_idx);
    TW_OWN(dev_alloc(ctx, &pair_cnt, (size_t)units), pair_cnt);
    TW_OWN(dev_alloc(ctx, &keep, (size_t)n + 1), keep);
    TW_OWN(dev_alloc(ctx, &rownnz, (size_t)n + 1), rownnz);
    TW_OWN(dev_alloc(ctx, &krow, (size_t)n + 1), krow);
    TW_OWN(dev_alloc(ctx, &eoff, (size_t)n + 1), eoff);
    TW_OWN(dev_alloc(ctx, &d_maxlen, 1), d_maxlen);
    if (units) TW_DO(upload_plain(ctx, d_text, text_utf16, sizeof(uint16_t) * (size_t)units));
    TW_DO(upload_plain(ctx, d_off, text_off, sizeof(int64_t) * (size_t)(n + 1)));
    TW_DO(upload_plain(ctx, d_created, created_ms, sizeof(int64_t) * (size_t)n));
    TW_DO(upload_plain(ctx, d_rc, retweet_cnt, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, d_fo, followers, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, d_fa, favourites, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, d_fr, friends, sizeof(int32_t) * (size_t)n));
    TW_DO(upload_plain(ctx, d_rt, is_retweet, (size_t)n));

    cudaMemsetAsync(d_maxlen, 0, sizeof(int), ctx->stream);
    k_feat_maxlen<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_off, n, d_maxlen);
    int maxlen = 0;
    cudaMemcpyAsync(&maxlen, d_maxlen, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { cleanup(); return fail(TWTML_E_CUDA, "tweet upload failed: %s", cudaGetErrorString(cudaGetLastError())); }
    if (maxlen > kMaxTerms + 1) { cleanup(); return fail(TWTML_E_INVALID_ARG, "a tweet has %d UTF-16 units; the featuriser takes at most %d", maxlen, kMaxTerms + 1); }

    const uint16_t* lower = nullptr;
    if (flags & TWTML_FLAG_LOWERCASE) TW_DO(lower_table(ctx, &lower));
    TweetCols tc{d_text, d_off, d_rc, d_fo, d_fa, d_fr, d_created, d_rt};
    cudaMemsetAsync(keep + n, 0, sizeof(long long), ctx->stream);
    cudaMemsetAsync(rownnz + n, 0, sizeof(long long), ctx->stream);
    const unsigned gridA = (unsigned)std::min<long long>((n + kFeatWarps - 1) / kFeatWarps, (long long)ctx->sm_count * 16);
    k_feat_terms<<<gridA, kFeatWarps * 32, 0, ctx->stream>>>(tc, n, rt_begin, rt_end, F, lower, pair_idx, pair_cnt, keep, rownnz);
    TW_DO(exclusive_sum(ctx, keep, krow, n + 1));
    TW_DO(exclusive_sum(ctx, rownnz, eoff, n + 1));
    long long tot[2] = {0, 0};
    cudaMemcpyAsync(&tot[0], krow + n, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream);
    cudaMemcpyAsync(&tot[1], eoff + n, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream);
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { cleanup(); return fail(TWTML_E_CUDA, "featuriser failed: %s", cudaGetErrorString(cudaGetLastError())); }
    const long long n_kept = tot[0], nnz = tot[1];
    TW_DO(new_csr(ctx, n_kept, (int64_t)F + 4, nnz, &b));
    const unsigned gridB = (unsigned)std::min<long long>((n + 3) / 4, (long long)ctx->sm_count * 16);
    k_feat_pack<<<gridB, 128, 0, ctx->stream>>>(tc, n, now_ms, F, pair_idx, pair_cnt, keep, krow, rownnz, eoff, b->rowptr, b->colidx, b->vals, b->labels);
    cudaMemcpyAsync(b->rowptr + n_kept, &tot[1], sizeof(long long), cudaMemcpyHostToDevice, ctx->stream);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
#undef TW_OWN
#undef TW_DO
    int rc = e == cudaSuccess ? 0 : fail(TWTML_E_CUDA, "featuriser pack failed: %s", cudaGetErrorString(e));
    if (!rc) rc = csr_finalize(ctx, b);
    if (rc) { batch_free_locked(b); return rc; }
    *out = b;
    return 0;
}

int32_t twtml_batch_export_csr(const twtml_batch* b, int64_t* rowptr, int32_t* colidx, double* vals, double* labels)
{
    if (!b || !rowptr) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (!b->csr) return fail(TWTML_E_INVALID_ARG, "twtml_batch_export_csr on a dense batch");
    twtml_ctx* ctx = b->ctx;
    Guard g(ctx);
    TW_CUDA(cudaMemcpyAsync(rowptr, b->rowptr, sizeof(int64_t) * (size_t)(b->n + 1), cudaMemcpyDeviceToHost, ctx->stream));
    if (b->nnz && colidx) TW_CUDA(cudaMemcpyAsync(colidx, b->colidx, sizeof(int32_t) * (size_t)b->nnz, cudaMemcpyDeviceToHost, ctx->stream));
    if (b->nnz && vals) TW_CUDA(cudaMemcpyAsync(vals, b->vals, sizeof(double) * (size_t)b->nnz, cudaMemcpyDeviceToHost, ctx->stream));
    if (b->n && labels) TW_CUDA(cudaMemcpyAsync(labels, b->labels, sizeof(double) * (size_t)b->n, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

} // extern "C"
