// twtml.cu -- the C ABI of include/twtml.h: contexts, models, dense micro-batches, and the
// predict / statistics / training entry points.  CSR batches and the tweet featuriser live in
// twtml_sparse.cu.  Host code here only validates, stages and enqueues: all arithmetic of the
// learning path runs in the kernels of dense.cuh / reduce.cuh / csr.cuh (no CPU fallback).
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <thread>
#include <unistd.h>

#include "dense.cuh"
#include "host.h"
#include "nccl_dyn.h"
#include "reduce.cuh"
#include "synth.cuh"

namespace twtml {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

const char* last_error_text() { return g_err; }

int DevBuf::reserve(size_t bytes)
{
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = (bytes + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        p = nullptr;
        return fail(TWTML_E_OOM, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e));
    }
    cap = want;
    return 0;
}

void DevBuf::release()
{
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}

#define TW_NCCL(expr)                                                                                     \
    do {                                                                                                  \
        ncclResult_t r__ = (expr);                                                                        \
        if (r__ != ncclSuccess) return fail(TWTML_E_NCCL, "%s: %s", #expr, nccl_api().GetErrorString(r__)); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// plans
// ---------------------------------------------------------------------------------------------
int dense_cfg(int64_t d, DenseCfg& c)
{
    const int64_t ldx = (d + 2) & ~(int64_t)1; // round_up(d + 1, 2)
    // column pairs that hold a feature; with d even the (label, pad) pair has no owner thread: only the
    // per-row threads read it, so narrow even widths (d = 128 -> 64 pairs) fill whole warps
    const int64_t c2 = (d + 1) / 2;
    if (c2 <= kIterThreads) c.cp = 1;
    else if (c2 <= 2 * kIterThreads) c.cp = 2;
    else if (c2 <= 4 * kIterThreads) c.cp = 4;
    else return fail(TWTML_E_INVALID_ARG, "dense batches support num_features <= %d; use the CSR form for wider rows", 8 * kIterThreads);
    c.c2 = (int)c2;
    if (c.cp == 1) {
        c.c2p = (c.c2 + 31) & ~31;
        c.rl = std::min(16, kIterThreads / c.c2p);
    } else {
        c.c2p = kIterThreads;
        c.rl = 1;
    }
    c.R = (4 / c.cp) * c.rl;
    c.stage_stride = (unsigned)(((size_t)c.R * ldx * 8 + 127) & ~(size_t)127);
    c.grid_per_sm = c.cp == 1 ? 2 : 1;
    const size_t budget = c.cp == 1 ? (228 * 1024 / 2 - 1024) : 227 * 1024;
    int s = (int)((budget - IterSmem::kStages) / c.stage_stride);
    c.nstages = std::min(8, s);
    if (c.nstages < 2) return fail(TWTML_E_INVALID_ARG, "row too wide for the dense pipeline");
    c.smem = IterSmem::kStages + (size_t)c.nstages * c.stage_stride;
    return 0;
}

void make_plan(int64_t n_local, int spr, int R, Plan& p, int64_t min_blocks)
{
    memset(&p, 0, sizeof(p));
    p.spr = spr;
    const int64_t nominal = (n_local + spr - 1) / spr;
    int64_t rb = 512;
    while (rb > 1 && rb * min_blocks > nominal) rb >>= 1; // a function of the row count only: keeps results grid-independent
    rb = std::max<int64_t>(1, rb / R) * R;
    p.rb = (int)rb;
    long long blk = 0;
    for (int s = 0; s <= spr; ++s) {
        p.shard_row0[s] = n_local * s / spr;
        if (s > 0) {
            const long long rows = p.shard_row0[s] - p.shard_row0[s - 1];
            blk += (rows + rb - 1) / rb;
        }
        p.shard_blk0[s] = blk;
    }
    for (int s = spr + 1; s <= kMaxShards; ++s) { p.shard_row0[s] = n_local; p.shard_blk0[s] = blk; }
    p.nblocks = blk;
}

// ---------------------------------------------------------------------------------------------
// iteration-kernel dispatch
// ---------------------------------------------------------------------------------------------
template <int CP, int MODE, bool RL1>
static int launch_iter_t(twtml_ctx* ctx, const DenseCfg& c, const DenseIterArgs& a)
{
    auto kern = k_dense_iter<CP, MODE, RL1>;
    constexpr int variant = (CP == 1 ? (RL1 ? 1 : 0) : (CP == 2 ? 2 : 3)) * 4 + MODE; // 16 kernel forms
    if (!(ctx->smem_attr_done & (1u << variant))) { // once per context (= per device) and kernel form
        TW_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        ctx->smem_attr_done |= 1u << variant;
    }
    long long grid = (long long)ctx->sm_count * c.grid_per_sm;
    if (grid > a.plan.nblocks) grid = a.plan.nblocks;
    if (grid < 1) return 0;
    kern<<<(unsigned)grid, kIterThreads, c.smem, ctx->stream>>>(a);
    TW_CUDA(cudaGetLastError());
    return 0;
}

template <int CP, bool RL1>
static int launch_iter_cp(twtml_ctx* ctx, const DenseCfg& c, const DenseIterArgs& a, int mode)
{
    switch (mode) {
    case MODE_TRAIN: return launch_iter_t<CP, MODE_TRAIN, RL1>(ctx, c, a);
    case MODE_TRAIN_STATS: return launch_iter_t<CP, MODE_TRAIN_STATS, RL1>(ctx, c, a);
    case MODE_PREDICT: return launch_iter_t<CP, MODE_PREDICT, RL1>(ctx, c, a);
    default: return launch_iter_t<CP, MODE_PREDICT_STATS, RL1>(ctx, c, a);
    }
}

static int launch_iter(twtml_ctx* ctx, const DenseCfg& c, const DenseIterArgs& a, int mode)
{
    if (c.cp == 1) return c.rl == 1 ? launch_iter_cp<1, true>(ctx, c, a, mode) : launch_iter_cp<1, false>(ctx, c, a, mode);
    if (c.cp == 2) return launch_iter_cp<2, true>(ctx, c, a, mode);
    return launch_iter_cp<4, true>(ctx, c, a, mode);
}

// sparse side (twtml_sparse.cu)
int csr_launch_iter(twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b, int mode, double* out_pred,
                    double* out_real, int iter_override);
int csr_partial_width(const twtml_batch* b);
int csr_reduce(twtml_ctx* ctx, const twtml_batch* b, double* shard_out, const TrainState* gate);

// ---------------------------------------------------------------------------------------------
// collectives
// ---------------------------------------------------------------------------------------------
static int allgather_inplace(twtml_ctx* ctx, void* buf, size_t count_per_rank, ncclDataType_t dt, size_t elem)
{
    if (ctx->world == 1) return 0;
    const char* send = static_cast<const char*>(buf) + (size_t)ctx->rank * count_per_rank * elem;
    TW_NCCL(nccl_api().AllGather(send, buf, count_per_rank, dt, ctx->comm, ctx->stream));
    ctx->prof.nccl_calls += 1;
    return 0;
}

// Map every rank's exchange buffer into this rank (collective; all ranks end with the same verdict).
struct PeerCard {
    cudaIpcMemHandle_t handle;
    long long pid;
    unsigned long long ptr;
    int dev;
    int ok;
};

// Collective: every rank offers `mine` (cudaMalloc'ed; null = "I could not allocate") and ends up with all ranks'
// buffers mapped in peer[] (CUDA IPC between processes, plain peer access between threads of one process), or with
// *all_ok = 0 on every rank when any rank failed.
static int map_peer_buffers(twtml_ctx* ctx, void* mine_ptr, double* peer[kMaxShards], bool ipc[kMaxShards], int* all_ok)
{
    const int W = ctx->world;
    PeerCard mine{};
    mine.pid = (long long)getpid();
    mine.dev = ctx->device;
    mine.ok = mine_ptr != nullptr;
    if (mine_ptr && cudaIpcGetMemHandle(&mine.handle, mine_ptr) != cudaSuccess) {
        cudaGetLastError();
        mine.ok = 0;
    }
    mine.ptr = (unsigned long long)mine_ptr;
    TW_TRY(ctx->gather.reserve(sizeof(PeerCard) * (size_t)W));
    char* g = ctx->gather.as<char>();
    TW_CUDA(cudaMemcpyAsync(g + sizeof(PeerCard) * ctx->rank, &mine, sizeof(mine), cudaMemcpyHostToDevice, ctx->stream));
    TW_TRY(allgather_inplace(ctx, g, sizeof(PeerCard), ncclInt8, 1));
    std::vector<PeerCard> cards((size_t)W);
    TW_CUDA(cudaMemcpyAsync(cards.data(), g, sizeof(PeerCard) * (size_t)W, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    int ok = 1;
    for (int r = 0; r < W; ++r) ok &= cards[r].ok;
    for (int r = 0; r < W && ok; ++r) {
        if (r == ctx->rank) { peer[r] = static_cast<double*>(mine_ptr); continue; }
        if (cards[r].pid == mine.pid) { // another thread of this process: plain peer access
            int can = 0;
            cudaDeviceCanAccessPeer(&can, ctx->device, cards[r].dev);
            cudaError_t e = can ? cudaDeviceEnablePeerAccess(cards[r].dev, 0) : cudaErrorInvalidDevice;
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            if (e != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
            peer[r] = reinterpret_cast<double*>(cards[r].ptr);
        } else {
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, cards[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
            peer[r] = static_cast<double*>(p);
            ipc[r] = true;
        }
    }
    // second round: a rank that could not map a peer vetoes the scheme for everybody
    int* flags = reinterpret_cast<int*>(g);
    TW_CUDA(cudaMemcpyAsync(flags + ctx->rank, &ok, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    TW_TRY(allgather_inplace(ctx, flags, 1, ncclInt32, sizeof(int)));
    std::vector<int> oks((size_t)W);
    TW_CUDA(cudaMemcpyAsync(oks.data(), flags, sizeof(int) * (size_t)W, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < W; ++r) ok &= oks[r];
    *all_ok = ok;
    return 0;
}

static int setup_peer_exchange(twtml_ctx* ctx)
{
    ctx->p2p_ok = false;
    const int W = ctx->world;
    if (W < 2 || W > kMaxShards) return 0;
    if (cudaMalloc(&ctx->xbuf, kXbufBytes) != cudaSuccess || cudaMemset(ctx->xbuf, 0, kXbufBytes) != cudaSuccess) {
        cudaGetLastError();
        if (ctx->xbuf) cudaFree(ctx->xbuf);
        ctx->xbuf = nullptr;
    }
    int ok = 0;
    TW_TRY(map_peer_buffers(ctx, ctx->xbuf, ctx->peer, ctx->peer_ipc, &ok));
    ctx->peers_mappable = ok != 0;
    ctx->p2p_ok = ok != 0 && kMaxShards % W == 0 && getenv("TWTML_NO_P2P") == nullptr;
    return 0;
}

// The wide exchange buffer (reduce.cuh WideExchange) for models of up to d columns.  Collective: called by every rank at
// the same point of a training call (the decision depends on the model width and on what earlier calls allocated, both
// equal on all ranks).
static int setup_wide_exchange(twtml_ctx* ctx, long long d)
{
    const long long want = (d + 2 + 15) & ~15ll;
    if (ctx->wx_dpad >= want) return 0;
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < kMaxShards; ++r) {
        if (ctx->wx_ipc[r] && ctx->wx_peer[r]) cudaIpcCloseMemHandle(ctx->wx_peer[r]);
        ctx->wx_peer[r] = nullptr;
        ctx->wx_ipc[r] = false;
    }
    // nobody frees a buffer a peer may still be reading: the last exchange on the old one ended before this point on every
    // rank, and the handle exchange below is a barrier before anyone touches the new one
    double* old = ctx->wxbuf;
    ctx->wxbuf = nullptr;
    ctx->wx_dpad = 0;
    const size_t bytes = wide_exchange_bytes(want);
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess || cudaMemset(p, 0, bytes) != cudaSuccess) {
        cudaGetLastError();
        if (p) cudaFree(p);
        p = nullptr;
    }
    int ok = 0;
    TW_TRY(map_peer_buffers(ctx, p, ctx->wx_peer, ctx->wx_ipc, &ok));
    if (old) cudaFree(old);
    if (!ok) {
        for (int r = 0; r < kMaxShards; ++r) {
            if (ctx->wx_ipc[r] && ctx->wx_peer[r]) cudaIpcCloseMemHandle(ctx->wx_peer[r]);
            ctx->wx_peer[r] = nullptr;
            ctx->wx_ipc[r] = false;
        }
        if (p) cudaFree(p);
        ctx->wx_failed = true; // every rank saw the same verdict: all fall back to the NCCL form from here on
        return 0;
    }
    ctx->wxbuf = static_cast<double*>(p);
    ctx->wx_dpad = want;
    return 0;
}

static void teardown_peer_exchange(twtml_ctx* ctx)
{
    for (int r = 0; r < kMaxShards; ++r)
        if (ctx->peer_ipc[r] && ctx->peer[r]) cudaIpcCloseMemHandle(ctx->peer[r]);
    if (ctx->xbuf) cudaFree(ctx->xbuf);
    ctx->xbuf = nullptr;
    for (int r = 0; r < kMaxShards; ++r)
        if (ctx->wx_ipc[r] && ctx->wx_peer[r]) cudaIpcCloseMemHandle(ctx->wx_peer[r]);
    if (ctx->wxbuf) cudaFree(ctx->wxbuf);
    ctx->wxbuf = nullptr;
}

int ctx_global_rows(twtml_ctx* ctx, int64_t n_local, int64_t* n_global, int64_t* row0)
{
    if (ctx->world == 1) { *n_global = n_local; *row0 = 0; return 0; }
    TW_TRY(ctx->gather.reserve(sizeof(long long) * ctx->world));
    long long* g = ctx->gather.as<long long>();
    long long v = n_local;
    TW_CUDA(cudaMemcpyAsync(g + ctx->rank, &v, sizeof(v), cudaMemcpyHostToDevice, ctx->stream));
    TW_TRY(allgather_inplace(ctx, g, 1, ncclInt64, sizeof(long long)));
    std::vector<long long> h(ctx->world);
    TW_CUDA(cudaMemcpyAsync(h.data(), g, sizeof(long long) * ctx->world, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    long long tot = 0, before = 0;
    for (int r = 0; r < ctx->world; ++r) { if (r < ctx->rank) before += h[r]; tot += h[r]; }
    *n_global = tot;
    *row0 = before;
    return 0;
}

int pool_alloc(twtml_ctx* ctx, void** p, size_t bytes, cudaStream_t stream)
{
    *p = nullptr;
    if (bytes == 0) bytes = 8;
    if (!stream) stream = ctx->stream;
    cudaError_t e = cudaMallocFromPoolAsync(p, bytes, ctx->pool, stream);
    if (e != cudaSuccess) { // give cached blocks back and try once more
        cudaGetLastError();
        cudaStreamSynchronize(stream);
        cudaMemPoolTrimTo(ctx->pool, 0);
        { std::lock_guard<std::mutex> lk(ctx->block_mu); for (auto& kv : ctx->block_free) cudaFree(kv.second); ctx->block_free.clear(); }
        e = cudaMallocFromPoolAsync(p, bytes, ctx->pool, stream);
    }
    if (e != cudaSuccess) {
        *p = nullptr;
        cudaGetLastError();
        return fail(TWTML_E_OOM, "device allocation of %zu bytes: %s", bytes, cudaGetErrorString(e));
    }
    return 0;
}

void pool_free(twtml_ctx* ctx, void* p, cudaStream_t stream)
{
    if (p) cudaFreeAsync(p, stream ? stream : ctx->stream);
}

// Sizes are rounded up to an eighth of their leading power of two (at least 2 MiB), so that the intervals of a stream,
// which differ a little in size, find each other's blocks; a block up to twice the wanted size is accepted.
static size_t block_size_class(size_t bytes)
{
    size_t lead = (size_t)1 << 21;
    while (lead * 2 <= bytes) lead *= 2;
    const size_t granule = std::max<size_t>((size_t)1 << 21, lead / 8);
    return (std::max<size_t>(bytes, 1) + granule - 1) / granule * granule;
}

static void block_trim(twtml_ctx* ctx) // caller holds block_mu
{
    for (auto& kv : ctx->block_free) cudaFree(kv.second);
    ctx->block_free.clear();
}

int block_alloc(twtml_ctx* ctx, void** p, size_t bytes)
{
    *p = nullptr;
    const size_t want = block_size_class(bytes);
    std::lock_guard<std::mutex> lk(ctx->block_mu);
    auto it = ctx->block_free.lower_bound(want);
    if (it != ctx->block_free.end() && it->first <= 2 * want) {
        *p = it->second;
        ctx->block_live[*p] = it->first;
        ctx->block_free.erase(it);
        return 0;
    }
    cudaSetDevice(ctx->device);
    cudaError_t e = cudaMalloc(p, want);
    if (e != cudaSuccess) { // give the idle blocks and the pool's cached memory back and try once more
        cudaGetLastError();
        block_trim(ctx);
        cudaMemPoolTrimTo(ctx->pool, 0);
        e = cudaMalloc(p, want);
    }
    if (e != cudaSuccess) {
        *p = nullptr;
        cudaGetLastError();
        return fail(TWTML_E_OOM, "device allocation of %zu bytes: %s", want, cudaGetErrorString(e));
    }
    ctx->block_live[*p] = want;
    return 0;
}

void block_release(twtml_ctx* ctx, void* p)
{
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(ctx->block_mu);
        auto it = ctx->block_live.find(p);
        if (it != ctx->block_live.end()) {
            ctx->block_free.emplace(it->second, p);
            ctx->block_live.erase(it);
            return;
        }
    }
    pool_free(ctx, p); // not one of ours: stream-ordered pool memory (dense batches)
}

static int ensure_pinned(IngestLane& lane, size_t bytes)
{
    if (lane.pin_bytes >= bytes) return 0;
    for (int i = 0; i < 2; ++i) {
        if (lane.pin[i]) { cudaEventSynchronize(lane.pin_ev[i]); cudaFreeHost(lane.pin[i]); }
        lane.pin[i] = nullptr;
    }
    lane.pin_bytes = 0;
    for (int i = 0; i < 2; ++i) TW_CUDA(cudaMallocHost(&lane.pin[i], bytes));
    lane.pin_bytes = bytes;
    return 0;
}

// Host -> device in chunks through the two pinned buffers: chunk k+1 is packed by `pack` while chunk
// k is in flight.  pack(dst, first_unit, n_units) fills dst with n_units * unit_bytes bytes.
template <class PackFn>
int staged_upload(IngestLane& lane, void* dst, size_t unit_bytes, int64_t n_units, PackFn pack)
{
    if (n_units <= 0) return 0;
    const size_t chunk_bytes = (size_t)32 << 20;
    TW_TRY(ensure_pinned(lane, std::max(chunk_bytes, unit_bytes)));
    const int64_t per = std::max<int64_t>(1, (int64_t)(lane.pin_bytes / unit_bytes));
    int k = 0;
    for (int64_t u0 = 0; u0 < n_units; u0 += per, ++k) {
        const int64_t nu = std::min(per, n_units - u0);
        const int bi = k & 1;
        TW_CUDA(cudaEventSynchronize(lane.pin_ev[bi])); // the last copy out of this buffer (maybe an earlier call's) is done
        pack(static_cast<char*>(lane.pin[bi]), u0, nu);
        TW_CUDA(cudaMemcpyAsync(static_cast<char*>(dst) + (size_t)u0 * unit_bytes, lane.pin[bi], (size_t)nu * unit_bytes,
                                cudaMemcpyHostToDevice, lane.stream));
        TW_CUDA(cudaEventRecord(lane.pin_ev[bi], lane.stream));
    }
    return 0;
}
static void parallel_rows(int64_t n, int64_t min_per_thread, const std::function<void(int64_t, int64_t)>& fn);
static bool page_locked(const void* p);

// Host -> device copy of a plain array: straight DMA when the caller's memory is page-locked, else through
// the pinned ring with the staging memcpy spread over host threads.
int upload_plain(twtml_ctx* ctx, IngestLane& lane, void* dst, const void* src, size_t bytes)
{
    (void)ctx;
    if (bytes == 0) return 0;
    if (page_locked(src)) {
        TW_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, lane.stream));
        return 0;
    }
    const char* s = static_cast<const char*>(src);
    return staged_upload(lane, dst, 1, (int64_t)bytes, [s](char* out, int64_t u0, int64_t nu) {
        parallel_rows(nu, 4 << 20, [&](int64_t lo, int64_t hi) { memcpy(out + lo, s + u0 + lo, (size_t)(hi - lo)); });
    });
}

static bool page_locked(const void* p)
{
    cudaPointerAttributes pa{};
    const bool yes = cudaPointerGetAttributes(&pa, p) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return yes;
}

static void parallel_rows(int64_t n, int64_t min_per_thread, const std::function<void(int64_t, int64_t)>& fn)
{
    int hw = (int)std::thread::hardware_concurrency();
    int nt = (int)std::min<int64_t>(std::max(1, std::min(hw, 16)), std::max<int64_t>(1, n / min_per_thread));
    if (nt <= 1) { fn(0, n); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < nt; ++i) th.emplace_back(fn, n * i / nt, n * (i + 1) / nt);
    for (auto& t : th) t.join();
}

__global__ void k_scatter_labels(double* __restrict__ X, long long ldx, int d, const double* __restrict__ labels, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    X[i * ldx + d] = labels[i];
    for (long long j = d + 1; j < ldx; ++j) X[i * ldx + j] = 0.0;
}

// rows of d doubles, back to back  ->  rows of ldx doubles [x_0 .. x_{d-1}, label, 0 pad]
__global__ void k_repitch(const double* __restrict__ flat, double* __restrict__ X, long long ldx, int d, const double* __restrict__ labels,
                          long long rows)
{
    const long long total = rows * ldx;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / ldx;
        const int j = (int)(i - r * ldx);
        X[i] = j < d ? flat[r * d + j] : (j == d ? labels[r] : 0.0);
    }
}

// Page-locked, densely packed host rows -> the batch's pitched rows.  A pitched (2-D) copy of 8 KB rows reaches ~45 GB/s
// on this link, a flat one ~55 GB/s: so the rows land flat in a small device buffer (two halves, alternating) and a kernel
// spreads each chunk to the row pitch and drops the labels in.  The copies are ordered on the lane's stream; the kernels run
// on the context's spread stream, beside the next chunk's copy (events per half: copied -> spread -> may be overwritten); the
// lane's stream ends up waiting for the last kernel, so whatever is ordered after this call on it sees all the rows.
// Rows [row_begin, row_begin + nrows) of batch b; X and labels_dev point at the first of them.
static const size_t kStageChunkBytes = (size_t)64 << 20;
static cudaError_t flat_dense_ingest(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b, int64_t row_begin, int64_t nrows, const double* X,
                                     const double* labels_dev)
{
    const int64_t n = nrows, d = b->d;
    cudaStream_t stream = lane.stream;
    const int64_t rows_per = std::max<int64_t>(1, (int64_t)(kStageChunkBytes / (sizeof(double) * (size_t)d)));
    const size_t half = sizeof(double) * (size_t)rows_per * (size_t)d;
    const size_t need = n <= rows_per ? sizeof(double) * (size_t)n * (size_t)d : 2 * half; // a small batch uses one half only
    if (lane.stage.reserve(need)) return cudaErrorMemoryAllocation; // (growing it goes through cudaFree, which waits for the device)
    cudaStream_t spread = ctx->spread_stream;
    cudaError_t e;
    // the labels (copied on `stream` before this call) and, for a zone used before, its last kernels: order the spread stream
    // behind `stream` once; every kernel below additionally waits for its own chunk's copy
    if ((e = cudaEventRecord(lane.stage_ev[0][0], stream)) != cudaSuccess) return e;
    if ((e = cudaStreamWaitEvent(spread, lane.stage_ev[0][0], 0)) != cudaSuccess) return e;
    int k = 0;
    int64_t chunk = 0;
    for (int64_t r0 = 0; r0 < n; r0 += rows_per, k ^= 1, ++chunk) {
        const int64_t nr = std::min(rows_per, n - r0);
        double* st = reinterpret_cast<double*>(static_cast<char*>(lane.stage.p) + (size_t)k * half);
        // this half was last read by the kernel of chunk - 2
        if (chunk >= 2 && (e = cudaStreamWaitEvent(stream, lane.stage_ev[k][1], 0)) != cudaSuccess) return e;
        if ((e = cudaMemcpyAsync(st, X + r0 * d, sizeof(double) * (size_t)nr * (size_t)d, cudaMemcpyHostToDevice, stream)) != cudaSuccess) return e;
        if ((e = cudaEventRecord(lane.stage_ev[k][0], stream)) != cudaSuccess) return e;
        if ((e = cudaStreamWaitEvent(spread, lane.stage_ev[k][0], 0)) != cudaSuccess) return e;
        const long long total = nr * b->ldx;
        const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, (long long)ctx->sm_count * 16);
        k_repitch<<<grid, 256, 0, spread>>>(st, b->X + (row_begin + r0) * b->ldx, b->ldx, (int)d, labels_dev + r0, nr);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if ((e = cudaEventRecord(lane.stage_ev[k][1], spread)) != cudaSuccess) return e;
    }
    // both halves' last kernels before anything later on `stream` (the next rows' copies into this zone included)
    for (int h = 0; h < 2; ++h)
        if (chunk > h && (e = cudaStreamWaitEvent(stream, lane.stage_ev[h][1], 0)) != cudaSuccess) return e;
    return cudaSuccess;
}

__global__ void k_gather_dense(const double* __restrict__ X, long long ldx, int d, double* __restrict__ out,
                               double* __restrict__ labels, long long n)
{
    for (long long i = blockIdx.x; i < n; i += gridDim.x) {
        for (int j = threadIdx.x; j < d; j += blockDim.x) out[i * d + j] = X[i * ldx + j];
        if (threadIdx.x == 0) labels[i] = X[i * ldx + d];
    }
}

static void stats_from_out10(const double* o, twtml_stats* s)
{
    s->batch = (int64_t)o[0];
    s->mse = o[1];
    s->real_stdev = o[2];
    s->pred_stdev = o[3];
    s->mse_raw = o[4];
    s->real_stdev_raw = o[5];
    s->pred_stdev_raw = o[6];
    s->real_mean = o[7];
    s->pred_mean = o[8];
    s->nonfinite = (int64_t)o[9];
}

static void stats_empty(twtml_stats* s)
{
    memset(s, 0, sizeof(*s));
    s->mse = s->real_stdev = s->pred_stdev = s->mse_raw = s->real_stdev_raw = s->pred_stdev_raw = s->real_mean = s->pred_mean = NAN;
}

static cudaEvent_t prof_event(twtml_ctx* ctx, size_t i)
{
    while (ctx->ev.size() <= i) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        ctx->ev.push_back(e);
    }
    return ctx->ev[i];
}

// One pass of the iteration kernel in `mode` over batch b (dense or CSR).
static int run_pass(twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b, int mode, double* out_pred, double* out_real,
                    int iter_override)
{
    if (b->csr) return csr_launch_iter(ctx, m, b, mode, out_pred, out_real, iter_override);
    DenseCfg c;
    TW_TRY(dense_cfg(b->d, c));
    DenseIterArgs a{};
    a.X = b->X;
    a.w = m->d_w;
    a.state = ctx->state.as<TrainState>();
    a.partials = ctx->partials.as<double>();
    a.stat_partials = ctx->stat_partials.as<double>();
    a.out_pred = out_pred;
    a.out_real = out_real;
    a.plan = b->plan;
    a.ldx = b->ldx;
    a.d = (int)b->d;
    a.c2 = c.c2;
    a.c2p = c.c2p;
    a.rl = c.rl;
    a.nstages = c.nstages;
    a.stage_stride = c.stage_stride;
    a.frac = m->frac;
    a.row0_global = (unsigned long long)b->row0_global;
    a.iter_override = iter_override;
    return launch_iter(ctx, c, a, mode);
}

static int partial_width(const twtml_batch* b) { return b->csr ? csr_partial_width(b) : (int)b->ldx + 2; }
static int loss_slot(const twtml_batch* b) { return b->csr ? (int)b->d : (int)b->ldx; }

static bool use_sliced(const twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b);

static int reserve_scratch(twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b)
{
    const int pv = partial_width(b);
    TW_TRY(ctx->partials.reserve(sizeof(double) * (size_t)std::max<long long>(1, b->plan.nblocks) * (b->csr && !b->narrow ? 2 : pv)));
    TW_TRY(ctx->shards.reserve(sizeof(double) * (size_t)kMaxShards * std::max(1, ctx->world) * pv));
    TW_TRY(ctx->stat_partials.reserve(sizeof(double) * (size_t)std::max<long long>(1, b->plan.nblocks) * kStatStride));
    TW_TRY(ctx->stat_shards.reserve(sizeof(double) * (size_t)kMaxShards * std::max(1, ctx->world) * kStatStride));
    TW_TRY(ctx->state.reserve(sizeof(TrainState) + 64)); // + the CTA counter of k_shard_reduce_update
    TW_TRY(ctx->loss_hist.reserve(sizeof(double) * (size_t)std::max(1, m->num_iter)));
    TW_TRY(ctx->out10.reserve(sizeof(double) * 16));
    TW_TRY(ctx->upd_parts.reserve(sizeof(double) * 3 * (size_t)((m->d + kWideCols - 1) / kWideCols)));
    if (use_sliced(ctx, m, b)) {
        const size_t slice_len = (size_t)((m->d + ctx->world - 1) / ctx->world);
        TW_TRY(ctx->slices.reserve(sizeof(double) * (2 * (size_t)ctx->world * slice_len + 2 * (size_t)ctx->world + 8)));
        TW_TRY(ctx->wprev.reserve(sizeof(double) * (size_t)m->d));
    }
    return 0;
}

// block partials -> shard vectors -> (all-gather) ; leaves every rank with all shard vectors
// wide CSR models on several ranks: all-to-all of gradient slices + sliced update + all-gather of w
static bool use_sliced(const twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b)
{
    return ctx->world > 1 && b->csr && m->d > 2 * kWideCols && getenv("TWTML_NO_SLICED") == nullptr;
}

static bool use_sliced_p2p(const twtml_ctx* ctx, const twtml_model* m, const twtml_batch* b)
{
    static const bool off = getenv("TWTML_NO_P2P") != nullptr || getenv("TWTML_NO_P2P_WIDE") != nullptr;
    return use_sliced(ctx, m, b) && ctx->peers_mappable && !ctx->wx_failed && !off;
}

static WideExchange wide_exchange(const twtml_ctx* ctx)
{
    WideExchange wx{};
    for (int r = 0; r < ctx->world; ++r) wx.peer[r] = ctx->wx_peer[r];
    wx.dpad = ctx->wx_dpad;
    wx.world = ctx->world;
    wx.rank = ctx->rank;
    return wx;
}

// the peer-memory form of sliced_update (reduce.cuh): this rank's gradient already sits in its exchange buffer
static int sliced_update_p2p(twtml_ctx* ctx, twtml_model* m, const twtml_batch* b, const UpdateArgs& ua)
{
    (void)b;
    const int W = ctx->world, d = (int)m->d, slice_len = (d + W - 1) / W;
    const WideExchange wx = wide_exchange(ctx);
    double* pairs = ctx->slices.as<double>();
    SliceArgs sa{};
    sa.w = m->d_w;
    sa.state = ua.state;
    sa.W = W;
    sa.slice_len = slice_len;
    sa.lo = std::min(d, ctx->rank * slice_len);
    sa.hi = std::min(d, sa.lo + slice_len);
    sa.hi_all = d;
    sa.updater = ua.updater;
    sa.step = ua.step;
    sa.reg = ua.reg;
    const int nb = std::max(1, (sa.hi - sa.lo + kWideCols - 1) / kWideCols);
    const int nparts = (d + kWideCols - 1) / kWideCols;
    k_wide_signal<<<1, 32, 0, ctx->stream>>>(wx, ua.state);
    k_update_slice_p2p<<<nb, 1024, 0, ctx->stream>>>(sa, wx, ua.state, pairs);
    k_norms_wide_p2p<<<nparts, 1024, 0, ctx->stream>>>(m->d_w, wx, d, ua.state, ctx->upd_parts.as<double>());
    k_slice_finish<<<1, 1024, 0, ctx->stream>>>(ua, pairs, W, ctx->upd_parts.as<double>(), nparts, wx_seq(ctx->wxbuf, ctx->wx_dpad));
    TW_CUDA(cudaGetLastError());
    ctx->prof.other_launches += 4;
    return 0;
}

static int sliced_update(twtml_ctx* ctx, twtml_model* m, const twtml_batch* b, const UpdateArgs& ua)
{
    const int W = ctx->world, d = (int)m->d, slice_len = (d + W - 1) / W;
    double* mine = ctx->shards.as<double>() + (size_t)ctx->rank * partial_width(b); // [g_0..g_{d-1}, lossSum, count] of this rank
    double* slices = ctx->slices.as<double>();
    double* pairs = slices + (size_t)W * slice_len;
    // slices travel from a staging copy padded to W * slice_len (the tail beyond d carries nothing)
    double* send = pairs + 2 * W + 2;
    TW_CUDA(cudaMemcpyAsync(send, mine, sizeof(double) * (size_t)d, cudaMemcpyDeviceToDevice, ctx->stream));
    TW_NCCL(nccl_api().GroupStart());
    for (int p = 0; p < W; ++p) {
        TW_NCCL(nccl_api().Send(send + (size_t)p * slice_len, (size_t)slice_len, ncclDouble, p, ctx->comm, ctx->stream));
        TW_NCCL(nccl_api().Recv(slices + (size_t)p * slice_len, (size_t)slice_len, ncclDouble, p, ctx->comm, ctx->stream));
        TW_NCCL(nccl_api().Send(mine + d, 2, ncclDouble, p, ctx->comm, ctx->stream));
        TW_NCCL(nccl_api().Recv(pairs + 2 * p, 2, ncclDouble, p, ctx->comm, ctx->stream));
    }
    TW_NCCL(nccl_api().GroupEnd());
    TW_CUDA(cudaMemcpyAsync(ctx->wprev.p, m->d_w, sizeof(double) * (size_t)d, cudaMemcpyDeviceToDevice, ctx->stream));
    SliceArgs sa{};
    sa.slices = slices;
    sa.pairs = pairs;
    sa.w = m->d_w;
    sa.state = ua.state;
    sa.W = W;
    sa.slice_len = slice_len;
    sa.lo = std::min(d, ctx->rank * slice_len);
    sa.hi = std::min(d, sa.lo + slice_len);
    sa.hi_all = d;
    sa.updater = ua.updater;
    sa.step = ua.step;
    sa.reg = ua.reg;
    const int nb = std::max(1, (sa.hi - sa.lo + kWideCols - 1) / kWideCols);
    k_update_slice<<<nb, 1024, 0, ctx->stream>>>(sa);
    TW_CUDA(cudaGetLastError());
    TW_NCCL(nccl_api().AllGather(m->d_w + (size_t)ctx->rank * slice_len, m->d_w, (size_t)slice_len, ncclDouble, ctx->comm, ctx->stream));
    const int nparts = (d + kWideCols - 1) / kWideCols;
    k_norms_wide<<<nparts, 1024, 0, ctx->stream>>>(m->d_w, ctx->wprev.as<double>(), d, ua.state, ctx->upd_parts.as<double>());
    k_slice_finish<<<1, 1024, 0, ctx->stream>>>(ua, pairs, W, ctx->upd_parts.as<double>(), nparts);
    TW_CUDA(cudaGetLastError());
    ctx->prof.other_launches += 3;
    ctx->prof.nccl_calls += 2;
    return 0;
}

static bool use_p2p(const twtml_ctx* ctx, const twtml_batch* b)
{
    static const bool csr_nccl = getenv("TWTML_NO_P2P_NARROW") != nullptr; // A/B switch: narrow CSR over NCCL as in round 1
    if (b->csr && (!b->narrow || csr_nccl)) return false;
    return ctx->p2p_ok && partial_width(b) <= kPvCap && b->plan.spr * ctx->world <= kMaxShards;
}

static int exchange_shards(twtml_ctx* ctx, const twtml_batch* b)
{
    if (use_p2p(ctx, b)) return 0; // k_shard_reduce_p2p already stored the vectors on every rank
    return allgather_inplace(ctx, ctx->shards.as<double>(), (size_t)b->plan.spr * partial_width(b), ncclDouble, sizeof(double));
}

static int reduce_local(twtml_ctx* ctx, const twtml_batch* b, bool gated, bool wide_p2p = false)
{
    const int pv = partial_width(b), spr = b->plan.spr;
    const TrainState* gate = gated ? ctx->state.as<TrainState>() : nullptr;
    double* mine = ctx->shards.as<double>() + (size_t)ctx->rank * spr * pv;
    const bool p2p = gated && use_p2p(ctx, b); // dense, or the 1004-column CSR model whose CTAs leave the same kind of vectors
    if (b->csr && !p2p) {
        if (wide_p2p) mine = wx_G(ctx->wxbuf); // the peers read the gradient where the column pass leaves it
        TW_TRY(csr_reduce(ctx, b, mine, gate));
    } else if (p2p) {
        PeerExchange px{};
        for (int r = 0; r < ctx->world; ++r) px.peer[r] = ctx->peer[r];
        px.world = ctx->world;
        px.rank = ctx->rank;
        dim3 grid((pv + 31) / 32, spr), block(32, 32);
        k_shard_reduce_p2p<<<grid, block, 0, ctx->stream>>>(ctx->partials.as<double>(), pv, b->plan, px, gate);
        TW_CUDA(cudaGetLastError());
        ctx->prof.other_launches += 1;
    } else {
        dim3 grid((pv + 31) / 32, spr), block(32, 32);
        k_shard_reduce<<<grid, block, 0, ctx->stream>>>(ctx->partials.as<double>(), pv, b->plan, mine, gate);
        TW_CUDA(cudaGetLastError());
        ctx->prof.other_launches += 1;
    }
    return 0;
}

static int reduce_stats(twtml_ctx* ctx, const twtml_batch* b)
{
    const int spr = b->plan.spr;
    double* mine = ctx->stat_shards.as<double>() + (size_t)ctx->rank * spr * kStatStride;
    k_stats_shard<<<spr, 256, 0, ctx->stream>>>(ctx->stat_partials.as<double>(), b->plan, mine);
    TW_CUDA(cudaGetLastError());
    TW_TRY(allgather_inplace(ctx, ctx->stat_shards.as<double>(), (size_t)spr * kStatStride, ncclDouble, sizeof(double)));
    k_stats_final<<<1, 32, 0, ctx->stream>>>(ctx->stat_shards.as<double>(), spr * ctx->world, ctx->out10.as<double>());
    TW_CUDA(cudaGetLastError());
    ctx->prof.other_launches += 2;
    return 0;
}

// ---- asynchronous ingest: the worker thread ---------------------------------------------------------
static void worker_main(twtml_ctx* ctx)
{
    cudaSetDevice(ctx->device);
    for (;;) {
        std::function<void()> job;
        {
            std::unique_lock<std::mutex> lk(ctx->qmu);
            ctx->qcv.wait(lk, [&] { return ctx->worker_stop || !ctx->jobs.empty(); });
            if (ctx->jobs.empty()) return; // stop requested and nothing left to do
            job = std::move(ctx->jobs.front());
            ctx->jobs.pop_front();
        }
        job();
    }
}

// Queue `job` (the ingest of batch b on lane 1) behind the context's earlier asynchronous ingests.  Its status
// and message end up in b->pending / b->pending_err; the first call that uses the batch collects them.
int submit_ingest(twtml_ctx* ctx, twtml_batch* b, std::function<int()> job)
{
    auto task = std::make_shared<std::packaged_task<int()>>([b, job]() {
        const int rc = job();
        if (rc) b->pending_err = last_error_text();
        return rc;
    });
    b->pending = task->get_future().share();
    b->has_pending = true;
    {
        std::lock_guard<std::mutex> lk(ctx->qmu);
        if (!ctx->worker.joinable()) ctx->worker = std::thread(worker_main, ctx);
        ctx->jobs.emplace_back([task]() { (*task)(); });
    }
    ctx->qcv.notify_one();
    return 0;
}

static void stop_worker(twtml_ctx* ctx)
{
    {
        std::lock_guard<std::mutex> lk(ctx->qmu);
        ctx->worker_stop = true;
    }
    ctx->qcv.notify_all();
    if (ctx->worker.joinable()) ctx->worker.join();
}

// Host side of "is this batch usable": the worker's verdict for an asynchronously ingested batch (blocks until it is
// in) and the rows of a batch that is being appended to.  Local: any rank may call it alone (export, ready, predict).
int settle_batch(twtml_ctx* ctx, const twtml_batch* cb)
{
    (void)ctx;
    twtml_batch* b = const_cast<twtml_batch*>(cb);
    if (b->has_pending) {
        const int rc = b->pending.get();
        if (rc) return fail(rc, "asynchronous ingest failed: %s", b->pending_err.c_str());
    }
    if (!b->csr && b->rows_appended < b->n)
        return fail(TWTML_E_INVALID_ARG, "batch used before all of its rows were appended (%lld of %lld)", (long long)b->rows_appended, (long long)b->n);
    return 0;
}

// On a multi-rank context, the row-count exchange an asynchronous ingest left for later.  A collective on the compute
// stream: only the collective compute calls (statistics, training) come here, where every rank arrives with the same
// batch in the same order.
static int resolve_global_rows(twtml_ctx* ctx, const twtml_batch* cb)
{
    twtml_batch* b = const_cast<twtml_batch*>(cb);
    if (b->global_deferred) {
        TW_TRY(ctx_global_rows(ctx, b->n, &b->n_global, &b->row0_global));
        b->global_deferred = false;
    }
    return 0;
}

// ... and the device side: a batch ingested asynchronously may still be in flight on the copy stream
static int wait_ready(twtml_ctx* ctx, const twtml_batch* b)
{
    TW_TRY(settle_batch(ctx, b));
    if (b->ready) TW_CUDA(cudaStreamWaitEvent(ctx->stream, b->ready, 0));
    return 0;
}

static int check_pair(const twtml_model* m, const twtml_batch* b, const char* what)
{
    if (!m || !b) return fail(TWTML_E_INVALID_ARG, "%s: null handle", what);
    if (m->ctx != b->ctx) return fail(TWTML_E_INVALID_ARG, "%s: model and batch belong to different contexts", what);
    if (!m->initialized)
        return fail(TWTML_E_NOT_INITIALIZED, "Model must be initialized before starting %s.", what);
    if (m->d != b->d)
        return fail(TWTML_E_SHAPE_MISMATCH, "%s: model has %lld features, batch has %lld", what, (long long)m->d, (long long)b->d);
    return 0;
}

static int fetch_rows(twtml_ctx* ctx, double* host, const double* dev, int64_t n)
{
    if (host && n > 0) TW_CUDA(cudaMemcpyAsync(host, dev, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}

static int do_predict_stats(const twtml_model* m, const twtml_batch* b, double* out_real, double* out_pred, twtml_stats* out)
{
    twtml_ctx* ctx = m->ctx;
    TW_TRY(wait_ready(ctx, b));
    TW_TRY(resolve_global_rows(ctx, b));
    if (b->n_global == 0) { stats_empty(out); return 0; }
    TW_TRY(reserve_scratch(ctx, m, b));
    double *dp = nullptr, *dr = nullptr;
    if (out_pred) { TW_TRY(ctx->tmp_a.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, b->n))); dp = ctx->tmp_a.as<double>(); }
    if (out_real) { TW_TRY(ctx->tmp_b.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, b->n))); dr = ctx->tmp_b.as<double>(); }
    TW_TRY(run_pass(ctx, m, b, MODE_PREDICT_STATS, dp, dr, 0));
    TW_TRY(reduce_stats(ctx, b));
    double o[10];
    TW_CUDA(cudaMemcpyAsync(o, ctx->out10.p, sizeof(o), cudaMemcpyDeviceToHost, ctx->stream));
    TW_TRY(fetch_rows(ctx, out_pred, dp, b->n));
    TW_TRY(fetch_rows(ctx, out_real, dr, b->n));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    stats_from_out10(o, out);
    return 0;
}

static int do_train(twtml_model* m, const twtml_batch* b, bool with_stats, double* out_real, double* out_pred, twtml_stats* out,
                    double* loss_history, int32_t cap, int32_t* iters_run)
{
    twtml_ctx* ctx = m->ctx;
    if (iters_run) *iters_run = 0;
    ctx->prof = twtml_profile{};
    TW_TRY(wait_ready(ctx, b));
    TW_TRY(resolve_global_rows(ctx, b));
    if (b->n_global == 0) { // rdd.isEmpty: both output ops skip the batch
        if (with_stats && out) stats_empty(out);
        return 0;
    }
    TW_TRY(reserve_scratch(ctx, m, b));
    double *dp = nullptr, *dr = nullptr;
    if (with_stats && out_pred) { TW_TRY(ctx->tmp_a.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, b->n))); dp = ctx->tmp_a.as<double>(); }
    if (with_stats && out_real) { TW_TRY(ctx->tmp_b.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, b->n))); dr = ctx->tmp_b.as<double>(); }

    TrainState* st = ctx->state.as<TrainState>();
    unsigned* fuse_counter = reinterpret_cast<unsigned*>(ctx->state.as<char>() + sizeof(TrainState));
    TW_CUDA(cudaMemsetAsync(fuse_counter, 0, sizeof(unsigned), ctx->stream));
    k_train_init<<<1, 1024, 0, ctx->stream>>>(st, m->d_w, (int)m->d, m->updater, m->reg, m->num_iter);
    TW_CUDA(cudaGetLastError());
    if (with_stats && m->num_iter < 1) { // no iteration to ride on
        TW_TRY(run_pass(ctx, m, b, MODE_PREDICT_STATS, dp, dr, 0));
        TW_TRY(reduce_stats(ctx, b));
    }
    const int pv = partial_width(b);
    UpdateArgs ua{};
    ua.shards = ctx->shards.as<double>();
    ua.w = m->d_w;
    ua.state = st;
    ua.loss_hist = ctx->loss_hist.as<double>();
    ua.nshards = b->plan.spr * ctx->world;
    ua.pv = pv;
    ua.d = (int)m->d;
    ua.ldx = loss_slot(b);
    ua.max_iter = m->num_iter;
    ua.updater = m->updater;
    ua.step = m->step;
    ua.reg = m->reg;
    ua.tol = m->tol;
    // One GPU: the host may stop enqueuing once the device has flagged the end of the loop (converged early);
    // with several ranks every rank must issue the same sequence of exchanges, so all iterations are enqueued.
    const bool may_stop_early = ctx->world == 1 && ctx->host_done_dev != nullptr;
    if (may_stop_early) { *ctx->host_done = 0; ua.host_done = ctx->host_done_dev; }

    bool wide_p2p = false;
    if (use_sliced_p2p(ctx, m, b)) { // collective decision + (first time, or a wider model) the buffer's set-up
        TW_TRY(setup_wide_exchange(ctx, m->d));
        wide_p2p = !ctx->wx_failed && ctx->wxbuf != nullptr;
    }
    static const bool no_fuse = getenv("TWTML_NO_FUSED_TAIL") != nullptr; // A/B switch
    const bool p2p = ctx->world > 1 && use_p2p(ctx, b);
    const bool fused_tail = (ctx->world == 1 || p2p) && !no_fuse && m->d <= 2 * kWideCols && (!b->csr || b->narrow);
    PeerExchange px{};
    if (p2p) {
        for (int r = 0; r < ctx->world; ++r) px.peer[r] = ctx->peer[r];
        px.world = ctx->world;
        px.rank = ctx->rank;
    }
    const bool prof = ctx->profiling;
    size_t ne = 0;
    if (prof) TW_CUDA(cudaEventRecord(prof_event(ctx, ne++), ctx->stream));
    int enqueued = 0;
    for (int i = 1; i <= m->num_iter; ++i) {
        if (may_stop_early && i > 2 && *reinterpret_cast<volatile int*>(ctx->host_done)) break;
        ++enqueued;
        const int mode = (i == 1 && with_stats) ? MODE_TRAIN_STATS : MODE_TRAIN;
        if (prof) TW_CUDA(cudaEventRecord(prof_event(ctx, ne++), ctx->stream));
        TW_TRY(run_pass(ctx, m, b, mode, dp, dr, 0));
        // the timed "iteration kernel": k_dense_iter / k_csr_narrow alone; wide CSR: row pass + column pass + fold
        const bool one_kernel = !b->csr || b->narrow;
        if (prof && one_kernel) TW_CUDA(cudaEventRecord(prof_event(ctx, ne++), ctx->stream));
        ctx->prof.iter_launches += 1;
        if (fused_tail) { // narrow model: shard reduce (+ the exchange over peer memory) + update in one launch
            const int spr = b->plan.spr;
            dim3 grid((pv + 31) / 32, spr);
            if (p2p) {
                ua.xbuf = ctx->xbuf;
                k_shard_reduce_update<true><<<grid, 1024, 0, ctx->stream>>>(ctx->partials.as<double>(), pv, b->plan, ctx->shards.as<double>(), fuse_counter, ua, px);
            } else {
                k_shard_reduce_update<false><<<grid, 1024, 0, ctx->stream>>>(ctx->partials.as<double>(), pv, b->plan, ctx->shards.as<double>(), fuse_counter, ua, px);
            }
            TW_CUDA(cudaGetLastError());
            ctx->prof.other_launches += 1;
            if (i == 1 && with_stats) TW_TRY(reduce_stats(ctx, b));
            continue;
        }
        TW_TRY(reduce_local(ctx, b, true, wide_p2p));
        if (prof && !one_kernel) TW_CUDA(cudaEventRecord(prof_event(ctx, ne++), ctx->stream));
        if (use_sliced(ctx, m, b)) {
            TW_TRY(wide_p2p ? sliced_update_p2p(ctx, m, b, ua) : sliced_update(ctx, m, b, ua));
            if (i == 1 && with_stats) TW_TRY(reduce_stats(ctx, b));
            continue;
        }
        TW_TRY(exchange_shards(ctx, b));
        if (use_p2p(ctx, b)) { // read this iteration's vectors from the local exchange buffer once every writer has signalled
            ua.xbuf = ctx->xbuf;
        }
        if (m->d <= 2 * kWideCols) {
            k_update<<<1, 1024, 0, ctx->stream>>>(ua);
            ctx->prof.other_launches += 1;
        } else {
            const int nparts = (int)((m->d + kWideCols - 1) / kWideCols);
            k_update_wide<<<nparts, 1024, 0, ctx->stream>>>(ua, ctx->upd_parts.as<double>());
            k_update_finish<<<1, 1024, 0, ctx->stream>>>(ua, ctx->upd_parts.as<double>(), nparts);
            ctx->prof.other_launches += 2;
        }
        TW_CUDA(cudaGetLastError());
        if (i == 1 && with_stats) TW_TRY(reduce_stats(ctx, b));
    }
    if (prof) TW_CUDA(cudaEventRecord(prof_event(ctx, ne++), ctx->stream));

    TrainState hs{};
    double o[10];
    std::vector<double> hist((size_t)std::max(1, m->num_iter));
    TW_CUDA(cudaMemcpyAsync(&hs, st, sizeof(hs), cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaMemcpyAsync(hist.data(), ctx->loss_hist.p, sizeof(double) * hist.size(), cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaMemcpyAsync(m->h_w.data(), m->d_w, sizeof(double) * (size_t)m->d, cudaMemcpyDeviceToHost, ctx->stream));
    if (with_stats) {
        TW_CUDA(cudaMemcpyAsync(o, ctx->out10.p, sizeof(o), cudaMemcpyDeviceToHost, ctx->stream));
        TW_TRY(fetch_rows(ctx, out_pred, dp, b->n));
        TW_TRY(fetch_rows(ctx, out_real, dr, b->n));
    }
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    if (hs.pad == kPeerTimeoutMark) return fail(TWTML_E_NCCL, "peer exchange timed out: a rank never delivered its shard vectors");
    if (with_stats && out) stats_from_out10(o, out);
    if (iters_run) *iters_run = hs.nhist;
    if (loss_history)
        for (int i = 0; i < hs.nhist && i < cap; ++i) loss_history[i] = hist[i];
    if (prof) {
        float ms = 0.f, tot = 0.f;
        for (int i = 0; i < enqueued; ++i) {
            cudaEventElapsedTime(&ms, ctx->ev[1 + 2 * i], ctx->ev[2 + 2 * i]);
            tot += ms;
        }
        ctx->prof.iter_kernel_ms = tot;
        cudaEventElapsedTime(&ctx->prof.loop_ms, ctx->ev[0], ctx->ev[ne - 1]);
        if (b->csr) ctx->prof.iter_kernel_bytes = b->n * 16 + b->nnz * 12 + 16 * b->d;
        else ctx->prof.iter_kernel_bytes = 8 * b->n * (b->d + 1) + 16 * b->d;
    }
    return 0;
}

// Everything a statistics / training call may allocate (scratch buffers, the per-row output staging, the wide exchange
// buffer), done ahead of the call.  A device group runs this on all members first: an allocation on one member must not
// fall into the time another member's kernels are already spinning on this member's flags.
int prepare_compute(twtml_model* m, const twtml_batch* b, bool per_row_outputs)
{
    twtml_ctx* ctx = m->ctx;
    Guard g(ctx);
    TW_TRY(wait_ready(ctx, b));
    TW_TRY(resolve_global_rows(ctx, b));
    if (b->n_global == 0) return 0;
    TW_TRY(reserve_scratch(ctx, m, b));
    if (per_row_outputs) {
        TW_TRY(ctx->tmp_a.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, b->n)));
        TW_TRY(ctx->tmp_b.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, b->n)));
    }
    if (use_sliced_p2p(ctx, m, b)) TW_TRY(setup_wide_exchange(ctx, m->d));
    return 0;
}

// caller holds the context lock (or is the only owner)
void batch_free_locked(twtml_batch* b)
{
    if (b->has_pending) b->pending.wait(); // the worker still owns its fields
    if (b->has_pending && b->pending.get() != 0) cudaStreamSynchronize(b->ctx->copy_stream); // a failed ingest: its kernels may still run, its event was never recorded
    if (b->ready) {
        cudaEventSynchronize(b->ready); // its own copy only; later batches may still be arriving
        cudaEventDestroy(b->ready);
    }
    cudaStreamSynchronize(b->ctx->stream);
    for (void* p : {(void*)b->X, (void*)b->rowptr, (void*)b->colidx, (void*)b->vals, (void*)b->labels, (void*)b->colptr, (void*)b->rowidx,
                    (void*)b->cvals, (void*)b->headbits, (void*)b->nzcol, (void*)b->win_first_nz, (void*)b->wkey, (void*)b->win_first_part,
                    (void*)b->win_last_part, (void*)b->diff, (void*)b->bounds, (void*)b->pcolidx, (void*)b->pvals})
        block_release(b->ctx, p); // (the compute stream and the batch's own ingest are done: the arrays are idle)
    b->ctx->live_handles -= 1;
    delete b;
}

} // namespace twtml

using namespace twtml;

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

int32_t twtml_abi_version(void) { return TWTML_ABI_VERSION; }

int32_t twtml_last_error(char* buf, int64_t cap)
{
    if (!buf || cap <= 0) return TWTML_E_INVALID_ARG;
    snprintf(buf, (size_t)cap, "%s", g_err);
    return 0;
}

static int32_t create_common(int32_t device, int32_t rank, int32_t world, const uint8_t* id, twtml_ctx** out)
{
    if (!out) return fail(TWTML_E_INVALID_ARG, "twtml_create: out is null");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(TWTML_E_CUDA, "no CUDA device (%s); libtwtml has no CPU fallback", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(TWTML_E_INVALID_ARG, "device %d out of range [0,%d)", device, ndev);
    cudaDeviceProp prop;
    TW_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(TWTML_E_CUDA, "device %d is sm_%d%d; libtwtml is built for sm_100a only", device, prop.major, prop.minor);
    TW_CUDA(cudaSetDevice(device));
    twtml_ctx* c = new twtml_ctx();
    c->device = device;
    c->rank = rank;
    c->world = world;
    c->sm_count = prop.multiProcessorCount;
    // 8 virtual shards in total when the rank count divides 8, so that 1/2/4/8 GPUs add the same
    // shard vectors in the same order; otherwise one shard per rank.
    c->spr = (world <= kMaxShards && kMaxShards % world == 0) ? kMaxShards / world : 1;
    c->nshards = c->spr * world;
    cudaError_t se = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (se != cudaSuccess) { delete c; return fail(TWTML_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(se)); }
    cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&c->spread_stream, cudaStreamNonBlocking);
    c->lane[0].stream = c->stream;
    c->lane[1].stream = c->copy_stream;
    for (auto& lane : c->lane) {
        for (auto& half : lane.stage_ev)
            for (auto& ev : half) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
        for (auto& ev : lane.pin_ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    }
    {
        cudaMemPoolProps props{};
        props.allocType = cudaMemAllocationTypePinned;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        cudaError_t pe = cudaMemPoolCreate(&c->pool, &props);
        if (pe != cudaSuccess) { cudaStreamDestroy(c->stream); delete c; return fail(TWTML_E_CUDA, "cudaMemPoolCreate: %s", cudaGetErrorString(pe)); }
        unsigned long long keep = ~0ull; // freed batch memory stays cached for the next micro-batch
        cudaMemPoolSetAttribute(c->pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    if (cudaHostAlloc(reinterpret_cast<void**>(&c->host_done), sizeof(int), cudaHostAllocMapped) == cudaSuccess) {
        *c->host_done = 0;
        if (cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->host_done_dev), c->host_done, 0) != cudaSuccess) c->host_done_dev = nullptr;
    }
    cudaGetLastError();
    if (world > 1) {
        if (!nccl_api().ok) { delete c; return fail(TWTML_E_NCCL, "libnccl.so.2 could not be loaded: %s", dlerror()); }
        ncclUniqueId uid;
        memcpy(&uid, id, sizeof(uid));
        ncclResult_t r = nccl_api().CommInitRank(&c->comm, world, uid, rank);
        if (r != ncclSuccess) { delete c; return fail(TWTML_E_NCCL, "ncclCommInitRank: %s", nccl_api().GetErrorString(r)); }
        int prc = setup_peer_exchange(c);
        if (prc) { delete c; return prc; }
    }
    *out = c;
    return 0;
}

int32_t twtml_create(int32_t device, twtml_ctx** out) { return create_common(device, 0, 1, nullptr, out); }

int32_t twtml_create_all(const int32_t* devices, int32_t n_devices, twtml_ctx** out) { return grp::create_all(devices, n_devices, out); }

int32_t twtml_device_count(const twtml_ctx* ctx, int32_t* n_devices)
{
    if (!ctx || !n_devices) return fail(TWTML_E_INVALID_ARG, "null argument");
    *n_devices = ctx->members.empty() ? 1 : (int32_t)ctx->members.size();
    return 0;
}

int32_t twtml_nccl_unique_id(uint8_t* out, int64_t cap)
{
    static_assert(sizeof(ncclUniqueId) == TWTML_NCCL_ID_BYTES, "ncclUniqueId size");
    if (!out || cap < TWTML_NCCL_ID_BYTES) return fail(TWTML_E_INVALID_ARG, "twtml_nccl_unique_id: need %d bytes", TWTML_NCCL_ID_BYTES);
    if (!nccl_api().ok) return fail(TWTML_E_NCCL, "libnccl.so.2 could not be loaded");
    ncclUniqueId uid;
    TW_NCCL(nccl_api().GetUniqueId(&uid));
    memcpy(out, &uid, sizeof(uid));
    return 0;
}

int32_t twtml_create_ranked(int32_t device, int32_t rank, int32_t world_size, const uint8_t* nccl_id, int64_t id_bytes, twtml_ctx** out)
{
    if (world_size < 1 || rank < 0 || rank >= world_size) return fail(TWTML_E_INVALID_ARG, "bad rank %d / world %d", rank, world_size);
    if (world_size > 1 && (!nccl_id || id_bytes < TWTML_NCCL_ID_BYTES)) return fail(TWTML_E_INVALID_ARG, "nccl_id must hold %d bytes", TWTML_NCCL_ID_BYTES);
    return create_common(device, rank, world_size, nccl_id, out);
}

int32_t twtml_destroy(twtml_ctx* ctx)
{
    if (!ctx) return 0;
    if (!ctx->members.empty()) return grp::destroy(ctx);
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        if (ctx->live_handles.load() != 0)
            return fail(TWTML_E_INVALID_ARG, "twtml_destroy: %d models/batches of this context are still alive; free them first", ctx->live_handles.load());
    }
    stop_worker(ctx);
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    if (ctx->spread_stream) cudaStreamSynchronize(ctx->spread_stream);
    teardown_peer_exchange(ctx);
    if (ctx->host_done) cudaFreeHost(ctx->host_done);
    if (ctx->comm) nccl_api().CommDestroy(ctx->comm);
    for (DevBuf* b : {&ctx->partials, &ctx->shards, &ctx->stat_partials, &ctx->stat_shards, &ctx->state, &ctx->loss_hist, &ctx->out10,
                      &ctx->tmp_a, &ctx->tmp_b, &ctx->gather, &ctx->lower, &ctx->upd_parts, &ctx->slices, &ctx->wprev})
        b->release();
    for (auto& lane : ctx->lane) {
        for (DevBuf* b : {&lane.scan_tmp, &lane.sort_tmp, &lane.labels_tmp, &lane.stage, &lane.feat_arena}) b->release();
        for (int i = 0; i < 2; ++i) {
            if (lane.pin[i]) cudaFreeHost(lane.pin[i]);
            if (lane.pin_ev[i]) cudaEventDestroy(lane.pin_ev[i]);
        }
        for (auto& half : lane.stage_ev)
            for (auto& ev : half)
                if (ev) cudaEventDestroy(ev);
    }
    for (cudaEvent_t e : ctx->ev) cudaEventDestroy(e);
    {
        std::lock_guard<std::mutex> lk(ctx->block_mu);
        for (auto& kv : ctx->block_free) cudaFree(kv.second);
        for (auto& kv : ctx->block_live) cudaFree(kv.first);
        ctx->block_free.clear();
        ctx->block_live.clear();
    }
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    cudaStreamDestroy(ctx->stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->spread_stream) cudaStreamDestroy(ctx->spread_stream);
    delete ctx;
    return 0;
}

int32_t twtml_rank(const twtml_ctx* ctx, int32_t* rank, int32_t* world_size)
{
    if (!ctx) return fail(TWTML_E_INVALID_ARG, "null context");
    if (rank) *rank = ctx->rank;
    if (world_size) *world_size = ctx->world;
    return 0;
}

int32_t twtml_profile_enable(twtml_ctx* ctx, int32_t on)
{
    if (!ctx) return fail(TWTML_E_INVALID_ARG, "null context");
    if (!ctx->members.empty()) return grp::profile_enable(ctx, on);
    ctx->profiling = on != 0;
    return 0;
}

int32_t twtml_profile_get(const twtml_ctx* ctx, twtml_profile* out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (!ctx->members.empty()) return grp::profile_get(ctx, out);
    *out = ctx->prof;
    return 0;
}

// ---- model ----------------------------------------------------------------------------------
int32_t twtml_model_create(twtml_ctx* ctx, int64_t num_features, twtml_model** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (num_features < 1 || num_features > (int64_t)1 << 30) return fail(TWTML_E_INVALID_ARG, "num_features %lld out of range", (long long)num_features);
    if (!ctx->members.empty()) return grp::model_create(ctx, num_features, out);
    Guard g(ctx);
    twtml_model* m = new twtml_model();
    m->ctx = ctx;
    m->d = num_features;
    m->h_w.assign((size_t)num_features, 0.0);
    // room for world equal column slices (the all-gather of a sliced update writes whole slices)
    cudaError_t e = cudaMalloc(&m->d_w, sizeof(double) * (size_t)(num_features + 2 + ctx->world));
    if (e != cudaSuccess) { delete m; return fail(TWTML_E_OOM, "cudaMalloc weights: %s", cudaGetErrorString(e)); }
    ctx->live_handles += 1;
    *out = m;
    return 0;
}

int32_t twtml_model_free(twtml_model* m)
{
    if (!m) return 0;
    if (!m->members.empty()) return grp::model_free(m);
    {
        Guard g(m->ctx);
        cudaStreamSynchronize(m->ctx->stream);
        cudaFree(m->d_w);
        m->ctx->live_handles -= 1;
    }
    delete m;
    return 0;
}

int32_t twtml_set_step_size(twtml_model* m, double v)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_step_size(mm, v));
    m->step = v;
    return 0;
}
int32_t twtml_set_num_iterations(twtml_model* m, int32_t v)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_num_iterations(mm, v));
    if (v < 0) return fail(TWTML_E_INVALID_ARG, "numIterations must be >= 0");
    m->num_iter = v;
    return 0;
}
int32_t twtml_set_mini_batch_fraction(twtml_model* m, double v)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_mini_batch_fraction(mm, v));
    m->frac = v;
    return 0;
}
int32_t twtml_set_convergence_tol(twtml_model* m, double v)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_convergence_tol(mm, v));
    if (!(v >= 0.0 && v <= 1.0)) return fail(TWTML_E_INVALID_ARG, "convergenceTol must be in [0, 1]"); // require() in GradientDescent [U]
    m->tol = v;
    return 0;
}
int32_t twtml_set_reg_param(twtml_model* m, double v)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_reg_param(mm, v));
    m->reg = v;
    return 0;
}
int32_t twtml_set_updater(twtml_model* m, int32_t kind)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_updater(mm, kind));
    if (kind < 0 || kind > 2) return fail(TWTML_E_INVALID_ARG, "updater must be 0 (Simple), 1 (SquaredL2) or 2 (L1)");
    m->updater = kind;
    return 0;
}

int32_t twtml_set_initial_weights(twtml_model* m, const double* w, int64_t d)
{
    if (!m || !w) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (d != m->d) return fail(TWTML_E_SHAPE_MISMATCH, "setInitialWeights: got %lld values for %lld features", (long long)d, (long long)m->d);
    if (!m->members.empty()) return grp::set_initial_weights(m, w, d);
    Guard g(m->ctx);
    memcpy(m->h_w.data(), w, sizeof(double) * (size_t)d);
    TW_CUDA(cudaMemcpyAsync(m->d_w, m->h_w.data(), sizeof(double) * (size_t)d, cudaMemcpyHostToDevice, m->ctx->stream));
    TW_CUDA(cudaStreamSynchronize(m->ctx->stream));
    m->initialized = true;
    return 0;
}

int32_t twtml_get_weights(const twtml_model* m, double* out, int64_t d)
{
    if (!m || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (!m->initialized) return fail(TWTML_E_NOT_INITIALIZED, "Model must be initialized before starting prediction.");
    if (d != m->d) return fail(TWTML_E_SHAPE_MISMATCH, "get_weights: buffer of %lld for %lld features", (long long)d, (long long)m->d);
    if (!m->members.empty()) return grp::get_weights(m, out, d);
    memcpy(out, m->h_w.data(), sizeof(double) * (size_t)d);
    return 0;
}

int32_t twtml_num_features(const twtml_model* m, int64_t* d)
{
    if (!m || !d) return fail(TWTML_E_INVALID_ARG, "null argument");
    *d = m->d;
    return 0;
}

// ---- dense batches --------------------------------------------------------------------------
// defer_global: leave the multi-rank row-count exchange to the batch's first use (asynchronous ingest must not touch
// the communicator, which belongs to the compute stream)
static int32_t new_dense(twtml_ctx* ctx, int64_t n, int64_t d, twtml_batch** out, bool defer_global = false)
{
    DenseCfg c;
    TW_TRY(dense_cfg(d, c));
    twtml_batch* b = new twtml_batch();
    b->ctx = ctx;
    b->n = n;
    b->d = d;
    b->ldx = (d + 2) & ~(int64_t)1;
    b->rows_appended = n;
    make_plan(n, ctx->spr, c.R, b->plan);
    if (n > 0) {
        int rc0 = pool_alloc(ctx, reinterpret_cast<void**>(&b->X), sizeof(double) * (size_t)n * (size_t)b->ldx);
        if (rc0) { delete b; return rc0; }
    }
    if (defer_global && ctx->world > 1) {
        b->global_deferred = true;
    } else {
        int rc = ctx_global_rows(ctx, n, &b->n_global, &b->row0_global);
        if (rc) { pool_free(ctx, b->X); delete b; return rc; }
    }
    ctx->live_handles += 1;
    *out = b;
    return 0;
}

// rows [row0, row0 + nrows) of a dense batch from host memory, on `lane` (enqueued; nothing waits here)
static int dense_rows_in(twtml_ctx* ctx, IngestLane& lane, twtml_batch* b, int64_t row0, int64_t nrows, const double* X, int64_t ld_x,
                         const double* labels, double* labels_dev)
{
    const int64_t d = b->d, ldx = b->ldx;
    if (page_locked(X)) { // page-locked caller memory: DMA straight from it, labels dropped in on the device
        TW_TRY(upload_plain(ctx, lane, labels_dev, labels, sizeof(double) * (size_t)nrows));
        if (ld_x == d) { // packed rows: flat copies + a device kernel to the row pitch
            cudaError_t e = flat_dense_ingest(ctx, lane, b, row0, nrows, X, labels_dev);
            if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? TWTML_E_OOM : TWTML_E_CUDA, "dense ingest: %s", cudaGetErrorString(e));
        } else {
            TW_CUDA(cudaMemcpy2DAsync(b->X + row0 * ldx, (size_t)ldx * 8, X, (size_t)ld_x * 8, (size_t)d * 8, (size_t)nrows, cudaMemcpyHostToDevice, lane.stream));
            k_scatter_labels<<<(unsigned)((nrows + 255) / 256), 256, 0, lane.stream>>>(b->X + row0 * ldx, ldx, (int)d, labels_dev, nrows);
            TW_CUDA(cudaGetLastError());
        }
        return 0;
    }
    // pageable: pack rows [x.., label, pad] into the pinned ring, chunk k+1 while chunk k copies
    return staged_upload(lane, b->X + row0 * ldx, (size_t)ldx * 8, nrows, [&](char* dst, int64_t r0, int64_t nr) {
        parallel_rows(nr, 2048, [&](int64_t lo, int64_t hi) {
            for (int64_t r = lo; r < hi; ++r) {
                double* row = reinterpret_cast<double*>(dst) + r * ldx;
                memcpy(row, X + (r0 + r) * ld_x, sizeof(double) * (size_t)d);
                row[d] = labels[r0 + r];
                for (int64_t j = d + 1; j < ldx; ++j) row[j] = 0.0;
            }
        });
    });
}

int32_t twtml_batch_from_dense(twtml_ctx* ctx, int64_t n, int64_t d, const double* X, int64_t ld_x, const double* labels, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d < 1 || (n > 0 && (!X || !labels)) || ld_x < d) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_dense: bad shape or null data");
    if (!ctx->members.empty()) return grp::from_dense(ctx, n, d, X, ld_x, labels, false, out);
    Guard g(ctx);
    twtml_batch* b = nullptr;
    TW_TRY(new_dense(ctx, n, d, &b));
    if (n > 0) {
        IngestLane& lane = ctx->lane[0];
        int rc = lane.labels_tmp.reserve(sizeof(double) * (size_t)n);
        if (!rc) rc = dense_rows_in(ctx, lane, b, 0, n, X, ld_x, labels, lane.labels_tmp.as<double>());
        if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(TWTML_E_CUDA, "upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        if (rc) { pool_free(ctx, b->X); ctx->live_handles -= 1; delete b; return rc; }
    }
    *out = b;
    return 0;
}

// A dense batch whose memory is reserved now and whose rows arrive later, on the copy stream.
static int32_t begin_dense_async(twtml_ctx* ctx, int64_t n, int64_t d, twtml_batch** out)
{
    twtml_batch* b = nullptr;
    TW_TRY(new_dense(ctx, n, d, &b, true));
    b->rows_appended = 0;
    if (int rc = pool_alloc(ctx, reinterpret_cast<void**>(&b->labels), sizeof(double) * (size_t)std::max<int64_t>(n, 1))) { batch_free_locked(b); return rc; }
    cudaError_t e = cudaEventCreateWithFlags(&b->ready, cudaEventDisableTiming);
    // the two buffers were allocated in compute-stream order: the copy stream may touch them only after that point
    if (e == cudaSuccess) e = cudaEventRecord(b->ready, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->copy_stream, b->ready, 0);
    if (e != cudaSuccess) {
        int rc = fail(TWTML_E_CUDA, "asynchronous ingest: %s", cudaGetErrorString(e));
        batch_free_locked(b);
        return rc;
    }
    *out = b;
    return 0;
}

static int32_t append_dense_async(twtml_ctx* ctx, twtml_batch* b, int64_t row0, int64_t nrows, const double* X, int64_t ld_x, const double* labels)
{
    std::lock_guard<std::mutex> lane_lock(ctx->lane[1].mu);
    TW_TRY(dense_rows_in(ctx, ctx->lane[1], b, row0, nrows, X, ld_x, labels, b->labels + row0));
    TW_CUDA(cudaEventRecord(b->ready, ctx->copy_stream));
    b->rows_appended += nrows;
    return 0;
}

int32_t twtml_batch_from_dense_async(twtml_ctx* ctx, int64_t n, int64_t d, const double* X, int64_t ld_x, const double* labels, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d < 1 || (n > 0 && (!X || !labels)) || ld_x < d) return fail(TWTML_E_INVALID_ARG, "twtml_batch_from_dense_async: bad shape or null data");
    if (!ctx->members.empty()) return grp::from_dense(ctx, n, d, X, ld_x, labels, true, out);
    // only page-locked host memory can be copied without a host-side staging pass
    if (n > 0 && (!page_locked(X) || !page_locked(labels))) return twtml_batch_from_dense(ctx, n, d, X, ld_x, labels, out);
    Guard g(ctx);
    twtml_batch* b = nullptr;
    TW_TRY(begin_dense_async(ctx, n, d, &b));
    if (n > 0) {
        if (int rc = append_dense_async(ctx, b, 0, n, X, ld_x, labels)) { batch_free_locked(b); return rc; }
    }
    *out = b;
    return 0;
}

int32_t twtml_batch_dense_begin(twtml_ctx* ctx, int64_t n, int64_t d, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d < 1) return fail(TWTML_E_INVALID_ARG, "twtml_batch_dense_begin: bad shape");
    if (!ctx->members.empty()) return grp::dense_begin(ctx, n, d, out);
    Guard g(ctx);
    return begin_dense_async(ctx, n, d, out);
}

int32_t twtml_batch_dense_append(twtml_batch* b, int64_t nrows, const double* X, int64_t ld_x, const double* labels)
{
    if (!b) return fail(TWTML_E_INVALID_ARG, "null batch");
    if (!b->members.empty()) {
        if (nrows < 0 || (nrows > 0 && (!X || !labels)) || ld_x < b->d) return fail(TWTML_E_INVALID_ARG, "twtml_batch_dense_append: bad shape or null data");
        return grp::dense_append(b, nrows, X, ld_x, labels);
    }
    if (b->csr || !b->ready) return fail(TWTML_E_INVALID_ARG, "twtml_batch_dense_append: not a batch from twtml_batch_dense_begin");
    if (nrows < 0 || (nrows > 0 && (!X || !labels)) || ld_x < b->d) return fail(TWTML_E_INVALID_ARG, "twtml_batch_dense_append: bad shape or null data");
    if (b->rows_appended + nrows > b->n)
        return fail(TWTML_E_INVALID_ARG, "twtml_batch_dense_append: %lld rows after %lld exceed the batch's %lld", (long long)nrows, (long long)b->rows_appended, (long long)b->n);
    if (nrows == 0) return 0;
    twtml_ctx* ctx = b->ctx;
    Guard g(ctx);
    return append_dense_async(ctx, b, b->rows_appended, nrows, X, ld_x, labels);
}

int32_t twtml_batch_ready(const twtml_batch* b, int32_t wait, int32_t* ready)
{
    if (!b) return fail(TWTML_E_INVALID_ARG, "null batch");
    if (!b->members.empty()) return grp::batch_ready(b, wait, ready);
    int r = 1;
    if (b->has_pending) { // asynchronous CSR / tweet ingest: the worker's verdict first
        if (!wait && b->pending.wait_for(std::chrono::seconds(0)) != std::future_status::ready) {
            if (ready) *ready = 0;
            return 0;
        }
        const int rc = b->pending.get();
        if (rc) return fail(rc, "asynchronous ingest failed: %s", b->pending_err.c_str());
    }
    if (b->ready) {
        Guard g(b->ctx);
        if (wait) TW_CUDA(cudaEventSynchronize(b->ready));
        else {
            cudaError_t e = cudaEventQuery(b->ready);
            if (e == cudaErrorNotReady) r = 0;
            else if (e != cudaSuccess) return fail(TWTML_E_CUDA, "cudaEventQuery: %s", cudaGetErrorString(e));
        }
    }
    if (ready) *ready = r;
    return 0;
}

int32_t twtml_batch_synth_dense(twtml_ctx* ctx, int64_t n, int64_t d, uint64_t seed, int64_t row0, twtml_batch** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (n < 0 || d < 1) return fail(TWTML_E_INVALID_ARG, "twtml_batch_synth_dense: bad shape");
    if (!ctx->members.empty()) return grp::synth_dense(ctx, n, d, seed, row0, out);
    Guard g(ctx);
    twtml_batch* b = nullptr;
    TW_TRY(new_dense(ctx, n, d, &b));
    if (n > 0) {
        const double inv = 1.0 / std::sqrt((double)d);
        const unsigned gx = (unsigned)std::min<int64_t>(n, (int64_t)ctx->sm_count * 64);
        k_synth_dense_x<<<gx, 256, 0, ctx->stream>>>(b->X, b->ldx, (int)d, seed, row0, inv, n);
        k_synth_dense_y<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(b->X, b->ldx, (int)d, seed, row0, inv, n);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { pool_free(ctx, b->X); ctx->live_handles -= 1; delete b; return fail(TWTML_E_CUDA, "synthetic generator: %s", cudaGetErrorString(e)); }
    }
    *out = b;
    return 0;
}

int32_t twtml_batch_rows(const twtml_batch* b, int64_t* n)
{
    if (!b || !n) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (!b->members.empty()) return grp::batch_info(b, n, nullptr, nullptr, nullptr);
    *n = b->n;
    return 0;
}

int32_t twtml_batch_info(const twtml_batch* b, int64_t* n, int64_t* d, int64_t* nnz, int32_t* is_csr)
{
    if (!b) return fail(TWTML_E_INVALID_ARG, "null batch");
    if (!b->members.empty()) return grp::batch_info(b, n, d, nnz, is_csr);
    if (nnz && b->has_pending) { // only the finished ingest knows how many entries a featurised batch has
        const int rc = b->pending.get();
        if (rc) return fail(rc, "asynchronous ingest failed: %s", b->pending_err.c_str());
    }
    if (n) *n = b->n;
    if (d) *d = b->d;
    if (nnz) *nnz = b->csr ? b->nnz : b->n * b->d;
    if (is_csr) *is_csr = b->csr ? 1 : 0;
    return 0;
}

int32_t twtml_batch_export_dense(const twtml_batch* b, double* X, double* labels)
{
    if (!b || !X || !labels) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (b->csr) return fail(TWTML_E_INVALID_ARG, "twtml_batch_export_dense on a CSR batch");
    if (!b->members.empty()) return grp::export_dense(b, X, labels);
    if (b->n == 0) return 0;
    twtml_ctx* ctx = b->ctx;
    Guard g(ctx);
    TW_TRY(wait_ready(ctx, b));
    TW_TRY(ctx->tmp_a.reserve(sizeof(double) * (size_t)b->n * (size_t)b->d));
    TW_TRY(ctx->tmp_b.reserve(sizeof(double) * (size_t)b->n));
    k_gather_dense<<<(unsigned)std::min<int64_t>(b->n, 65535), 256, 0, ctx->stream>>>(b->X, b->ldx, (int)b->d, ctx->tmp_a.as<double>(), ctx->tmp_b.as<double>(), b->n);
    TW_CUDA(cudaGetLastError());
    TW_CUDA(cudaMemcpyAsync(X, ctx->tmp_a.p, sizeof(double) * (size_t)b->n * (size_t)b->d, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaMemcpyAsync(labels, ctx->tmp_b.p, sizeof(double) * (size_t)b->n, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int32_t twtml_batch_free(twtml_batch* b)
{
    if (!b) return 0;
    if (!b->members.empty()) return grp::batch_free(b);
    Guard g(b->ctx);
    batch_free_locked(b);
    return 0;
}

// ---- compute --------------------------------------------------------------------------------
int32_t twtml_predict(const twtml_model* m, const twtml_batch* b, double* out_pred)
{
    TW_TRY(check_pair(m, b, "prediction"));
    if (!out_pred && b->n > 0) return fail(TWTML_E_INVALID_ARG, "twtml_predict: out_pred is null");
    if (!m->members.empty()) return grp::predict(m, b, out_pred);
    twtml_ctx* ctx = m->ctx;
    Guard g(ctx);
    TW_TRY(wait_ready(ctx, b));
    if (b->n == 0) return 0;
    TW_TRY(reserve_scratch(ctx, m, b));
    TW_TRY(ctx->tmp_a.reserve(sizeof(double) * (size_t)b->n));
    TW_TRY(run_pass(ctx, m, b, MODE_PREDICT, ctx->tmp_a.as<double>(), nullptr, 0));
    TW_TRY(fetch_rows(ctx, out_pred, ctx->tmp_a.as<double>(), b->n));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int32_t twtml_predict_stats(const twtml_model* m, const twtml_batch* b, double* out_real, double* out_pred_rounded, twtml_stats* out)
{
    TW_TRY(check_pair(m, b, "prediction"));
    if (!out) return fail(TWTML_E_INVALID_ARG, "twtml_predict_stats: out is null");
    if (!m->members.empty()) return grp::predict_stats(m, b, out_real, out_pred_rounded, out);
    Guard g(m->ctx);
    return do_predict_stats(m, b, out_real, out_pred_rounded, out);
}

int32_t twtml_train_on_batch(twtml_model* m, const twtml_batch* b, double* loss_history, int32_t cap, int32_t* iters_run)
{
    TW_TRY(check_pair(m, b, "training"));
    if (!m->members.empty()) return grp::train(m, b, false, nullptr, nullptr, nullptr, loss_history, cap, iters_run);
    Guard g(m->ctx);
    return do_train(m, b, false, nullptr, nullptr, nullptr, loss_history, cap, iters_run);
}

int32_t twtml_step(twtml_model* m, const twtml_batch* b, double* out_real, double* out_pred_rounded, twtml_stats* out,
                   double* loss_history, int32_t cap, int32_t* iters_run)
{
    TW_TRY(check_pair(m, b, "training"));
    if (!out) return fail(TWTML_E_INVALID_ARG, "twtml_step: out is null");
    if (!m->members.empty()) return grp::train(m, b, true, out_real, out_pred_rounded, out, loss_history, cap, iters_run);
    Guard g(m->ctx);
    return do_train(m, b, true, out_real, out_pred_rounded, out, loss_history, cap, iters_run);
}

} // extern "C"
