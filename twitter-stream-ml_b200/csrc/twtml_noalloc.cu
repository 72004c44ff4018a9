// twtml_noalloc.cu -- twtml_predict_stats_static: the batch statistics for callers that must not
// allocate GPU memory.
//
// Who that is: a Spark executor (or any host application) that has already handed the whole GPU to
// its own memory pool, so a cudaMalloc from inside this library would either fail or carve memory
// out from under that pool; and a caller on a per-batch deadline, for whom cudaMalloc / cudaFree
// are unwelcome because they can stall until all work already queued on the device has finished.
// twtml_predict_stats grows per-context scratch buffers on demand (cudaMalloc on first use and
// whenever a larger batch arrives); this entry point never does.
//
// How: the scratch the statistics pass needs -- one 8-double tuple per row block, then one per
// shard, then the ten result values -- lives in ONE file-scope device array of fixed size below,
// which every GPU that loads libtwtml.so carries until the process exits (kScratchDoubles * 8
// bytes, about 1 MiB).  The inputs (model weights, batch) are already resident; the only output is
// the caller's twtml_stats struct.
//
// Rules that follow from sharing one array:
//  * limit: the batch may have at most kMaxBlocks row blocks (twtml_batch plans use blocks of up to
//    512 rows, so up to 8.4 million rows per GPU); larger batches are refused with
//    TWTML_E_INVALID_ARG, as are CSR batches and multi-rank contexts.
//  * every call overwrites the array, and it keeps the previous call's contents in between: the
//    part that is read without being fully rewritten first (shard tuples + results) is cleared at
//    the start of each call; the per-block tuples are all rewritten before they are read.
//  * two calls whose GPU work overlaps (two contexts on the same device from two threads) race on
//    the array.  Serialise them.
//  * the first call on each device resolves the array's device address (cudaGetSymbolAddress).
//    Make that first call before starting any stream capture on that device.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

#include "dense.cuh"
#include "host.h"
#include "reduce.cuh"

namespace twtml {

constexpr int kMaxBlocks = 16384;
constexpr size_t kScratchDoubles = (size_t)kMaxBlocks * kStatStride + (size_t)kMaxShards * kStatStride + 16;

// In host code this name is only a handle for the runtime, not a device address: it is never
// passed to a kernel.  setup_device() asks the runtime for the real address.
static __device__ __align__(256) double g_static_scratch[kScratchDoubles];

constexpr int kMaxDevices = 64;
static std::mutex g_setup_mu;
static double* g_scratch_addr[kMaxDevices] = {nullptr};
static unsigned g_smem_attr_done[kMaxDevices] = {0}; // per device: kernel forms whose shared-memory limit is already raised

// once per device, with that device current
static int setup_device(int device, double** out)
{
    if (device < 0 || device >= kMaxDevices) return fail(TWTML_E_INVALID_ARG, "device index %d not supported by the static-scratch entry point", device);
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (!g_scratch_addr[device]) {
        double* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch);
        if (e != cudaSuccess || !p) return fail(TWTML_E_CUDA, "cudaGetSymbolAddress(static scratch): %s", cudaGetErrorString(e));
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return 0;
}

template <int CP, bool RL1>
static int launch_stats_pass(twtml_ctx* ctx, const DenseCfg& c, const DenseIterArgs& a)
{
    auto kern = k_dense_iter<CP, MODE_PREDICT_STATS, RL1>;
    constexpr unsigned form = 1u << (CP == 1 ? (RL1 ? 1 : 0) : (CP == 2 ? 2 : 3));
    {   // once per device and kernel form (set-up work, like the address look-up)
        std::lock_guard<std::mutex> lk(g_setup_mu);
        if (!(g_smem_attr_done[ctx->device] & form)) {
            TW_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
            g_smem_attr_done[ctx->device] |= form;
        }
    }
    long long grid = std::min<long long>((long long)ctx->sm_count * c.grid_per_sm, a.plan.nblocks);
    kern<<<(unsigned)grid, kIterThreads, c.smem, ctx->stream>>>(a);
    TW_CUDA(cudaGetLastError());
    return 0;
}

} // namespace twtml

using namespace twtml;

extern "C" int32_t twtml_predict_stats_static(const twtml_model* m, const twtml_batch* b, twtml_stats* out)
{
    if (!m || !b || !out) return fail(TWTML_E_INVALID_ARG, "twtml_predict_stats_static: null argument");
    if (m->ctx != b->ctx) return fail(TWTML_E_INVALID_ARG, "twtml_predict_stats_static: model and batch belong to different contexts");
    if (!m->initialized) return fail(TWTML_E_NOT_INITIALIZED, "Model must be initialized before starting prediction.");
    if (m->d != b->d) return fail(TWTML_E_SHAPE_MISMATCH, "twtml_predict_stats_static: model has %lld features, batch has %lld", (long long)m->d, (long long)b->d);
    twtml_ctx* ctx = m->ctx;
    if (ctx->world != 1 || !ctx->members.empty()) return fail(TWTML_E_INVALID_ARG, "twtml_predict_stats_static: single-GPU contexts only");
    if (b->csr) return fail(TWTML_E_INVALID_ARG, "twtml_predict_stats_static: dense batches only");
    if (b->plan.nblocks > kMaxBlocks)
        return fail(TWTML_E_INVALID_ARG, "twtml_predict_stats_static: batch has %lld row blocks, the static scratch holds %d (about %lld rows); use twtml_predict_stats",
                    b->plan.nblocks, kMaxBlocks, (long long)kMaxBlocks * 512);
    Guard g(ctx);
    if (b->n == 0) {
        memset(out, 0, sizeof(*out));
        out->mse = out->real_stdev = out->pred_stdev = out->mse_raw = out->real_stdev_raw = out->pred_stdev_raw = out->real_mean = out->pred_mean = NAN;
        return 0;
    }
    if (b->ready) TW_CUDA(cudaStreamWaitEvent(ctx->stream, b->ready, 0)); // batch still arriving (asynchronous ingest)
    double* scratch = nullptr;
    TW_TRY(setup_device(ctx->device, &scratch));
    double* block_tuples = scratch;
    double* shard_tuples = scratch + (size_t)kMaxBlocks * kStatStride;
    double* result10 = shard_tuples + (size_t)kMaxShards * kStatStride;
    // the tail of the array still holds the previous call's shard tuples and results
    TW_CUDA(cudaMemsetAsync(shard_tuples, 0, sizeof(double) * ((size_t)kMaxShards * kStatStride + 16), ctx->stream));

    DenseCfg c;
    TW_TRY(dense_cfg(b->d, c));
    DenseIterArgs a{};
    a.X = b->X;
    a.w = m->d_w;
    a.stat_partials = block_tuples;
    a.plan = b->plan;
    a.ldx = b->ldx;
    a.d = (int)b->d;
    a.c2 = c.c2;
    a.c2p = c.c2p;
    a.rl = c.rl;
    a.nstages = c.nstages;
    a.stage_stride = c.stage_stride;
    a.frac = 1.0;
    if (c.cp == 1) TW_TRY(c.rl == 1 ? (launch_stats_pass<1, true>(ctx, c, a)) : (launch_stats_pass<1, false>(ctx, c, a)));
    else if (c.cp == 2) TW_TRY((launch_stats_pass<2, true>(ctx, c, a)));
    else TW_TRY((launch_stats_pass<4, true>(ctx, c, a)));
    k_stats_shard<<<b->plan.spr, 256, 0, ctx->stream>>>(block_tuples, b->plan, shard_tuples);
    k_stats_final<<<1, 32, 0, ctx->stream>>>(shard_tuples, b->plan.spr, result10);
    TW_CUDA(cudaGetLastError());
    double o[10];
    TW_CUDA(cudaMemcpyAsync(o, result10, sizeof(o), cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    out->batch = (int64_t)o[0];
    out->mse = o[1];
    out->real_stdev = o[2];
    out->pred_stdev = o[3];
    out->mse_raw = o[4];
    out->real_stdev_raw = o[5];
    out->pred_stdev_raw = o[6];
    out->real_mean = o[7];
    out->pred_mean = o[8];
    out->nonfinite = (int64_t)o[9];
    return 0;
}
