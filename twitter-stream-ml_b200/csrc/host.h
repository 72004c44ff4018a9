// host.h -- host-side structures behind the opaque handles of include/twtml.h.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h> // types only: libnccl is dlopen'ed at run time (nccl_dyn.h), never linked

#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <unordered_map>
#include <future>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/twtml.h"
#include "common.cuh"

namespace twtml {

void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define TW_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t e__ = (expr);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return fail(e__ == cudaErrorMemoryAllocation ? TWTML_E_OOM : TWTML_E_CUDA, "%s: %s (%s:%d)", \
                        #expr, cudaGetErrorString(e__), __FILE__, __LINE__);                            \
    } while (0)

#define TW_TRY(expr)            \
    do {                        \
        int rc__ = (expr);      \
        if (rc__ != 0) return rc__; \
    } while (0)

// a cudaMalloc'ed buffer that only ever grows
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);
    void release();
    template <class T> T* as() const { return static_cast<T*>(p); }
};

// Everything one ingest needs besides the batch it fills: a stream and the scratch that goes with it.  A context has
// two lanes so that the ingest of micro-batch t+1 never shares a buffer with the compute (or the ingest) of batch t:
// lane 0 works on the compute stream for the blocking twtml_batch_from_* calls, lane 1 on the copy stream for the
// *_async ones.
struct IngestLane {
    std::mutex mu;                             // held while work is being enqueued on the lane (caller thread or the ingest worker)
    cudaStream_t stream = nullptr;
    DevBuf scan_tmp, sort_tmp;                 // CUB temporaries (prefix sums, the transpose's radix sort)
    DevBuf feat_arena;                         // the featuriser's scratch arrays (tweet columns, terms, scans): kept from interval to interval
    DevBuf labels_tmp;                         // labels on their way into a dense batch's label column
    DevBuf stage;                              // flat landing zone of page-locked dense rows (two halves)
    cudaEvent_t stage_ev[2][2] = {};           // [half][0: chunk copied, 1: chunk spread]
    void* pin[2] = {nullptr, nullptr};         // pinned staging ring for pageable caller memory
    size_t pin_bytes = 0;
    cudaEvent_t pin_ev[2] = {nullptr, nullptr};
};

// One worker thread per member context of a device group (twtml_create_all): the members' calls of one operation run
// side by side, as they must -- each contains the exchange with the other members.
struct MemberThreads;

} // namespace twtml

struct twtml_ctx {
    // a device group (twtml_create_all): this handle only fans calls out to its members, one ranked context per GPU
    std::vector<twtml_ctx*> members;
    twtml::MemberThreads* threads = nullptr;
    int device = 0, rank = 0, world = 1, sm_count = 148;
    int spr = twtml::kMaxShards;  // shards per rank
    int nshards = twtml::kMaxShards;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr; // asynchronous ingest (twtml_batch_from_dense_async)
    cudaMemPool_t pool = nullptr;       // batch memory: stream-ordered, kept across batches (no OS round trip, no device-wide sync)
    // The arrays of CSR batches: blocks kept by this context and handed out again only once nothing on the device can still
    // touch them (batch_free_locked returns them after the compute stream and the batch's own ingest are done), so a block
    // carries no stream ordering at all.  The ingest lane allocates these arrays and the compute side frees them; through
    // the stream-ordered pool that cross-stream reuse made the lane's allocation wait for whatever the compute stream
    // had queued -- the training call beside it (twtml.cu block_alloc).
    std::mutex block_mu;
    std::multimap<size_t, void*> block_free;          // size -> idle block
    std::unordered_map<void*, size_t> block_live;     // block in use -> size
    ncclComm_t comm = nullptr;
    std::mutex mu;
    std::atomic<int> live_handles{0}; // models + batches created on this context and not yet freed
    // peer-memory exchange of the shard vectors (reduce.cuh PeerExchange); falls back to NCCL when not set up
    bool p2p_ok = false;
    bool peers_mappable = false;            // every rank can map every rank's memory (the dense scheme additionally needs world | 8)
    // wide-model exchange buffer (reduce.cuh WideExchange), set up at the first training call that needs it
    double* wxbuf = nullptr;
    long long wx_dpad = 0;
    double* wx_peer[twtml::kMaxShards] = {};
    bool wx_ipc[twtml::kMaxShards] = {};
    bool wx_failed = false;
    double* xbuf = nullptr;                 // this rank's exchange buffer
    double* peer[twtml::kMaxShards] = {};   // every rank's buffer as mapped here (peer[rank] == xbuf)
    bool peer_ipc[twtml::kMaxShards] = {};  // mapped with cudaIpcOpenMemHandle (to be closed)
    int* host_done = nullptr;     // mapped pinned word: the update kernel sets it when a training loop has ended
    int* host_done_dev = nullptr; // ... its device alias
    unsigned smem_attr_done = 0; // dense iteration-kernel forms whose dynamic shared-memory limit is raised
    unsigned narrow_attr_done = 0; // ... and the single-pass CSR kernel's forms
    // scratch, reused across calls
    twtml::DevBuf partials, shards, stat_partials, stat_shards, state, loss_hist, out10, tmp_a, tmp_b, gather, lower, upd_parts, slices, wprev;
    twtml::IngestLane lane[2];                 // [0] blocking ingest on the compute stream, [1] asynchronous ingest on the copy stream
    cudaStream_t spread_stream = nullptr;      // k_repitch runs here, beside the next chunk's copy
    // asynchronous CSR / tweet ingest: jobs run in order on one worker thread per context (started on first use), so the
    // host-side waits an ingest needs (sizes that only the device knows) never block the thread that drives the compute
    std::thread worker;
    std::mutex qmu;
    std::condition_variable qcv;
    std::deque<std::function<void()>> jobs;
    bool worker_stop = false;
    // profiling
    bool profiling = false;
    std::vector<cudaEvent_t> ev;
    twtml_profile prof{};
};

struct twtml_model {
    std::vector<twtml_model*> members; // group model: one replica per member context (they stay bit-identical)
    twtml_ctx* ctx = nullptr;
    int64_t d = 0;
    double step = 0.1, frac = 1.0, tol = 0.001, reg = 0.0;
    int num_iter = 50, updater = 0;
    bool initialized = false;
    double* d_w = nullptr; // device weights, d rounded up
    std::vector<double> h_w;
};

struct twtml_batch {
    std::vector<twtml_batch*> members;            // group batch: contiguous row blocks, one per member context
    std::vector<std::vector<int64_t>> owned_i64;  // ... and the re-based row pointers / text offsets the members read
    twtml_ctx* ctx = nullptr;
    bool csr = false;
    int64_t n = 0, d = 0, nnz = 0;
    int64_t n_global = 0, row0_global = 0;
    cudaEvent_t ready = nullptr; // set by asynchronous ingest: recorded on the copy stream when the rows are in HBM
    // asynchronous ingest through the context's worker thread (CSR, tweets): the job's verdict; fields below n / d are
    // the worker's until it is in
    std::shared_future<int> pending;
    bool has_pending = false;
    std::string pending_err;
    bool global_deferred = false; // multi-rank: n_global / row0_global still to be exchanged (done at first use, on the compute stream)
    int64_t rows_appended = 0;    // twtml_batch_dense_begin / _append: rows delivered so far
    // dense
    double* X = nullptr;
    int64_t ldx = 0;
    // csr (+ its transpose for the column pass)
    long long* rowptr = nullptr;
    int* colidx = nullptr;
    double* vals = nullptr;
    double* labels = nullptr;
    long long* colptr = nullptr;
    int* rowidx = nullptr;
    double* cvals = nullptr;
    unsigned* headbits = nullptr;  // window index of the transpose (csr.cuh CscView)
    int* nzcol = nullptr;
    long long* win_first_nz = nullptr;
    int* wkey = nullptr;
    double* win_first_part = nullptr;
    double* win_last_part = nullptr;
    long long nwin = 0;
    // narrow feature spaces (csr_narrow.cuh): no transpose, only the first row of every warp of a fixed grid
    bool narrow = false;
    int nwarps = 0;
    long long* bounds = nullptr;
    int* pcolidx = nullptr;        // the entries in the order the single-pass kernel streams them (k_narrow_permute), or null
    double* pvals = nullptr;
    double* diff = nullptr;        // residuals of the current iteration, n
    twtml::Plan plan{};
};

namespace twtml {
// shared between twtml.cu and twtml_sparse.cu
int upload_plain(twtml_ctx* ctx, IngestLane& lane, void* dst, const void* src, size_t bytes);
const char* last_error_text();
int submit_ingest(twtml_ctx* ctx, twtml_batch* b, std::function<int()> job);
int settle_batch(twtml_ctx* ctx, const twtml_batch* b);
int ctx_global_rows(twtml_ctx* ctx, int64_t n_local, int64_t* n_global, int64_t* row0);
void make_plan(int64_t n_local, int spr, int R, Plan& p, int64_t min_blocks = 64);
void batch_free_locked(twtml_batch* b);
int prepare_compute(twtml_model* m, const twtml_batch* b, bool per_row_outputs);
// batch memory from the context's pool, ordered on the compute stream
int pool_alloc(twtml_ctx* ctx, void** p, size_t bytes, cudaStream_t stream = nullptr);
void pool_free(twtml_ctx* ctx, void* p, cudaStream_t stream = nullptr);
int block_alloc(twtml_ctx* ctx, void** p, size_t bytes);   // for memory that is freed with block_release only
void block_release(twtml_ctx* ctx, void* p);               // p must be idle on the device; pool memory is passed on to pool_free
// launch shape of the dense iteration kernel for a feature width
struct DenseCfg {
    int cp, c2, c2p, rl, R, nstages, grid_per_sm;
    unsigned stage_stride;
    size_t smem;
};
int dense_cfg(int64_t d, DenseCfg& c);
// device groups (twtml_group.cu)
namespace grp {
int create_all(const int32_t* devices, int32_t n, twtml_ctx** out);
int destroy(twtml_ctx* g);
int profile_enable(twtml_ctx* g, int32_t on);
int profile_get(const twtml_ctx* g, twtml_profile* out);
int model_create(twtml_ctx* g, int64_t d, twtml_model** out);
int model_free(twtml_model* m);
int set_initial_weights(twtml_model* m, const double* w, int64_t d);
int get_weights(const twtml_model* m, double* out, int64_t d);
int from_dense(twtml_ctx* g, int64_t n, int64_t d, const double* X, int64_t ld_x, const double* labels, bool async, twtml_batch** out);
int dense_begin(twtml_ctx* g, int64_t n, int64_t d, twtml_batch** out);
int dense_append(twtml_batch* b, int64_t nrows, const double* X, int64_t ld_x, const double* labels);
int from_csr(twtml_ctx* g, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals, const double* labels,
             bool async, twtml_batch** out);
struct TweetCols {
    const uint16_t* text_utf16;
    const int64_t* text_off;
    const int32_t *retweet_cnt, *followers, *favourites, *friends;
    const int64_t* created_ms;
    const uint8_t* is_retweet;
};
int from_tweets(twtml_ctx* g, const TweetCols& t, int64_t n, int64_t now_ms, int32_t rt_begin, int32_t rt_end, int32_t F, uint32_t flags,
                bool async, twtml_batch** out);
int synth_dense(twtml_ctx* g, int64_t n, int64_t d, uint64_t seed, int64_t row0, twtml_batch** out);
int synth_csr(twtml_ctx* g, int64_t n, int32_t d_text, uint64_t seed, int64_t row0, twtml_batch** out);
int batch_ready(const twtml_batch* b, int32_t wait, int32_t* ready);
int batch_info(const twtml_batch* b, int64_t* n, int64_t* d, int64_t* nnz, int32_t* is_csr);
int export_csr(const twtml_batch* b, int64_t* rowptr, int32_t* colidx, double* vals, double* labels);
int export_dense(const twtml_batch* b, double* X, double* labels);
int batch_free(twtml_batch* b);
int predict(const twtml_model* m, const twtml_batch* b, double* out_pred);
int predict_stats(const twtml_model* m, const twtml_batch* b, double* out_real, double* out_pred, twtml_stats* out);
int train(twtml_model* m, const twtml_batch* b, bool with_stats, double* out_real, double* out_pred, twtml_stats* out, double* loss_history,
          int32_t cap, int32_t* iters_run);
} // namespace grp
struct Guard {
    std::lock_guard<std::mutex> lk;
    explicit Guard(twtml_ctx* c) : lk(c->mu) { cudaSetDevice(c->device); }
};
} // namespace twtml
