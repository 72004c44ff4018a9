// twtml_group.cu -- device groups: one handle, one host process, several GPUs (twtml_create_all).
//
// The reference is ONE JVM whose SparkContext spreads an RDD's partitions over local[N] worker threads
// (LinearRegression.scala:36).  A host that wants all the GPUs of a box behind the same three calls
// (model.trainOn, predict, latestModel) gets them through a group context: it owns one ranked context per
// device (the same row-sharded ranks twtml_create_ranked makes, here as threads of this process, exchanging
// their shard vectors through peer memory), cuts every micro-batch into contiguous row blocks -- rank r
// gets rows [n r / G, n (r + 1) / G), so concatenating the ranks' outputs keeps the row order -- and runs
// the members' calls side by side on one worker thread each (every compute call contains the exchange with
// the other members, so they cannot run one after the other).  Nothing here computes: it only splits
// arrays, forwards calls and puts outputs back together.  With n % 8 == 0 the result is bit-identical to
// the single-GPU one (same virtual shards, same order; DESIGN.md "Determinism").
#include <algorithm>
#include <cstring>
#include <memory>

#include "host.h"

namespace twtml {

struct MemberThreads {
    struct Slot {
        std::thread th;
        std::mutex mu;
        std::condition_variable cv;
        std::function<int()> fn;
        bool has = false, done = false, stop = false;
        int rc = 0;
        std::string err;
    };
    std::vector<std::unique_ptr<Slot>> slots;

    explicit MemberThreads(int n)
    {
        for (int i = 0; i < n; ++i) {
            slots.emplace_back(new Slot());
            Slot* s = slots.back().get();
            s->th = std::thread([s] {
                for (;;) {
                    std::function<int()> fn;
                    {
                        std::unique_lock<std::mutex> lk(s->mu);
                        s->cv.wait(lk, [&] { return s->has || s->stop; });
                        if (s->stop) return;
                        fn = std::move(s->fn);
                        s->has = false;
                    }
                    const int rc = fn();
                    {
                        std::lock_guard<std::mutex> lk(s->mu);
                        s->rc = rc;
                        s->err = rc ? last_error_text() : "";
                        s->done = true;
                    }
                    s->cv.notify_all();
                }
            });
        }
    }
    ~MemberThreads()
    {
        for (auto& s : slots) {
            {
                std::lock_guard<std::mutex> lk(s->mu);
                s->stop = true;
            }
            s->cv.notify_all();
            s->th.join();
        }
    }
    // fn(r) on member r's thread, all at once; the first failure (lowest rank) is reported on the caller's thread
    int run(const std::function<int(int)>& fn)
    {
        for (size_t r = 0; r < slots.size(); ++r) {
            Slot* s = slots[r].get();
            {
                std::lock_guard<std::mutex> lk(s->mu);
                s->fn = [fn, r] { return fn((int)r); };
                s->done = false;
                s->has = true;
            }
            s->cv.notify_all();
        }
        int rc = 0;
        std::string err;
        for (auto& sp : slots) {
            Slot* s = sp.get();
            std::unique_lock<std::mutex> lk(s->mu);
            s->cv.wait(lk, [&] { return s->done; });
            if (s->rc && !rc) { rc = s->rc; err = s->err; }
        }
        return rc ? fail(rc, "%s", err.c_str()) : 0;
    }
};

namespace grp {

static int G(const twtml_ctx* g) { return (int)g->members.size(); }
static int64_t lo_of(int64_t n, int r, int g) { return n * r / g; }

int create_all(const int32_t* devices, int32_t n, twtml_ctx** out)
{
    if (!out) return fail(TWTML_E_INVALID_ARG, "twtml_create_all: out is null");
    *out = nullptr;
    if (!devices || n < 1 || n > kMaxShards) return fail(TWTML_E_INVALID_ARG, "twtml_create_all: 1 to %d devices", kMaxShards);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return fail(TWTML_E_INVALID_ARG, "twtml_create_all: device %d listed twice", devices[i]);
    if (n == 1) return twtml_create(devices[0], out);
    uint8_t id[TWTML_NCCL_ID_BYTES];
    TW_TRY(twtml_nccl_unique_id(id, sizeof(id)));
    twtml_ctx* g = new twtml_ctx();
    g->members.assign((size_t)n, nullptr);
    g->threads = new MemberThreads(n);
    g->world = 1; // to its user the group is one context
    g->device = devices[0];
    std::vector<int> dev(devices, devices + n);
    // the members' communicator is built by all ranks at once (ncclCommInitRank blocks until every rank has joined)
    int rc = g->threads->run([&](int r) { return twtml_create_ranked(dev[r], r, n, id, sizeof(id), &g->members[r]); });
    if (rc) {
        std::string msg = last_error_text();
        for (twtml_ctx* m : g->members)
            if (m) twtml_destroy(m);
        delete g->threads;
        delete g;
        return fail(rc, "%s", msg.c_str());
    }
    *out = g;
    return 0;
}

int destroy(twtml_ctx* g)
{
    if (g->live_handles.load() != 0)
        return fail(TWTML_E_INVALID_ARG, "twtml_destroy: %d models/batches of this context are still alive; free them first", g->live_handles.load());
    int rc = g->threads->run([&](int r) { return twtml_destroy(g->members[r]); });
    delete g->threads;
    delete g;
    return rc;
}

int profile_enable(twtml_ctx* g, int32_t on)
{
    for (twtml_ctx* m : g->members) TW_TRY(twtml_profile_enable(m, on));
    return 0;
}

int profile_get(const twtml_ctx* g, twtml_profile* out)
{
    TW_TRY(twtml_profile_get(g->members[0], out));
    for (size_t r = 1; r < g->members.size(); ++r) { // the slowest member sets the pace
        twtml_profile p;
        TW_TRY(twtml_profile_get(g->members[r], &p));
        out->iter_kernel_ms = std::max(out->iter_kernel_ms, p.iter_kernel_ms);
        out->loop_ms = std::max(out->loop_ms, p.loop_ms);
    }
    return 0;
}

// ---- model ------------------------------------------------------------------------------------
int model_create(twtml_ctx* g, int64_t d, twtml_model** out)
{
    twtml_model* m = new twtml_model();
    m->ctx = g;
    m->d = d;
    m->members.assign(g->members.size(), nullptr);
    for (size_t r = 0; r < g->members.size(); ++r) {
        int rc = twtml_model_create(g->members[r], d, &m->members[r]);
        if (rc) {
            for (twtml_model* mm : m->members) twtml_model_free(mm);
            delete m;
            return rc;
        }
    }
    g->live_handles += 1;
    *out = m;
    return 0;
}

int model_free(twtml_model* m)
{
    for (twtml_model* mm : m->members) twtml_model_free(mm);
    m->ctx->live_handles -= 1;
    delete m;
    return 0;
}

int set_initial_weights(twtml_model* m, const double* w, int64_t d)
{
    for (twtml_model* mm : m->members) TW_TRY(twtml_set_initial_weights(mm, w, d));
    m->initialized = true;
    return 0;
}

int get_weights(const twtml_model* m, double* out, int64_t d) { return twtml_get_weights(m->members[0], out, d); }

// ---- batches ----------------------------------------------------------------------------------
static twtml_batch* shell(twtml_ctx* g, bool csr, int64_t d)
{
    twtml_batch* b = new twtml_batch();
    b->ctx = g;
    b->csr = csr;
    b->d = d;
    b->members.assign(g->members.size(), nullptr);
    g->live_handles += 1;
    return b;
}

static void finish(twtml_batch* b)
{
    b->n = b->nnz = 0;
    for (twtml_batch* mb : b->members) { b->n += mb->n; }
    b->n_global = b->n;
}

int batch_free(twtml_batch* b)
{
    for (twtml_batch* mb : b->members) twtml_batch_free(mb);
    b->ctx->live_handles -= 1;
    delete b;
    return 0;
}

static int fail_batch(twtml_batch* b, int rc)
{
    std::string msg = last_error_text();
    batch_free(b);
    return fail(rc, "%s", msg.c_str());
}

int from_dense(twtml_ctx* g, int64_t n, int64_t d, const double* X, int64_t ld_x, const double* labels, bool async, twtml_batch** out)
{
    twtml_batch* b = shell(g, false, d);
    const int W = G(g);
    auto one = [&](int r) {
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W);
        const double* Xr = X ? X + lo * ld_x : nullptr;
        const double* yr = labels ? labels + lo : nullptr;
        return async ? twtml_batch_from_dense_async(g->members[r], hi - lo, d, Xr, ld_x, yr, &b->members[r])
                     : twtml_batch_from_dense(g->members[r], hi - lo, d, Xr, ld_x, yr, &b->members[r]);
    };
    // side by side: the blocking form (and the asynchronous one, when it has to fall back to it for pageable memory)
    // exchanges the members' row counts inside the call
    int rc = g->threads->run(one);
    if (rc) return fail_batch(b, rc);
    finish(b);
    *out = b;
    return 0;
}

int dense_begin(twtml_ctx* g, int64_t n, int64_t d, twtml_batch** out)
{
    twtml_batch* b = shell(g, false, d);
    const int W = G(g);
    for (int r = 0; r < W; ++r) {
        int rc = twtml_batch_dense_begin(g->members[r], lo_of(n, r + 1, W) - lo_of(n, r, W), d, &b->members[r]);
        if (rc) return fail_batch(b, rc);
    }
    finish(b);
    b->rows_appended = 0;
    *out = b;
    return 0;
}

int dense_append(twtml_batch* b, int64_t nrows, const double* X, int64_t ld_x, const double* labels)
{
    if (b->rows_appended + nrows > b->n)
        return fail(TWTML_E_INVALID_ARG, "twtml_batch_dense_append: %lld rows after %lld exceed the batch's %lld", (long long)nrows, (long long)b->rows_appended, (long long)b->n);
    // the appended rows [a, a + nrows) of the whole batch go to the members whose blocks they fall into
    const int64_t a = b->rows_appended;
    int64_t first = 0;
    for (twtml_batch* mb : b->members) {
        const int64_t lo = std::max(a, first), hi = std::min(a + nrows, first + mb->n);
        if (hi > lo) TW_TRY(twtml_batch_dense_append(mb, hi - lo, X + (lo - a) * ld_x, ld_x, labels + (lo - a)));
        first += mb->n;
    }
    b->rows_appended += nrows;
    return 0;
}

int from_csr(twtml_ctx* g, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals, const double* labels, bool async,
             twtml_batch** out)
{
    twtml_batch* b = shell(g, true, d);
    const int W = G(g);
    b->owned_i64.resize((size_t)W);
    for (int r = 0; r < W; ++r) { // each member reads its rows through row pointers that start at 0
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W);
        std::vector<int64_t>& rp = b->owned_i64[(size_t)r];
        rp.resize((size_t)(hi - lo + 1));
        for (int64_t i = lo; i <= hi; ++i) rp[(size_t)(i - lo)] = rowptr[i] - rowptr[lo];
    }
    auto one = [&](int r) {
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W), e0 = rowptr[lo];
        const int64_t* rp = b->owned_i64[(size_t)r].data();
        const int32_t* ci = colidx ? colidx + e0 : nullptr;
        const double* va = vals ? vals + e0 : nullptr;
        const double* la = labels ? labels + lo : nullptr;
        return async ? twtml_batch_from_csr_async(g->members[r], hi - lo, d, rp, ci, va, la, &b->members[r])
                     : twtml_batch_from_csr(g->members[r], hi - lo, d, rp, ci, va, la, &b->members[r]);
    };
    int rc = 0;
    if (async) {
        for (int r = 0; r < W && !rc; ++r) rc = one(r);
    } else {
        rc = g->threads->run(one);
    }
    if (rc) return fail_batch(b, rc);
    finish(b);
    *out = b;
    return 0;
}

int from_tweets(twtml_ctx* g, const TweetCols& t, int64_t n, int64_t now_ms, int32_t rt_begin, int32_t rt_end, int32_t F, uint32_t flags, bool async,
                twtml_batch** out)
{
    twtml_batch* b = shell(g, true, (int64_t)F + 4);
    const int W = G(g);
    b->owned_i64.resize((size_t)W);
    for (int r = 0; r < W; ++r) {
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W);
        std::vector<int64_t>& off = b->owned_i64[(size_t)r];
        off.resize((size_t)(hi - lo + 1));
        for (int64_t i = lo; i <= hi; ++i) off[(size_t)(i - lo)] = n > 0 ? t.text_off[i] - t.text_off[lo] : 0;
    }
    auto one = [&](int r) {
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W);
        const uint16_t* text = (n > 0 && t.text_utf16) ? t.text_utf16 + t.text_off[lo] : t.text_utf16;
        auto fn = async ? twtml_batch_from_tweets_async : twtml_batch_from_tweets;
        auto at = [&](auto* p) { return p ? p + lo : p; };
        return fn(g->members[r], text, b->owned_i64[(size_t)r].data(), at(t.retweet_cnt), at(t.followers), at(t.favourites), at(t.friends),
                  at(t.created_ms), at(t.is_retweet), hi - lo, now_ms, rt_begin, rt_end, F, flags, &b->members[r]);
    };
    int rc = 0;
    if (async) {
        for (int r = 0; r < W && !rc; ++r) rc = one(r);
    } else {
        rc = g->threads->run(one);
    }
    if (rc) return fail_batch(b, rc);
    finish(b);
    *out = b;
    return 0;
}

int synth_dense(twtml_ctx* g, int64_t n, int64_t d, uint64_t seed, int64_t row0, twtml_batch** out)
{
    twtml_batch* b = shell(g, false, d);
    const int W = G(g);
    int rc = g->threads->run([&](int r) {
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W);
        return twtml_batch_synth_dense(g->members[r], hi - lo, d, seed, row0 + lo, &b->members[r]);
    });
    if (rc) return fail_batch(b, rc);
    finish(b);
    *out = b;
    return 0;
}

int synth_csr(twtml_ctx* g, int64_t n, int32_t d_text, uint64_t seed, int64_t row0, twtml_batch** out)
{
    twtml_batch* b = shell(g, true, (int64_t)d_text + 4);
    const int W = G(g);
    int rc = g->threads->run([&](int r) {
        const int64_t lo = lo_of(n, r, W), hi = lo_of(n, r + 1, W);
        return twtml_batch_synth_csr(g->members[r], hi - lo, d_text, seed, row0 + lo, &b->members[r]);
    });
    if (rc) return fail_batch(b, rc);
    finish(b);
    *out = b;
    return 0;
}

int batch_ready(const twtml_batch* b, int32_t wait, int32_t* ready)
{
    int all = 1;
    for (twtml_batch* mb : b->members) {
        int32_t r = 1;
        TW_TRY(twtml_batch_ready(mb, wait, &r));
        all &= r;
    }
    if (ready) *ready = all;
    return 0;
}

int batch_info(const twtml_batch* b, int64_t* n, int64_t* d, int64_t* nnz, int32_t* is_csr)
{
    int64_t tn = 0, tz = 0;
    for (twtml_batch* mb : b->members) {
        int64_t a = 0, z = 0;
        TW_TRY(twtml_batch_info(mb, &a, nullptr, nnz ? &z : nullptr, nullptr));
        tn += a;
        tz += z;
    }
    if (n) *n = tn;
    if (d) *d = b->d;
    if (nnz) *nnz = tz;
    if (is_csr) *is_csr = b->csr ? 1 : 0;
    return 0;
}

int export_dense(const twtml_batch* b, double* X, double* labels)
{
    int64_t first = 0;
    for (twtml_batch* mb : b->members) {
        if (mb->n > 0) TW_TRY(twtml_batch_export_dense(mb, X + first * b->d, labels + first));
        first += mb->n;
    }
    return 0;
}

int export_csr(const twtml_batch* b, int64_t* rowptr, int32_t* colidx, double* vals, double* labels)
{
    int64_t first = 0, e0 = 0;
    rowptr[0] = 0;
    for (twtml_batch* mb : b->members) {
        int64_t n = 0, z = 0;
        TW_TRY(twtml_batch_info(mb, &n, nullptr, &z, nullptr));
        std::vector<int64_t> rp((size_t)n + 1);
        TW_TRY(twtml_batch_export_csr(mb, rp.data(), colidx ? colidx + e0 : nullptr, vals ? vals + e0 : nullptr, labels ? labels + first : nullptr));
        for (int64_t i = 0; i <= n; ++i) rowptr[first + i] = rp[(size_t)i] + e0;
        first += n;
        e0 += z;
    }
    return 0;
}

// ---- compute ----------------------------------------------------------------------------------
static std::vector<int64_t> row_offsets(const twtml_batch* b)
{
    std::vector<int64_t> off;
    int64_t first = 0;
    for (twtml_batch* mb : b->members) { off.push_back(first); first += mb->n; }
    return off;
}

int predict(const twtml_model* m, const twtml_batch* b, double* out_pred)
{
    const std::vector<int64_t> off = row_offsets(b);
    return m->ctx->threads->run([&](int r) { return twtml_predict(m->members[r], b->members[r], out_pred ? out_pred + off[r] : nullptr); });
}

int predict_stats(const twtml_model* m, const twtml_batch* b, double* out_real, double* out_pred, twtml_stats* out)
{
    const std::vector<int64_t> off = row_offsets(b);
    std::vector<twtml_stats> st(b->members.size());
    // allocations first, on every member, before any member's kernels start waiting for the others
    TW_TRY(m->ctx->threads->run([&](int r) { return prepare_compute(m->members[r], b->members[r], out_real || out_pred); }));
    TW_TRY(m->ctx->threads->run([&](int r) {
        return twtml_predict_stats(m->members[r], b->members[r], out_real ? out_real + off[r] : nullptr, out_pred ? out_pred + off[r] : nullptr, &st[r]);
    }));
    *out = st[0]; // every member holds the statistics of the whole batch
    return 0;
}

int train(twtml_model* m, const twtml_batch* b, bool with_stats, double* out_real, double* out_pred, twtml_stats* out, double* loss_history,
          int32_t cap, int32_t* iters_run)
{
    const std::vector<int64_t> off = row_offsets(b);
    const size_t W = b->members.size();
    std::vector<twtml_stats> st(W);
    std::vector<int32_t> it(W, 0);
    std::vector<std::vector<double>> hist(W, std::vector<double>((size_t)std::max(cap, 1)));
    TW_TRY(m->ctx->threads->run([&](int r) { return prepare_compute(m->members[r], b->members[r], with_stats && (out_real || out_pred)); }));
    TW_TRY(m->ctx->threads->run([&](int r) {
        if (with_stats)
            return twtml_step(m->members[r], b->members[r], out_real ? out_real + off[r] : nullptr, out_pred ? out_pred + off[r] : nullptr, &st[r],
                              hist[r].data(), cap, &it[r]);
        return twtml_train_on_batch(m->members[r], b->members[r], hist[r].data(), cap, &it[r]);
    }));
    if (with_stats && out) *out = st[0];
    if (iters_run) *iters_run = it[0];
    if (loss_history)
        for (int i = 0; i < it[0] && i < cap; ++i) loss_history[i] = hist[0][(size_t)i];
    return 0;
}

} // namespace grp
} // namespace twtml
