// twtml.hpp -- header-only C++ host facade over the C ABI of twtml.h, named after the Scala API the
// reference drives (LinearRegression.scala:28-32,53-66,86): StreamingLinearRegressionWithSGD with
// setStepSize / setNumIterations / setMiniBatchFraction / setInitialWeights / setConvergenceTol,
// latestModel(), trainOn, predictOnValues; LinearRegressionModel.  The reference's host language is
// Scala (JVM, absent from the build image; the JNA/Scala twin is in INTEGRATION.md); this is the
// compiled-language host side that is built and tested here (tests/cpp/driver_loop.cpp).
// It only marshals: every number comes from libtwtml.so.
#pragma once
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "twtml.h"

namespace twtml {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& m) : std::runtime_error("libtwtml " + std::to_string(c) + ": " + m), code(c) {}
};
// what MLlib throws from StreamingLinearAlgorithm when the model has no initial weights [U]
struct IllegalArgumentException : std::invalid_argument {
    using std::invalid_argument::invalid_argument;
};

inline void check(int rc)
{
    if (rc == 0) return;
    char buf[512];
    twtml_last_error(buf, sizeof(buf));
    if (rc == TWTML_E_NOT_INITIALIZED) throw IllegalArgumentException(buf);
    throw Error(rc, buf);
}

class Context {
public:
    explicit Context(int device = 0) { check(twtml_create(device, &h_)); }
    Context(int device, int rank, int world, const uint8_t* nccl_id) { check(twtml_create_ranked(device, rank, world, nccl_id, TWTML_NCCL_ID_BYTES, &h_)); }
    // one handle over several GPUs of this process (twtml_create_all): the single-JVM shape of the reference
    explicit Context(const std::vector<int>& devices)
    {
        std::vector<int32_t> d(devices.begin(), devices.end());
        check(twtml_create_all(d.data(), (int32_t)d.size(), &h_));
    }
    int deviceCount() const { int32_t n = 0; check(twtml_device_count(h_, &n)); return n; }
    ~Context() { twtml_destroy(h_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    twtml_ctx* handle() const { return h_; }

private:
    twtml_ctx* h_ = nullptr;
};

// One micro-batch (the RDD[LabeledPoint] of one interval) resident on the GPU.
class Batch {
public:
    static Batch fromDense(Context& c, const double* X, int64_t ld, const double* labels, int64_t n, int64_t d)
    {
        Batch b;
        check(twtml_batch_from_dense(c.handle(), n, d, X, ld, labels, &b.h_));
        return b;
    }
    static Batch fromCsr(Context& c, int64_t n, int64_t d, const int64_t* rowptr, const int32_t* colidx, const double* vals, const double* labels)
    {
        Batch b;
        check(twtml_batch_from_csr(c.handle(), n, d, rowptr, colidx, vals, labels, &b.h_));
        return b;
    }
    Batch(Batch&& o) noexcept : h_(o.h_) { o.h_ = nullptr; }
    Batch& operator=(Batch&& o) noexcept { std::swap(h_, o.h_); return *this; }
    ~Batch() { twtml_batch_free(h_); }
    int64_t count() const { int64_t n = 0; check(twtml_batch_rows(h_, &n)); return n; }
    bool isEmpty() const { return count() == 0; }
    twtml_batch* handle() const { return h_; }

private:
    Batch() = default;
    twtml_batch* h_ = nullptr;
};

struct LinearRegressionModel {
    std::vector<double> weights;
    double intercept = 0.0;
};

class StreamingLinearRegressionWithSGD {
public:
    explicit StreamingLinearRegressionWithSGD(Context& c, double stepSize = 0.1, int numIterations = 50, double miniBatchFraction = 1.0)
        : ctx_(c), step_(stepSize), iters_(numIterations), frac_(miniBatchFraction) {}
    ~StreamingLinearRegressionWithSGD() { twtml_model_free(m_); }
    StreamingLinearRegressionWithSGD(const StreamingLinearRegressionWithSGD&) = delete;

    StreamingLinearRegressionWithSGD& setStepSize(double v) { step_ = v; return push(); }
    StreamingLinearRegressionWithSGD& setNumIterations(int v) { iters_ = v; return push(); }
    StreamingLinearRegressionWithSGD& setMiniBatchFraction(double v) { frac_ = v; return push(); }
    StreamingLinearRegressionWithSGD& setConvergenceTol(double v) { tol_ = v; return push(); }
    StreamingLinearRegressionWithSGD& setInitialWeights(const std::vector<double>& w)
    {
        twtml_model_free(m_);
        m_ = nullptr;
        check(twtml_model_create(ctx_.handle(), (int64_t)w.size(), &m_));
        check(twtml_set_initial_weights(m_, w.data(), (int64_t)w.size()));
        d_ = (int64_t)w.size();
        return push();
    }
    LinearRegressionModel latestModel() const
    {
        require("prediction");
        LinearRegressionModel lm;
        lm.weights.resize((size_t)d_);
        check(twtml_get_weights(m_, lm.weights.data(), d_));
        return lm;
    }
    // StreamingLinearAlgorithm.trainOn for one interval: a non-empty batch replaces the weights
    void trainOn(const Batch& rdd)
    {
        require("training");
        lossHistory.assign((size_t)(iters_ > 0 ? iters_ : 1), 0.0);
        int32_t it = 0;
        check(twtml_train_on_batch(m_, rdd.handle(), lossHistory.data(), (int32_t)lossHistory.size(), &it));
        lossHistory.resize((size_t)it);
    }
    // predictOnValues: w.x per row, row order kept
    std::vector<double> predictOnValues(const Batch& rdd) const
    {
        require("prediction");
        std::vector<double> p((size_t)rdd.count());
        check(twtml_predict(m_, rdd.handle(), p.data()));
        return p;
    }
    // the stats output op of LinearRegression.scala:53-66
    twtml_stats stats(const Batch& rdd, std::vector<double>* real = nullptr, std::vector<double>* pred = nullptr) const
    {
        require("prediction");
        const size_t n = (size_t)rdd.count();
        if (real) real->assign(n, 0.0);
        if (pred) pred->assign(n, 0.0);
        twtml_stats s{};
        check(twtml_predict_stats(m_, rdd.handle(), real ? real->data() : nullptr, pred ? pred->data() : nullptr, &s));
        return s;
    }
    std::vector<double> lossHistory; // stochasticLossHistory of the last trainOn

private:
    void require(const char* what) const
    {
        if (!m_) throw IllegalArgumentException(std::string("Model must be initialized before starting ") + what + ".");
    }
    StreamingLinearRegressionWithSGD& push()
    {
        if (m_) {
            check(twtml_set_step_size(m_, step_));
            check(twtml_set_num_iterations(m_, iters_));
            check(twtml_set_mini_batch_fraction(m_, frac_));
            check(twtml_set_convergence_tol(m_, tol_));
        }
        return *this;
    }
    Context& ctx_;
    twtml_model* m_ = nullptr;
    int64_t d_ = 0;
    double step_;
    int iters_;
    double frac_, tol_ = 0.001;
};

} // namespace twtml
