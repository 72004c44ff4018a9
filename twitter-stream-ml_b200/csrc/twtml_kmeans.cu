// twtml_kmeans.cu -- the reference's second app (KMeans.scala): one batch interval of
//     val scaledData = new StandardScaler(false, true).fit(rdd).transform(rdd)            (:96)
//     model.latestModel.update(scaledData, model.decayFactor, model.timeUnit)             (:98)
//     val pred = model.latestModel.predict(scaledData).toArray                            (:105)
// with StreamingKMeans(k = 3, halfLife 5 batches, random centres of dimension 2, :69-73) on (retweets, followers).
// MLlib 1.6.1 semantics [U]: StandardScaler / MultivariateOnlineSummarizer (unbiased column variance), KMeans.findClosest
// with MLUtils.fastSquaredDistance, StreamingKMeansModel.update (discount, lambda rule, split of a dying cluster).
// The data are tiny next to the regression path (2 columns); what matters here is that every number the app would read
// comes from the GPU in a fixed summation order: per-warp accumulators in shared memory, blocks folded in block order.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "host.h"

struct twtml_kmeans {
    twtml_ctx* ctx = nullptr;
    int k = 0, dim = 0, time_unit = 0; // 0 "batches", 1 "points"
    double decay = 1.0;
    bool initialized = false;
    double* d_centers = nullptr; // k x dim, then k weights, then dim std, then dim 1/std
    twtml::DevBuf x, scaled, assign, moments, sums;
};

namespace twtml {

constexpr int kKmMaxK = 16, kKmMaxDim = 16; // the reference uses k = 3, dim = 2
constexpr int kKmThreads = 256, kKmWarps = kKmThreads / 32;

// ---- column moments over ALL values (zeros included): (n, mean, M2) per column, Chan merges in a fixed order.
// MultivariateOnlineSummarizer [U] runs Welford over the non-zero values and corrects for the zeros afterwards; the
// quantity is the same unbiased variance, so the two agree to rounding.
__device__ __forceinline__ void chan(double& n1, double& mu1, double& m1, double n2, double mu2, double m2)
{
    if (n2 == 0.0) return;
    if (n1 == 0.0) { n1 = n2; mu1 = mu2; m1 = m2; return; }
    const double n = n1 + n2, delta = mu2 - mu1;
    mu1 = mu1 + delta * n2 / n;
    m1 = m1 + m2 + delta * delta * n1 * n2 / n;
    n1 = n;
}

static __global__ void __launch_bounds__(kKmThreads) k_km_moments(const double* __restrict__ X, long long n, int dim, double* __restrict__ part)
{
    __shared__ double sn[kKmThreads], smu[kKmThreads], sm2[kKmThreads];
    const long long lo = n * blockIdx.x / gridDim.x, hi = n * (blockIdx.x + 1) / gridDim.x;
    for (int j = 0; j < dim; ++j) {
        double cn = 0.0, mu = 0.0, m2 = 0.0;
        for (long long r = lo + threadIdx.x; r < hi; r += kKmThreads) { // Welford, rows of this thread in order
            const double v = X[r * dim + j];
            cn += 1.0;
            const double dl = v - mu;
            mu += dl / cn;
            m2 += dl * (v - mu);
        }
        sn[threadIdx.x] = cn; smu[threadIdx.x] = mu; sm2[threadIdx.x] = m2;
        __syncthreads();
        for (int s = kKmThreads / 2; s >= 1; s >>= 1) { // fixed tree
            if (threadIdx.x < s) chan(sn[threadIdx.x], smu[threadIdx.x], sm2[threadIdx.x], sn[threadIdx.x + s], smu[threadIdx.x + s], sm2[threadIdx.x + s]);
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            double* o = part + ((long long)blockIdx.x * dim + j) * 3;
            o[0] = sn[0]; o[1] = smu[0]; o[2] = sm2[0];
        }
        __syncthreads();
    }
}

// blocks in block order -> std_j = sqrt(M2 / (n - 1)) (0 when n <= 1), inv_j = 1 / std_j or 0  [U] StandardScalerModel.transform
static __global__ void k_km_std(const double* __restrict__ part, int nblocks, int dim, double* __restrict__ std_out, double* __restrict__ inv_out)
{
    const int j = threadIdx.x;
    if (j >= dim) return;
    double cn = 0.0, mu = 0.0, m2 = 0.0;
    for (int b = 0; b < nblocks; ++b) chan(cn, mu, m2, part[((long long)b * dim + j) * 3], part[((long long)b * dim + j) * 3 + 1], part[((long long)b * dim + j) * 3 + 2]);
    double var = cn - 1.0 > 0.0 ? m2 / (cn - 1.0) : 0.0;
    if (var < 0.0) var = 0.0;
    const double sd = sqrt(var);
    std_out[j] = sd;
    inv_out[j] = sd != 0.0 ? 1.0 / sd : 0.0;
}

// ---- assignment (+ optional scaling, + optional per-cluster sums) ---------------------------------------------------
// KMeans.findClosest [U]: centres in index order, examined only if (||c|| - ||x||)^2 < best, replace only when strictly
// closer; MLUtils.fastSquaredDistance [U]: ||a||^2 + ||b||^2 - 2 a.b when its error bound is below 1e-6, else sum (a-b)^2.
// Unfused arithmetic, like the JVM.
struct KmArgs {
    const double* X;       // n x dim, packed
    const double* centers; // k x dim
    const double* inv_std; // dim, or null: no scaling
    double* scaled_out;    // n x dim or null
    int* assign_out;       // n or null
    double* part;          // gridDim.x x k x (dim + 1) or null (predict only)
    long long n;
    int k, dim;
};

static __global__ void __launch_bounds__(kKmThreads) k_km_assign(const KmArgs a)
{
    __shared__ double sc[kKmMaxK * kKmMaxDim], scn[kKmMaxK], sinv[kKmMaxDim];
    __shared__ double acc[kKmWarps][kKmMaxK * (kKmMaxDim + 1)];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, k = a.k, dim = a.dim, pw = dim + 1;
    for (int i = t; i < k * dim; i += kKmThreads) sc[i] = a.centers[i];
    for (int i = t; i < dim; i += kKmThreads) sinv[i] = a.inv_std ? a.inv_std[i] : 1.0;
    for (int i = t; i < kKmWarps * kKmMaxK * (kKmMaxDim + 1); i += kKmThreads) (&acc[0][0])[i] = 0.0;
    __syncthreads();
    if (t < k) {
        double s = 0.0;
        for (int j = 0; j < dim; ++j) s = __dadd_rn(s, __dmul_rn(sc[t * dim + j], sc[t * dim + j]));
        scn[t] = sqrt(s);
    }
    __syncthreads();
    const long long lo = a.n * blockIdx.x / gridDim.x, hi = a.n * (blockIdx.x + 1) / gridDim.x;
    const double eps = 2.220446049250313e-16;
    for (long long base = lo; base < hi; base += kKmThreads) { // chunks of one row per thread; all threads stay in the loop
        const long long r = base + t;
        const bool in = r < hi;
        double x[kKmMaxDim];
        double xn = 0.0;
        int best_c = 0;
        if (in) {
            for (int j = 0; j < dim; ++j) {
                double v = a.X[r * dim + j];
                if (a.inv_std) v = __dmul_rn(v, sinv[j]);
                x[j] = v;
                xn = __dadd_rn(xn, __dmul_rn(v, v));
                if (a.scaled_out) a.scaled_out[r * dim + j] = v;
            }
            xn = sqrt(xn);
            double best = __longlong_as_double(0x7ff0000000000000ll);
            for (int c = 0; c < k; ++c) {
                double lb = __dadd_rn(scn[c], -xn);
                lb = __dmul_rn(lb, lb);
                if (lb < best) {
                    const double ssn = __dadd_rn(__dmul_rn(scn[c], scn[c]), __dmul_rn(xn, xn)), nd = __dadd_rn(scn[c], -xn);
                    const double bound1 = __ddiv_rn(__dmul_rn(__dmul_rn(2.0, eps), ssn), __dadd_rn(__dmul_rn(nd, nd), eps));
                    double dist = 0.0;
                    if (bound1 < 1e-6) {
                        double dot = 0.0;
                        for (int j = 0; j < dim; ++j) dot = __dadd_rn(dot, __dmul_rn(sc[c * dim + j], x[j]));
                        dist = __dadd_rn(ssn, -__dmul_rn(2.0, dot));
                    } else {
                        for (int j = 0; j < dim; ++j) {
                            const double df = __dadd_rn(sc[c * dim + j], -x[j]);
                            dist = __dadd_rn(dist, __dmul_rn(df, df));
                        }
                    }
                    if (dist < best) { best = dist; best_c = c; }
                }
            }
            if (a.assign_out) a.assign_out[r] = best_c;
        }
        if (a.part) { // per cluster: the warp's 32 rows folded by a fixed butterfly, then added to the warp's accumulator
            for (int c = 0; c < k; ++c) {
                const bool mine = in && best_c == c;
                for (int j = 0; j <= dim; ++j) {
                    double v = mine ? (j < dim ? x[j] : 1.0) : 0.0;
                    for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
                    if (lane == 0) acc[warp][c * pw + j] += v;
                }
            }
        }
    }
    if (a.part) {
        __syncthreads();
        for (int i = t; i < k * pw; i += kKmThreads) {
            double s = 0.0;
            for (int w = 0; w < kKmWarps; ++w) s += acc[w][i]; // warp order
            a.part[(long long)blockIdx.x * k * pw + i] = s;
        }
    }
}

// StreamingKMeansModel.update [U] from the blocks' (sum, count) partials, one thread (k x dim numbers)
static __global__ void k_km_update(const double* __restrict__ part, int nblocks, int k, int dim, double* __restrict__ centers,
                                   double* __restrict__ weights, double decay, int time_unit, double n_points)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int pw = dim + 1;
    const double discount = time_unit == 1 ? pow(decay, n_points) : decay;
    for (int c = 0; c < k; ++c) weights[c] = __dmul_rn(weights[c], discount);
    for (int c = 0; c < k; ++c) {
        double cnt = 0.0;
        for (int b = 0; b < nblocks; ++b) cnt += part[((long long)b * k + c) * pw + dim];
        if (cnt == 0.0) continue;
        const double updated = __dadd_rn(weights[c], cnt);
        const double lambda = cnt / fmax(updated, 1e-16);
        weights[c] = updated;
        for (int j = 0; j < dim; ++j) {
            double sum = 0.0;
            for (int b = 0; b < nblocks; ++b) sum += part[((long long)b * k + c) * pw + j]; // block order
            double v = __dmul_rn(centers[c * dim + j], __dadd_rn(1.0, -lambda));
            v = __dadd_rn(v, __dmul_rn(lambda / cnt, sum));
            centers[c * dim + j] = v;
        }
    }
    int largest = 0, smallest = 0;
    for (int c = 1; c < k; ++c) {
        if (weights[c] > weights[largest]) largest = c;
        if (weights[c] < weights[smallest]) smallest = c;
    }
    if (weights[smallest] < __dmul_rn(1e-8, weights[largest])) { // a dying cluster: split the largest one
        const double w = __dadd_rn(weights[largest], weights[smallest]) / 2.0;
        weights[largest] = w;
        weights[smallest] = w;
        for (int j = 0; j < dim; ++j) {
            const double x = centers[largest * dim + j];
            const double p = __dmul_rn(1e-14, fmax(fabs(x), 1.0));
            centers[largest * dim + j] = __dadd_rn(x, p);
            centers[smallest * dim + j] = __dadd_rn(x, -p);
        }
    }
}

// standard normals for setRandomCenters: Box-Muller over the library's counter-based generator (MLlib draws them from
// XORShiftRandom(seed) [U], which cannot be reproduced without the JVM's seed hashing: documented difference)
static __global__ void k_km_random(double* centers, int count, unsigned long long seed)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const double u1 = fmax(u01(seed, 2ull * i), 1e-300), u2 = u01(seed, 2ull * i + 1);
    centers[i] = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

static int km_blocks(const twtml_ctx* ctx, long long n) { return (int)std::max<long long>(1, std::min<long long>((n + kKmThreads - 1) / kKmThreads, (long long)ctx->sm_count * 4)); }

// rows from the caller (stride ld) -> packed n x dim on the device
static int km_upload(twtml_kmeans* m, const double* X, int64_t n, int64_t ld)
{
    twtml_ctx* ctx = m->ctx;
    TW_TRY(m->x.reserve(sizeof(double) * (size_t)std::max<int64_t>(1, n) * m->dim));
    if (n == 0) return 0;
    if (ld == m->dim) return upload_plain(ctx, ctx->lane[0], m->x.p, X, sizeof(double) * (size_t)n * m->dim);
    std::vector<double> packed((size_t)n * m->dim);
    for (int64_t r = 0; r < n; ++r) memcpy(&packed[(size_t)r * m->dim], X + r * ld, sizeof(double) * (size_t)m->dim);
    TW_TRY(upload_plain(ctx, ctx->lane[0], m->x.p, packed.data(), sizeof(double) * packed.size()));
    TW_CUDA(cudaStreamSynchronize(ctx->stream)); // `packed` dies here
    return 0;
}

} // namespace twtml

using namespace twtml;

extern "C" {

int32_t twtml_kmeans_create(twtml_ctx* ctx, int32_t k, int32_t dim, twtml_kmeans** out)
{
    if (!ctx || !out) return fail(TWTML_E_INVALID_ARG, "null argument");
    if (!ctx->members.empty() || ctx->world != 1) return fail(TWTML_E_INVALID_ARG, "twtml_kmeans_create: single-GPU contexts only");
    if (k < 1 || k > kKmMaxK || dim < 1 || dim > kKmMaxDim) return fail(TWTML_E_INVALID_ARG, "twtml_kmeans_create: 1 <= k <= %d, 1 <= dim <= %d", kKmMaxK, kKmMaxDim);
    Guard g(ctx);
    twtml_kmeans* m = new twtml_kmeans();
    m->ctx = ctx;
    m->k = k;
    m->dim = dim;
    cudaError_t e = cudaMalloc(&m->d_centers, sizeof(double) * (size_t)(k * dim + k + 2 * dim));
    if (e != cudaSuccess) { delete m; return fail(TWTML_E_OOM, "cudaMalloc centres: %s", cudaGetErrorString(e)); }
    ctx->live_handles += 1;
    *out = m;
    return 0;
}

int32_t twtml_kmeans_free(twtml_kmeans* m)
{
    if (!m) return 0;
    {
        Guard g(m->ctx);
        cudaStreamSynchronize(m->ctx->stream);
        cudaFree(m->d_centers);
        for (DevBuf* b : {&m->x, &m->scaled, &m->assign, &m->moments, &m->sums}) b->release();
        m->ctx->live_handles -= 1;
    }
    delete m;
    return 0;
}

int32_t twtml_kmeans_set_decay_factor(twtml_kmeans* m, double a)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    if (!(a >= 0.0)) return fail(TWTML_E_INVALID_ARG, "decayFactor must be >= 0");
    m->decay = a;
    return 0;
}

int32_t twtml_kmeans_set_time_unit(twtml_kmeans* m, int32_t time_unit)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    if (time_unit != 0 && time_unit != 1) return fail(TWTML_E_INVALID_ARG, "timeUnit must be 0 (batches) or 1 (points)");
    m->time_unit = time_unit;
    return 0;
}

int32_t twtml_kmeans_set_half_life(twtml_kmeans* m, double half_life, int32_t time_unit)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    if (time_unit != 0 && time_unit != 1) return fail(TWTML_E_INVALID_ARG, "timeUnit must be 0 (batches) or 1 (points)");
    if (!(half_life > 0.0)) return fail(TWTML_E_INVALID_ARG, "halfLife must be > 0");
    m->decay = std::exp(std::log(0.5) / half_life); // StreamingKMeans.setHalfLife [U]
    m->time_unit = time_unit;
    return 0;
}

int32_t twtml_kmeans_set_centers(twtml_kmeans* m, const double* centers, const double* weights)
{
    if (!m || !centers || !weights) return fail(TWTML_E_INVALID_ARG, "null argument");
    Guard g(m->ctx);
    TW_CUDA(cudaMemcpy(m->d_centers, centers, sizeof(double) * (size_t)m->k * m->dim, cudaMemcpyHostToDevice));
    TW_CUDA(cudaMemcpy(m->d_centers + m->k * m->dim, weights, sizeof(double) * (size_t)m->k, cudaMemcpyHostToDevice));
    m->initialized = true;
    return 0;
}

int32_t twtml_kmeans_set_random_centers(twtml_kmeans* m, double weight, uint64_t seed)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    Guard g(m->ctx);
    k_km_random<<<1, 1024, 0, m->ctx->stream>>>(m->d_centers, m->k * m->dim, seed);
    TW_CUDA(cudaGetLastError());
    std::vector<double> w((size_t)m->k, weight);
    TW_CUDA(cudaMemcpyAsync(m->d_centers + m->k * m->dim, w.data(), sizeof(double) * w.size(), cudaMemcpyHostToDevice, m->ctx->stream));
    TW_CUDA(cudaStreamSynchronize(m->ctx->stream));
    m->initialized = true;
    return 0;
}

int32_t twtml_kmeans_get_centers(const twtml_kmeans* m, double* centers, double* weights)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    if (!m->initialized) return fail(TWTML_E_NOT_INITIALIZED, "Initial cluster centers must be set before starting predictions");
    Guard g(m->ctx);
    if (centers) TW_CUDA(cudaMemcpy(centers, m->d_centers, sizeof(double) * (size_t)m->k * m->dim, cudaMemcpyDeviceToHost));
    if (weights) TW_CUDA(cudaMemcpy(weights, m->d_centers + m->k * m->dim, sizeof(double) * (size_t)m->k, cudaMemcpyDeviceToHost));
    return 0;
}

static int32_t km_run(twtml_kmeans* m, const double* X, int64_t n, int64_t ld, bool standardize, bool update, double* std_out,
                      double* scaled_out, int32_t* assign_out)
{
    if (!m) return fail(TWTML_E_INVALID_ARG, "null model");
    if (!m->initialized) return fail(TWTML_E_NOT_INITIALIZED, "Initial cluster centers must be set before starting %s", update ? "training" : "predictions");
    if (n < 0 || (n > 0 && !X) || ld < m->dim) return fail(TWTML_E_INVALID_ARG, "twtml_kmeans: bad shape or null data");
    twtml_ctx* ctx = m->ctx;
    Guard g(ctx);
    const int k = m->k, dim = m->dim;
    double* centers = m->d_centers;
    double* weights = centers + k * dim;
    double* d_std = weights + k;
    double* d_inv = d_std + dim;
    if (n == 0) { // `if (rdd.count > 0)` (KMeans.scala:94): nothing happens for an empty interval
        if (std_out) for (int j = 0; j < dim; ++j) std_out[j] = 0.0;
        return 0;
    }
    TW_TRY(km_upload(m, X, n, ld));
    const int nb = km_blocks(ctx, n);
    if (standardize) {
        TW_TRY(m->moments.reserve(sizeof(double) * (size_t)nb * dim * 3));
        k_km_moments<<<nb, kKmThreads, 0, ctx->stream>>>(m->x.as<double>(), n, dim, m->moments.as<double>());
        k_km_std<<<1, 32, 0, ctx->stream>>>(m->moments.as<double>(), nb, dim, d_std, d_inv);
    }
    KmArgs a{};
    a.X = m->x.as<double>();
    a.centers = centers;
    a.inv_std = standardize ? d_inv : nullptr;
    a.n = n;
    a.k = k;
    a.dim = dim;
    if (scaled_out || update) { // the updated model predicts on the scaled data: keep it
        TW_TRY(m->scaled.reserve(sizeof(double) * (size_t)n * dim));
        a.scaled_out = m->scaled.as<double>();
    }
    if (update) {
        TW_TRY(m->sums.reserve(sizeof(double) * (size_t)nb * k * (dim + 1)));
        a.part = m->sums.as<double>();
        k_km_assign<<<nb, kKmThreads, 0, ctx->stream>>>(a);
        k_km_update<<<1, 32, 0, ctx->stream>>>(m->sums.as<double>(), nb, k, dim, centers, weights, m->decay, m->time_unit, (double)n);
        a.X = m->scaled.as<double>(); // predict(scaledData) with the UPDATED centres (KMeans.scala:105)
        a.inv_std = nullptr;
        a.scaled_out = nullptr;
        a.part = nullptr;
    }
    if (assign_out || !update) {
        TW_TRY(m->assign.reserve(sizeof(int) * (size_t)n));
        a.assign_out = m->assign.as<int>();
        k_km_assign<<<nb, kKmThreads, 0, ctx->stream>>>(a);
    }
    TW_CUDA(cudaGetLastError());
    if (std_out && standardize) TW_CUDA(cudaMemcpyAsync(std_out, d_std, sizeof(double) * (size_t)dim, cudaMemcpyDeviceToHost, ctx->stream));
    if (scaled_out) TW_CUDA(cudaMemcpyAsync(scaled_out, m->scaled.p, sizeof(double) * (size_t)n * dim, cudaMemcpyDeviceToHost, ctx->stream));
    if (assign_out) TW_CUDA(cudaMemcpyAsync(assign_out, m->assign.p, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    TW_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int32_t twtml_kmeans_update(twtml_kmeans* m, const double* X, int64_t n, int64_t ld, int32_t standardize, double* std_out, double* scaled_out,
                            int32_t* assign_out)
{
    return km_run(m, X, n, ld, standardize != 0, true, std_out, scaled_out, assign_out);
}

int32_t twtml_kmeans_predict(twtml_kmeans* m, const double* X, int64_t n, int64_t ld, int32_t* assign_out)
{
    if (!assign_out && n > 0) return fail(TWTML_E_INVALID_ARG, "twtml_kmeans_predict: assign_out is null");
    return km_run(m, X, n, ld, false, false, nullptr, nullptr, assign_out);
}

} // extern "C"
