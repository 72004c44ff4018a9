/*
 * twtml_oracle_kmeans.c -- CPU restatement of the reference's second, half-disabled app: one batch
 * interval of KMeans.scala:93-105 (StandardScaler(false, true).fit(rdd).transform(rdd), then
 * StreamingKMeansModel.update, then predict with the updated model).
 *
 * TEST INFRASTRUCTURE ONLY (same rules as twtml_oracle.c) and PARITY UNPINNED: the reference has no test
 * on this path; the arithmetic is MLlib 1.6.1's ([U]: mllib/feature/StandardScaler.scala,
 * mllib/stat/MultivariateOnlineSummarizer.scala, mllib/clustering/{StreamingKMeans,KMeans}.scala,
 * mllib/util/MLUtils.scala fastSquaredDistance), restated from the published sources, unverifiable here.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TW_EXPORT __attribute__((visibility("default")))

/* MultivariateOnlineSummarizer [U]: per column, Welford over the NON-ZERO values only (add() skips zeros),
 * then   variance_j = (M2n_j + mean_j^2 * nnz_j * (n - nnz_j) / n) / (n - 1)   for n > 1, else 0,
 * the unbiased sample variance over all n values.  StandardScaler.fit keeps std_j = sqrt(variance_j). */
TW_EXPORT void twtml_oracle_column_std(const double* X, int64_t n, int64_t ld, int32_t dim, double* std_out)
{
    for (int32_t j = 0; j < dim; ++j) {
        double mean = 0.0, m2n = 0.0, nnz = 0.0;
        for (int64_t r = 0; r < n; ++r) {
            const double v = X[r * ld + j];
            if (v != 0.0) {
                const double prev = mean, diff = v - prev;
                mean = prev + diff / (nnz + 1.0);
                m2n += (v - mean) * diff;
                nnz += 1.0;
            }
        }
        double var = 0.0;
        if ((double)n - 1.0 > 0.0) {
            var = (m2n + mean * mean * nnz * ((double)n - nnz) / (double)n) / ((double)n - 1.0);
            if (var < 0.0) var = 0.0;
        }
        std_out[j] = sqrt(var);
    }
}

/* StandardScalerModel.transform with withMean = false, withStd = true [U]: value * (1 / std) where std != 0, else 0. */
TW_EXPORT void twtml_oracle_scale(const double* X, int64_t n, int64_t ld, int32_t dim, const double* std, double* out)
{
    for (int64_t r = 0; r < n; ++r)
        for (int32_t j = 0; j < dim; ++j) out[r * dim + j] = X[r * ld + j] * (std[j] != 0.0 ? 1.0 / std[j] : 0.0);
}

static double dotd(const double* a, const double* b, int32_t d)
{
    double s = 0.0;
    for (int32_t j = 0; j < d; ++j) s += a[j] * b[j];
    return s;
}

/* MLUtils.fastSquaredDistance [U] for dense vectors (precision 1e-6): the expanded form when its error bound allows,
 * else the exact sum of squared differences. */
static double fast_sqdist(const double* v1, double norm1, const double* v2, double norm2, int32_t d)
{
    const double eps = 2.220446049250313e-16; /* MLUtils.EPSILON: the loop `while (1.0 + eps / 2.0 != 1.0) eps /= 2.0` */
    const double ssn = norm1 * norm1 + norm2 * norm2, nd = norm1 - norm2;
    const double bound1 = 2.0 * eps * ssn / (nd * nd + eps);
    if (bound1 < 1e-6) return ssn - 2.0 * dotd(v1, v2, d);
    double s = 0.0;
    for (int32_t j = 0; j < d; ++j) {
        const double df = v1[j] - v2[j];
        s += df * df;
    }
    return s;
}

/* KMeans.findClosest [U]: centers in index order, a center is examined only if (||c|| - ||x||)^2 < best so far, and
 * replaces the best only when strictly closer (ties keep the lower index). */
static int32_t find_closest(const double* centers, const double* cnorm, int32_t k, int32_t d, const double* x, double xnorm)
{
    double best = INFINITY;
    int32_t bi = 0;
    for (int32_t c = 0; c < k; ++c) {
        double lb = cnorm[c] - xnorm;
        lb = lb * lb;
        if (lb < best) {
            const double dist = fast_sqdist(centers + (int64_t)c * d, cnorm[c], x, xnorm, d);
            if (dist < best) {
                best = dist;
                bi = c;
            }
        }
    }
    return bi;
}

TW_EXPORT void twtml_oracle_kmeans_predict(const double* centers, int32_t k, int32_t dim, const double* X, int64_t n, int64_t ld,
                                           int32_t* out)
{
    double* cn = (double*)malloc(sizeof(double) * (size_t)k);
    for (int32_t c = 0; c < k; ++c) cn[c] = sqrt(dotd(centers + (int64_t)c * dim, centers + (int64_t)c * dim, dim));
    for (int64_t r = 0; r < n; ++r) out[r] = find_closest(centers, cn, k, dim, X + r * ld, sqrt(dotd(X + r * ld, X + r * ld, dim)));
    free(cn);
}

/* StreamingKMeansModel.update(data, decayFactor, timeUnit) [U]; time_unit 0 = "batches", 1 = "points".
 * centers k x dim and weights k are updated in place.  X is the (already scaled) data, n x dim with stride ld. */
TW_EXPORT void twtml_oracle_kmeans_update(double* centers, double* weights, int32_t k, int32_t dim, const double* X, int64_t n,
                                          int64_t ld, double decay, int32_t time_unit)
{
    int32_t* lab = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    twtml_oracle_kmeans_predict(centers, k, dim, X, n, ld, lab);
    double* sums = (double*)calloc((size_t)k * dim, sizeof(double));
    double* cnt = (double*)calloc((size_t)k, sizeof(double));
    for (int64_t r = 0; r < n; ++r) {
        for (int32_t j = 0; j < dim; ++j) sums[(int64_t)lab[r] * dim + j] += X[r * ld + j];
        cnt[lab[r]] += 1.0;
    }
    double discount = decay;
    if (time_unit == 1) discount = pow(decay, (double)n);
    for (int32_t c = 0; c < k; ++c) weights[c] *= discount; /* BLAS.scal(discount, clusterWeights) */
    for (int32_t c = 0; c < k; ++c) {
        if (cnt[c] == 0.0) continue; /* pointStats holds only the clusters that received points */
        const double updated = weights[c] + cnt[c];
        const double lambda = cnt[c] / (updated > 1e-16 ? updated : 1e-16);
        weights[c] = updated;
        for (int32_t j = 0; j < dim; ++j) { /* scal(1 - lambda, centroid); axpy(lambda / count, sum, centroid) */
            double v = centers[(int64_t)c * dim + j] * (1.0 - lambda);
            v += (lambda / cnt[c]) * sums[(int64_t)c * dim + j];
            centers[(int64_t)c * dim + j] = v;
        }
    }
    /* a dying cluster: split the largest one */
    int32_t largest = 0, smallest = 0;
    for (int32_t c = 1; c < k; ++c) {
        if (weights[c] > weights[largest]) largest = c;   /* maxBy / minBy keep the first of equals */
        if (weights[c] < weights[smallest]) smallest = c;
    }
    if (weights[smallest] < 1e-8 * weights[largest]) {
        const double w = (weights[largest] + weights[smallest]) / 2.0;
        weights[largest] = w;
        weights[smallest] = w;
        for (int32_t j = 0; j < dim; ++j) {
            const double x = centers[(int64_t)largest * dim + j];
            const double p = 1e-14 * (fabs(x) > 1.0 ? fabs(x) : 1.0);
            centers[(int64_t)largest * dim + j] = x + p;
            centers[(int64_t)smallest * dim + j] = x - p;
        }
    }
    free(lab);
    free(sums);
    free(cnt);
}
