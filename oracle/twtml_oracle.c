/*
 * twtml_oracle.c -- CPU restatement of twtml-spark's per-micro-batch learning path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under twitter-stream-ml_b200/ may link, import or
 * call this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs do, and there only as the checker / the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference (QilinGu/twitter-stream-ml) ships no test, golden vector
 * or fixture for this path (its two suites cover config parsing and the web JSON
 * round-trip only), and its arithmetic lives in an un-vendored dependency,
 * org.apache.spark:spark-mllib_2.10:1.6.1 (spark/build.sbt:15,22-24), which cannot be
 * built or run here (no JVM).  This file therefore restates
 *   - the reference's own Scala sources line by line (cited as file:line under
 *     /root/reference/spark/src/main/scala/com/giorgioinf/twtml/spark/), and
 *   - the published MLlib 1.6.1 algorithms (cited as [U] + upstream class name),
 * and is pinned only by hand-derived known answers (tests/golden/) and by an
 * independently written numpy twin (oracle/oracle_np.py).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp).  -ffp-contract=off
 * matters: the JVM never fuses a*b+c, so neither may this file.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TW_EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------
 * Utils.round  (Utils.scala:4-5)
 *   BigDecimal(number).setScale(0, HALF_UP).toDouble
 * Scala 2.10 builds the BigDecimal from Double.toString(number); a decimal string that
 * reads exactly k.5 can only come from the double k+0.5, so HALF_UP on the string equals
 * HALF_UP on the binary value: truncate, then step away from zero when the (exactly
 * representable) remainder is >= 0.5.  BigDecimal has no negative zero, hence "+ 0.0".
 * NaN / Infinity make the reference throw NumberFormatException; here they pass through
 * and callers test isfinite().
 * ---------------------------------------------------------------------------------- */
TW_EXPORT double twtml_oracle_round(double x)
{
    if (!isfinite(x)) return x;
    double r = trunc(x);
    double rem = x - r;                    /* exact */
    if (fabs(rem) >= 0.5) r += (x < 0.0) ? -1.0 : 1.0;
    return r + 0.0;
}

/* ------------------------------------------------------------------------------------
 * HashingTF  [U] mllib/feature/HashingTF.scala, core/util/Utils.scala nonNegativeMod
 *   indexOf(term) = nonNegativeMod(term.##, numFeatures)
 * term is a Scala String of 1 or 2 UTF-16 units, so term.## = String.hashCode
 *   = s[0]*31^(n-1) + ... + s[n-1]  in int32 arithmetic.
 * ---------------------------------------------------------------------------------- */
static int32_t java_string_hash(const uint16_t* s, int64_t n)
{
    uint32_t h = 0;
    for (int64_t i = 0; i < n; ++i) h = 31u * h + (uint32_t)s[i];
    return (int32_t)h;
}

static int32_t non_negative_mod(int32_t x, int32_t m)
{
    int32_t r = x % m;
    return r < 0 ? r + m : r;
}

TW_EXPORT int32_t twtml_oracle_hash_term(const uint16_t* s, int64_t n, int32_t num_features)
{
    return non_negative_mod(java_string_hash(s, n), num_features);
}

/* MllibHelper.filtrate / retweetInterval (MllibHelper.scala:84-94): isRetweet &&
 * begin <= retweetedStatus.retweetCount <= end, both ends inclusive. */
TW_EXPORT int32_t twtml_oracle_filtrate(uint8_t is_retweet, int32_t retweet_count,
                                        int32_t rt_begin, int32_t rt_end)
{
    return is_retweet && retweet_count >= rt_begin && retweet_count <= rt_end;
}

/* ------------------------------------------------------------------------------------
 * MllibHelper.featurize over a batch of tweets  (MllibHelper.scala:42-56,58-71,73-82)
 *
 * text_utf16/text_off: the retweeted status text AFTER java String.toLowerCase
 * (MllibHelper.scala:43-45), as UTF-16 code units.  The NFD/accent-strip result is
 * computed and discarded by the reference (:49-51), so nothing of it appears here.
 * text.sliding(2) (:54): L >= 2 -> L-1 two-unit windows; L == 1 -> one one-unit window;
 * L == 0 -> none.  HashingTF.transform counts occurrences per bucket and
 * Vectors.sparse sorts by index.  featurizeNumbers (:58-71): three int counts * 1e-12
 * and (now - created) * 1e-14 (Math.pow(10,-12)/(10,-14) taken as the doubles nearest
 * 1e-12 / 1e-14; a JVM whose pow is 1 ulp off cannot be checked here).  featurize (:73-82)
 * appends indices num_text..num_text+3 with the four numbers, stored even when 0.0;
 * label = retweetCount.toDouble (:81).
 *
 * Rows failing filtrate are dropped (LinearRegression.scala:44-46), order preserved.
 * Two-call protocol: with colidx == NULL only rowptr/labels sizes are computed.
 * Returns the number of kept rows; *nnz_out receives total nnz.
 * ---------------------------------------------------------------------------------- */
TW_EXPORT int64_t twtml_oracle_featurize(
    const uint16_t* text_utf16, const int64_t* text_off,
    const int32_t* retweet_cnt, const int32_t* followers, const int32_t* favourites,
    const int32_t* friends, const int64_t* created_ms, const uint8_t* is_retweet,
    int64_t n, int64_t now_ms, int32_t rt_begin, int32_t rt_end, int32_t num_text_features,
    int64_t* rowptr, int32_t* colidx, double* vals, double* labels, int64_t* nnz_out)
{
    const int32_t F = num_text_features;
    int32_t* counts = (int32_t*)calloc((size_t)F, sizeof(int32_t));
    int64_t kept = 0, nnz = 0;
    if (rowptr) rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (!twtml_oracle_filtrate(is_retweet[i], retweet_cnt[i], rt_begin, rt_end)) continue;
        const uint16_t* t = text_utf16 + text_off[i];
        int64_t L = text_off[i + 1] - text_off[i];
        if (L == 1) counts[twtml_oracle_hash_term(t, 1, F)] += 1;
        for (int64_t k = 0; k + 1 < L; ++k) counts[twtml_oracle_hash_term(t + k, 2, F)] += 1;
        for (int32_t b = 0; b < F; ++b) {
            if (counts[b]) {
                if (colidx) { colidx[nnz] = b; vals[nnz] = (double)counts[b]; }
                ++nnz;
                counts[b] = 0;
            }
        }
        if (colidx) {
            colidx[nnz + 0] = F + 0; vals[nnz + 0] = (double)followers[i] * 1e-12;
            colidx[nnz + 1] = F + 1; vals[nnz + 1] = (double)favourites[i] * 1e-12;
            colidx[nnz + 2] = F + 2; vals[nnz + 2] = (double)friends[i] * 1e-12;
            colidx[nnz + 3] = F + 3; vals[nnz + 3] = (double)(now_ms - created_ms[i]) * 1e-14;
        }
        nnz += 4;
        if (labels) labels[kept] = (double)retweet_cnt[i];
        ++kept;
        if (rowptr) rowptr[kept] = nnz;
    }
    free(counts);
    if (nnz_out) *nnz_out = nnz;
    return kept;
}

/* ------------------------------------------------------------------------------------
 * Counter-based stand-in for RDD.sample(false, frac, 42 + i)  [U] RDD.scala,
 * RandomSampler.scala.  Spark seeds an XORShiftRandom per partition, so its draw is a
 * function of the partitioning and cannot be reproduced outside Spark; what is kept is
 * the contract (identity for frac >= 1, Bernoulli(frac) per row keyed by 42 + i).  The
 * draw here is splitmix64 over (seed, global row index) so that it does not depend on how
 * rows are split over partitions or GPUs.
 * ---------------------------------------------------------------------------------- */
static inline uint64_t splitmix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

static inline double u01(uint64_t seed, uint64_t idx)
{
    return (double)(splitmix64(seed ^ idx) >> 11) * 0x1.0p-53;
}

TW_EXPORT int32_t twtml_oracle_sampled(int32_t iter, uint64_t global_row, double frac)
{
    if (frac >= 1.0) return 1;
    if (frac <= 0.0) return 0;
    uint64_t seed = splitmix64((uint64_t)(42 + iter));
    return u01(seed, global_row) < frac;
}

/* ------------------------------------------------------------------------------------
 * Synthetic dense micro-batch (measurement tool; not in the
 * reference).  X[r][j] = (2u-1) * inv_sqrt_d, u = u01(seed, r*d + j);  w*[j] the same
 * with seed+1 keyed by j;  y[r] = sum_j X[r][j]*w*[j] (index order, unfused) +
 * 0.01*(2u'-1), u' = u01(seed+2, r).  r is the GLOBAL row index.
 * ---------------------------------------------------------------------------------- */
TW_EXPORT void twtml_oracle_synth_dense(int64_t n, int64_t d, uint64_t seed, int64_t row0,
                                        double* X, int64_t ldx, double* y, int threads)
{
    const double inv_sqrt_d = 1.0 / sqrt((double)d);
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < n; ++i) {
        uint64_t r = (uint64_t)(row0 + i);
        double acc = 0.0;
        for (int64_t j = 0; j < d; ++j) {
            double x = (2.0 * u01(seed, r * (uint64_t)d + (uint64_t)j) - 1.0) * inv_sqrt_d;
            double ws = (2.0 * u01(seed + 1, (uint64_t)j) - 1.0) * inv_sqrt_d;
            if (X) X[i * ldx + j] = x;
            acc = acc + x * ws;
        }
        y[i] = acc + 0.01 * (2.0 * u01(seed + 2, r) - 1.0);
    }
}

/* Synthetic CSR micro-batch (config 5, distribution B = uniform columns): row r has
 * k = 1 + (h % 40) text entries drawn as distinct sorted columns in [0, D_text), TF count
 * 1 + (h' % 3), then the four numeric slots D_text..D_text+3 (values ~ 1e-6 scale),
 * label in [100, 1000].  Deterministic in (seed, global row).  Two-call protocol. */
static int cmp_i32(const void* a, const void* b)
{
    int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
    return (x > y) - (x < y);
}

TW_EXPORT int64_t twtml_oracle_synth_csr(int64_t n, int32_t d_text, uint64_t seed, int64_t row0,
                                         int64_t* rowptr, int32_t* colidx, double* vals,
                                         double* labels)
{
    int64_t nnz = 0;
    int32_t tmp[64];
    rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        uint64_t r = (uint64_t)(row0 + i);
        uint64_t h = splitmix64(seed ^ (r * 0x100000001b3ull));
        int k = 1 + (int)(h % 40);
        int m = 0;
        for (int t = 0; t < k; ++t) {
            int32_t c = (int32_t)(splitmix64(h + (uint64_t)t) % (uint64_t)d_text);
            tmp[m++] = c;
        }
        qsort(tmp, (size_t)m, sizeof(int32_t), cmp_i32);
        int u = 0;
        for (int t = 0; t < m; ++t)
            if (t == 0 || tmp[t] != tmp[t - 1]) tmp[u++] = tmp[t];
        for (int t = 0; t < u; ++t) {
            if (colidx) {
                colidx[nnz] = tmp[t];
                vals[nnz] = (double)(1 + (int)(splitmix64(h ^ (uint64_t)tmp[t]) % 3));
            }
            ++nnz;
        }
        for (int t = 0; t < 4; ++t) {
            if (colidx) {
                colidx[nnz] = d_text + t;
                vals[nnz] = u01(seed + 7 + (uint64_t)t, r) * 1e-6;
            }
            ++nnz;
        }
        if (labels) labels[i] = 100.0 + floor(u01(seed + 3, r) * 901.0);
        rowptr[i + 1] = nnz;
    }
    return nnz;
}

/* ------------------------------------------------------------------------------------
 * BLAS level-1 as MLlib 1.6.1 runs it  [U] mllib/linalg/BLAS.scala
 *   dense.dense  -> f2jBLAS.ddot: unrolled by 5, but "dtemp + a0 + a1 + ..." evaluates
 *                   left to right, i.e. plain index order;
 *   sparse.dense -> sum += values(k) * w(indices(k)) in stored order;
 *   axpy         -> y(i) += a * x(i), each element independent.
 * ---------------------------------------------------------------------------------- */
static inline double dot_dense(const double* x, const double* w, int64_t d)
{
    double s = 0.0;
    for (int64_t j = 0; j < d; ++j) s = s + x[j] * w[j];
    return s;
}

static inline double dot_sparse(const int32_t* idx, const double* v, int64_t nnz, const double* w)
{
    double s = 0.0;
    for (int64_t k = 0; k < nnz; ++k) s = s + v[k] * w[idx[k]];
    return s;
}

/* Accumulator of treeAggregate in runMiniBatchSGD: (gradSum[d], lossSum, count). */
typedef struct {
    double* g;
    double loss;
    int64_t cnt;
} agg_t;

/* Updaters  [U] mllib/optimization/Updater.scala.  kind 0 = SimpleUpdater (what
 * LinearRegressionWithSGD, hence the reference, uses; regParam ignored), 1 =
 * SquaredL2Updater, 2 = L1Updater (both outside the reference path).
 * g is the averaged gradient.  Returns regVal. */
static double updater_compute(int kind, double* w, const double* g, int64_t d, double step,
                              int iter, double reg)
{
    double this_step = step / sqrt((double)iter);
    if (kind == 0) {
        for (int64_t j = 0; j < d; ++j) w[j] = w[j] + (-this_step) * g[j];
        return 0.0;
    } else if (kind == 1) {
        double shrink = 1.0 - this_step * reg;
        double nrm2 = 0.0;
        for (int64_t j = 0; j < d; ++j) {
            w[j] = w[j] * shrink;
            w[j] = w[j] + (-this_step) * g[j];
            nrm2 += w[j] * w[j];
        }
        double nrm = sqrt(nrm2);
        return 0.5 * reg * nrm * nrm;
    } else {
        double shrinkage = reg * this_step;
        double l1 = 0.0;
        for (int64_t j = 0; j < d; ++j) {
            w[j] = w[j] + (-this_step) * g[j];
            double a = fabs(w[j]) - shrinkage;
            double sg = (w[j] > 0.0) - (w[j] < 0.0);
            w[j] = sg * (a > 0.0 ? a : 0.0);
            l1 += fabs(w[j]);
        }
        return l1 * reg;
    }
}

/* RDD.treeAggregate combine order  [U] core/rdd/RDD.scala: with P partition results and
 * depth 2, scale = max(ceil(sqrt(P)), 2); while P > scale + P/scale the partials are
 * folded by key (i % (P/scale)); the survivors are reduced in index order. */
static void tree_combine(agg_t* a, int P, int64_t d)
{
    int scale = (int)ceil(sqrt((double)P));
    if (scale < 2) scale = 2;
    int cur = P;
    while (cur > scale + cur / scale) {
        int next = cur / scale;
        for (int i = next; i < cur; ++i) {
            agg_t* dst = &a[i % next];
            for (int64_t j = 0; j < d; ++j) dst->g[j] += a[i].g[j];
            dst->loss += a[i].loss;
            dst->cnt += a[i].cnt;
        }
        cur = next;
    }
    for (int i = 1; i < cur; ++i) {
        for (int64_t j = 0; j < d; ++j) a[0].g[j] += a[i].g[j];
        a[0].loss += a[i].loss;
        a[0].cnt += a[i].cnt;
    }
}

/* ------------------------------------------------------------------------------------
 * GradientDescent.runMiniBatchSGD  [U] mllib/optimization/GradientDescent.scala, reached
 * from StreamingLinearAlgorithm.trainOn (called at LinearRegression.scala:86) through
 * GeneralizedLinearAlgorithm.run (no intercept, no scaling for LinearRegressionWithSGD).
 *
 *   n == 0 -> weights returned unchanged, no iterations.
 *   regVal = updater.compute(w, 0, 0, 1, reg)._2
 *   for i = 1 .. numIterations while !converged:
 *      (gradSum, lossSum, m) = sample(frac, 42+i).treeAggregate(seqOp, combOp)
 *         seqOp: LeastSquaresGradient.compute: diff = dot(x,w) - y; axpy(diff, x, grad);
 *                loss += diff*diff/2; count += 1
 *      if m > 0: lossHistory += lossSum/m + regVal
 *                (w, regVal) = updater.compute(w, gradSum/m, stepSize, i, reg)
 *                converged = (an earlier iteration also had m > 0: previousWeights != None)
 *                            && ||w_prev - w|| < tol * max(||w||, 1)
 *   The iteration counter restarts at 1 on every micro-batch.
 *
 * Rows come either dense (X, ldx) or CSR (rowptr/colidx/vals) -- exactly one non-NULL.
 * partitions = number of RDD partitions (contiguous row ranges); threads = OpenMP
 * threads folding them.  row0 = global index of the first row (sampler key).
 * Returns 0, or -1 on bad arguments.
 * ---------------------------------------------------------------------------------- */
static int32_t sgd_core(
    const double* X, int64_t ldx,
    const int64_t* rowptr, const int32_t* colidx, const double* vals,
    const double* y, int64_t n, int64_t d,
    double* w, double step_size, int32_t num_iterations, double mini_batch_fraction,
    double convergence_tol, double reg_param, int32_t updater_kind,
    int32_t partitions, int32_t threads, int64_t row0,
    double* loss_history, int32_t* iters_run, int synth, uint64_t synth_seed)
{
    if ((!synth && (X == NULL) == (rowptr == NULL)) || d <= 0 || n < 0) return -1;
    if (iters_run) *iters_run = 0;
    if (n == 0) return 0;
    int P = partitions < 1 ? 1 : partitions;
    if ((int64_t)P > n) P = (int)n;
    if (threads < 1) threads = 1;

    agg_t* agg = (agg_t*)malloc(sizeof(agg_t) * (size_t)P);
    for (int p = 0; p < P; ++p) agg[p].g = (double*)malloc(sizeof(double) * (size_t)d);
    double* gavg = (double*)calloc((size_t)d, sizeof(double));
    double* wprev = (double*)malloc(sizeof(double) * (size_t)d);

    double reg_val = updater_compute(updater_kind, w, gavg, d, 0.0, 1, reg_param);
    int converged = 0, i = 1, nhist = 0;
    while (!converged && i <= num_iterations) {
#pragma omp parallel for schedule(static) num_threads(threads)
        for (int p = 0; p < P; ++p) {
            agg_t* a = &agg[p];
            memset(a->g, 0, sizeof(double) * (size_t)d);
            a->loss = 0.0;
            a->cnt = 0;
            int64_t lo = (int64_t)p * n / P, hi = (int64_t)(p + 1) * n / P;
            double* xrow = synth ? (double*)malloc(sizeof(double) * (size_t)d) : NULL;
            for (int64_t r = lo; r < hi; ++r) {
                if (!twtml_oracle_sampled(i, (uint64_t)(row0 + r), mini_batch_fraction)) continue;
                double diff;
                if (synth) { /* the row is regenerated from its counters instead of being stored */
                    double yr;
                    twtml_oracle_synth_dense(1, d, synth_seed, row0 + r, xrow, d, &yr, 1);
                    diff = dot_dense(xrow, w, d) - yr;
                    for (int64_t j = 0; j < d; ++j) a->g[j] += diff * xrow[j];
                } else if (X) {
                    const double* x = X + r * ldx;
                    diff = dot_dense(x, w, d) - y[r];
                    for (int64_t j = 0; j < d; ++j) a->g[j] += diff * x[j];
                } else {
                    int64_t b = rowptr[r], e = rowptr[r + 1];
                    diff = dot_sparse(colidx + b, vals + b, e - b, w) - y[r];
                    for (int64_t k = b; k < e; ++k) a->g[colidx[k]] += diff * vals[k];
                }
                a->loss += diff * diff / 2.0;
                a->cnt += 1;
            }
            free(xrow);
        }
        tree_combine(agg, P, d);
        int64_t m = agg[0].cnt;
        if (m > 0) {
            const int had_previous = nhist >= 1; /* previousWeights != None: set by an earlier iteration with m > 0 */
            loss_history[nhist++] = agg[0].loss / (double)m + reg_val;
            memcpy(wprev, w, sizeof(double) * (size_t)d);
            for (int64_t j = 0; j < d; ++j) gavg[j] = agg[0].g[j] / (double)m;
            reg_val = updater_compute(updater_kind, w, gavg, d, step_size, i, reg_param);
            if (had_previous) {
                double dn = 0.0, wn = 0.0;
                for (int64_t j = 0; j < d; ++j) {
                    double df = wprev[j] - w[j];
                    dn += df * df;
                    wn += w[j] * w[j];
                }
                dn = sqrt(dn);
                wn = sqrt(wn);
                converged = dn < convergence_tol * (wn > 1.0 ? wn : 1.0);
            }
        }
        ++i;
    }
    if (iters_run) *iters_run = nhist;
    for (int p = 0; p < P; ++p) free(agg[p].g);
    free(agg);
    free(gavg);
    free(wprev);
    return 0;
}

TW_EXPORT int32_t twtml_oracle_sgd(
    const double* X, int64_t ldx,
    const int64_t* rowptr, const int32_t* colidx, const double* vals,
    const double* y, int64_t n, int64_t d,
    double* w, double step_size, int32_t num_iterations, double mini_batch_fraction,
    double convergence_tol, double reg_param, int32_t updater_kind,
    int32_t partitions, int32_t threads, int64_t row0,
    double* loss_history, int32_t* iters_run)
{
    return sgd_core(X, ldx, rowptr, colidx, vals, y, n, d, w, step_size, num_iterations, mini_batch_fraction,
                    convergence_tol, reg_param, updater_kind, partitions, threads, row0, loss_history, iters_run, 0, 0);
}

/* The same on the synthetic dense micro-batch of twtml_oracle_synth_dense(seed, row0 ..), with every
 * row regenerated on the fly: lets the 2^24 x 1004 configuration (134.75 GB) be checked on a host
 * that could never hold it. */
TW_EXPORT int32_t twtml_oracle_sgd_synth(
    int64_t n, int64_t d, uint64_t seed, int64_t row0,
    double* w, double step_size, int32_t num_iterations, double convergence_tol,
    int32_t partitions, int32_t threads, double* loss_history, int32_t* iters_run)
{
    return sgd_core(NULL, 0, NULL, NULL, NULL, NULL, n, d, w, step_size, num_iterations, 1.0, convergence_tol, 0.0, 0,
                    partitions, threads, row0, loss_history, iters_run, 1, seed);
}

/* ------------------------------------------------------------------------------------
 * StatCounter  [U] core/util/StatCounter.scala: merge(value) is Welford's update,
 * merge(other) is Chan's; stdev = sqrt(m2 / n) (population).
 * ---------------------------------------------------------------------------------- */
typedef struct {
    int64_t n;
    double mu, m2;
} statc_t;

static void stat_add(statc_t* s, double v)
{
    double delta = v - s->mu;
    s->n += 1;
    s->mu += delta / (double)s->n;
    s->m2 += delta * (v - s->mu);
}

static void stat_merge(statc_t* a, const statc_t* b)
{
    if (b->n == 0) return;
    if (a->n == 0) { *a = *b; return; }
    double delta = b->mu - a->mu;
    int64_t n = a->n + b->n;
    if (b->n * 10 < a->n) a->mu = a->mu + (delta * (double)b->n) / (double)n;
    else if (a->n * 10 < b->n) a->mu = b->mu - (delta * (double)a->n) / (double)n;
    else a->mu = (a->mu * (double)a->n + b->mu * (double)b->n) / (double)n;
    a->m2 += b->m2 + (delta * delta * (double)a->n * (double)b->n) / (double)n;
    a->n = n;
}

/* ------------------------------------------------------------------------------------
 * The predict + statistics block of the driver loop (LinearRegression.scala:53-66):
 *   p_r   = round(latestModel.predict(x_r))          (:56-58; predict = w.x + 0.0 [U])
 *   batch = count                                     (:59)
 *   realStdev = round(stdev(y)), predStdev = round(stdev(p))       (:63-64)
 *   mse   = round(mean((y - p)^2))                    (:65; math.pow(v - p, 2))
 * The weights are those BEFORE this batch is trained on (:53 is registered before :86).
 * out[0..8] = batch, mse, realStdev, predStdev (rounded), then the raw mse, stdevs and
 * the two means.  pred_raw (nullable) receives w.x unrounded (= predictOnValues [U]).
 * ---------------------------------------------------------------------------------- */
TW_EXPORT int32_t twtml_oracle_predict_stats(
    const double* X, int64_t ldx,
    const int64_t* rowptr, const int32_t* colidx, const double* vals,
    const double* y, int64_t n, int64_t d, const double* w, int32_t partitions,
    double* pred_raw, double* pred_rounded, double* out9)
{
    if ((X == NULL) == (rowptr == NULL)) return -1;
    int P = partitions < 1 ? 1 : partitions;
    if ((int64_t)P > n) P = n > 0 ? (int)n : 1;
    statc_t real = {0, 0, 0}, pred = {0, 0, 0}, se = {0, 0, 0};
    for (int p = 0; p < P; ++p) {
        statc_t r1 = {0, 0, 0}, p1 = {0, 0, 0}, s1 = {0, 0, 0};
        int64_t lo = (int64_t)p * n / P, hi = (int64_t)(p + 1) * n / P;
        for (int64_t r = lo; r < hi; ++r) {
            double raw;
            if (X) raw = dot_dense(X + r * ldx, w, d);
            else {
                int64_t b = rowptr[r];
                raw = dot_sparse(colidx + b, vals + b, rowptr[r + 1] - b, w);
            }
            raw = raw + 0.0; /* + intercept */
            double pr = twtml_oracle_round(raw);
            if (pred_raw) pred_raw[r] = raw;
            if (pred_rounded) pred_rounded[r] = pr;
            double e = y[r] - pr;
            stat_add(&r1, y[r]);
            stat_add(&p1, pr);
            stat_add(&s1, e * e);
        }
        stat_merge(&real, &r1);
        stat_merge(&pred, &p1);
        stat_merge(&se, &s1);
    }
    double rs = n > 0 ? sqrt(real.m2 / (double)real.n) : NAN;
    double ps = n > 0 ? sqrt(pred.m2 / (double)pred.n) : NAN;
    double mse = n > 0 ? se.mu : NAN;
    out9[0] = (double)n;
    out9[1] = twtml_oracle_round(mse);
    out9[2] = twtml_oracle_round(rs);
    out9[3] = twtml_oracle_round(ps);
    out9[4] = mse;
    out9[5] = rs;
    out9[6] = ps;
    out9[7] = real.mu;
    out9[8] = pred.mu;
    return 0;
}

TW_EXPORT int32_t twtml_oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
