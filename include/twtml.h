/*
 * twtml.h -- C ABI of libtwtml.so: the per-micro-batch learning path of twtml-spark
 * (QilinGu/twitter-stream-ml) on NVIDIA B200 (sm_100a).
 *
 * The reference has no FFI seam on this path: its driver (LinearRegression.scala) calls the
 * Scala API of org.apache.spark.mllib.regression.StreamingLinearRegressionWithSGD (MLlib
 * 1.6.1, un-vendored; spark/build.sbt:15) and its own MllibHelper.  Each entry point below
 * names the reference call site it stands behind; paths are relative to
 * spark/src/main/scala/com/giorgioinf/twtml/spark/ in the reference tree, "[U]" marks a
 * class of MLlib 1.6.1.  INTEGRATION.md shows the JNA binding and the Scala facade that keep
 * those call sites source-compatible.
 *
 * Conventions
 *  - plain C, no C++ types, no exceptions; loadable with JNA / ctypes / dlopen (no jni.h).
 *  - every function returns 0 (TWTML_OK) or a negative twtml_status; the message of the last
 *    failure on the calling thread is read with twtml_last_error().
 *  - every call is blocking: results are on the host when it returns.  Streams are internal.
 *    A context may be used from any host thread (the Spark JobScheduler does not pin output
 *    ops to one OS thread), but from one thread at a time per context.
 *  - the caller owns every pointer it passes; the library copies inputs during the call and
 *    keeps no caller pointer afterwards.  Opaque handles are released only by their
 *    _free/_destroy function.
 *  - there is NO CPU fallback: without an sm_100 device twtml_create fails with TWTML_E_CUDA.
 *  - all floating point is IEEE binary64 (MLlib vectors are Array[Double]); indices int32,
 *    row counts / offsets int64.
 */
#ifndef TWTML_H
#define TWTML_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden; only this ABI is exported */
#endif

#define TWTML_ABI_VERSION 1

typedef enum twtml_status {
    TWTML_OK = 0,
    TWTML_E_INVALID_ARG = -1,
    TWTML_E_NOT_INITIALIZED = -2, /* model used before set_initial_weights: the facade turns this into
                                     IllegalArgumentException("Model must be initialized before starting
                                     training." / "... prediction.")  [U] StreamingLinearAlgorithm */
    TWTML_E_CUDA = -3,
    TWTML_E_NCCL = -4,
    TWTML_E_OOM = -5,
    TWTML_E_SHAPE_MISMATCH = -6,
    TWTML_E_NONFINITE = -7        /* Utils.round on NaN/Infinity throws NumberFormatException (Utils.scala:4-5) */
} twtml_status;

typedef struct twtml_ctx twtml_ctx;     /* one GPU (one rank of the row-sharded job) */
typedef struct twtml_model twtml_model; /* StreamingLinearRegressionWithSGD state: weights + hyper-parameters */
typedef struct twtml_batch twtml_batch; /* one micro-batch (one RDD[LabeledPoint]) resident in HBM */

/* Numbers the driver loop reports per batch (LinearRegression.scala:59-65, sent on by
 * SessionStats.scala:22-33).  mse / real_stdev / pred_stdev are rounded with Utils.round like
 * the reference; the *_raw fields are the same quantities before that last rounding. */
typedef struct twtml_stats {
    int64_t batch;         /* rdd.count                                  (:59) */
    double mse;            /* round(mean((label - round(pred))^2))       (:65) */
    double real_stdev;     /* round(stdev(label)), population            (:63) */
    double pred_stdev;     /* round(stdev(round(pred))), population      (:64) */
    double mse_raw;
    double real_stdev_raw;
    double pred_stdev_raw;
    double real_mean;
    double pred_mean;
    int64_t nonfinite;     /* rows whose prediction was NaN/Inf (the reference would throw) */
} twtml_stats;

/* Device-side timing of the last twtml_train_on_batch / twtml_step on a context, taken with
 * CUDA events on the library's own stream (measurement aid, no reference counterpart). */
typedef struct twtml_profile {
    int32_t iter_launches;   /* fused SGD-iteration kernel launches */
    int32_t other_launches;  /* reduce / update / stats / cast kernels */
    int32_t nccl_calls;
    float iter_kernel_ms;    /* sum of the iteration-kernel durations */
    float loop_ms;           /* first iteration launch .. last update kernel end */
    int64_t iter_kernel_bytes; /* algorithmic bytes one iteration-kernel launch reads+writes */
} twtml_profile;

/* ---- context ------------------------------------------------------------------------------ */

int32_t twtml_abi_version(void);

/* Copies the calling thread's last error message (NUL-terminated, truncated to cap). */
int32_t twtml_last_error(char* buf, int64_t cap);

/* Single-GPU context on CUDA device `device`.  Stands where the reference creates its
 * SparkContext (LinearRegression.scala:36): the thing the batches and the model live on. */
int32_t twtml_create(int32_t device, twtml_ctx** out);

/* One rank of a row-sharded job: world_size processes (or threads), one GPU each, every rank
 * holding a contiguous block of each micro-batch's rows.  Replaces the executor side of
 * RDD.treeAggregate + SparkContext.broadcast inside GradientDescent.runMiniBatchSGD [U]: per
 * iteration the ranks exchange (gradSum[d], lossSum, count) partials with one NCCL all-gather
 * and add them in a fixed order, so every rank applies the identical update.
 * nccl_id: TWTML_NCCL_ID_BYTES bytes from twtml_nccl_unique_id() on rank 0, handed to the
 * other ranks by the host application.  Collective: all ranks must call it together. */
#define TWTML_NCCL_ID_BYTES 128
int32_t twtml_nccl_unique_id(uint8_t* out, int64_t cap);
int32_t twtml_create_ranked(int32_t device, int32_t rank, int32_t world_size, const uint8_t* nccl_id,
                            int64_t id_bytes, twtml_ctx** out);
/* One handle that drives several GPUs of this box from ONE host process -- the shape of the reference,
 * which is a single JVM whose SparkContext (LinearRegression.scala:36) spreads every RDD over local[N]
 * worker threads.  The group cuts each micro-batch into n_devices contiguous row blocks, keeps a
 * bit-identical replica of every model on each device and runs the per-iteration exchange between them
 * through peer memory; the caller uses the returned context exactly like a single-GPU one (same
 * entry points; output arrays hold all rows, in the caller's row order).  1 <= n_devices <= 8; when
 * n_devices divides 8 and the batch has a multiple of 8 rows the results equal the single-GPU ones
 * bit for bit.  twtml_predict_stats_static is not offered on groups. */
int32_t twtml_create_all(const int32_t* devices, int32_t n_devices, twtml_ctx** out);
int32_t twtml_device_count(const twtml_ctx* ctx, int32_t* n_devices);
int32_t twtml_destroy(twtml_ctx* ctx);
int32_t twtml_rank(const twtml_ctx* ctx, int32_t* rank, int32_t* world_size);

int32_t twtml_profile_enable(twtml_ctx* ctx, int32_t on);
int32_t twtml_profile_get(const twtml_ctx* ctx, twtml_profile* out);

/* ---- model: StreamingLinearRegressionWithSGD [U], built at LinearRegression.scala:28-32 --- */

/* new StreamingLinearRegressionWithSGD(): stepSize 0.1, numIterations 50, miniBatchFraction
 * 1.0 (class defaults [U]); convergenceTol 0.001, regParam 0.0 (GradientDescent defaults [U]);
 * weights unset until twtml_set_initial_weights. */
int32_t twtml_model_create(twtml_ctx* ctx, int64_t num_features, twtml_model** out);
int32_t twtml_model_free(twtml_model* m);
int32_t twtml_set_step_size(twtml_model* m, double v);          /* .setStepSize          (:30) */
int32_t twtml_set_num_iterations(twtml_model* m, int32_t v);    /* .setNumIterations     (:29) */
int32_t twtml_set_mini_batch_fraction(twtml_model* m, double v);/* .setMiniBatchFraction (:31) */
int32_t twtml_set_convergence_tol(twtml_model* m, double v);    /* .setConvergenceTol [U] */
/* regParam and the updater: the reference path uses SimpleUpdater (updater 0), which ignores
 * regParam [U].  1 = SquaredL2Updater, 2 = L1Updater [U] are offered beyond the reference. */
int32_t twtml_set_reg_param(twtml_model* m, double v);
int32_t twtml_set_updater(twtml_model* m, int32_t kind);
int32_t twtml_set_initial_weights(twtml_model* m, const double* w, int64_t d); /* .setInitialWeights (:32) */
/* latestModel().weights (:57).  TWTML_E_NOT_INITIALIZED if never set. */
int32_t twtml_get_weights(const twtml_model* m, double* out, int64_t d);
int32_t twtml_num_features(const twtml_model* m, int64_t* d);

/* ---- micro-batch ingest --------------------------------------------------------------------- */

/* Raw tweets -> filtered, featurised batch, on the GPU.  Replaces
 * .filter(MllibHelper.filtrate).map(MllibHelper.featurize) (LinearRegression.scala:44-46;
 * MllibHelper.scala:42-56,58-71,73-82,84-94) with HashingTF(num_text_features) [U].
 *  text_utf16/text_off[n+1]: retweetedStatus.getText as UTF-16 code units, ALREADY lower-cased
 *      by the caller (java String.toLowerCase, MllibHelper.scala:43-45), unless
 *      TWTML_FLAG_LOWERCASE asks the library to apply the simple one-to-one Unicode BMP
 *      lower-case map (differs from Java only where Java's mapping is context- or
 *      length-changing: U+0130, final sigma, and supplementary-plane letters).
 *  now_ms: System.currentTimeMillis at MllibHelper.scala:61, injected.
 *  rt_begin/rt_end: numRetweetBegin/End (MllibHelper.scala:15-16), both inclusive.
 * The batch is CSR of width num_text_features + 4; rows failing the filter are dropped,
 * order kept.  n may be 0.
 * Only tweets that pass the filter are featurised (featurize runs after filtrate, :44-46), and the
 * filter does not look at the text.  When at most half of the tweets pass and text_utf16 is
 * page-locked host memory the device can address (cudaHostAlloc / cudaHostRegister), the text column
 * is therefore not copied: the featuriser reads the kept tweets' text where it lies, over the host
 * link.  Same result, byte for byte; pageable text is copied as before. */
#define TWTML_FLAG_LOWERCASE 1u
int32_t twtml_batch_from_tweets(twtml_ctx* ctx, const uint16_t* text_utf16, const int64_t* text_off,
                                const int32_t* retweet_cnt, const int32_t* followers,
                                const int32_t* favourites, const int32_t* friends,
                                const int64_t* created_ms, const uint8_t* is_retweet, int64_t n,
                                int64_t now_ms, int32_t rt_begin, int32_t rt_end,
                                int32_t num_text_features, uint32_t flags, twtml_batch** out);

/* An RDD[LabeledPoint] of SparseVectors, packed by the caller (what trainOn receives at
 * LinearRegression.scala:86).  colidx strictly increasing within a row, < d. */
int32_t twtml_batch_from_csr(twtml_ctx* ctx, int64_t n, int64_t d, const int64_t* rowptr,
                             const int32_t* colidx, const double* vals, const double* labels,
                             twtml_batch** out);

/* The same with DenseVectors: X row-major, row stride ld_x doubles (>= d). */
int32_t twtml_batch_from_dense(twtml_ctx* ctx, int64_t n, int64_t d, const double* X, int64_t ld_x,
                               const double* labels, twtml_batch** out);

/* The same, returning while the rows are still on their way: the copy runs on the library's copy
 * stream and overlaps whatever the context is computing, which is how batch t+1 is ingested under
 * the SGD of batch t (north_star: "packed once into pinned host memory and copied asynchronously";
 * the reference's receiver likewise fills the next RDD while foreachRDD works on the current one,
 * LinearRegression.scala:44-47,53,86).  X and labels must be page-locked host memory and stay
 * untouched until the batch has been used by a compute call or twtml_batch_ready(b, 1, ..) has
 * returned; with pageable memory the call simply behaves like twtml_batch_from_dense.  Works on
 * multi-rank contexts too: the ranks' row counts are then exchanged by the first statistics or training
 * call that uses the batch (the collective calls, which every rank makes in the same order anyway). */
int32_t twtml_batch_from_dense_async(twtml_ctx* ctx, int64_t n, int64_t d, const double* X, int64_t ld_x,
                                     const double* labels, twtml_batch** out);

/* A dense micro-batch that arrives in pieces -- the way a Spark Streaming receiver stores the blocks
 * of one batch interval before the interval's RDD exists (TwitterUtils.createStream ...,
 * LinearRegression.scala:44).  _begin reserves HBM for n rows of d features; every _append copies the
 * next nrows rows (same conventions as twtml_batch_from_dense_async: asynchronous from page-locked
 * memory, which must stay untouched until twtml_batch_ready(b, 1, ..) returns; staged and blocking
 * from pageable memory).  The batch can be used once all n rows have been appended.  A host can thus
 * stream a batch far larger than its own page-locked buffer (2^24 x 1004 rows are 135 GB). */
int32_t twtml_batch_dense_begin(twtml_ctx* ctx, int64_t n, int64_t d, twtml_batch** out);
int32_t twtml_batch_dense_append(twtml_batch* b, int64_t nrows, const double* X, int64_t ld_x,
                                 const double* labels);

/* twtml_batch_from_csr / twtml_batch_from_tweets that return at once: validation, copies, the
 * featuriser and the transpose of wide models run on a worker thread of the context and on its copy
 * stream, beside whatever the context computes (same reference behaviour as above: the receiver fills
 * the RDD of interval t+1 while foreachRDD works on interval t).  The caller's arrays must stay
 * untouched until the batch has been used by a compute call or twtml_batch_ready(b, 1, ..) has
 * returned.  twtml_batch_rows is valid at once (for tweets the host applies MllibHelper.filtrate
 * itself); twtml_batch_info's nnz waits for the featuriser.  A failure of the deferred work is
 * reported by the first call that uses the batch.  Bit-identical to the blocking calls. */
int32_t twtml_batch_from_csr_async(twtml_ctx* ctx, int64_t n, int64_t d, const int64_t* rowptr,
                                   const int32_t* colidx, const double* vals, const double* labels,
                                   twtml_batch** out);
int32_t twtml_batch_from_tweets_async(twtml_ctx* ctx, const uint16_t* text_utf16, const int64_t* text_off,
                                      const int32_t* retweet_cnt, const int32_t* followers,
                                      const int32_t* favourites, const int32_t* friends,
                                      const int64_t* created_ms, const uint8_t* is_retweet, int64_t n,
                                      int64_t now_ms, int32_t rt_begin, int32_t rt_end,
                                      int32_t num_text_features, uint32_t flags, twtml_batch** out);
/* *ready = 1 once the batch's rows (all that were appended so far) are in HBM; wait != 0 blocks
 * until then. */
int32_t twtml_batch_ready(const twtml_batch* b, int32_t wait, int32_t* ready);

/* Synthetic micro-batches generated in HBM (measurement tool for SURVEY.md 8d configs 2-5; no
 * reference counterpart).  Bit-identical to oracle/twtml_oracle.c's generators for the same
 * (seed, global first row). */
int32_t twtml_batch_synth_dense(twtml_ctx* ctx, int64_t n, int64_t d, uint64_t seed, int64_t row0,
                                twtml_batch** out);
int32_t twtml_batch_synth_csr(twtml_ctx* ctx, int64_t n, int32_t d_text, uint64_t seed, int64_t row0,
                              twtml_batch** out);

int32_t twtml_batch_rows(const twtml_batch* b, int64_t* n);          /* rdd.count (:59) */
int32_t twtml_batch_info(const twtml_batch* b, int64_t* n, int64_t* d, int64_t* nnz, int32_t* is_csr);
/* Copy a CSR batch back (featuriser parity checks).  Buffers sized from twtml_batch_info. */
int32_t twtml_batch_export_csr(const twtml_batch* b, int64_t* rowptr, int32_t* colidx, double* vals,
                               double* labels);
/* Copy a dense batch back: X row-major n x d (ld = d), labels n. */
int32_t twtml_batch_export_dense(const twtml_batch* b, double* X, double* labels);
int32_t twtml_batch_free(twtml_batch* b);

/* ---- compute ---------------------------------------------------------------------------------- */

/* predictOnValues / predictOn [U] (mapValues(latestModel.predict)): out_pred[n] = w.x + 0.0,
 * unrounded, row order kept.  TWTML_E_NOT_INITIALIZED before set_initial_weights. */
int32_t twtml_predict(const twtml_model* m, const twtml_batch* b, double* out_pred);

/* The stats foreachRDD of the driver loop (LinearRegression.scala:53-66) with the CURRENT
 * weights: out_real[n] = labels (:61), out_pred_rounded[n] = Utils.round(predict) (:56-58,
 * :62) -- both nullable -- and the reported numbers.  With ranks, out arrays hold this rank's
 * rows and *out holds the statistics of the whole batch. */
int32_t twtml_predict_stats(const twtml_model* m, const twtml_batch* b, double* out_real,
                            double* out_pred_rounded, twtml_stats* out);

/* The statistics of twtml_predict_stats (no per-row arrays) for callers that must not allocate GPU
 * memory.  NO COUNTERPART IN THE REFERENCE: it exists for hosts that have given the whole GPU to
 * their own memory pool, or that cannot afford the device-wide stall a cudaMalloc / cudaFree can
 * cause in the middle of a batch interval.  twtml_predict_stats grows per-context scratch with
 * cudaMalloc on demand; this call uses one fixed array compiled into the library (about 1 MiB,
 * resident on every GPU that loads it) and allocates nothing.  Limits: dense batches on a
 * single-GPU context with at most 16384 row blocks (about 8.4 million rows); anything larger is
 * refused with TWTML_E_INVALID_ARG.  All calls on one device share that array: calls whose GPU work
 * overlaps (two contexts, two threads) race and must be serialised by the caller.  The first call on
 * a device looks up the array's address; make it before starting any stream capture there. */
int32_t twtml_predict_stats_static(const twtml_model* m, const twtml_batch* b, twtml_stats* out);

/* One trainOn step (LinearRegression.scala:86 -> StreamingLinearAlgorithm.trainOn ->
 * GeneralizedLinearAlgorithm.run -> GradientDescent.runMiniBatchSGD [U]) on one micro-batch;
 * the model's weights are replaced by the result.  An empty batch leaves them untouched.
 * loss_history[cap] (nullable) receives stochasticLossHistory [U]; *iters_run the iterations
 * actually executed (convergence can stop early). */
int32_t twtml_train_on_batch(twtml_model* m, const twtml_batch* b, double* loss_history, int32_t cap,
                             int32_t* iters_run);

/* Both output ops of one batch interval in reference order -- stats with the pre-training
 * weights, then training (LinearRegression.scala:53 is registered before :86) -- with the
 * statistics taken from the tile the first SGD iteration loads, so the batch is read once
 * less than by the two calls above.  Same results as predict_stats followed by train. */
int32_t twtml_step(twtml_model* m, const twtml_batch* b, double* out_real, double* out_pred_rounded,
                   twtml_stats* out, double* loss_history, int32_t cap, int32_t* iters_run);

/* ---- the reference's second app: StreamingKMeans on (retweets, followers), KMeans.scala:69-113 ------------- */

typedef struct twtml_kmeans twtml_kmeans; /* StreamingKMeans [U] state: k centres of dimension dim + their weights */

/* new StreamingKMeans().setK(k) (KMeans.scala:69-70); 1 <= k, dim <= 16; single-GPU contexts. */
int32_t twtml_kmeans_create(twtml_ctx* ctx, int32_t k, int32_t dim, twtml_kmeans** out);
int32_t twtml_kmeans_free(twtml_kmeans* m);
int32_t twtml_kmeans_set_decay_factor(twtml_kmeans* m, double a);                          /* .setDecayFactor (:71) */
/* the timeUnit argument of StreamingKMeansModel.update (:98): 0 = "batches", 1 = "points" (discount = decay^points). */
int32_t twtml_kmeans_set_time_unit(twtml_kmeans* m, int32_t time_unit);
/* .setHalfLife(halfLife, timeUnit) (:72): decayFactor = exp(ln(0.5) / halfLife); time_unit 0 = "batches", 1 = "points". */
int32_t twtml_kmeans_set_half_life(twtml_kmeans* m, double half_life, int32_t time_unit);
/* .setInitialCenters(centers, weights): centers k x dim row-major. */
int32_t twtml_kmeans_set_centers(twtml_kmeans* m, const double* centers, const double* weights);
/* .setRandomCenters(dim, weight, seed) (:73): standard-normal centres, all weights = weight.  MLlib draws them from
 * XORShiftRandom(seed); this library uses its own counter-based generator, so the centres differ from a JVM run's. */
int32_t twtml_kmeans_set_random_centers(twtml_kmeans* m, double weight, uint64_t seed);
/* latestModel.clusterCenters / clusterWeights (:101); either output may be null. */
int32_t twtml_kmeans_get_centers(const twtml_kmeans* m, double* centers, double* weights);
/* One batch interval of KMeans.scala:93-105 on n points (X row-major, stride ld >= dim):
 *   standardize != 0: scaledData = new StandardScaler(false, true).fit(rdd).transform(rdd)   (:96; std_out[dim] receives
 *       the fitted column standard deviations, unbiased, 0 for a single point; a column with std 0 scales to 0.0)
 *   model.latestModel.update(scaledData, decayFactor, timeUnit)                              (:98)
 *   assign_out[n] = model.latestModel.predict(scaledData) with the UPDATED centres           (:105)
 * scaled_out[n x dim], std_out, assign_out are nullable.  n == 0 leaves the model untouched (`if (rdd.count > 0)`, :94). */
int32_t twtml_kmeans_update(twtml_kmeans* m, const double* X, int64_t n, int64_t ld, int32_t standardize,
                            double* std_out, double* scaled_out, int32_t* assign_out);
/* latestModel.predict(points): index of the closest centre per point (ties: the lower index). */
int32_t twtml_kmeans_predict(twtml_kmeans* m, const double* X, int64_t n, int64_t ld, int32_t* assign_out);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* TWTML_H */
