package com.giorgioinf.twtml.gpu

import com.sun.jna.Pointer
import com.sun.jna.ptr.{IntByReference, PointerByReference}
import org.apache.spark.mllib.linalg.{DenseVector, SparseVector, Vector, Vectors}
import org.apache.spark.mllib.regression.{LabeledPoint, LinearRegressionModel}
import org.apache.spark.rdd.RDD
import org.apache.spark.streaming.dstream.DStream
import scala.reflect.ClassTag

/** Source-compatible stand-in for org.apache.spark.mllib.regression.StreamingLinearRegressionWithSGD (MLlib 1.6.1) at
  * the reference's call sites (LinearRegression.scala:28-32,57,86).  The batches of a DStream are collected on the
  * driver (the reference runs local[N]: driver and executors share one JVM), packed as CSR and handed to libtwtml.so;
  * the weights live on the GPU between batches and are mirrored into `weights` after every training step, so
  * latestModel() returns a genuine MLlib LinearRegressionModel.
  */
class StreamingLinearRegressionWithSGD(private var stepSize: Double, private var numIterations: Int,
    private var miniBatchFraction: Double) extends Serializable {

  def this() = this(0.1, 50, 1.0)                                     // class defaults of MLlib 1.6.1

  private var convergenceTol: Double = 0.001                          // GradientDescent default
  private var weights: Array[Double] = null
  @transient private var ctx: Pointer = null
  @transient private var model: Pointer = null
  @transient var lastLossHistory: Array[Double] = Array.empty

  def setStepSize(v: Double): this.type = { stepSize = v; push(); this }                     // LinearRegression.scala:30
  def setNumIterations(v: Int): this.type = { numIterations = v; push(); this }              // :29
  def setMiniBatchFraction(v: Double): this.type = { miniBatchFraction = v; push(); this }   // :31
  def setConvergenceTol(v: Double): this.type = { convergenceTol = v; push(); this }
  def setInitialWeights(w: Vector): this.type = {                                            // :32
    weights = w.toArray.clone()
    if (model != null) { Twtml.lib.twtml_model_free(model); model = null }
    this
  }

  def latestModel(): LinearRegressionModel = {                                               // :57
    require("prediction")
    new LinearRegressionModel(Vectors.dense(weights.clone()), 0.0)
  }

  /** StreamingLinearAlgorithm.trainOn: registers the output operation and returns (:86). */
  def trainOn(data: DStream[LabeledPoint]): Unit = {
    require("training")
    data.foreachRDD { (rdd: RDD[LabeledPoint], _) => if (!rdd.isEmpty) trainBatch(rdd.collect()) }
  }

  def predictOnValues[K: ClassTag](data: DStream[(K, Vector)]): DStream[(K, Double)] = {
    require("prediction")
    data.transform { rdd =>
      val kv = rdd.collect()
      val preds = predictBatch(kv.map(_._2))
      rdd.sparkContext.parallelize(kv.map(_._1).zip(preds).toSeq, 1)
    }
  }

  def predictOn(data: DStream[Vector]): DStream[Double] = predictOnValues(data.map(v => (0, v))).map(_._2)

  /** The stats foreachRDD of LinearRegression.scala:53-66 in one call: (stats, real[], round(pred)[]). */
  def stats(points: Array[LabeledPoint]): (TwtmlStats, Array[Double], Array[Double]) = {
    require("prediction")
    val b = pack(points.map(_.features), points.map(_.label))
    try {
      val st = new TwtmlStats
      val real = new Array[Double](points.length)
      val pred = new Array[Double](points.length)
      Twtml.check(Twtml.lib.twtml_predict_stats(native(), b, real, pred, st))
      if (st.nonfinite > 0) throw new NumberFormatException("Utils.round on a non-finite prediction")   // Utils.scala:4-5
      (st, real, pred)
    } finally Twtml.lib.twtml_batch_free(b)
  }

  // ---- plumbing -------------------------------------------------------------------------------------------------
  private def require(what: String): Unit =
    if (weights == null) throw new IllegalArgumentException(s"Model must be initialized before starting $what.")

  private def native(): Pointer = {
    if (ctx == null) { val p = new PointerByReference; Twtml.check(Twtml.lib.twtml_create(0, p)); ctx = p.getValue }
    if (model == null) {
      val p = new PointerByReference
      Twtml.check(Twtml.lib.twtml_model_create(ctx, weights.length.toLong, p))
      model = p.getValue
      Twtml.check(Twtml.lib.twtml_set_initial_weights(model, weights, weights.length.toLong))
      push()
    }
    model
  }

  private def push(): Unit = if (model != null) {
    Twtml.check(Twtml.lib.twtml_set_step_size(model, stepSize))
    Twtml.check(Twtml.lib.twtml_set_num_iterations(model, numIterations))
    Twtml.check(Twtml.lib.twtml_set_mini_batch_fraction(model, miniBatchFraction))
    Twtml.check(Twtml.lib.twtml_set_convergence_tol(model, convergenceTol))
  }

  /** Vectors -> CSR (dense vectors become full rows); labels alongside. */
  private def pack(xs: Array[Vector], labels: Array[Double]): Pointer = {
    native()
    val d = weights.length
    val rowptr = new Array[Long](xs.length + 1)
    val idx = Array.newBuilder[Int]
    val vals = Array.newBuilder[Double]
    var nnz = 0L
    var r = 0
    while (r < xs.length) {
      xs(r) match {
        case sv: SparseVector => idx ++= sv.indices; vals ++= sv.values; nnz += sv.indices.length
        case dv: DenseVector  => idx ++= (0 until d); vals ++= dv.values; nnz += d
      }
      r += 1
      rowptr(r) = nnz
    }
    val out = new PointerByReference
    Twtml.check(Twtml.lib.twtml_batch_from_csr(ctx, xs.length.toLong, d.toLong, rowptr, idx.result(), vals.result(), labels, out))
    out.getValue
  }

  private def trainBatch(points: Array[LabeledPoint]): Unit = {
    val b = pack(points.map(_.features), points.map(_.label))
    try {
      val hist = new Array[Double](math.max(numIterations, 1))
      val iters = new IntByReference
      Twtml.check(Twtml.lib.twtml_train_on_batch(native(), b, hist, hist.length, iters))
      lastLossHistory = hist.take(iters.getValue)
      Twtml.check(Twtml.lib.twtml_get_weights(model, weights, weights.length.toLong))
    } finally Twtml.lib.twtml_batch_free(b)
  }

  private def predictBatch(xs: Array[Vector]): Array[Double] = {
    if (xs.isEmpty) return Array.empty
    val b = pack(xs, new Array[Double](xs.length))
    try {
      val out = new Array[Double](xs.length)
      Twtml.check(Twtml.lib.twtml_predict(native(), b, out))
      out
    } finally Twtml.lib.twtml_batch_free(b)
  }
}
