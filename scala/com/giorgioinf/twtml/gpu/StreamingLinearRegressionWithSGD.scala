package com.giorgioinf.twtml.gpu

import com.sun.jna.Pointer
import com.sun.jna.ptr.{IntByReference, PointerByReference}
import org.apache.spark.mllib.linalg.{DenseVector, SparseVector, Vector, Vectors}
import org.apache.spark.mllib.regression.{LabeledPoint, LinearRegressionModel}
import org.apache.spark.rdd.RDD
import org.apache.spark.streaming.dstream.DStream
import scala.reflect.ClassTag

/** Source-compatible stand-in for org.apache.spark.mllib.regression.StreamingLinearRegressionWithSGD (MLlib 1.6.1) at
  * the reference's call sites (LinearRegression.scala:28-32,57,86).  Every batch of a DStream is packed PARTITION-WISE on
  * the executors into primitive arrays (no boxing, no per-row objects), the parts are laid end to end in direct
  * ByteBuffers and handed to libtwtml.so with ONE native call per batch: twtml_batch_from_dense when the batch holds
  * DenseVectors, twtml_batch_from_csr when it holds SparseVectors (the reference's featuriser emits those,
  * MllibHelper.scala:76-81).  The weights live on the GPU(s) between batches and are mirrored into `weights` after every
  * training step, so latestModel() returns a genuine MLlib LinearRegressionModel.
  *
  * devices: the GPUs of this box to use; more than one goes through twtml_create_all (one handle, rows split in
  * contiguous blocks, bit-identical model replicas).  The reference runs local[N] -- driver and executors are one JVM --
  * which is what lets the packed parts reach the driver without leaving the process.
  */
class StreamingLinearRegressionWithSGD(private var stepSize: Double, private var numIterations: Int,
    private var miniBatchFraction: Double, private val devices: Array[Int]) extends Serializable {

  def this(stepSize: Double, numIterations: Int, miniBatchFraction: Double) = this(stepSize, numIterations, miniBatchFraction, Array(0))
  def this() = this(0.1, 50, 1.0)                                     // class defaults of MLlib 1.6.1
  def this(devices: Array[Int]) = this(0.1, 50, 1.0, devices)

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
    val d = weights.length
    data.foreachRDD { (rdd: RDD[LabeledPoint], _) =>
      val parts = rdd.mapPartitions(it => Iterator.single(Packer.pack(it.map(p => (p.label, p.features)), d)), true).collect()
      if (parts.exists(_.rows > 0)) trainBatch(parts)                                        // if (!rdd.isEmpty) ... [U]
    }
  }

  def predictOnValues[K: ClassTag](data: DStream[(K, Vector)]): DStream[(K, Double)] = {
    require("prediction")
    val d = weights.length
    data.transform { rdd =>
      val keys = rdd.map(_._1).collect()
      val parts = rdd.mapPartitions(it => Iterator.single(Packer.pack(it.map(kv => (0.0, kv._2)), d)), true).collect()
      val preds = predictBatch(parts)
      rdd.sparkContext.parallelize(keys.zip(preds).toSeq, 1)                                 // mapValues keeps the row order [U]
    }
  }

  def predictOn(data: DStream[Vector]): DStream[Double] = predictOnValues(data.map(v => (0, v))).map(_._2)

  /** The stats foreachRDD of LinearRegression.scala:53-66 in one call: (stats, real[], round(pred)[]). */
  def stats(rdd: RDD[LabeledPoint]): (TwtmlStats, Array[Double], Array[Double]) = {
    require("prediction")
    val d = weights.length
    val parts = rdd.mapPartitions(it => Iterator.single(Packer.pack(it.map(p => (p.label, p.features)), d)), true).collect()
    val n = parts.map(_.rows).sum
    val b = upload(parts)
    try {
      val st = new TwtmlStats
      val real = new Array[Double](n)
      val pred = new Array[Double](n)
      Twtml.check(Twtml.lib.twtml_predict_stats(native(), b, real, pred, st))
      if (st.nonfinite > 0) throw new NumberFormatException("Utils.round on a non-finite prediction")   // Utils.scala:4-5
      (st, real, pred)
    } finally Twtml.lib.twtml_batch_free(b)
  }

  // ---- plumbing -------------------------------------------------------------------------------------------------
  private def require(what: String): Unit =
    if (weights == null) throw new IllegalArgumentException(s"Model must be initialized before starting $what.")

  private def native(): Pointer = {
    if (ctx == null) {
      val p = new PointerByReference
      if (devices.length == 1) Twtml.check(Twtml.lib.twtml_create(devices(0), p))
      else Twtml.check(Twtml.lib.twtml_create_all(devices, devices.length, p))
      ctx = p.getValue
    }
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

  /** The partitions' packed rows, end to end in direct buffers (what the library DMAs or stages from), one native call. */
  private def upload(parts: Array[PackedPart]): Pointer = {
    native()
    val d = weights.length
    val n = parts.map(_.rows.toLong).sum
    val out = new PointerByReference
    val labels = Packer.direct(8L * n).asDoubleBuffer()
    parts.foreach(p => labels.put(p.labels))
    if (parts.forall(_.dense)) {                                   // DenseVectors: row-major n x d, no index arrays at all
      val x = Packer.direct(8L * n * d).asDoubleBuffer()
      parts.foreach(p => x.put(p.vals))
      Twtml.check(Twtml.lib.twtml_batch_from_dense(ctx, n, d.toLong, x, d.toLong, labels, out))
    } else {
      val csr = parts.map(p => if (p.dense) Packer.denseToCsr(p, d) else p)
      val nnz = csr.map(_.vals.length.toLong).sum
      val rowptr = Packer.direct(8L * (n + 1)).asLongBuffer()
      val idx = Packer.direct(4L * nnz).asIntBuffer()
      val vals = Packer.direct(8L * nnz).asDoubleBuffer()
      var base = 0L
      rowptr.put(0L)
      csr.foreach { p =>
        var r = 1
        while (r <= p.rows) { rowptr.put(base + p.rowptr(r)); r += 1 }               // re-based onto the whole batch
        idx.put(p.colidx); vals.put(p.vals)
        base += p.vals.length
      }
      Twtml.check(Twtml.lib.twtml_batch_from_csr(ctx, n, d.toLong, rowptr, idx, vals, labels, out))
    }
    out.getValue
  }

  private def trainBatch(parts: Array[PackedPart]): Unit = {
    val b = upload(parts)
    try {
      val hist = new Array[Double](math.max(numIterations, 1))
      val iters = new IntByReference
      Twtml.check(Twtml.lib.twtml_train_on_batch(native(), b, hist, hist.length, iters))
      lastLossHistory = hist.take(iters.getValue)
      Twtml.check(Twtml.lib.twtml_get_weights(model, weights, weights.length.toLong))
    } finally Twtml.lib.twtml_batch_free(b)
  }

  private def predictBatch(parts: Array[PackedPart]): Array[Double] = {
    val n = parts.map(_.rows).sum
    if (n == 0) return Array.empty
    val b = upload(parts)
    try {
      val out = new Array[Double](n)
      Twtml.check(Twtml.lib.twtml_predict(native(), b, out))
      out
    } finally Twtml.lib.twtml_batch_free(b)
  }
}

/** One partition's rows as primitive arrays.  dense: vals holds rows x d values row-major and rowptr / colidx are empty;
  * otherwise CSR with rowptr starting at 0.  Serializable so that collect() can carry it (in local[N] it never leaves the JVM). */
case class PackedPart(rows: Int, dense: Boolean, labels: Array[Double], rowptr: Array[Long], colidx: Array[Int], vals: Array[Double])

object Packer {
  import java.nio.{ByteBuffer, ByteOrder}
  import scala.collection.mutable.ArrayBuilder

  def direct(bytes: Long): ByteBuffer = ByteBuffer.allocateDirect(math.max(bytes, 8L).toInt).order(ByteOrder.nativeOrder())

  /** Runs on the executor that holds the partition: one pass, specialised builders (ArrayBuilder.ofDouble / ofInt / ofLong
    * append primitives, nothing is boxed).  A partition of DenseVectors stays dense; the first SparseVector switches the
    * partition to CSR (rows seen so far are re-emitted with explicit indices). */
  def pack(it: Iterator[(Double, Vector)], d: Int): PackedPart = {
    val labels = new ArrayBuilder.ofDouble
    val vals = new ArrayBuilder.ofDouble
    val idx = new ArrayBuilder.ofInt
    val rowptr = new ArrayBuilder.ofLong
    var rows = 0
    var nnz = 0L
    var dense = true
    rowptr += 0L
    while (it.hasNext) {
      val (y, x) = it.next()
      labels += y
      x match {
        case dv: DenseVector if dense => vals ++= dv.values
        case v =>
          if (dense) {                                            // first sparse row: give the dense rows so far their indices
            var r = 0
            while (r < rows) { var j = 0; while (j < d) { idx += j; j += 1 }; nnz += d; rowptr += nnz; r += 1 }
            dense = false
          }
          v match {
            case sv: SparseVector => idx ++= sv.indices; vals ++= sv.values; nnz += sv.indices.length
            case dv: DenseVector  => var j = 0; while (j < d) { idx += j; j += 1 }; vals ++= dv.values; nnz += d
          }
          rowptr += nnz
      }
      rows += 1
    }
    if (dense) PackedPart(rows, true, labels.result(), Array.empty, Array.empty, vals.result())
    else PackedPart(rows, false, labels.result(), rowptr.result(), idx.result(), vals.result())
  }

  /** A dense partition inside a batch that also has sparse ones. */
  def denseToCsr(p: PackedPart, d: Int): PackedPart = {
    val rowptr = Array.tabulate[Long](p.rows + 1)(r => r.toLong * d)
    val idx = new Array[Int](p.rows * d)
    var k = 0
    while (k < idx.length) { idx(k) = k % d; k += 1 }
    p.copy(dense = false, rowptr = rowptr, colidx = idx)
  }
}
