package com.giorgioinf.twtml.gpu

import com.sun.jna.{Library, Native, Pointer, Structure}
import com.sun.jna.ptr.{IntByReference, LongByReference, PointerByReference}

/** include/twtml.h, one method per entry point used by the facade; every call returns 0 or a negative twtml_status. */
trait TwtmlNative extends Library {
  def twtml_abi_version(): Int
  def twtml_last_error(buf: Array[Byte], cap: Long): Int
  def twtml_create(device: Int, out: PointerByReference): Int
  def twtml_create_all(devices: Array[Int], nDevices: Int, out: PointerByReference): Int
  def twtml_device_count(ctx: Pointer, n: IntByReference): Int
  def twtml_nccl_unique_id(out: Array[Byte], cap: Long): Int
  def twtml_create_ranked(device: Int, rank: Int, world: Int, id: Array[Byte], idBytes: Long, out: PointerByReference): Int
  def twtml_destroy(ctx: Pointer): Int
  def twtml_model_create(ctx: Pointer, numFeatures: Long, out: PointerByReference): Int
  def twtml_model_free(m: Pointer): Int
  def twtml_set_step_size(m: Pointer, v: Double): Int
  def twtml_set_num_iterations(m: Pointer, v: Int): Int
  def twtml_set_mini_batch_fraction(m: Pointer, v: Double): Int
  def twtml_set_convergence_tol(m: Pointer, v: Double): Int
  def twtml_set_initial_weights(m: Pointer, w: Array[Double], d: Long): Int
  def twtml_get_weights(m: Pointer, out: Array[Double], d: Long): Int
  def twtml_batch_from_tweets(ctx: Pointer, textUtf16: Array[Char], textOff: Array[Long], retweetCnt: Array[Int],
      followers: Array[Int], favourites: Array[Int], friends: Array[Int], createdMs: Array[Long], isRetweet: Array[Byte],
      n: Long, nowMs: Long, rtBegin: Int, rtEnd: Int, numTextFeatures: Int, flags: Int, out: PointerByReference): Int
  // the packed batch travels in direct NIO buffers (JNA passes a direct Buffer as the address of its memory: no copy on
  // the way into the call)
  def twtml_batch_from_csr(ctx: Pointer, n: Long, d: Long, rowptr: java.nio.LongBuffer, colidx: java.nio.IntBuffer,
      vals: java.nio.DoubleBuffer, labels: java.nio.DoubleBuffer, out: PointerByReference): Int
  def twtml_batch_from_dense(ctx: Pointer, n: Long, d: Long, x: java.nio.DoubleBuffer, ldX: Long, labels: java.nio.DoubleBuffer,
      out: PointerByReference): Int
  def twtml_batch_from_csr_async(ctx: Pointer, n: Long, d: Long, rowptr: java.nio.LongBuffer, colidx: java.nio.IntBuffer,
      vals: java.nio.DoubleBuffer, labels: java.nio.DoubleBuffer, out: PointerByReference): Int
  def twtml_batch_from_dense_async(ctx: Pointer, n: Long, d: Long, x: java.nio.DoubleBuffer, ldX: Long, labels: java.nio.DoubleBuffer,
      out: PointerByReference): Int
  def twtml_batch_ready(b: Pointer, waitFor: Int, ready: IntByReference): Int
  def twtml_batch_rows(b: Pointer, n: LongByReference): Int
  def twtml_batch_free(b: Pointer): Int
  def twtml_predict(m: Pointer, b: Pointer, outPred: Array[Double]): Int
  def twtml_predict_stats(m: Pointer, b: Pointer, outReal: Array[Double], outPredRounded: Array[Double], out: TwtmlStats): Int
  def twtml_train_on_batch(m: Pointer, b: Pointer, lossHistory: Array[Double], cap: Int, itersRun: IntByReference): Int
  def twtml_step(m: Pointer, b: Pointer, outReal: Array[Double], outPredRounded: Array[Double], out: TwtmlStats,
      lossHistory: Array[Double], cap: Int, itersRun: IntByReference): Int
}

/** struct twtml_stats, field order as in the header. */
class TwtmlStats extends Structure {
  var batch: Long = 0L
  var mse: Double = 0.0
  var real_stdev: Double = 0.0
  var pred_stdev: Double = 0.0
  var mse_raw: Double = 0.0
  var real_stdev_raw: Double = 0.0
  var pred_stdev_raw: Double = 0.0
  var real_mean: Double = 0.0
  var pred_mean: Double = 0.0
  var nonfinite: Long = 0L
  override protected def getFieldOrder: java.util.List[String] = java.util.Arrays.asList("batch", "mse", "real_stdev",
    "pred_stdev", "mse_raw", "real_stdev_raw", "pred_stdev_raw", "real_mean", "pred_mean", "nonfinite")
}

object Twtml {
  lazy val lib: TwtmlNative = Native.loadLibrary("twtml", classOf[TwtmlNative]).asInstanceOf[TwtmlNative]

  /** 0 -> ok; TWTML_E_NOT_INITIALIZED (-2) -> MLlib's own IllegalArgumentException and message; else RuntimeException. */
  def check(rc: Int): Unit = if (rc != 0) {
    val buf = new Array[Byte](512)
    lib.twtml_last_error(buf, 512L)
    val msg = new String(buf.takeWhile(_ != 0), "UTF-8")
    if (rc == -2) throw new IllegalArgumentException(msg) else throw new RuntimeException(s"libtwtml $rc: $msg")
  }
}
