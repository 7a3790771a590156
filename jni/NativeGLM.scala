/*
 * Scala side of the drop-in: the natives of jni/photon_ml_jni.c and the objective-function classes that put them
 * behind the reference's own plug-in API (ObjectiveFunction / DiffFunction / TwiceDiffFunction).
 *
 * Not compiled in the build image (no JVM, no Spark jars); tests/test_jni_text.py keeps the @native list below and
 * the C shim in step.  Written against the reference at photon-api/src/main/scala/com/linkedin/photon/ml:
 *   function/glm/DistributedGLMLossFunction.scala:49-170      -> NativeDistributedGLMLossFunction
 *   function/glm/SingleNodeGLMLossFunction.scala:42-160       -> NativeSingleNodeGLMLossFunction
 *   function/svm/DistributedSmoothedHingeLossFunction.scala:36-130, SingleNodeSmoothedHingeLossFunction.scala:34-115
 *                                                              -> the two NativeSmoothedHinge classes
 *   algorithm/RandomEffectCoordinate.scala:108-119            -> NativeRandomEffectSolver (one batched call per partition)
 *
 * Deployment shape: one executor JVM per GPU, one partition of the training RDD per executor (coalesce to the number
 * of GPUs).  An executor keeps ONE NativeExecutor (context, peer handle, packed partitions).  The driver-side objective
 * runs a Spark job whose tasks call the natives on "their" executor; the merge that treeAggregate did on the driver
 * happens between the GPUs (pml_peer_*), so every task returns the finished value and ITS SLICE of the vector.
 */
package com.linkedin.photon.ml.function.glm

import scala.collection.mutable

import breeze.linalg.{DenseVector, Vector}
import org.apache.spark.{SparkContext, TaskContext}
import org.apache.spark.broadcast.Broadcast
import org.apache.spark.rdd.RDD

import com.linkedin.photon.ml.data.LabeledPoint
import com.linkedin.photon.ml.function._
import com.linkedin.photon.ml.function.svm.SmoothedHingeLossFunction
import com.linkedin.photon.ml.normalization.NormalizationContext
import com.linkedin.photon.ml.optimization.RegularizationType
import com.linkedin.photon.ml.optimization.game.GLMOptimizationConfiguration

object NativeGLM {                       // 1:1 with jni/photon_ml_jni.c, which is 1:1 with include/photon_ml_b200.h
  System.loadLibrary("photon_ml_b200_jni")

  val LOGISTIC = 0; val POISSON = 1; val SQUARED = 2; val SMOOTHED_HINGE = 3

  @native def ctxCreate(device: Int): Long
  @native def ctxDestroy(ctx: Long): Unit
  @native def datasetCreate(ctx: Long, n: Long, d: Int, rowPtr: Array[Long], col: Array[Int], value: Array[Double],
                            labels: Array[Double], offsets: Array[Double], weights: Array[Double]): Long
  @native def datasetDestroy(ds: Long): Unit
  @native def setOffsets(ds: Long, offsets: Array[Double]): Unit
  @native def setWeights(ds: Long, weights: Array[Double]): Unit
  @native def setNormalization(ds: Long, factors: Array[Double], shifts: Array[Double], interceptId: Int): Unit
  @native def warmUp(ds: Long, loss: Int, d: Int, n: Long): Unit
  @native def valueAndGradient(ds: Long, loss: Int, coef: Array[Double], outGrad: Array[Double]): Double
  @native def hessianVector(ds: Long, loss: Int, coef: Array[Double], tag: Long, v: Array[Double], out: Array[Double]): Unit
  @native def hessianDiagonal(ds: Long, loss: Int, coef: Array[Double], out: Array[Double]): Unit
  @native def scores(ds: Long, coef: Array[Double], outScores: Array[Double]): Unit
  // random-effect coordinates: all entities of a partition in one call (pml_batch_*)
  @native def batchCreate(ctx: Long, entRowPtr: Array[Long], entDim: Array[Int], rowPtr: Array[Long], colLocal: Array[Int],
                          value: Array[Double], labels: Array[Double], offsets: Array[Double], weights: Array[Double]): Long
  @native def batchDestroy(batch: Long): Unit
  @native def batchValueAndGradient(batch: Long, loss: Int, coefs: Array[Double], outValues: Array[Double],
                                    outGrads: Array[Double]): Unit
  @native def batchHessianVector(batch: Long, loss: Int, coefs: Array[Double], tag: Long, vs: Array[Double],
                                 outHvs: Array[Double]): Unit
  @native def batchSetOffsets(batch: Long, offsets: Array[Double]): Unit
  @native def batchSetActive(batch: Long, active: Array[Byte]): Unit
  // several executors, one GPU each: broadcast + treeAggregate replaced by the exchange over NVLink (pml_peer_*)
  @native def peerCreate(ctx: Long, rank: Int, world: Int, d: Int): Long
  @native def peerHandle(peer: Long): Array[Byte]                       // 64 bytes, collected by the driver ...
  @native def peerConnect(peer: Long, allHandles: Array[Byte]): Unit    // ... and sent back in rank order
  @native def peerSlice(peer: Long, out: Array[Long]): Unit             // out(0) = lo, out(1) = hi
  @native def valueAndGradientAllReduce(ds: Long, peer: Long, loss: Int, coef: Array[Double], outGrad: Array[Double],
                                        gatherOutput: Boolean): Double
  @native def hessianVectorAllReduce(ds: Long, peer: Long, loss: Int, coef: Array[Double], tag: Long, v: Array[Double],
                                     out: Array[Double], gatherOutput: Boolean): Unit
  @native def hessianDiagonalAllReduce(ds: Long, peer: Long, loss: Int, coef: Array[Double], out: Array[Double],
                                       gatherOutput: Boolean): Unit
  @native def peerDestroy(peer: Long): Unit

  def lossId(f: PointwiseLossFunction): Int = f match {
    case LogisticLossFunction => LOGISTIC
    case PoissonLossFunction => POISSON
    case SquaredLossFunction => SQUARED
    case other => throw new IllegalArgumentException(s"no native kernel for $other")
  }
}

/** Flattens LabeledPoints to the CSR arrays datasetCreate takes (lib/data/LabeledPoint.scala:32-36). */
private[glm] final case class Csr(n: Long, d: Int, rowPtr: Array[Long], col: Array[Int], value: Array[Double],
                                  labels: Array[Double], offsets: Array[Double], weights: Array[Double])

private[glm] object Csr {
  def apply(points: Iterable[LabeledPoint]): Csr = {
    val rowPtr = mutable.ArrayBuilder.make[Long]; rowPtr += 0L
    val col = mutable.ArrayBuilder.make[Int]; val value = mutable.ArrayBuilder.make[Double]
    val labels = mutable.ArrayBuilder.make[Double]; val offsets = mutable.ArrayBuilder.make[Double]
    val weights = mutable.ArrayBuilder.make[Double]
    var nnz = 0L; var n = 0L; var d = 0
    points.foreach { p =>
      d = p.features.length
      p.features.activeIterator.foreach { case (j, x) => col += j; value += x; nnz += 1 }   // stored entries, any order
      rowPtr += nnz; labels += p.label; offsets += p.offset; weights += p.weight; n += 1
    }
    Csr(n, d, rowPtr.result(), col.result(), value.result(), labels.result(), offsets.result(), weights.result())
  }
}

/** One per executor JVM: the GPU context, the peer handle, and the partitions packed so far (keyed by RDD id and
  * partition index; a re-offset RDD -- Coordinate.updateModel, lib/algorithm/Coordinate.scala:59-62 -- is recognised
  * by its parent's key and only has its offsets replaced). */
object NativeExecutor {
  private var ctx = 0L
  private var peer = 0L
  private var world = 1
  private val packed = mutable.HashMap.empty[(Int, Int), Long]
  private val normOf = mutable.HashMap.empty[Long, NormalizationContext]

  def context(device: Int): Long = synchronized { if (ctx == 0L) ctx = NativeGLM.ctxCreate(device); ctx }

  def handleFor(device: Int, rank: Int, worldSize: Int, d: Int): Array[Byte] = synchronized {
    if (peer == 0L) { peer = NativeGLM.peerCreate(context(device), rank, worldSize, d); world = worldSize }
    NativeGLM.peerHandle(peer)
  }

  def connect(allHandles: Array[Byte]): Unit = synchronized { NativeGLM.peerConnect(peer, allHandles) }

  def dataset(rddId: Int, offsetsOnlyOf: Option[Int], points: Iterator[LabeledPoint], loss: Int, device: Int): Long =
    synchronized {
      val part = TaskContext.getPartitionId()
      packed.get((rddId, part)) match {
        case Some(ds) => ds
        case None =>
          val ds = offsetsOnlyOf.flatMap(parent => packed.get((parent, part))) match {
            case Some(parentDs) =>                        // same rows, new offsets: no re-pack
              NativeGLM.setOffsets(parentDs, points.map(_.offset).toArray)
              parentDs
            case None =>
              val c = Csr(points.toIterable)
              val created = NativeGLM.datasetCreate(context(device), c.n, c.d, c.rowPtr, c.col, c.value, c.labels,
                c.offsets, c.weights)
              NativeGLM.warmUp(created, loss, c.d, c.n)    // lazy buffers exist before arrays get pinned
              created
          }
          packed((rddId, part)) = ds
          ds
      }
    }

  def normalize(ds: Long, norm: NormalizationContext): Unit = synchronized {
    if (!normOf.get(ds).exists(_ eq norm)) {
      NativeGLM.setNormalization(ds, norm.factors.map(_.toArray).orNull, norm.shifts.map(_.toArray).orNull,
        norm.interceptId.getOrElse(-1))
      normOf(ds) = norm
    }
  }

  def peerHandleValue: Long = peer
  def slice: (Int, Int) = { val s = new Array[Long](2); NativeGLM.peerSlice(peer, s); (s(0).toInt, s(1).toInt) }

  def release(rddId: Int): Unit = synchronized {
    packed.keys.filter(_._1 == rddId).toList.foreach { k => packed.remove(k).foreach(NativeGLM.datasetDestroy) }
  }
}

/** What one task hands back: the objective value (the same on every rank) and this rank's slice of the vector. */
private[glm] final case class Slice(lo: Int, values: Array[Double], value: Double)

private[glm] object Assemble {
  def apply(d: Int, parts: Array[Slice]): (Double, Vector[Double]) = {
    val out = new Array[Double](d)
    parts.foreach(s => System.arraycopy(s.values, 0, out, s.lo, s.values.length))
    (parts.head.value, DenseVector(out))
  }
}

/** Drop-in for DistributedGLMLossFunction: same supertype, same Data / Coefficients types, same signatures. */
protected[ml] class NativeDistributedGLMLossFunction private[glm] (loss: Int, sc: SparkContext, treeAggregateDepth: Int)
  extends DistributedObjectiveFunction(sc, treeAggregateDepth) with TwiceDiffFunction {

  // one GPU per executor: executor i uses device (i mod GPUs per host); the mapping is deployment configuration
  protected def deviceOf(partition: Int): Int = partition % sc.getConf.getInt("spark.photon.gpusPerHost", 8)

  private def evaluate(input: RDD[LabeledPoint], normalizationContext: Broadcast[NormalizationContext])
                      (body: (Long, Long) => Slice): Array[Slice] = {
    val id = input.id
    val parent = input.dependencies.headOption.map(_.rdd.id)          // a re-offset copy points at the packed RDD
    val lossId = loss
    val dev: Int => Int = deviceOf
    input.mapPartitionsWithIndex { (part, points) =>
      val ds = NativeExecutor.dataset(id, parent, points, lossId, dev(part))
      NativeExecutor.normalize(ds, normalizationContext.value)
      Iterator.single(body(ds, NativeExecutor.peerHandleValue))
    }.collect()
  }

  override protected[ml] def calculate(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): (Double, Vector[Double]) = {
    val lossId = loss
    val parts = evaluate(input, normalizationContext) { (ds, peer) =>
      val w = coefficients.value.toArray
      val g = new Array[Double](w.length)
      val f = NativeGLM.valueAndGradientAllReduce(ds, peer, lossId, w, g, gatherOutput = false)
      val (lo, hi) = NativeExecutor.slice
      Slice(lo, java.util.Arrays.copyOfRange(g, lo, hi), f)
    }
    Assemble(coefficients.value.length, parts)
  }

  override protected[ml] def value(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): Double =
    calculate(input, coefficients, normalizationContext)._1

  override protected[ml] def gradient(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): Vector[Double] =
    calculate(input, coefficients, normalizationContext)._2

  override protected[ml] def hessianVector(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      multiplyVector: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): Vector[Double] = {
    val lossId = loss
    // the Broadcast's id names the coefficient vector: TRON's CG loop passes the same one up to 20 times
    // (lib/optimization/TRON.scala:294-301), and the kernels then reuse the cached curvature of every row
    val tag = coefficients.id + 1
    val parts = evaluate(input, normalizationContext) { (ds, peer) =>
      val w = coefficients.value.toArray
      val out = new Array[Double](w.length)
      NativeGLM.hessianVectorAllReduce(ds, peer, lossId, w, tag, multiplyVector.value.toArray, out, gatherOutput = false)
      val (lo, hi) = NativeExecutor.slice
      Slice(lo, java.util.Arrays.copyOfRange(out, lo, hi), 0.0)
    }
    Assemble(coefficients.value.length, parts)._2
  }

  override protected[ml] def hessianDiagonal(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]]): Vector[Double] = {
    val id = input.id
    val parent = input.dependencies.headOption.map(_.rdd.id)
    val lossId = loss
    val dev: Int => Int = deviceOf
    val parts = input.mapPartitionsWithIndex { (part, points) =>
      val ds = NativeExecutor.dataset(id, parent, points, lossId, dev(part))
      val w = coefficients.value.toArray
      val out = new Array[Double](w.length)
      NativeGLM.hessianDiagonalAllReduce(ds, NativeExecutor.peerHandleValue, lossId, w, out, gatherOutput = false)
      val (lo, hi) = NativeExecutor.slice
      Iterator.single(Slice(lo, java.util.Arrays.copyOfRange(out, lo, hi), 0.0))
    }.collect()
    Assemble(coefficients.value.length, parts)._2
  }
}

object NativeDistributedGLMLossFunction {
  // same shape as DistributedGLMLossFunction.apply (:151-170): L2 is still mixed in on the Scala side, and only for
  // RegularizationType.L2 (ELASTIC_NET gets no L2 term there either)
  def apply(sc: SparkContext, configuration: GLMOptimizationConfiguration, treeAggregateDepth: Int)
           (singleLossFunction: PointwiseLossFunction): NativeDistributedGLMLossFunction = {
    val loss = NativeGLM.lossId(singleLossFunction)
    val regularizationContext = configuration.regularizationContext
    regularizationContext.regularizationType match {
      case RegularizationType.L2 =>
        new NativeDistributedGLMLossFunction(loss, sc, treeAggregateDepth) with L2RegularizationTwiceDiff {
          l2RegWeight = regularizationContext.getL2RegularizationWeight(configuration.regularizationWeight)
        }
      case _ => new NativeDistributedGLMLossFunction(loss, sc, treeAggregateDepth)
    }
  }
}

/** Drop-in for DistributedSmoothedHingeLossFunction (function/svm/DistributedSmoothedHingeLossFunction.scala:36-103):
  * a DiffFunction only, and it ignores the normalization context, as the reference's calculateAt does. */
protected[ml] class NativeDistributedSmoothedHingeLossFunction private[glm] (sc: SparkContext, treeAggregateDepth: Int)
  extends DistributedObjectiveFunction(sc, treeAggregateDepth) with DiffFunction {

  protected def deviceOf(partition: Int): Int = partition % sc.getConf.getInt("spark.photon.gpusPerHost", 8)

  override protected[ml] def calculate(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): (Double, Vector[Double]) = {
    val id = input.id
    val parent = input.dependencies.headOption.map(_.rdd.id)
    val dev: Int => Int = deviceOf
    val parts = input.mapPartitionsWithIndex { (part, points) =>
      val ds = NativeExecutor.dataset(id, parent, points, NativeGLM.SMOOTHED_HINGE, dev(part))
      val w = coefficients.value.toArray
      val g = new Array[Double](w.length)
      val f = NativeGLM.valueAndGradientAllReduce(ds, NativeExecutor.peerHandleValue, NativeGLM.SMOOTHED_HINGE, w, g,
        gatherOutput = false)
      val (lo, hi) = NativeExecutor.slice
      Iterator.single(Slice(lo, java.util.Arrays.copyOfRange(g, lo, hi), f))
    }.collect()
    Assemble(coefficients.value.length, parts)
  }

  override protected[ml] def value(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): Double =
    calculate(input, coefficients, normalizationContext)._1

  override protected[ml] def gradient(
      input: RDD[LabeledPoint],
      coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): Vector[Double] =
    calculate(input, coefficients, normalizationContext)._2
}

object NativeDistributedSmoothedHingeLossFunction {
  def apply(sc: SparkContext, configuration: GLMOptimizationConfiguration, treeAggregateDepth: Int)
      : NativeDistributedSmoothedHingeLossFunction = {
    val regularizationContext = configuration.regularizationContext
    regularizationContext.regularizationType match {
      case RegularizationType.L2 =>
        new NativeDistributedSmoothedHingeLossFunction(sc, treeAggregateDepth) with L2RegularizationDiff {
          l2RegWeight = regularizationContext.getL2RegularizationWeight(configuration.regularizationWeight)
        }
      case _ => new NativeDistributedSmoothedHingeLossFunction(sc, treeAggregateDepth)
    }
  }
}

/** Drop-in for SingleNodeGLMLossFunction (function/glm/SingleNodeGLMLossFunction.scala:42-125): Data is an
  * Iterable[LabeledPoint] held by ONE JVM, Coefficients a plain Vector.  The Iterable is packed on first sight
  * (identity-keyed) on this JVM's GPU; no exchange is involved. */
protected[ml] class NativeSingleNodeGLMLossFunction private[glm] (loss: Int, device: Int)
  extends SingleNodeObjectiveFunction with TwiceDiffFunction {

  private val packed = new java.util.IdentityHashMap[Iterable[LabeledPoint], java.lang.Long]()
  private val normOf = mutable.HashMap.empty[Long, NormalizationContext]
  private var lastCoef: Vector[Double] = _
  private var tag = 0L

  private def handle(input: Iterable[LabeledPoint], norm: Broadcast[NormalizationContext]): Long = synchronized {
    var ds = packed.get(input)
    if (ds == null) {
      val c = Csr(input)
      ds = NativeGLM.datasetCreate(NativeExecutor.context(device), c.n, c.d, c.rowPtr, c.col, c.value, c.labels,
        c.offsets, c.weights)
      NativeGLM.warmUp(ds, loss, c.d, c.n)
      packed.put(input, ds)
    }
    if (norm != null && !normOf.get(ds).exists(_ eq norm.value)) {
      val n = norm.value
      NativeGLM.setNormalization(ds, n.factors.map(_.toArray).orNull, n.shifts.map(_.toArray).orNull,
        n.interceptId.getOrElse(-1))
      normOf(ds) = n
    }
    ds
  }

  override protected[ml] def calculate(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): (Double, Vector[Double]) = {
    val g = new Array[Double](coefficients.length)
    val f = NativeGLM.valueAndGradient(handle(input, normalizationContext), loss, coefficients.toArray, g)
    (f, DenseVector(g))
  }

  override protected[ml] def value(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): Double =
    NativeGLM.valueAndGradient(handle(input, normalizationContext), loss, coefficients.toArray, null)

  override protected[ml] def gradient(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): Vector[Double] =
    calculate(input, coefficients, normalizationContext)._2

  override protected[ml] def hessianVector(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      multiplyVector: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): Vector[Double] = {
    if (!(coefficients eq lastCoef)) { lastCoef = coefficients; tag += 1 }   // TRON's CG loop passes the same object
    val out = new Array[Double](coefficients.length)
    NativeGLM.hessianVector(handle(input, normalizationContext), loss, coefficients.toArray, tag, multiplyVector.toArray, out)
    DenseVector(out)
  }

  override protected[ml] def hessianDiagonal(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double]): Vector[Double] = {
    val out = new Array[Double](coefficients.length)
    NativeGLM.hessianDiagonal(handle(input, null), loss, coefficients.toArray, out)
    DenseVector(out)
  }

  def close(): Unit = synchronized {
    val it = packed.values().iterator()
    while (it.hasNext) NativeGLM.datasetDestroy(it.next())
    packed.clear(); normOf.clear()
  }
}

object NativeSingleNodeGLMLossFunction {
  // same shape as SingleNodeGLMLossFunction.apply (:139-160)
  def apply(configuration: GLMOptimizationConfiguration, device: Int = 0)
           (singleLossFunction: PointwiseLossFunction): NativeSingleNodeGLMLossFunction = {
    val loss = NativeGLM.lossId(singleLossFunction)
    val regularizationContext = configuration.regularizationContext
    regularizationContext.regularizationType match {
      case RegularizationType.L2 =>
        new NativeSingleNodeGLMLossFunction(loss, device) with L2RegularizationTwiceDiff {
          l2RegWeight = regularizationContext.getL2RegularizationWeight(configuration.regularizationWeight)
        }
      case _ => new NativeSingleNodeGLMLossFunction(loss, device)
    }
  }
}

/** Drop-in for SingleNodeSmoothedHingeLossFunction (function/svm/SingleNodeSmoothedHingeLossFunction.scala:34-115). */
protected[ml] class NativeSingleNodeSmoothedHingeLossFunction private[glm] (device: Int)
  extends SingleNodeObjectiveFunction with DiffFunction {

  private val inner = new NativeSingleNodeGLMLossFunction(NativeGLM.SMOOTHED_HINGE, device)

  override protected[ml] def calculate(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): (Double, Vector[Double]) =
    inner.calculate(input, coefficients, null)        // the hinge objective ignores normalization

  override protected[ml] def value(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): Double =
    inner.value(input, coefficients, null)

  override protected[ml] def gradient(
      input: Iterable[LabeledPoint],
      coefficients: Vector[Double],
      normalizationContext: Broadcast[NormalizationContext]): Vector[Double] =
    inner.gradient(input, coefficients, null)
}

object NativeSingleNodeSmoothedHingeLossFunction {
  def apply(configuration: GLMOptimizationConfiguration, device: Int = 0): NativeSingleNodeSmoothedHingeLossFunction = {
    val regularizationContext = configuration.regularizationContext
    regularizationContext.regularizationType match {
      case RegularizationType.L2 =>
        new NativeSingleNodeSmoothedHingeLossFunction(device) with L2RegularizationDiff {
          l2RegWeight = regularizationContext.getL2RegularizationWeight(configuration.regularizationWeight)
        }
      case _ => new NativeSingleNodeSmoothedHingeLossFunction(device)
    }
  }
}

/** The random-effect counterpart (RandomEffectCoordinate.updateModel, algorithm/RandomEffectCoordinate.scala:108-119):
  * instead of one SingleNodeOptimizationProblem.run per entity inside mapValues, the entities of a partition are
  * packed as ONE block-diagonal batch and every optimizer step evaluates all of them in one call.  The per-entity
  * optimizer then runs in lock step over `coefs` (the entities' coefficient vectors, concatenated); entities that
  * have converged are switched off with batchSetActive and stop costing bandwidth. */
final class NativeRandomEffectSolver(device: Int, loss: Int, entities: IndexedSeq[Iterable[LabeledPoint]]) {
  private val parts = entities.map(Csr(_))
  val entDim: Array[Int] = parts.map(_.d).toArray
  val coefPtr: Array[Long] = entDim.scanLeft(0L)(_ + _)
  private val entRowPtr = parts.scanLeft(0L)(_ + _.n).toArray
  private val rowPtr = {
    val b = mutable.ArrayBuilder.make[Long]; b += 0L; var base = 0L
    parts.foreach { c => c.rowPtr.drop(1).foreach(p => b += base + p); base += c.rowPtr.last }
    b.result()
  }
  private val batch = NativeGLM.batchCreate(NativeExecutor.context(device), entRowPtr, entDim, rowPtr,
    parts.flatMap(_.col).toArray, parts.flatMap(_.value).toArray, parts.flatMap(_.labels).toArray,
    parts.flatMap(_.offsets).toArray, parts.flatMap(_.weights).toArray)

  /** values(e), grads(coefPtr(e) until coefPtr(e + 1)) for every entity e */
  def valueAndGradient(coefs: Array[Double], values: Array[Double], grads: Array[Double]): Unit =
    NativeGLM.batchValueAndGradient(batch, loss, coefs, values, grads)

  def hessianVector(coefs: Array[Double], tag: Long, vs: Array[Double], out: Array[Double]): Unit =
    NativeGLM.batchHessianVector(batch, loss, coefs, tag, vs, out)

  def setOffsets(offsets: Array[Double]): Unit = NativeGLM.batchSetOffsets(batch, offsets)
  def setActive(active: Array[Byte]): Unit = NativeGLM.batchSetActive(batch, active)
  def close(): Unit = NativeGLM.batchDestroy(batch)
}
