// Scala side of the drop-in -- NOT compiled in this image (no JVM). See INTEGRATION.md section 3.
package com.linkedin.photon.ml.function.glm

object NativeGLM {                       // 1:1 with include/photon_ml_b200.h
  System.loadLibrary("photon_ml_b200_jni")
  @native def ctxCreate(device: Int): Long
  @native def datasetCreate(ctx: Long, n: Long, d: Int, rowPtr: Array[Long], col: Array[Int], value: Array[Double],
                            labels: Array[Double], offsets: Array[Double], weights: Array[Double]): Long
  @native def setOffsets(ds: Long, offsets: Array[Double]): Unit
  @native def setWeights(ds: Long, weights: Array[Double]): Unit
  @native def setNormalization(ds: Long, factors: Array[Double], shifts: Array[Double], interceptId: Int): Unit
  @native def valueAndGradient(ds: Long, loss: Int, coef: Array[Double], outGrad: Array[Double]): Double
  @native def hessianVector(ds: Long, loss: Int, coef: Array[Double], tag: Long, v: Array[Double], out: Array[Double]): Unit
  @native def hessianDiagonal(ds: Long, loss: Int, coef: Array[Double], out: Array[Double]): Unit
  @native def datasetDestroy(ds: Long): Unit
  // random-effect coordinates: all entities of a partition in one call (pml_batch_*)
  @native def batchCreate(ctx: Long, entRowPtr: Array[Long], entDim: Array[Int], rowPtr: Array[Long], colLocal: Array[Int],
                          value: Array[Double], labels: Array[Double], offsets: Array[Double], weights: Array[Double]): Long
  @native def batchValueAndGradient(batch: Long, loss: Int, coefs: Array[Double], outValues: Array[Double], outGrads: Array[Double]): Unit
  @native def batchHessianVector(batch: Long, loss: Int, coefs: Array[Double], tag: Long, vs: Array[Double], outHvs: Array[Double]): Unit
  @native def batchSetOffsets(batch: Long, offsets: Array[Double]): Unit
  @native def batchSetActive(batch: Long, active: Array[Byte]): Unit
  @native def batchDestroy(batch: Long): Unit
  // several executors, one GPU each: the merge of treeAggregate over NVLink peer memory (pml_peer_*)
  @native def peerCreate(ctx: Long, rank: Int, world: Int, d: Int): Long
  @native def peerHandle(peer: Long): Array[Byte]                       // 64 bytes, collected by the driver ...
  @native def peerConnect(peer: Long, allHandles: Array[Byte]): Unit    // ... and sent back in rank order
  @native def valueAndGradientAllReduce(ds: Long, peer: Long, loss: Int, coef: Array[Double], outGrad: Array[Double]): Double
  @native def hessianVectorAllReduce(ds: Long, peer: Long, loss: Int, coef: Array[Double], tag: Long, v: Array[Double], out: Array[Double]): Unit
  @native def peerDestroy(peer: Long): Unit
}

/** Drop-in for DistributedGLMLossFunction (api/function/glm/DistributedGLMLossFunction.scala:49-137):
  * same supertype, same Data / Coefficients types, same method signatures. */
protected[ml] class NativeDistributedGLMLossFunction private (loss: Int, sc: SparkContext, depth: Int)
  extends DistributedObjectiveFunction(sc, depth) with TwiceDiffFunction {

  // RDD identity -> packed device handle; first call collects + packs, later calls reuse it.  When only offsets
  // changed (Coordinate.updateModel, lib/algorithm/Coordinate.scala:59-62) the shim calls setOffsets instead.
  private val packed = new java.util.WeakHashMap[RDD[LabeledPoint], java.lang.Long]()
  private def handle(input: RDD[LabeledPoint], norm: Broadcast[NormalizationContext]): Long = ...

  override protected[ml] def calculate(input: RDD[LabeledPoint], coefficients: Broadcast[Vector[Double]],
      normalizationContext: Broadcast[NormalizationContext]): (Double, Vector[Double]) = {
    val g = new Array[Double](coefficients.value.length)
    val f = NativeGLM.valueAndGradient(handle(input, normalizationContext), loss, coefficients.value.toArray, g)
    (f, DenseVector(g))
  }
  override protected[ml] def value(...) = calculate(...)._1          // or outGrad = null: value only
  override protected[ml] def gradient(...) = calculate(...)._2
  override protected[ml] def hessianVector(input, coefficients, multiplyVector, normalizationContext) = {
    val out = new Array[Double](coefficients.value.length)
    // coefficients.id identifies the coefficient vector: TRON's CG loop passes the same Broadcast up to 20 times
    NativeGLM.hessianVector(handle(input, normalizationContext), loss, coefficients.value.toArray, coefficients.id + 1,
      multiplyVector.value.toArray, out)
    DenseVector(out)
  }
  override protected[ml] def hessianDiagonal(input, coefficients) = ...
}

object NativeDistributedGLMLossFunction {
  // same shape as DistributedGLMLossFunction.apply (:151-170): L2 is still mixed in on the Scala side
  def apply(sc: SparkContext, configuration: GLMOptimizationConfiguration, treeAggregateDepth: Int)
           (singleLossFunction: PointwiseLossFunction): NativeDistributedGLMLossFunction = {
    val loss = singleLossFunction match {
      case LogisticLossFunction => 0; case PoissonLossFunction => 1; case SquaredLossFunction => 2 }
    configuration.regularizationContext.regularizationType match {
      case RegularizationType.L2 =>
        new NativeDistributedGLMLossFunction(loss, sc, treeAggregateDepth) with L2RegularizationTwiceDiff {
          l2RegWeight = configuration.regularizationContext.getL2RegularizationWeight(configuration.regularizationWeight) }
      case _ => new NativeDistributedGLMLossFunction(loss, sc, treeAggregateDepth)
    }
  }
}
