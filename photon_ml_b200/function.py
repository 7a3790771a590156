"""Host-side mirror of the reference's objective-function interface, backed by the CUDA library.

Same names, argument meaning and error behaviour as
  lib/function/ObjectiveFunction.scala:25-73, DiffFunction.scala:25-52, TwiceDiffFunction.scala:25-52,
  api/function/glm/DistributedGLMLossFunction.scala:49-171, SingleNodeGLMLossFunction.scala:42-153,
  api/function/svm/DistributedSmoothedHingeLossFunction.scala:84-103,
  lib/function/L2Regularization.scala:25-184, lib/normalization/NormalizationContext.scala:39-111
(paths under photon-lib / photon-api src/main/scala/com/linkedin/photon/ml).  `Data` is a packed
photon_ml_b200.Dataset instead of an RDD / Iterable of LabeledPoint; `Coefficients` is a numpy
vector.  In a JVM deployment these classes are the Scala shim of INTEGRATION.md; here they let the
tests and the optimizer drivers read like the reference's.
"""
import numpy as np

from . import abi


class NormalizationContext:
    """(factors, shifts, interceptId); NormalizationContext.scala:39-47."""

    def __init__(self, factors=None, shifts=None, interceptId=None):
        if shifts is not None and interceptId is None:
            raise abi.IllegalArgumentError("Shift without intercept is illegal.")
        if factors is not None and shifts is not None and len(factors) != len(shifts):
            raise abi.IllegalArgumentError("Factors and shifts vectors should have the same size")
        self.factors = None if factors is None else np.asarray(factors, dtype=np.float64)
        self.shifts = None if shifts is None else np.asarray(shifts, dtype=np.float64)
        self.interceptId = interceptId

    def modelToOriginalSpace(self, coef):   # :64-80
        if self.factors is None and self.shifts is None and not isinstance(coef, np.ndarray):
            return coef.clone()              # device tensors pass through an empty context untouched
        out = np.array(coef, dtype=np.float64)
        if self.factors is not None:
            out = out * self.factors
        if self.shifts is not None:
            out[self.interceptId] -= out.dot(self.shifts)
        return out

    def modelToTransformedSpace(self, coef):  # :97-111
        if self.factors is None and self.shifts is None and not isinstance(coef, np.ndarray):
            return coef.clone()
        out = np.array(coef, dtype=np.float64)
        if self.shifts is not None:
            out[self.interceptId] += out.dot(self.shifts)
        if self.factors is not None:
            out = out / self.factors
        return out


def NoNormalization():
    return NormalizationContext(None, None, None)


import itertools

# Hessian-vector cache tags: one process-wide counter, so that two objectives evaluating the same data set can
# never present the same tag for different coefficients (a data set remembers only the last tag it saw)
_TAGS = itertools.count(1)


def _apply_normalization(data, ctx):
    """The reference passes the NormalizationContext on every call; the packed data set keeps what was set last.  The
    record of what is set lives ON THE DATA SET (not in the objective, not keyed by id()): several objectives may
    take turns on one data set, and a new data set must never inherit another one's record."""
    empty = ctx.factors is None and ctx.shifts is None
    applied = getattr(data, "_applied_norm", None)
    if empty:
        if applied is not None:
            data.set_normalization(None, None, None)     # clear whatever an earlier caller left behind
        return
    if applied is not ctx:
        data.set_normalization(ctx.factors, ctx.shifts, ctx.interceptId)
        data._applied_norm = ctx


class _NativeObjective:
    """value / gradient / calculate (+ hessianVector / hessianDiagonal) on a packed Dataset."""

    twice_differentiable = True

    def __init__(self, loss):
        self.loss = loss
        self._tags = {}

    def _apply(self, data, ctx):
        _apply_normalization(data, ctx)

    def domainDimension(self, data):
        return data.d

    def value(self, data, coefficients, normalizationContext):
        return self.calculate(data, coefficients, normalizationContext, want_gradient=False)[0]

    def gradient(self, data, coefficients, normalizationContext):
        return self.calculate(data, coefficients, normalizationContext)[1]

    def calculate(self, data, coefficients, normalizationContext, want_gradient=True):
        self._apply(data, normalizationContext)
        return data.value_and_gradient(self.loss, coefficients, want_gradient=want_gradient)

    def hessianVector(self, data, coefficients, multiplyVector, normalizationContext):
        if not self.twice_differentiable:
            raise abi.IllegalArgumentError("not a TwiceDiffFunction")
        self._apply(data, normalizationContext)
        # the coefficient vector's identity is the cache tag (TRON's CG loop keeps passing the same one)
        key = id(coefficients)
        if self._tags.get("key") != key or not np.array_equal(self._tags.get("copy"), coefficients):
            self._tags = {"key": key, "copy": np.array(coefficients, copy=True), "tag": next(_TAGS)}
        return data.hessian_vector(self.loss, coefficients, multiplyVector, tag=self._tags["tag"])

    def hessianDiagonal(self, data, coefficients):
        if not self.twice_differentiable:
            raise abi.IllegalArgumentError("not a TwiceDiffFunction")
        return data.hessian_diagonal(self.loss, coefficients)


class L2RegularizationMixin:
    """L2Regularization / L2RegularizationDiff / L2RegularizationTwiceDiff, stacked on the host side
    exactly as in the reference (the kernels return the unregularised sums)."""

    l2RegWeight = 0.0

    def calculate(self, data, coefficients, normalizationContext, want_gradient=True):
        value, grad = super().calculate(data, coefficients, normalizationContext, want_gradient)
        c = np.asarray(coefficients, dtype=np.float64)
        value = value + self.l2RegWeight * c.dot(c) / 2          # :69-70
        if grad is not None:
            grad = grad + c * self.l2RegWeight                    # :133
        return value, grad

    def hessianVector(self, data, coefficients, multiplyVector, normalizationContext):
        hv = super().hessianVector(data, coefficients, multiplyVector, normalizationContext)
        return hv + np.asarray(multiplyVector, dtype=np.float64) * self.l2RegWeight   # :151-157,176

    def hessianDiagonal(self, data, coefficients):
        return super().hessianDiagonal(data, coefficients) + self.l2RegWeight          # :167-168,183


class DistributedGLMLossFunction(_NativeObjective):
    """api/function/glm/DistributedGLMLossFunction.scala; `loss` is one of abi.LOGISTIC / POISSON / SQUARED."""

    def __init__(self, loss, treeAggregateDepth=1):
        if treeAggregateDepth <= 0:   # DistributedObjectiveFunction.scala:39
            raise abi.IllegalArgumentError("Tree aggregate depth must be greater than 0: %d" % treeAggregateDepth)
        super().__init__(loss)

    @staticmethod
    def create(loss, l2RegWeight=None, treeAggregateDepth=1):
        """The factory of :151-170: L2 is mixed in only when the regularization type is L2."""
        if l2RegWeight is None:
            return DistributedGLMLossFunction(loss, treeAggregateDepth)
        obj = _DistributedGLMLossFunctionWithL2(loss, treeAggregateDepth)
        obj.l2RegWeight = float(l2RegWeight)
        return obj


class _DistributedGLMLossFunctionWithL2(L2RegularizationMixin, DistributedGLMLossFunction):
    pass


SingleNodeGLMLossFunction = DistributedGLMLossFunction   # same contract over a local data set


class DistributedSmoothedHingeLossFunction(_NativeObjective):
    """api/function/svm/DistributedSmoothedHingeLossFunction.scala: DiffFunction only; ignores normalization."""

    twice_differentiable = False

    def __init__(self, treeAggregateDepth=1):
        super().__init__(abi.SMOOTHED_HINGE)

    @staticmethod
    def create(l2RegWeight=None, treeAggregateDepth=1):
        if l2RegWeight is None:
            return DistributedSmoothedHingeLossFunction(treeAggregateDepth)
        obj = _HingeWithL2(treeAggregateDepth)
        obj.l2RegWeight = float(l2RegWeight)
        return obj


class _HingeWithL2(L2RegularizationMixin, DistributedSmoothedHingeLossFunction):
    pass


class DeviceGLMLossFunction:
    """The same objective with coefficients, gradients and Hessian-vector products as torch tensors on the
    GPU (the *_dev entry points of the C ABI): an optimizer driven through it keeps its whole state in device
    memory and only reads scalars back.  L2 is applied on the device when l2RegWeight is given.  Run the
    optimizer inside `with objective.stream():` so that torch's own kernels are ordered on the context's stream."""

    twice_differentiable = True

    def __init__(self, ctx, loss, l2RegWeight=None):
        import torch
        self._torch = torch
        self.ctx, self.loss = ctx, loss
        self.l2RegWeight = None if l2RegWeight is None else float(l2RegWeight)
        self._ext = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", ctx.device))
        self._tag, self._tag_src, self._tag_version = 0, None, -1

    def stream(self):
        return self._torch.cuda.stream(self._ext)

    def _apply(self, data, ctx):
        _apply_normalization(data, ctx)

    def calculate(self, data, coefficients, normalizationContext, want_gradient=True):
        torch = self._torch
        self._apply(data, normalizationContext)
        value = torch.empty(1, dtype=torch.float64, device=coefficients.device)
        grad = torch.empty_like(coefficients) if want_gradient else None
        data.value_and_gradient_dev(self.loss, coefficients.data_ptr(), value.data_ptr(),
                                    grad.data_ptr() if want_gradient else None)
        f = float(value)
        if self.l2RegWeight is not None:
            f += self.l2RegWeight * float(coefficients.dot(coefficients)) / 2
            if want_gradient:
                grad += coefficients * self.l2RegWeight
        return f, grad

    def hessianVector(self, data, coefficients, multiplyVector, normalizationContext):
        torch = self._torch
        self._apply(data, normalizationContext)
        # TRON passes the same tensor object for a whole CG loop; a tensor updated in place (its version counter
        # moves) is a new coefficient vector
        if self._tag_src is not coefficients or self._tag_version != coefficients._version:
            self._tag_src, self._tag_version = coefficients, coefficients._version
            self._tag = next(_TAGS)
        hv = torch.empty_like(coefficients)
        data.hessian_vector_dev(self.loss, coefficients.data_ptr(), multiplyVector.data_ptr(), hv.data_ptr(), tag=self._tag)
        if self.l2RegWeight is not None:
            hv += multiplyVector * self.l2RegWeight
        return hv


class ShardedDeviceGLMLossFunction(DeviceGLMLossFunction):
    """DistributedGLMLossFunction over rows sharded across GPUs, one process per GPU: `data` is THIS rank's shard and
    every evaluation ends in the exchange over peer memory (abi.Peer: pml_value_and_gradient_allreduce_dev /
    pml_hessian_vector_allreduce_dev), which leaves bit-identical results on all ranks -- so every rank can run the
    same optimizer on the same numbers (SPMD) and all ranks hold the same model, as Spark's driver + broadcast do
    (DistributedObjectiveFunction.scala:57-58,73).  The normalization context must be the same on every rank."""

    def __init__(self, ctx, loss, peer, l2RegWeight=None):
        super().__init__(ctx, loss, l2RegWeight)
        self.peer = peer
        result = self._torch.as_tensor(peer.result_array(), device=self._torch.device("cuda", ctx.device))
        self._vector, self._value = result[:peer.d], result[peer.d:peer.d + 1]   # the library's result buffer

    def calculate(self, data, coefficients, normalizationContext, want_gradient=True):
        self._apply(data, normalizationContext)
        self.peer.value_and_gradient(data, self.loss, coefficients.data_ptr())
        f = float(self._value)
        grad = self._vector.clone() if want_gradient else None
        if self.l2RegWeight is not None:
            f += self.l2RegWeight * float(coefficients.dot(coefficients)) / 2
            if want_gradient:
                grad += coefficients * self.l2RegWeight
        return f, grad

    def hessianVector(self, data, coefficients, multiplyVector, normalizationContext):
        self._apply(data, normalizationContext)
        if self._tag_src is not coefficients or self._tag_version != coefficients._version:
            self._tag_src, self._tag_version = coefficients, coefficients._version
            self._tag = next(_TAGS)
        self.peer.hessian_vector(data, self.loss, coefficients.data_ptr(), multiplyVector.data_ptr(), tag=self._tag)
        hv = self._vector.clone()
        if self.l2RegWeight is not None:
            hv += multiplyVector * self.l2RegWeight
        return hv
