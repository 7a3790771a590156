// pointwise.cuh -- device-side pointwise losses and small reduction helpers shared by the
// kernel translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/photon_ml_b200.h"

namespace pml {

// ------------------------------------------------------------------------------------------
// Pointwise losses.  The branch structure is the reference's; __dmul_rn / __dsub_rn keep nvcc
// from contracting a*b+c into an FMA where Scala rounds twice.
// ------------------------------------------------------------------------------------------

// api/function/glm/LogisticLossFunction.scala:54
__device__ __forceinline__ double sigmoid(double z) { return 1.0 / (1.0 + exp(-z)); }

// (l, l'):  LogisticLossFunction.scala:70-77, PoissonLossFunction.scala:40-43,
// SquaredLossFunction.scala:42-45, svm/SmoothedHingeLossFunction.scala:40-63.
__device__ __forceinline__ void loss_and_dz(int loss, double z, double y, double &l, double &dl)
{
    switch (loss) {
    case PML_LOSS_LOGISTIC: {
        // The reference takes (log1pExp(-z), -sigmoid(-z)) for positive labels and (log1pExp(z), sigmoid(z))
        // otherwise: both are (log1pExp(m), +-sigmoid(m)) with m = -+z.  Written without the branches -- a warp
        // mixing labels or signs would otherwise run every exp/log1p twice, and fp64 transcendentals are what
        // bounds the per-row epilogue -- and with one exp: e = exp(-|m|) serves log1pExp (lib/util/MathUtils.scala:34-40:
        // x > 0 ? x + log1p(exp(-x)) : log1p(exp(x))) exactly as in the reference
        // and gives sigmoid(m) = 1/(1+e) for m > 0 (the reference's expression) or e/(1+e) (equal to the
        // reference's 1/(1+exp(-m)) to an ulp).
        const bool positive = y > 0.5;  // MathConst.POSITIVE_RESPONSE_THRESHOLD, lib/constants/MathConst.scala:23
        const double m = positive ? -z : z;
        const double e = exp(-fabs(m));
        const double lp = log1p(e);
        l = m > 0 ? m + lp : lp;
        const double sg = (m > 0 ? 1.0 : e) / (1.0 + e);
        dl = positive ? -sg : sg;
        break;
    }
    case PML_LOSS_POISSON: {
        double pred = exp(z);  // no overflow guard, as in the reference
        l = __dsub_rn(pred, __dmul_rn(z, y));
        dl = pred - y;
        break;
    }
    case PML_LOSS_SQUARED: {
        double delta = z - y;
        l = __dmul_rn(delta, delta) / 2.0;
        dl = delta;
        break;
    }
    default: {  // smoothed hinge; note `label < 0.5`, not `label > 0.5`
        double yy = (y < 0.5) ? -1.0 : 1.0;
        double t = yy * z;
        double lv, dv;
        if (t <= 0) lv = 0.5 - t;
        else if (t < 1) lv = __dmul_rn(__dmul_rn(0.5, 1.0 - t), 1.0 - t);
        else lv = 0.0;
        if (t < 0) dv = -1.0;
        else if (t < 1) dv = t - 1.0;
        else dv = 0.0;
        l = lv;
        dl = dv * yy;
        break;
    }
    }
}

// l'':  LogisticLossFunction.scala:86-89, PoissonLossFunction.scala:52, SquaredLossFunction.scala:54
__device__ __forceinline__ double dzz_loss(int loss, double z)
{
    switch (loss) {
    case PML_LOSS_LOGISTIC: {
        double s = sigmoid(z);
        return __dmul_rn(s, 1.0 - s);
    }
    case PML_LOSS_POISSON:
        return exp(z);
    case PML_LOSS_SQUARED:
        return 1.0;
    default:
        return __longlong_as_double(0x7ff8000000000000LL);  // not twice differentiable
    }
}

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum (a, b) over the CTA in a fixed order; result valid in every thread.  `red` holds
// 2*warps doubles.  Safe to call repeatedly (leading barrier).
template <int WARPS>
__device__ __forceinline__ void block_sum2(double &a, double &b, double *red)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    a = warp_sum(a);
    b = warp_sum(b);
    __syncthreads();
    if (lane == 0) {
        red[2 * w] = a;
        red[2 * w + 1] = b;
    }
    __syncthreads();
    double sa = 0.0, sb = 0.0;
#pragma unroll
    for (int i = 0; i < WARPS; ++i) {
        sa += red[2 * i];
        sb += red[2 * i + 1];
    }
    a = sa;
    b = sb;
}

__device__ __forceinline__ double widen(float v) { return (double)v; }
__device__ __forceinline__ double widen(double v) { return v; }

// Feature values as stored: float, double, or nothing at all when every stored value is exactly 1.0 (indicator
// features: a9a, hashed Criteo-style categories, GLMix id features) -- the kernels then read only the indices.
struct One {};
template <typename VT>
__device__ __forceinline__ double load_value(const VT *__restrict__ val, int64_t k)
{
    return widen(__ldcs(val + k));
}
template <>
__device__ __forceinline__ double load_value<One>(const One *__restrict__, int64_t)
{
    return 1.0;
}

}  // namespace pml
