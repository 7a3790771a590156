"""Test-side objectives with the photon_ml_b200.function interface: the CPU oracle wrapped as an
ObjectiveFunction (so the same optimizer drivers can run on it), and the quadratic test objective of
the reference's optimizer tests (lib-integ/optimization/IntegTestObjective.scala:33-41: sum (x_i - 4)^2)."""
import numpy as np

import oracle


class OracleGLMObjective:
    def __init__(self, loss, l2RegWeight=0.0):
        self.loss, self.l2RegWeight = loss, l2RegWeight
        self.twice_differentiable = loss != oracle.SMOOTHED_HINGE

    def calculate(self, data, coefficients, ctx, want_gradient=True):
        value, grad = oracle.value_and_gradient(data, self.loss, coefficients, ctx.factors, ctx.shifts, ctx.interceptId)
        c = np.asarray(coefficients)
        return value + self.l2RegWeight * c.dot(c) / 2, grad + c * self.l2RegWeight

    def hessianVector(self, data, coefficients, v, ctx):
        return oracle.hessian_vector(data, self.loss, coefficients, v, ctx.factors, ctx.shifts, ctx.interceptId) \
            + np.asarray(v) * self.l2RegWeight


class QuadraticObjective:
    CENTROID = 4.0
    twice_differentiable = True

    def calculate(self, data, coefficients, ctx, want_gradient=True):
        delta = np.asarray(coefficients) - self.CENTROID
        return float(delta.dot(delta)), 2.0 * delta

    def hessianVector(self, data, coefficients, v, ctx):
        return 2.0 * np.asarray(v)
