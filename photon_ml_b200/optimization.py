"""Optimizer drivers for the convergence metric: restatements of the reference's
callers of the objective, so that "wall time to convergence" and iteration counts can be measured
without a JVM.  In a JVM deployment the unchanged Scala optimizers drive the C ABI instead.

  Optimizer   lib/optimization/Optimizer.scala:37-245   -- exact restatement (tolerances from the state at
                                                            zero coefficients, convergence reasons, loop)
  TRON        lib/optimization/TRON.scala:80-339        -- exact restatement (in-tree algorithm)
  LBFGS/OWLQN lib/optimization/LBFGS.scala:39-156, OWLQN.scala:40-86 are adaptors over Breeze 0.11.2
              (org.scalanlp:breeze_2.11:0.11.2, NOT in the reference tree).  The classes below are a
              Breeze-STYLE L-BFGS (two-loop recursion, m corrections, strong-Wolfe line search with
              c1 = 1e-4, c2 = 0.9, first step 1/|g|) and OWL-QN (pseudo-gradient, orthant projection,
              backtracking) restated from the published algorithms.  NOT Breeze-verified: iteration
              counts are comparable between objectives run under this same driver (GPU vs CPU oracle),
              not against Breeze.

The objective passed in follows photon_ml_b200.function: calculate(data, coef, normCtx) ->
(value, gradient) and, for TRON, hessianVector(data, coef, v, normCtx).
"""
import math
import time

import numpy as np


def _norm(x):
    """2-norm of a numpy vector or a torch tensor (the drivers work on either: with torch tensors on the GPU the
    whole optimizer state stays in device memory and only scalars come back)."""
    if isinstance(x, np.ndarray):
        return float(np.linalg.norm(x))
    return float(x.norm())


def _zeros_like(x):
    return np.zeros_like(x) if isinstance(x, np.ndarray) else x.new_zeros(x.shape)


def _copy(x):
    return np.array(x, dtype=np.float64) if isinstance(x, np.ndarray) else x.clone()


def _where(cond, a, b):
    """np.where / torch.where with python scalars allowed on either side."""
    if isinstance(cond, np.ndarray):
        return np.where(cond, a, b)
    import torch
    like = a if torch.is_tensor(a) else (b if torch.is_tensor(b) else None)
    if not torch.is_tensor(a):
        a = torch.full_like(like, a) if like is not None else torch.full(cond.shape, a, dtype=torch.float64, device=cond.device)
    if not torch.is_tensor(b):
        b = torch.full_like(a, b)
    return torch.where(cond, a, b)


def _sign(x):
    return np.sign(x) if isinstance(x, np.ndarray) else x.sign()


def _abs_sum(x):
    return float(np.abs(x).sum()) if isinstance(x, np.ndarray) else float(x.abs().sum())


EPSILON = 1e-12   # MathConst.EPSILON, lib/constants/MathConst.scala:21


def compute_variances(objectiveFunction, data, coefficients):
    """Distributed / SingleNodeOptimizationProblem.computeVariances (api/optimization/DistributedOptimizationProblem.scala:
    81-95, SingleNodeOptimizationProblem.scala:57-67): 1 / (hessianDiagonal_j + EPSILON) -- the one caller of
    TwiceDiffFunction.hessianDiagonal; None for an objective that is not twice differentiable."""
    if not getattr(objectiveFunction, "twice_differentiable", False):
        return None
    return 1.0 / (objectiveFunction.hessianDiagonal(data, coefficients) + EPSILON)


class OptimizerState:
    def __init__(self, coefficients, loss, gradient, iter):
        self.coefficients, self.loss, self.gradient, self.iter = coefficients, loss, gradient, iter


MAX_ITERATIONS = "MaxIterations"
OBJECTIVE_NOT_IMPROVING = "ObjectiveNotImproving"
FUNCTION_VALUES_CONVERGED = "FunctionValuesConverged"
GRADIENT_CONVERGED = "GradientConverged"


def project_coefficients(coefficients, constraintMap):
    """OptimizationUtils.projectCoefficientsToSubspace, lib/optimization/OptimizationUtils.scala:56-70."""
    if not constraintMap:
        return coefficients
    out = np.array(coefficients, copy=True)
    for key, (lo, hi) in constraintMap.items():
        out[key] = min(max(out[key], lo), hi)
    return out


class Optimizer:
    """Optimizer.scala:37-245."""

    def __init__(self, relTolerance, maxNumIterations, normalizationContext, constraintMap=None, isTrackingState=True):
        self.relTolerance = relTolerance
        self.maxNumIterations = maxNumIterations
        self.normalizationContext = normalizationContext
        self.constraintMap = constraintMap
        self.isTrackingState = isTrackingState
        self.clearOptimizerInnerState()
        self.evaluations = 0          # calculate() calls (value+gradient evaluations)
        self.hessian_vector_evaluations = 0

    def clearOptimizerInnerState(self):
        self.currentState = None
        self.previousState = None
        self.trackedStates = [] if self.isTrackingState else None
        self.lossAbsTolerance = 0.0
        self.gradientAbsTolerance = 0.0

    def _setAbsTolerances(self, state):          # :68-71
        self.lossAbsTolerance = state.loss * self.relTolerance
        self.gradientAbsTolerance = _norm(state.gradient) * self.relTolerance

    def _updateCurrentState(self, state):        # :73-90
        if self.trackedStates is not None and state is not self.currentState:
            self.trackedStates.append((state.iter, state.loss, time.perf_counter()))
        self.previousState = self.currentState
        self.currentState = state

    def getConvergenceReason(self):              # :136-150
        cur, prev = self.currentState, self.previousState
        if cur is None:
            return None
        if cur.iter >= self.maxNumIterations:
            return MAX_ITERATIONS
        if cur.iter == prev.iter:
            return OBJECTIVE_NOT_IMPROVING
        if abs(cur.loss - prev.loss) <= self.lossAbsTolerance:
            return FUNCTION_VALUES_CONVERGED
        if _norm(cur.gradient) <= self.gradientAbsTolerance:
            return GRADIENT_CONVERGED
        return None

    def isDone(self):
        return self.getConvergenceReason() is not None

    def calculateState(self, objectiveFunction, coefficients, data, iter=0):
        value, gradient = objectiveFunction.calculate(data, coefficients, self.normalizationContext)
        self.evaluations += 1
        return OptimizerState(coefficients, value, gradient, iter)

    def optimize(self, objectiveFunction, initialCoefficients, data):   # :172-196
        normalized = self.normalizationContext.modelToTransformedSpace(initialCoefficients)
        self.clearOptimizerInnerState()
        self._setAbsTolerances(self.calculateState(objectiveFunction, _zeros_like(normalized), data))
        initialState = self.calculateState(objectiveFunction, normalized, data)
        self.init(objectiveFunction, initialState, data)
        self._updateCurrentState(initialState)
        while True:
            self._updateCurrentState(self.runOneIteration(objectiveFunction, self.currentState, data))
            if self.isDone():
                break
        self.convergenceReason = self.getConvergenceReason()
        return self.currentState.coefficients, self.currentState.loss


# ---------------------------------------------------------------------------------------------
class TRON(Optimizer):
    """TRON.scala:80-339 (trust-region Newton with truncated conjugate gradient)."""

    DEFAULT_MAX_NUM_FAILURE = 5
    DEFAULT_TOLERANCE = 1.0e-5
    DEFAULT_MAX_ITER = 15
    MAX_CG_ITERATIONS = 20

    def __init__(self, normalizationContext, maxNumImprovementFailures=5, tolerance=1.0e-5, maxNumIterations=15,
                 constraintMap=None, isTrackingState=True):
        super().__init__(tolerance, maxNumIterations, normalizationContext, constraintMap, isTrackingState)
        self.maxNumImprovementFailures = maxNumImprovementFailures
        self.eta0, self.eta1, self.eta2 = 1e-4, 0.25, 0.75
        self.sigma1, self.sigma2, self.sigma3 = 0.25, 0.5, 4.0
        self.delta = float("inf")
        self.cg_iterations = []

    def init(self, objectiveFunction, initState, data):          # :112-113
        self.delta = _norm(initState.gradient)

    def clearOptimizerInnerState(self):
        super().clearOptimizerInnerState()
        self.delta = float("inf")

    def runOneIteration(self, objectiveFunction, currState, data):   # :152-254
        prevCoefficients = currState.coefficients
        prevValue, prevGradient, prevIter = currState.loss, currState.gradient, currState.iter
        isFirstIteration = currState.iter == 0
        improved, failures, finalState = False, 0, currState
        eta0, eta1, eta2 = self.eta0, self.eta1, self.eta2
        sigma1, sigma2, sigma3 = self.sigma1, self.sigma2, self.sigma3
        while True:
            cgIter, step, residual = self._truncatedCG(objectiveFunction, prevGradient, self.delta, data, prevCoefficients)
            self.cg_iterations.append(cgIter)
            updated = prevCoefficients + step
            gs = float(prevGradient.dot(step))
            predictedReduction = -0.5 * (gs - float(step.dot(residual)))
            value, gradient = objectiveFunction.calculate(data, updated, self.normalizationContext)
            self.evaluations += 1
            actualReduction = prevValue - value
            stepNorm = _norm(step)
            if isFirstIteration:
                self.delta = min(self.delta, stepNorm)
            if value - prevValue - gs <= 0:
                alpha = sigma3
            else:
                alpha = max(sigma1, -0.5 * (gs / (value - prevValue - gs)))
            if actualReduction < eta0 * predictedReduction:
                self.delta = min(max(alpha, sigma1) * stepNorm, sigma2 * self.delta)
            elif actualReduction < eta1 * predictedReduction:
                self.delta = max(sigma1 * self.delta, min(alpha * stepNorm, sigma2 * self.delta))
            elif actualReduction < eta2 * predictedReduction:
                self.delta = max(sigma1 * self.delta, min(alpha * stepNorm, sigma3 * self.delta))
            else:
                self.delta = max(self.delta, min(alpha * stepNorm, sigma3 * self.delta))
            if actualReduction > eta0 * predictedReduction:
                improved = True
                finalState = OptimizerState(project_coefficients(updated, self.constraintMap), value, gradient, prevIter + 1)
            else:
                failures += 1
            if improved or failures >= self.maxNumImprovementFailures:
                break
        return finalState

    def _truncatedCG(self, objectiveFunction, gradient, truncationBoundary, data, coefficients):   # :279-339
        step = _zeros_like(gradient)
        residual = gradient * -1.0
        direction = _copy(residual)
        tol = 0.1 * _norm(gradient)
        iteration, done = 0, False
        rTr = float(residual.dot(residual))
        while iteration < self.MAX_CG_ITERATIONS and not done:
            if _norm(residual) <= tol:
                done = True
            else:
                iteration += 1
                Hd = objectiveFunction.hessianVector(data, coefficients, direction, self.normalizationContext)
                self.hessian_vector_evaluations += 1
                alpha = rTr / float(direction.dot(Hd))
                step += direction * alpha
                if _norm(step) > truncationBoundary:
                    alpha = -alpha
                    step += direction * alpha
                    std = float(step.dot(direction))
                    sts = float(step.dot(step))
                    dtd = float(direction.dot(direction))
                    dsq = truncationBoundary * truncationBoundary
                    rad = math.sqrt(std * std + dtd * (dsq - sts))
                    if std >= 0:
                        alpha = (dsq - sts) / (std + rad)
                    else:
                        alpha = (rad - std) / dtd
                    step += direction * alpha
                    alpha = -alpha
                    residual += Hd * alpha
                    done = True
                else:
                    alpha = -alpha
                    residual += Hd * alpha
                    rnew = float(residual.dot(residual))
                    beta = rnew / rTr
                    direction = direction * beta + residual
                    rTr = rnew
        return iteration, step, residual


# ---------------------------------------------------------------------------------------------
def _strong_wolfe(phi, f0, g0, init_step, c1=1e-4, c2=0.9, max_iter=10, max_zoom=10):
    """Bracketing + zoom line search for the strong Wolfe conditions (Nocedal & Wright alg. 3.5/3.6).
    phi(t) -> (value, directional derivative, payload)."""
    def zoom(lo, f_lo, d_lo, hi, f_hi, d_hi, payload_lo):
        for _ in range(max_zoom):
            # cubic interpolation between lo and hi, safeguarded to the middle 80 %
            d1 = d_lo + d_hi - 3 * (f_lo - f_hi) / (lo - hi)
            disc = d1 * d1 - d_lo * d_hi
            if disc >= 0:
                d2 = math.copysign(math.sqrt(disc), hi - lo)
                t = hi - (hi - lo) * (d_hi + d2 - d1) / (d_hi - d_lo + 2 * d2)
            else:
                t = 0.5 * (lo + hi)
            a, b = min(lo, hi), max(lo, hi)
            if not (a + 0.1 * (b - a) <= t <= b - 0.1 * (b - a)):
                t = 0.5 * (lo + hi)
            f_t, d_t, payload = phi(t)
            if f_t > f0 + c1 * t * g0 or f_t >= f_lo:
                hi, f_hi, d_hi = t, f_t, d_t
            else:
                if abs(d_t) <= -c2 * g0:
                    return t, f_t, payload
                if d_t * (hi - lo) >= 0:
                    hi, f_hi, d_hi = lo, f_lo, d_lo
                lo, f_lo, d_lo, payload_lo = t, f_t, d_t, payload
        return lo, f_lo, payload_lo

    t_prev, f_prev, d_prev, payload_prev = 0.0, f0, g0, None
    t = init_step
    for i in range(max_iter):
        f_t, d_t, payload = phi(t)
        if f_t > f0 + c1 * t * g0 or (i > 0 and f_t >= f_prev):
            return zoom(t_prev, f_prev, d_prev, t, f_t, d_t, payload_prev)
        if abs(d_t) <= -c2 * g0:
            return t, f_t, payload
        if d_t >= 0:
            return zoom(t, f_t, d_t, t_prev, f_prev, d_prev, payload)
        t_prev, f_prev, d_prev, payload_prev = t, f_t, d_t, payload
        t *= 1.5
    return t_prev, f_prev, payload_prev


class LBFGS(Optimizer):
    """LBFGS.scala:39-156 with a Breeze-style L-BFGS underneath (see the module docstring)."""

    DEFAULT_MAX_ITER = 100
    DEFAULT_NUM_CORRECTIONS = 10
    DEFAULT_TOLERANCE = 1.0e-7

    def __init__(self, normalizationContext, numCorrections=10, tolerance=1.0e-7, maxNumIterations=100,
                 constraintMap=None, isTrackingState=True):
        super().__init__(tolerance, maxNumIterations, normalizationContext, constraintMap, isTrackingState)
        self.m = numCorrections
        self.tolerance = tolerance

    l1RegWeight = 0.0

    def init(self, objectiveFunction, initState, data):
        # breezeOptimizer.iterations(f, x0) followed by next(): Breeze evaluates f(x0) itself (LBFGS.scala:64-65)
        value, gradient = objectiveFunction.calculate(data, initState.coefficients, self.normalizationContext)
        self.evaluations += 1
        self._x = _copy(initState.coefficients)
        self._f, self._g = value, gradient
        self._S, self._Y = [], []
        self._inner_iter = 0
        self._inner_converged = False

    # -- pieces OWLQN overrides ---------------------------------------------------------------
    def _adjusted(self, x, f, g):
        return f, g

    def _direction(self, adj_g):
        q = -_copy(adj_g)
        alphas = []
        for s, y in zip(reversed(self._S), reversed(self._Y)):
            rho = 1.0 / float(y.dot(s))
            a = rho * float(s.dot(q))
            alphas.append((a, rho, s, y))
            q -= a * y
        if self._S:
            s, y = self._S[-1], self._Y[-1]
            q *= float(s.dot(y)) / float(y.dot(y))
        for a, rho, s, y in reversed(alphas):
            b = rho * float(y.dot(q))
            q += (a - b) * s
        return q

    def _line_search(self, objectiveFunction, data, x, f, g, adj_f, adj_g, direction):
        gd = float(adj_g.dot(direction))

        def phi(t):
            xt = x + t * direction
            ft, gt = objectiveFunction.calculate(data, xt, self.normalizationContext)
            self.evaluations += 1
            return ft, float(gt.dot(direction)), (xt, ft, gt)

        init = 1.0 / _norm(direction) if self._inner_iter == 0 else 1.0
        _, _, payload = _strong_wolfe(phi, f, gd, init)
        return payload

    # -----------------------------------------------------------------------------------------
    def runOneIteration(self, objectiveFunction, currState, data):
        if self._inner_converged:
            return currState                   # "LBFGS is converged": the same state object comes back
        x, f, g = self._x, self._f, self._g
        adj_f, adj_g = self._adjusted(x, f, g)
        if _norm(adj_g) <= max(self.tolerance * abs(adj_f), 1e-8):
            self._inner_converged = True       # already stationary (Breeze tests convergence on the initial state too)
            return currState
        direction = self._direction(adj_g)
        if float(adj_g.dot(direction)) >= 0:   # not a descent direction: drop the history
            self._S, self._Y = [], []
            direction = -adj_g
        payload = self._line_search(objectiveFunction, data, x, f, g, adj_f, adj_g, direction)
        if payload is None:
            self._inner_converged = True
            return currState
        xn, fn, gn = payload
        s, y = xn - x, gn - g
        if float(s.dot(y)) > 1e-10 * _norm(s) * _norm(y):
            self._S.append(s)
            self._Y.append(y)
            if len(self._S) > self.m:
                self._S.pop(0)
                self._Y.pop(0)
        self._x, self._f, self._g = xn, fn, gn
        self._inner_iter += 1
        adj_fn, adj_gn = self._adjusted(xn, fn, gn)
        # Breeze's own stopping rule: |adjusted gradient| <= max(tolerance * |adjusted value|, 1e-8)
        if _norm(adj_gn) <= max(self.tolerance * abs(adj_fn), 1e-8) or self._inner_iter >= self.maxNumIterations:
            self._inner_converged = True
        return OptimizerState(project_coefficients(xn, self.constraintMap), adj_fn, adj_gn, currState.iter + 1)


class OWLQN(LBFGS):
    """OWLQN.scala:40-86: L1 weight applied inside the optimizer (constant for every coordinate)."""

    def __init__(self, l1RegWeight, normalizationContext, numCorrections=10, tolerance=1.0e-7, maxNumIterations=100,
                 constraintMap=None, isTrackingState=True):
        super().__init__(normalizationContext, numCorrections, tolerance, maxNumIterations, constraintMap, isTrackingState)
        self.l1RegWeight = float(l1RegWeight)

    def _adjusted(self, x, f, g):
        l1 = self.l1RegWeight
        adj_f = f + l1 * _abs_sum(x)
        right, left = g + l1, g - l1
        pg = _where(x > 0, right, _where(x < 0, left, _where(left > 0, left, _where(right < 0, right, 0.0))))
        return adj_f, pg

    def _direction(self, adj_g):
        d = super()._direction(adj_g)
        return _where(d * adj_g < 0, d, 0.0)       # keep only components that descend along the pseudo-gradient

    def _line_search(self, objectiveFunction, data, x, f, g, adj_f, adj_g, direction):
        orthant = _where(x != 0, _sign(x), -_sign(adj_g))
        gd = float(adj_g.dot(direction))
        t = 1.0 / _norm(direction) if self._inner_iter == 0 else 1.0
        for _ in range(30):
            xt = x + t * direction
            xt = _where(_sign(xt) == orthant, xt, 0.0)
            ft, gt = objectiveFunction.calculate(data, xt, self.normalizationContext)
            self.evaluations += 1
            adj_ft = ft + self.l1RegWeight * _abs_sum(xt)
            if adj_ft <= adj_f + 1e-4 * float(adj_g.dot(xt - x)):
                return xt, ft, gt
            t *= 0.5
        return None
