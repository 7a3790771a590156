"""Lock-step TRON for all entities of a random-effect coordinate, with every vector in device memory.

The reference solves one small problem per entity inside a `mapValues`
(api/algorithm/RandomEffectCoordinate.scala:108-119): SingleNodeOptimizationProblem.run ->
Optimizer.optimize (lib/optimization/Optimizer.scala:172-196) -> TRON.runOneIteration / truncatedCG
(lib/optimization/TRON.scala:152-254, 279-339) over a SingleNodeGLMLossFunction with the L2 mix-in.

Here the same control flow runs for ALL entities of a pml.Batch at once.  Every scalar of the single-entity algorithm
(delta, alpha, beta, rTr, the tolerances, the iteration and failure counters, the convergence reason) becomes a
vector with one slot per entity, every branch becomes a mask, and every vector operation is the batched kernel of the
library: pml_batch_value_and_gradient_dev / pml_batch_hessian_vector_dev for the objective,
pml_batch_segment_dot_dev / pml_batch_segment_axpby_dev for Breeze's dot and axpy.  An entity's iterates are what a
run of photon_ml_b200.optimization.TRON on that entity alone produces (same operations in the same order; only the
summation order inside dot products may differ in the last bits) -- tests/test_gpu_batched_tron.py checks exactly
that.  torch is used for the per-entity scalar bookkeeping (E-sized element-wise ops) and as the owner of the buffers.
"""
import numpy as np

from .function import _TAGS

MAX_ITERATIONS, OBJECTIVE_NOT_IMPROVING, FUNCTION_VALUES_CONVERGED, GRADIENT_CONVERGED = 0, 1, 2, 3
REASONS = ["MaxIterations", "ObjectiveNotImproving", "FunctionValuesConverged", "GradientConverged"]


class BatchedTRON:
    MAX_CG_ITERATIONS = 20      # TRON.scala:97-98

    def __init__(self, batch, loss, l2RegWeight=0.0, tolerance=1.0e-5, maxNumIterations=15, maxNumImprovementFailures=5,
                 skip_converged=True, fused_cg=True):
        import torch
        self.torch = torch
        self.batch, self.loss, self.l2 = batch, loss, float(l2RegWeight)
        self.tol, self.max_iter, self.max_fail = tolerance, maxNumIterations, maxNumImprovementFailures
        self.eta0, self.eta1, self.eta2 = 1e-4, 0.25, 0.75
        self.sigma1, self.sigma2, self.sigma3 = 0.25, 0.5, 4.0
        ctx = batch.ctx
        self.device = torch.device("cuda", ctx.device)
        self.stream = torch.cuda.ExternalStream(ctx.stream, device=self.device)
        self.E, self.D = batch.n_entities, batch.d_total
        self.evaluations = 0
        self.hessian_vector_evaluations = 0
        self._tag = 0
        self.skip_converged, self._masked, self._mask8 = bool(skip_converged), False, None
        # fused_cg: one kernel per CG iteration for the per-entity vector algebra (pml_batch_cg_step_dev) instead of
        # three dots and up to five axpy launches; same rounding, so the iterates are bit-identical either way.
        # 10^6 entities: 0.038 s per solve against 0.044 s unfused (profiles/r02_re_solve_profile.json).
        self.fused_cg = bool(fused_cg)

    # ---- the library's batched kernels -------------------------------------------------------------------------
    def _new(self, n, dtype=None):
        return self.torch.empty(n, dtype=dtype or self.torch.float64, device=self.device)

    def _download_counts(self, *tensors):
        """Per-entity counters to numpy (int32) in ONE copy through a pinned buffer kept on the batch: three pageable
        int64 copies of 10^6 entities cost several milliseconds, a fifth of a whole lock-step solve."""
        torch = self.torch
        stacked = torch.stack([t.to(torch.int32) for t in tensors])
        host = getattr(self.batch, "_counts_host", None)
        if host is None or host.numel() < stacked.numel():
            host = torch.empty(stacked.numel(), dtype=torch.int32, pin_memory=True)
            self.batch._counts_host = host
        view = host[:stacked.numel()].view(stacked.shape)
        # (a blocking copy: a non-blocking one ties the pinned block to the context's stream inside torch's host
        # allocator, which then records an event on that stream when the block is freed -- after Context.close())
        view.copy_(stacked)
        return [view[i].numpy().copy() for i in range(len(tensors))]

    def _dot(self, x, y):
        out = self._new(self.E)
        self.batch.segment_dot_dev(x.data_ptr(), y.data_ptr(), out.data_ptr())
        return out

    def _axpby(self, a, x, c, y, out, active=None):
        """out = a[e] x + c[e] y per entity (None = 1), only where active."""
        self.batch.segment_axpby_dev(None if a is None else a.data_ptr(), x.data_ptr(), None if c is None else c.data_ptr(),
                                     None if y is None else y.data_ptr(), out.data_ptr(),
                                     None if active is None else active.data_ptr())

    def _only(self, mask):
        """Entities outside `mask` (a bool tensor; None = everybody) drop out of the next evaluations: their results
        come back as 0 and -- the point -- their rows are no longer streamed (pml_batch_set_active_dev).  In the
        reference a converged entity's solve has simply ended (RandomEffectCoordinate.scala:112-119)."""
        if not self.skip_converged:
            return
        if mask is None:
            if self._masked:
                self.batch.set_active_dev(None)
                self._masked = False
            return
        self._mask8 = mask.to(self.torch.uint8)      # kept alive until the next mask replaces it
        self.batch.set_active_dev(self._mask8.data_ptr())
        self._masked = True

    def _objective(self, w, only=None):
        """(value[E], gradient[D]) with the L2 term of L2RegularizationDiff (lib/function/L2Regularization.scala:69-125)."""
        self._only(only)
        f, g = self._new(self.E), self._new(self.D)
        self.batch.value_and_gradient_dev(self.loss, w.data_ptr(), f.data_ptr(), g.data_ptr())
        self.evaluations += 1
        if self.l2 != 0.0:
            f = f + self._dot(w, w) * (self.l2 / 2)
            self._axpby(self._l2_per_entity(), w, None, g, g)   # g += l2 * w: product rounded, then the sum
        return f, g

    def _hessian_vector(self, w, v, regularize=True):
        hv = self._new(self.D)
        self.batch.hessian_vector_dev(self.loss, w.data_ptr(), v.data_ptr(), hv.data_ptr(), tag=self._tag)
        self.hessian_vector_evaluations += 1
        if self.l2 != 0.0 and regularize:
            self._axpby(self._l2_per_entity(), v, None, hv, hv)
        return hv

    def _l2_per_entity(self):
        if getattr(self, "_l2_vec", None) is None:
            self._l2_vec = self.torch.full((self.E,), self.l2, dtype=self.torch.float64, device=self.device)
        return self._l2_vec

    # ---- TRON.truncatedCG for the entities in `mask` -------------------------------------------------------------
    def _truncated_cg_fused(self, w, g, delta, mask):
        """The same loop with the per-entity vector algebra of an iteration done by ONE kernel (pml_batch_cg_step_dev:
        alpha, step, trust-region boundary, residual, beta, direction -- rounding for rounding what the unfused loop
        below does with three dots and up to five axpy launches, each a pass over all entities' coefficients)."""
        torch = self.torch
        step = torch.zeros(self.D, dtype=torch.float64, device=self.device)
        residual = -g
        direction = residual.clone()
        cg_tol = 0.1 * self._dot(g, g).sqrt()
        rTr = self._dot(residual, residual)
        delta = delta.contiguous()
        iters = torch.zeros(self.E, dtype=torch.int64, device=self.device)
        over8 = torch.zeros(self.E, dtype=torch.uint8, device=self.device)
        done = ~mask
        for _ in range(self.MAX_CG_ITERATIONS):
            done = done | (rTr.sqrt() <= cg_tol)
            if bool(done.all()):
                break
            run = ~done
            iters += run
            self._only(run)
            Hd = self._hessian_vector(w, direction, regularize=False)
            run8 = run.to(torch.uint8)
            self.batch.cg_step_dev(run8.data_ptr(), self.l2, direction.data_ptr(), Hd.data_ptr(), step.data_ptr(),
                                   residual.data_ptr(), rTr.data_ptr(), delta.data_ptr(), over8.data_ptr())
            done = done | over8.to(torch.bool)
        return iters, step, residual

    def _truncated_cg(self, w, g, delta, mask):
        if self.fused_cg:
            return self._truncated_cg_fused(w, g, delta, mask)
        torch = self.torch
        step = torch.zeros(self.D, dtype=torch.float64, device=self.device)
        residual = -g
        direction = residual.clone()
        cg_tol = 0.1 * self._dot(g, g).sqrt()
        rTr = self._dot(residual, residual)
        iters = torch.zeros(self.E, dtype=torch.int64, device=self.device)
        done = ~mask
        for _ in range(self.MAX_CG_ITERATIONS):
            done = done | (rTr.sqrt() <= cg_tol)     # rTr is residual . residual of the current residual
            if bool(done.all()):
                break
            run = ~done
            iters += run
            self._only(run)                          # within one coefficient tag the set only ever shrinks
            Hd = self._hessian_vector(w, direction)
            alpha = rTr / self._dot(direction, Hd)
            run8 = run.to(torch.uint8)
            self._axpby(alpha, direction, None, step, step, run8)                  # step += direction * alpha
            over = run & (self._dot(step, step).sqrt() > delta)
            inside = run & ~over
            if bool(over.any()):
                # the step left the trust region: back to the boundary along `direction`, then stop (TRON.scala:311-327)
                over8 = over.to(torch.uint8)
                self._axpby(-alpha, direction, None, step, step, over8)
                std, sts, dtd = self._dot(step, direction), self._dot(step, step), self._dot(direction, direction)
                dsq = delta * delta
                rad = (std * std + dtd * (dsq - sts)).sqrt()
                alpha_b = torch.where(std >= 0, (dsq - sts) / (std + rad), (rad - std) / dtd)
                self._axpby(alpha_b, direction, None, step, step, over8)
                self._axpby(-alpha_b, Hd, None, residual, residual, over8)
                done = done | over
            if bool(inside.any()):
                in8 = inside.to(torch.uint8)
                self._axpby(-alpha, Hd, None, residual, residual, in8)             # residual -= Hd * alpha
                rnew = self._dot(residual, residual)
                beta = rnew / rTr
                self._axpby(beta, direction, None, residual, direction, in8)       # direction = direction * beta + residual
                rTr = torch.where(inside, rnew, rTr)
        return iters, step, residual

    # ---- Optimizer.optimize + TRON.runOneIteration, one slot per entity ------------------------------------------
    def optimize(self, initial=None):
        """Returns (coefficients[D], values[E], info) -- torch tensors on the batch's device; info holds per-entity
        iteration counts and convergence reasons (numpy) and the numbers of batched evaluations."""
        torch = self.torch
        with torch.cuda.stream(self.stream):
            w = torch.zeros(self.D, dtype=torch.float64, device=self.device) if initial is None else \
                torch.as_tensor(initial, dtype=torch.float64, device=self.device).clone()
            self._tag = next(_TAGS)
            f0, g0 = self._objective(torch.zeros_like(w))           # tolerances from the state at zero (Optimizer.scala:68-71)
            loss_tol = f0 * self.tol
            grad_tol = self._dot(g0, g0).sqrt() * self.tol
            if initial is None:
                f, g = f0.clone(), g0.clone()
                self.evaluations += 1                                # the reference evaluates the initial state separately
            else:
                f, g = self._objective(w)
            delta = self._dot(g, g).sqrt()                           # TRON.init (:112-113)
            it = torch.zeros(self.E, dtype=torch.int64, device=self.device)
            done = torch.zeros(self.E, dtype=torch.bool, device=self.device)
            reason = torch.full((self.E,), -1, dtype=torch.int64, device=self.device)
            cg_total = torch.zeros(self.E, dtype=torch.int64, device=self.device)
            while not bool(done.all()):
                active = ~done
                prev_f, prev_it = f.clone(), it.clone()
                first = it == 0
                failures = torch.zeros(self.E, dtype=torch.int64, device=self.device)
                attempting = active.clone()
                while bool(attempting.any()):
                    self._tag = next(_TAGS)                                   # new coefficient vector for the curvature cache
                    cg_iters, step, residual = self._truncated_cg(w, g, delta, attempting)
                    cg_total += cg_iters
                    w_try = w + step
                    gs = self._dot(g, step)
                    predicted = -0.5 * (gs - self._dot(step, residual))
                    f_try, g_try = self._objective(w_try, only=attempting)
                    actual = f - f_try
                    step_norm = self._dot(step, step).sqrt()
                    delta = torch.where(attempting & first, torch.minimum(delta, step_norm), delta)
                    denom = f_try - f - gs
                    alpha = torch.where(denom <= 0, torch.full_like(denom, self.sigma3),
                                        torch.clamp(-0.5 * (gs / denom), min=self.sigma1))
                    a_sn = alpha * step_norm
                    new_delta = torch.where(
                        actual < self.eta0 * predicted,
                        torch.minimum(torch.clamp(alpha, min=self.sigma1) * step_norm, self.sigma2 * delta),
                        torch.where(actual < self.eta1 * predicted,
                                    torch.maximum(self.sigma1 * delta, torch.minimum(a_sn, self.sigma2 * delta)),
                                    torch.where(actual < self.eta2 * predicted,
                                                torch.maximum(self.sigma1 * delta, torch.minimum(a_sn, self.sigma3 * delta)),
                                                torch.maximum(delta, torch.minimum(a_sn, self.sigma3 * delta)))))
                    delta = torch.where(attempting, new_delta, delta)
                    accepted = attempting & (actual > self.eta0 * predicted)
                    acc8 = accepted.to(torch.uint8)
                    self._axpby(None, w_try, None, None, w, acc8)    # w, g <- the trial point where it was accepted
                    self._axpby(None, g_try, None, None, g, acc8)
                    f = torch.where(accepted, f_try, f)
                    it = it + accepted
                    failures = failures + (attempting & ~accepted)
                    attempting = attempting & ~accepted & (failures < self.max_fail)
                # Optimizer.getConvergenceReason (:136-150), in its order
                gnorm = self._dot(g, g).sqrt()
                r = torch.full((self.E,), -1, dtype=torch.int64, device=self.device)
                r = torch.where(gnorm <= grad_tol, torch.full_like(r, GRADIENT_CONVERGED), r)
                r = torch.where((f - prev_f).abs() <= loss_tol, torch.full_like(r, FUNCTION_VALUES_CONVERGED), r)
                r = torch.where(it == prev_it, torch.full_like(r, OBJECTIVE_NOT_IMPROVING), r)
                r = torch.where(it >= self.max_iter, torch.full_like(r, MAX_ITERATIONS), r)
                newly = active & (r >= 0)
                reason = torch.where(newly, r, reason)
                done = done | newly
            self._only(None)
            its, reasons, cgs = self._download_counts(it, reason, cg_total)
            info = {"iterations": its, "reasons": reasons, "cg_iterations": cgs,
                    "evaluations": self.evaluations, "hessian_vector_evaluations": self.hessian_vector_evaluations}
        return w, f, info


def reason_names(codes):
    return [REASONS[int(c)] for c in np.asarray(codes)]
