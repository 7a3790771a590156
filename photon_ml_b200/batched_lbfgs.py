"""Lock-step L-BFGS / OWL-QN over all entities of a batch (the optimizer GAME defaults to for its coordinates:
api/util/GameTestUtils.scala:92, api/optimization/OptimizerFactory.scala:46-62; driver lib/optimization/LBFGS.scala:59-107,
OWLQN.scala:81-85).

Every scalar of the single-problem driver (optimization.LBFGS, the Breeze-style restatement) becomes a tensor with one
slot per entity, every branch a mask, every vector operation a batched kernel over the concatenated per-entity vectors
(pml_batch_segment_dot_dev / pml_batch_segment_axpby_dev), so that entity e goes through exactly the iterates a
single-problem run on entity e would.  The strong-Wolfe line search is a per-entity state machine (bracketing, then
zoom): entities sit in different phases at different step lengths, and each round evaluates all trial points with ONE
batched value+gradient call; entities whose search has ended are masked out and their rows are no longer streamed.
The correction pairs are kept as 2 x m concatenated vectors; every entity has its own ring in them (an entity that
skips an update -- curvature test -- or drops its history keeps its own list), and the two-loop recursion is one kernel
(pml_batch_lbfgs_direction_dev).  fused_direction=False is the same algebra as 2m dots and 2m + 1 axpys over lists
ordered oldest to newest, invalid slots carrying rho = 0; both give the same bits.

Like optimization.LBFGS this is NOT Breeze-verified (Breeze 0.11.2 is not in the reference tree): counts are
comparable with optimization.LBFGS / OWLQN on the same objective, entity by entity (tests/test_gpu_batched_lbfgs.py).
"""
import numpy as np

from .batched import (FUNCTION_VALUES_CONVERGED, GRADIENT_CONVERGED, MAX_ITERATIONS, OBJECTIVE_NOT_IMPROVING, REASONS,
                      BatchedTRON)

__all__ = ["BatchedLBFGS", "REASONS"]


class BatchedLBFGS(BatchedTRON):
    """l1RegWeight > 0 turns it into OWL-QN (pseudo-gradient, orthant projection, backtracking line search)."""

    C1, C2, LS_MAX_ITER, LS_MAX_ZOOM = 1e-4, 0.9, 10, 10     # optimization._strong_wolfe

    def __init__(self, batch, loss, l2RegWeight=0.0, l1RegWeight=0.0, numCorrections=10, tolerance=1.0e-7,
                 maxNumIterations=100, skip_converged=True, fused_direction=True):
        super().__init__(batch, loss, l2RegWeight=l2RegWeight, tolerance=tolerance, maxNumIterations=maxNumIterations,
                         skip_converged=skip_converged)
        self.m, self.l1 = int(numCorrections), float(l1RegWeight)
        # fused_direction: the two-loop recursion is ONE kernel (pml_batch_lbfgs_direction_dev) over per-entity rings of
        # correction pairs instead of 2m dots and 2m + 1 axpys, and a push writes one slot instead of shifting 2(m - 1)
        # vectors; same rounding, so the iterates are bit-identical either way (the ABI takes up to 16 pairs).
        self.fused_direction = bool(fused_direction) and self.m <= 16
        self.entity_evaluations = None

    # ---- small helpers ------------------------------------------------------------------------------------------
    def _zeros(self, n, dtype=None):
        return self.torch.zeros(n, dtype=dtype or self.torch.float64, device=self.device)

    def _per_coef(self, per_entity):
        """A per-entity quantity spread over the coefficients of each entity."""
        if getattr(self, "_ent_of", None) is None:
            t = self.torch
            ptr = t.as_tensor(np.asarray(self.batch.coef_ptr), device=self.device)
            self._ent_of = t.repeat_interleave(t.arange(self.E, device=self.device), ptr[1:] - ptr[:-1])
        return per_entity[self._ent_of]

    def _masked_copy(self, src, dst, mask):
        self._axpby(None, src, None, None, dst, mask.to(self.torch.uint8))

    def _evaluate(self, x, mask):
        """Objective (with the L2 term) at x for the entities in mask; counts per-entity evaluations."""
        if mask is None:                      # everybody
            f, g = self._objective(x, only=None)
            self.entity_evaluations += 1
        else:
            f, g = self._objective(x, only=mask)
            self.entity_evaluations += mask.to(self.torch.int64)
        return f, g

    def _adjusted(self, x, f, g):
        """OWLQN._adjusted: value + l1 |x|_1 and the pseudo-gradient (the identity for plain L-BFGS)."""
        if self.l1 == 0.0:
            return f, g
        t = self.torch
        l1 = self.l1
        ones = t.ones_like(x)
        adj_f = f + l1 * self._dot(x.abs(), ones)
        right, left = g + l1, g - l1
        zero = t.zeros_like(g)
        pg = t.where(x > 0, right, t.where(x < 0, left, t.where(left > 0, left, t.where(right < 0, right, zero))))
        return adj_f, pg

    # ---- L-BFGS two-loop recursion over per-entity histories (optimization.LBFGS._direction) ----------------------
    def _direction(self, adj_g, go):
        t = self.torch
        if self.fused_direction:
            q = self._zeros(self.D)
            self.batch.lbfgs_direction_dev(go.to(t.uint8).data_ptr(), self.m, self._S.data_ptr(), self._Y.data_ptr(),
                                           self._rho.data_ptr(), self._head.data_ptr(), self._len.data_ptr(),
                                           adj_g.data_ptr(), q.data_ptr())
            if self.l1 != 0.0:
                q = t.where(q * adj_g < 0, q, t.zeros_like(q))
            return q
        q = -adj_g
        alphas = [None] * self.m
        for i in range(self.m - 1, -1, -1):              # newest (slot m-1) to oldest; rho = 0 where the slot is empty
            a = self._rho[i] * self._dot(self._S[i], q)
            alphas[i] = a
            self._axpby(None, q, -a, self._Y[i], q)      # q -= a * y
        has = self._len > 0
        sy, yy = self._dot(self._S[self.m - 1], self._Y[self.m - 1]), self._dot(self._Y[self.m - 1], self._Y[self.m - 1])
        scale = t.where(has, sy / yy, t.ones_like(sy))
        self._axpby(scale, q, None, None, q)             # q *= s.y / y.y of the newest pair
        for i in range(self.m):
            b = self._rho[i] * self._dot(self._Y[i], q)
            self._axpby(None, q, alphas[i] - b, self._S[i], q)   # q += (a - b) * s
        if self.l1 != 0.0:                               # OWLQN._direction: keep components descending along the pseudo-gradient
            q = t.where(q * adj_g < 0, q, t.zeros_like(q))
        return q

    def _push_pair(self, s, y, mask, sy):
        """Append (s, y) to the history of the entities in mask (sy = s . y per entity, which the caller has): their
        slots shift down by one, the newest goes last -- or, with rings, the head moves on by one."""
        t = self.torch
        m8 = mask.to(t.uint8)
        if self.fused_direction:
            rho_new = 1.0 / sy
            self.batch.lbfgs_push_dev(m8.data_ptr(), self.m, rho_new.data_ptr(), s.data_ptr(), y.data_ptr(),
                                      self._S.data_ptr(), self._Y.data_ptr(), self._rho.data_ptr(), self._head.data_ptr(),
                                      self._len.data_ptr())
            return
        for i in range(self.m - 1):
            self._axpby(None, self._S[i + 1], None, None, self._S[i], m8)
            self._axpby(None, self._Y[i + 1], None, None, self._Y[i], m8)
            self._rho[i] = t.where(mask, self._rho[i + 1], self._rho[i])
        self._axpby(None, s, None, None, self._S[self.m - 1], m8)
        self._axpby(None, y, None, None, self._Y[self.m - 1], m8)
        self._rho[self.m - 1] = t.where(mask, 1.0 / sy, self._rho[self.m - 1])
        self._len = t.where(mask, t.clamp(self._len + 1, max=self.m), self._len)

    def _drop_history(self, mask):
        t = self.torch
        if self.fused_direction:
            self._len.masked_fill_(mask, 0)
            return
        for i in range(self.m):
            self._rho[i] = t.where(mask, t.zeros_like(self._rho[i]), self._rho[i])
        self._len = t.where(mask, t.zeros_like(self._len), self._len)

    # ---- strong-Wolfe line search, one state machine per entity (optimization._strong_wolfe) ----------------------
    def _strong_wolfe(self, x, f0, direction, g0, init, mask):
        """Returns (ok[E], t[E], f[E], g[D]): for the entities in mask the accepted step, value and gradient."""
        t_ = self.torch
        E, c1, c2 = self.E, self.C1, self.C2
        BRACKET, ZOOM, DONE = 0, 1, 2
        phase = t_.where(mask, t_.full((E,), BRACKET, device=self.device), t_.full((E,), DONE, device=self.device))
        step = init.clone()
        t_prev, f_prev, d_prev = self._zeros(E), f0.clone(), g0.clone()
        has_prev = t_.zeros(E, dtype=t_.bool, device=self.device)
        g_prev = self._zeros(self.D)
        lo, f_lo, d_lo = self._zeros(E), self._zeros(E), self._zeros(E)
        hi, f_hi, d_hi = self._zeros(E), self._zeros(E), self._zeros(E)
        has_lo = t_.zeros(E, dtype=t_.bool, device=self.device)
        g_lo = self._zeros(self.D)
        i_br = t_.zeros(E, dtype=t_.int64, device=self.device)
        i_zm = t_.zeros(E, dtype=t_.int64, device=self.device)
        ok = t_.zeros(E, dtype=t_.bool, device=self.device)
        res_t, res_f, res_g = self._zeros(E), self._zeros(E), self._zeros(self.D)
        xt = x.clone()        # trial points; an entity's slice is rewritten while it is searching and ignored afterwards

        def finish(who, tt, ff, gg, valid):
            nonlocal ok, res_t, res_f, phase
            ok = t_.where(who, valid, ok)
            res_t, res_f = t_.where(who, tt, res_t), t_.where(who, ff, res_f)
            self._masked_copy(gg, res_g, who & valid)
            phase = t_.where(who, t_.full_like(phase, DONE), phase)

        while True:
            running = phase != DONE
            if not bool(running.any()):
                break
            zooming = phase == ZOOM
            # the zoom phase picks its trial point by safeguarded cubic interpolation between lo and hi
            d1 = d_lo + d_hi - 3 * (f_lo - f_hi) / (lo - hi)
            disc = d1 * d1 - d_lo * d_hi
            d2 = t_.copysign(t_.sqrt(t_.clamp(disc, min=0.0)), hi - lo)
            cubic = hi - (hi - lo) * (d_hi + d2 - d1) / (d_hi - d_lo + 2 * d2)
            mid = 0.5 * (lo + hi)
            tz = t_.where(disc >= 0, cubic, mid)
            a, b = t_.minimum(lo, hi), t_.maximum(lo, hi)
            inside = (a + 0.1 * (b - a) <= tz) & (tz <= b - 0.1 * (b - a))
            tz = t_.where(inside, tz, mid)
            step = t_.where(zooming, tz, step)
            # phi(step) for everybody still searching: one batched evaluation
            self._axpby(None, x, step, direction, xt, running.to(t_.uint8))
            f_t, g_t = self._evaluate(xt, running)
            d_t = self._dot(g_t, direction)
            armijo_fail = f_t > f0 + c1 * step * g0
            curv_ok = d_t.abs() <= -c2 * g0
            # ---- bracketing phase
            br = running & ~zooming
            to_zoom_a = br & (armijo_fail | ((i_br > 0) & (f_t >= f_prev)))      # zoom(prev, t), payload of prev
            br_done = br & ~to_zoom_a & curv_ok
            to_zoom_b = br & ~to_zoom_a & ~curv_ok & (d_t >= 0)                 # zoom(t, prev), payload of t
            br_next = br & ~to_zoom_a & ~br_done & ~to_zoom_b
            # ---- zoom phase
            zm_hi = zooming & (armijo_fail | (f_t >= f_lo))
            zm_done = zooming & ~zm_hi & curv_ok
            zm_lo = zooming & ~zm_hi & ~curv_ok
            flip = zm_lo & (d_t * (hi - lo) >= 0)
            # results that are final now
            finish(br_done | zm_done, step, f_t, g_t, t_.ones_like(ok))
            # zoom updates (order as in the reference: hi <- lo where flipped, then lo <- t)
            hi, f_hi, d_hi = t_.where(zm_hi, step, hi), t_.where(zm_hi, f_t, f_hi), t_.where(zm_hi, d_t, d_hi)
            hi, f_hi, d_hi = t_.where(flip, lo, hi), t_.where(flip, f_lo, f_hi), t_.where(flip, d_lo, d_hi)
            lo, f_lo, d_lo = t_.where(zm_lo, step, lo), t_.where(zm_lo, f_t, f_lo), t_.where(zm_lo, d_t, d_lo)
            self._masked_copy(g_t, g_lo, zm_lo)
            has_lo = has_lo | zm_lo
            i_zm = i_zm + (zooming & ~zm_done).to(t_.int64)
            out_of_zoom = zooming & ~zm_done & (i_zm >= self.LS_MAX_ZOOM)
            finish(out_of_zoom, lo, f_lo, g_lo, has_lo)
            # entering zoom from the bracket
            lo = t_.where(to_zoom_a, t_prev, t_.where(to_zoom_b, step, lo))
            f_lo = t_.where(to_zoom_a, f_prev, t_.where(to_zoom_b, f_t, f_lo))
            d_lo = t_.where(to_zoom_a, d_prev, t_.where(to_zoom_b, d_t, d_lo))
            hi = t_.where(to_zoom_a, step, t_.where(to_zoom_b, t_prev, hi))
            f_hi = t_.where(to_zoom_a, f_t, t_.where(to_zoom_b, f_prev, f_hi))
            d_hi = t_.where(to_zoom_a, d_t, t_.where(to_zoom_b, d_prev, d_hi))
            self._masked_copy(g_prev, g_lo, to_zoom_a)
            self._masked_copy(g_t, g_lo, to_zoom_b)
            has_lo = t_.where(to_zoom_a, has_prev, t_.where(to_zoom_b, t_.ones_like(has_lo), has_lo))
            phase = t_.where(to_zoom_a | to_zoom_b, t_.full_like(phase, ZOOM), phase)
            # bracket goes on: remember this point, lengthen the step
            t_prev, f_prev, d_prev = t_.where(br_next, step, t_prev), t_.where(br_next, f_t, f_prev), t_.where(br_next, d_t, d_prev)
            self._masked_copy(g_t, g_prev, br_next)
            has_prev = has_prev | br_next
            step = t_.where(br_next, step * 1.5, step)
            i_br = i_br + br_next.to(t_.int64)
            out_of_bracket = br_next & (i_br >= self.LS_MAX_ITER)
            finish(out_of_bracket, t_prev, f_prev, g_prev, has_prev)
        return ok, res_t, res_f, res_g

    def _backtracking(self, x, adj_f, adj_g, direction, init, mask):
        """OWLQN._line_search: orthant projection and Armijo backtracking on the adjusted value."""
        t_ = self.torch
        orthant = t_.where(x != 0, x.sign(), -adj_g.sign())
        ones = t_.ones_like(x)
        step = init.clone()
        running = mask.clone()
        ok = t_.zeros(self.E, dtype=t_.bool, device=self.device)
        res_x, res_f, res_g = x.clone(), self._zeros(self.E), self._zeros(self.D)
        for _ in range(30):
            if not bool(running.any()):
                break
            xt = x + self._per_coef(step) * direction
            xt = t_.where(xt.sign() == orthant, xt, t_.zeros_like(xt))
            f_t, g_t = self._evaluate(xt, running)
            adj_ft = f_t + self.l1 * self._dot(xt.abs(), ones)
            accept = running & (adj_ft <= adj_f + 1e-4 * self._dot(adj_g, xt - x))
            self._masked_copy(xt, res_x, accept)
            self._masked_copy(g_t, res_g, accept)
            res_f = t_.where(accept, f_t, res_f)
            ok = ok | accept
            running = running & ~accept
            step = t_.where(running, step * 0.5, step)
        return ok, res_x, res_f, res_g

    # ---- Optimizer.optimize + LBFGS.init / runOneIteration, one slot per entity ------------------------------------
    def optimize(self, initial=None):
        """Returns (coefficients[D], values[E], info).  info: per-entity iterations, reasons and evaluation counts
        (numpy), and the number of batched evaluations."""
        t = self.torch
        with t.cuda.stream(self.stream):
            E, D = self.E, self.D
            everybody = None
            self.entity_evaluations = t.zeros(E, dtype=t.int64, device=self.device)
            x = self._zeros(D) if initial is None else t.as_tensor(initial, dtype=t.float64, device=self.device).clone()
            # Optimizer.optimize (:172-196): tolerances from the state at zero, then the initial state, then
            # LBFGS.init evaluates the initial point once more (Breeze does it itself, LBFGS.scala:64-65)
            f0, g0 = self._evaluate(self._zeros(D), everybody)
            loss_tol = f0 * self.tol
            grad_tol = self._dot(g0, g0).sqrt() * self.tol
            cur_f, cur_g = self._evaluate(x, everybody)                 # currentState: (loss, gradient) as the Optimizer sees them
            f, g = self._evaluate(x, everybody)                         # the L-BFGS inner state
            if self.fused_direction:
                self._S, self._Y, self._rho = self._zeros(self.m * D), self._zeros(self.m * D), self._zeros(self.m * E)
                self._head = t.zeros(E, dtype=t.int32, device=self.device)
                self._len = t.zeros(E, dtype=t.int32, device=self.device)
            else:
                self._S = [self._zeros(D) for _ in range(self.m)]
                self._Y = [self._zeros(D) for _ in range(self.m)]
                self._rho = [self._zeros(E) for _ in range(self.m)]
                self._len = t.zeros(E, dtype=t.int64, device=self.device)
            inner_iter = t.zeros(E, dtype=t.int64, device=self.device)
            inner_converged = t.zeros(E, dtype=t.bool, device=self.device)
            it = t.zeros(E, dtype=t.int64, device=self.device)
            done = t.zeros(E, dtype=t.bool, device=self.device)
            reason = t.full((E,), -1, dtype=t.int64, device=self.device)
            # Per-entity sums depend only on the entity's own numbers, so what an iteration has computed for the state
            # it leaves is what the next one would compute again: the adjusted state and its squared gradient norm are
            # carried over, and |cur_g|^2 follows cur_g entity by entity.
            adj_f, adj_g = self._adjusted(x, f, g)
            adj_gg = self._dot(adj_g, adj_g)
            cur_gg = self._dot(cur_g, cur_g)
            first_pass = True
            while not bool(done.all()):
                active = ~done
                prev_f, prev_it = cur_f.clone(), it.clone()
                stationary = adj_gg.sqrt() <= t.clamp(self.tol * adj_f.abs(), min=1e-8)
                inner_converged = inner_converged | (active & stationary)
                go = active & ~inner_converged                          # the others hand back the same state
                if bool(go.any()):
                    direction = self._direction(adj_g, go)
                    gd = self._dot(adj_g, direction)
                    not_descent = go & (gd >= 0)
                    if bool(not_descent.any()):
                        self._drop_history(not_descent)
                        self._masked_copy(-adj_g, direction, not_descent)
                        gd = self._dot(adj_g, direction)
                    # (inner_iter is 0 only in the first pass: an entity either moves or stops for good)
                    init = 1.0 / self._dot(direction, direction).sqrt() if first_pass else t.ones_like(gd)
                    if self.l1 == 0.0:
                        ok, step, fn, gn = self._strong_wolfe(x, f, direction, gd, init, go)
                        xn = x.clone()
                        self._axpby(None, x, step, direction, xn, (go & ok).to(t.uint8))
                    else:
                        ok, xn, fn, gn = self._backtracking(x, adj_f, adj_g, direction, init, go)
                    inner_converged = inner_converged | (go & ~ok)      # the search found nothing: "converged", same state
                    moved = go & ok
                    s, y = xn - x, gn - g
                    sy = self._dot(s, y)
                    keep = moved & (sy > 1e-10 * self._dot(s, s).sqrt() * self._dot(y, y).sqrt())
                    if bool(keep.any()):
                        self._push_pair(s, y, keep, sy)
                    x = xn                                              # xn is x wherever the entity has not moved
                    self._masked_copy(gn, g, moved)
                    f = t.where(moved, fn, f)
                    inner_iter = inner_iter + moved.to(t.int64)
                    adj_f, adj_g = self._adjusted(x, f, g)
                    adj_gg = self._dot(adj_g, adj_g)
                    stop = adj_gg.sqrt() <= t.clamp(self.tol * adj_f.abs(), min=1e-8)
                    inner_converged = inner_converged | (moved & (stop | (inner_iter >= self.max_iter)))
                    cur_f = t.where(moved, adj_f, cur_f)
                    self._masked_copy(adj_g, cur_g, moved)
                    cur_gg = t.where(moved, adj_gg, cur_gg)
                    it = it + moved.to(t.int64)
                    first_pass = False
                # Optimizer.getConvergenceReason (:136-150), in its order
                gnorm = cur_gg.sqrt()
                r = t.full((E,), -1, dtype=t.int64, device=self.device)
                r = t.where(gnorm <= grad_tol, t.full_like(r, GRADIENT_CONVERGED), r)
                r = t.where((cur_f - prev_f).abs() <= loss_tol, t.full_like(r, FUNCTION_VALUES_CONVERGED), r)
                r = t.where(it == prev_it, t.full_like(r, OBJECTIVE_NOT_IMPROVING), r)
                r = t.where(it >= self.max_iter, t.full_like(r, MAX_ITERATIONS), r)
                newly = active & (r >= 0)
                reason = t.where(newly, r, reason)
                done = done | newly
            self._only(None)
            its, reasons, evals = self._download_counts(it, reason, self.entity_evaluations)
            info = {"iterations": its, "reasons": reasons, "entity_evaluations": evals, "evaluations": self.evaluations}
            self._S = self._Y = self._rho = None
        return x, cur_f, info
