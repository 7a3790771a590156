"""Coordinate descent over a fixed-effect coordinate and random-effect coordinates with every per-row vector in
device memory -- the callers either side of the hot path, restated so a GAME pass can be
measured without a JVM.

Reference flow (lib/algorithm/CoordinateDescent.scala:138-260, lib/algorithm/Coordinate.scala:59-62): for every pass,
for every coordinate in the update sequence: partial score = sum of the OTHER coordinates' scores; the coordinate adds
it to its offsets (DataSet.addScoresToOffsets: api/data/FixedEffectDataSet.scala:45-54, LocalDataSet.scala:82-89),
re-trains from its current model, and is scored again (FixedEffectCoordinate.scala:158-165,
RandomEffectCoordinate.scala:157-187).  There the scores are RDDs joined by row id; here each coordinate keeps its
rows in its own order with a row map to the global order, and the exchange is pml_scores_dev ->
pml_offsets_from_scores_dev -> pml_dataset_set_offsets_dev.

  FixedEffect(dataset, ...)            one GLM over all rows: TRON with the optimizer state in device memory
  RandomEffect(batch, row_map, ...)    one small GLM per entity, rows grouped by entity: lock-step batched TRON
"""
import time

from . import abi
from .batched import BatchedTRON
from .function import DeviceGLMLossFunction, NoNormalization, ShardedDeviceGLMLossFunction
from .optimization import TRON


class FixedEffect:
    """FixedEffectCoordinate (api/algorithm/FixedEffectCoordinate.scala).  `row_map[i]` = global row of local row i
    (None = the coordinate's rows are in global order); `base_offset` = the data's own offsets, local order."""

    def __init__(self, dataset, loss, l2RegWeight=0.0, base_offset=None, row_map=None, tolerance=1e-5, maxNumIterations=15,
                 peer=None):
        import torch
        self.torch = torch
        self.ds, self.loss, self.l2 = dataset, loss, l2RegWeight
        self.ctx = dataset.ctx
        self.device = torch.device("cuda", self.ctx.device)
        self.base_offset, self.row_map = base_offset, row_map
        self.tolerance, self.max_iter = tolerance, maxNumIterations
        # peer: the rows are one rank's shard (one process per GPU); every evaluation is merged over peer memory and
        # every rank trains the same model on the same numbers
        self.objective = (DeviceGLMLossFunction(self.ctx, loss, l2RegWeight=l2RegWeight or None) if peer is None else
                          ShardedDeviceGLMLossFunction(self.ctx, loss, peer, l2RegWeight=l2RegWeight or None))
        self.model = torch.zeros(dataset.d, dtype=torch.float64, device=self.device)
        self.n = dataset.n

    def train(self):
        opt = TRON(NoNormalization(), tolerance=self.tolerance, maxNumIterations=self.max_iter, isTrackingState=False)
        self.model, value = opt.optimize(self.objective, self.model, self.ds)
        return float(value), {"iterations": opt.currentState.iter, "reason": opt.convergenceReason,
                              "evaluations": opt.evaluations, "hessian_vector_evaluations": opt.hessian_vector_evaluations}

    def score_local(self):
        out = self.torch.empty(self.n, dtype=self.torch.float64, device=self.device)
        self.ds.scores_dev(self.model.data_ptr(), out.data_ptr())
        return out

    def set_offsets(self, offsets):
        self.ds.set_offsets_dev(offsets.data_ptr())

    def regularization_term(self):
        return 0.5 * self.l2 * float(self.model.dot(self.model)) if self.l2 else 0.0


class RandomEffect:
    """RandomEffectCoordinate (api/algorithm/RandomEffectCoordinate.scala): `batch` holds the rows grouped by entity,
    `row_map[i]` = global row of the batch's row i."""

    def __init__(self, batch, row_map, loss, l2RegWeight=0.0, base_offset=None, tolerance=1e-5, maxNumIterations=15,
                 optimizer="TRON"):
        import torch
        self.torch = torch
        self.batch, self.loss, self.l2 = batch, loss, l2RegWeight
        self.ctx = batch.ctx
        self.device = torch.device("cuda", self.ctx.device)
        self.base_offset, self.row_map = base_offset, row_map
        self.tolerance, self.max_iter = tolerance, maxNumIterations
        self.model = torch.zeros(batch.d_total, dtype=torch.float64, device=self.device)
        self.n = batch.dataset.n
        self.optimizer = optimizer     # "TRON" or "LBFGS" (OptimizerFactory.scala:46-62; GAME's tests default to LBFGS)

    def train(self):
        if self.optimizer == "LBFGS":
            from .batched_lbfgs import BatchedLBFGS
            solver = BatchedLBFGS(self.batch, self.loss, l2RegWeight=self.l2, tolerance=self.tolerance,
                                  maxNumIterations=self.max_iter)
        else:
            solver = BatchedTRON(self.batch, self.loss, l2RegWeight=self.l2, tolerance=self.tolerance,
                                 maxNumIterations=self.max_iter)
        self.model, values, info = solver.optimize(self.model)
        return float(values.sum()), {"mean_iterations": float(info["iterations"].mean()),
                                     "evaluations": info["evaluations"],
                                     "hessian_vector_evaluations": info.get("hessian_vector_evaluations", 0)}

    def score_local(self):
        out = self.torch.empty(self.n, dtype=self.torch.float64, device=self.device)
        self.batch.dataset.scores_dev(self.model.data_ptr(), out.data_ptr())
        return out

    def set_offsets(self, offsets):
        self.batch.dataset.set_offsets_dev(offsets.data_ptr())

    def regularization_term(self):
        return 0.5 * self.l2 * float(self.model.dot(self.model)) if self.l2 else 0.0


class CoordinateDescent:
    """CoordinateDescent.run (lib/algorithm/CoordinateDescent.scala:138-260) without validation: `coordinates` is an
    ordered {name: FixedEffect | RandomEffect} (the update sequence), `n` the number of global rows."""

    def __init__(self, coordinates, n, descent_iterations=1, all_reduce_sum=None, exchange_scores=None):
        """all_reduce_sum(float) -> float: with one process per GPU, sums a rank-local scalar over the ranks (the logged
        objective then counts every rank's random-effect terms; a fixed effect trained through the peer exchange is
        global already).  None = single process."""
        import torch
        self.all_reduce_sum = all_reduce_sum
        # exchange_scores(tensor[n]): in-place sum over the ranks of a global score vector of which every rank has
        # filled in the rows it owns (zeros elsewhere) -- the one place where per-row data crosses ranks when the
        # coordinates' rows are partitioned differently (the reference joins score RDDs by row id:
        # CoordinateDescent.scala:189,218).  None = every coordinate's rows live in this process.
        self.exchange_scores = exchange_scores
        self.torch = torch
        self.coordinates, self.n, self.passes = dict(coordinates), int(n), descent_iterations
        first = next(iter(self.coordinates.values()))
        self.ctx, self.device = first.ctx, first.device
        self.stream = torch.cuda.ExternalStream(self.ctx.stream, device=self.device)
        self.history = []

    def _to_global(self, coord, local_scores):
        if coord.row_map is None:
            return local_scores
        # rows the coordinate does not cover (dropped by its active-data bounds) score 0, as in the reference
        out = self.torch.zeros(self.n, dtype=self.torch.float64, device=self.device)
        out[coord.row_map] = local_scores
        return out

    def run(self):
        torch = self.torch
        with torch.cuda.stream(self.stream):
            scores = {name: torch.zeros(self.n, dtype=torch.float64, device=self.device) for name in self.coordinates}
            for p in range(self.passes):
                for name, coord in self.coordinates.items():
                    t0 = time.perf_counter()
                    partial = torch.zeros(self.n, dtype=torch.float64, device=self.device)
                    for other, s in scores.items():
                        if other != name:
                            partial += s
                    offsets = torch.empty(coord.n, dtype=torch.float64, device=self.device)
                    abi.offsets_from_scores_dev(self.ctx, coord.n,
                                                None if coord.base_offset is None else coord.base_offset.data_ptr(),
                                                partial.data_ptr(),
                                                None if coord.row_map is None else coord.row_map.data_ptr(),
                                                offsets.data_ptr())
                    coord.set_offsets(offsets)
                    objective, info = coord.train()
                    scores[name] = self._to_global(coord, coord.score_local())
                    if self.exchange_scores is not None:
                        self.exchange_scores(scores[name])
                    self.ctx.synchronize()
                    # the coordinate's objective = training loss of the full model + its own regularization term;
                    # add the other coordinates' terms for the value CoordinateDescent logs
                    def everywhere(x, c):     # a rank-local term of a coordinate whose rows / entities are sharded
                        local = self.all_reduce_sum is not None and getattr(c, "objective", None).__class__.__name__ != \
                            "ShardedDeviceGLMLossFunction"
                        return self.all_reduce_sum(x) if local else x
                    total = everywhere(objective, coord) + sum(everywhere(c.regularization_term(), c)
                                                               for o, c in self.coordinates.items() if o != name)
                    self.history.append({"pass": p, "coordinate": name, "objective": total,
                                         "seconds": time.perf_counter() - t0, **info})
        return {name: c.model for name, c in self.coordinates.items()}, self.history
