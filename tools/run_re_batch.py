"""The per-entity batch of bench.py on its own: one batched evaluation with its roofline, and the lock-step solve
with and without skipping the rows of converged entities."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import photon_ml_b200 as pml  # noqa: E402

ctx = pml.Context(0)
print(json.dumps(bench.random_effect_batch(ctx), indent=1))
