import sys, json, time
sys.path.insert(0, "/root/repo")
import bench, photon_ml_b200 as pml
ctx = pml.Context(0)
out = bench.random_effect_batch(ctx)
print(json.dumps(out, indent=1))
