"""One table from the per-N bench lines kept under gpurun_out/ or profiles/ (r02_bench_nN.json, r02_c2_nN.json, r02_c4_nN.json)."""
import glob
import json
import os
import sys

d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
for tag, title in (("bench", "configs[1] value+gradient evals/s (strong scaling)"), ("c2", "configs[2] evals/s (weak scaling, 25 M rows per GPU)"),
                   ("c3", "configs[3] GLMix coordinate descent, rows x passes per second (strong scaling)"),
                   ("c4", "configs[4] per-member problems solved/s (strong scaling) and the full coordinate descent")):
    print(title)
    base = None
    for n in (1, 2, 4, 8):
        path = os.path.join(d, "r02_%s_n%d.json" % (tag, n))
        if not os.path.exists(path) or os.path.getsize(path) == 0:
            continue
        j = json.loads(open(path).read().strip().splitlines()[-1])
        v = j["value"]
        if base is None:
            base = (n, v)
        ideal = base[1] * (n / base[0] if j["scaling"] == "strong" else 1.0)
        extra = ""
        if "e2e" in j:
            extra = " e2e %.1f (%.3f ms)" % (j["e2e"]["value"], j["e2e"]["ms_per_step"])
        if tag == "c2":
            extra += " rows/s %.3g" % j["rows_per_s"]
        if tag == "c3":
            extra += " wall %.3f s for %d passes over %d rows" % (j["wall_s"], j["steps"], j["config"]["rows_total"])
        if tag == "c4" and j.get("coordinate_descent"):
            cd = j["coordinate_descent"]
            extra += "; fixed (%.3g entries) + per-member coordinate descent: %d passes over %d rows in %.3f s" % (
                cd["fixed_effect_entries_total"], cd["passes"], cd["rows_total"], cd["wall_s"])
        print("  N=%d: %.1f %s, %.4f ms/step, efficiency %.3f%s" % (n, v, j["unit"], j["ms_per_step"], v / ideal, extra))
