"""Per-kernel times (CUDA events) of a scene after it has been stepped for a while: tools/late_profile.py WORKLOAD STEPS"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
L = bench.capi.load_product()
name, before = sys.argv[1], int(sys.argv[2])
scene = bench.make_workload(name, L)
p, pos, vel = scene[:3]
ctx = L.create(p, pos, vel, **bench.EXTRA.get(name, {}))
done = 0
while done < before:
    ctx.run_steps(100); done += 100
    if ctx.out_of_grid_count():
        print("diverged after", done); sys.exit(1)
import numpy as np
print("pressure>0 frac", float((ctx.read("pressure") > 0).mean()), "after", done, "steps")
c0 = ctx.debug_counters()
ctx.profile(True)
ms = ctx.run_steps(30) / 30
pr = ctx.profile_report()
print(name, "ms/step %.4f" % ms, {k: round(v["ms"] / 30, 4) for k, v in pr.items()})
c1 = ctx.debug_counters()
print("counters: needy last search %d (%.1f%%), hist-first mode %d, fallbacks to the plain scan per step %.0f (%.2f%%)" % (c1[0], 100.0 * c1[0] / len(pos), c1[1], (c1[3] - c0[3]) / 30.0, 100.0 * (c1[3] - c0[3]) / 30.0 / len(pos)))
nm = ctx.read("neighborMap")
print("mean neighbours", float((nm[:, 0] >= 0).reshape(-1, 32).sum(1).mean()))
