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
ctx.profile(True)
ms = ctx.run_steps(30) / 30
pr = ctx.profile_report()
print(name, "ms/step %.4f" % ms, {k: round(v["ms"] / 30, 4) for k, v in pr.items()})
nm = ctx.read("neighborMap")
print("mean neighbours", float((nm[:, 0] >= 0).reshape(-1, 32).sum(1).mean()))
