#!/bin/bash
mkdir -p gpurun_out
for h2 in 101 50 30 10; do
SPH_HIST2_PCT=$h2 timeout 200 python - > gpurun_out/b16_cfgb_$h2.log 2>&1 <<'PY'
import sys; sys.path.insert(0, '.')
import bench
L = bench.capi.load_product()
for name in ("config_b", "config_a"):
    scene = bench.make_workload(name, L)
    p, pos, vel = scene[:3]
    ctx = L.create(p, pos, vel, **bench.EXTRA.get(name, {}))
    ctx.run_steps(20)
    best = min(ctx.run_steps(50) / 50 for _ in range(5))
    c = ctx.debug_counters()
    print(name, "ms/step %.4f" % best, "needy %.1f%% mode %d" % (100.0 * c[0] / len(pos), c[1]))
PY
done
grep -h "ms/step" gpurun_out/b16_cfgb_*.log
