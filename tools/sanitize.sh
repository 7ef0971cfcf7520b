#!/bin/bash
# compute-sanitizer evidence (SURVEY.md §5 "race detection", §8a.2 item 11): memcheck and racecheck over
#   the smoke scene (3 fused steps), one worm step (elastic + membranes + muscles), one 2-slab peer-memory step.
# Logs -> gpurun_out/r02_sanitizer_{memcheck,racecheck}.log (copied to profiles/ by hand).
mkdir -p gpurun_out
cat > /tmp/san_driver.py <<'P'
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import common
import __graft_entry__ as g
g.smoke()
P_ = common.product_lib()
cfg = common.config_worm(P_)
a = common.create(P_, cfg)
a.set_muscles(common.muscle_signal(0)); a.step(); a.close()
print("worm step OK")
slab = importlib.import_module(common.PKG + ".slab")
pos, vel = common.synthetic_block(12, 12, 24, (20, 20, 64), seed=2)
p = common.fill_counts(P_.default_params(box=(10, 10, 32)), pos); p.hash_mask = 0xFFFFFFFF
parts, caps = slab.split_fluid(p, pos, 2)
ranks = [slab.SlabRank(P_, p, pos[o], vel[o], o.astype(np.uint32), z0, z1, r, 2, **caps) for r, (z0, z1, o) in enumerate(parts)]
tr = slab.PeerTransport.local(ranks)
for _ in range(2): slab.slab_step(tr)
for r in ranks: r.sync(); r.close()
print("slab steps OK")
P
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python /tmp/san_driver.py > gpurun_out/r02_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?" >> gpurun_out/r02_sanitizer_$tool.log
  tail -6 gpurun_out/r02_sanitizer_$tool.log
done
