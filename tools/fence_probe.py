"""Probe (1 GPU): one slab rank without neighbours; does a D2H copy in flight slow the linked step down?"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common
slab = importlib.import_module(common.PKG + ".slab")
scenes = importlib.import_module(common.PKG + ".scenes")
P = common.product_lib()
sc = scenes.slab_scene("fluid_8m", 1, 0)
p = P.default_params(box=sc["box"]); p.hash_mask = 0xFFFFFFFF
p.particle_count = len(sc["gid"]); p.num_liquid = sc["n_liquid"]; p.num_boundary = len(sc["gid"]) - sc["n_liquid"]
r = slab.SlabRank(P, p, sc["position"], sc["velocity"], sc["gid"], sc["z0"], sc["z1"], 0, 1, halo_cap=sc["halo_cap"], mig_cap=sc["mig_cap"])
tr = slab.PeerTransport.local([r])
for _ in range(3): slab.slab_step(tr)
r.ctx.stream_timer_start()
for _ in range(5): slab.slab_step(tr)
print("device only  %.3f ms/step" % (r.ctx.stream_timer_stop() / 5))
bufs = [P.pinned_array((r.capacity, 4)), P.pinned_array((r.capacity, 4))]
r.read_owned_begin(bufs[0]); r.ctx.read_end()
r.sync(); t0 = time.perf_counter()
for s in range(6):
    slab.slab_step(tr); r.ctx.read_end(); r.read_owned_begin(bufs[s & 1])
r.ctx.read_end(); r.sync()
print("pipelined    %.3f ms/step" % ((time.perf_counter() - t0) / 6 * 1e3))
r.ctx.profile(True)
for s in range(6):
    slab.slab_step(tr); r.ctx.read_end(); r.read_owned_begin(bufs[s & 1])
r.ctx.read_end()
pr = r.ctx.profile_report()
print({k: round(v["ms"] / 6, 3) for k, v in pr.items()})
