"""Probe (GPU): pipelined read-back with 2 local ranks on one GPU, loopback protocol vs peer-memory links."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402

slab = importlib.import_module(common.PKG + ".slab")
scenes = importlib.import_module(common.PKG + ".scenes")
P = common.product_lib()
world = 2
name = sys.argv[1] if len(sys.argv) > 1 else "fluid_2m"
for kind in ("loop", "peer"):
    ranks = []
    for r in range(world):
        sc = scenes.slab_scene(name, world, r)
        p = P.default_params(box=sc["box"])
        p.hash_mask = 0xFFFFFFFF
        p.particle_count = len(sc["gid"]); p.num_liquid = sc["n_liquid"]; p.num_boundary = len(sc["gid"]) - sc["n_liquid"]
        ranks.append(slab.SlabRank(P, p, sc["position"], sc["velocity"], sc["gid"], sc["z0"], sc["z1"], r, world,
                                   halo_cap=sc["halo_cap"], mig_cap=sc["mig_cap"]))
    tr = slab.PeerTransport.local(ranks) if kind == "peer" else slab.LoopbackTransport(ranks)
    for _ in range(3):
        slab.slab_step(tr)
    for r in ranks:
        r.sync()
    t0 = time.perf_counter()
    for _ in range(5):
        slab.slab_step(tr)
    for r in ranks:
        r.sync()
    dev = (time.perf_counter() - t0) / 5 * 1e3
    bufs = [[P.pinned_array((r.capacity, 4)), P.pinned_array((r.capacity, 4))] for r in ranks]
    for i, r in enumerate(ranks):
        r.read_owned_begin(bufs[i][0]); r.ctx.read_end()
    t0 = time.perf_counter()
    marks = []
    for s in range(6):
        a = time.perf_counter()
        slab.slab_step(tr)
        b = time.perf_counter()
        for i, r in enumerate(ranks):
            r.ctx.read_end()
        c = time.perf_counter()
        for i, r in enumerate(ranks):
            r.read_owned_begin(bufs[i][s & 1])
        marks.append(((b - a) * 1e3, (c - b) * 1e3))
    for r in ranks:
        r.ctx.read_end()
    pip = (time.perf_counter() - t0) / 6 * 1e3
    t0 = time.perf_counter()
    for s in range(4):
        slab.slab_step(tr)
        for i, r in enumerate(ranks):
            r.read_owned_begin(bufs[i][0]); r.ctx.read_end()
    blk = (time.perf_counter() - t0) / 4 * 1e3
    print(kind, name, "device %.2f ms  pipelined %.2f ms  blocking %.2f ms" % (dev, pip, blk), "marks (step call, read_end wait):", [(round(x, 2), round(y, 2)) for x, y in marks])
    for r in ranks:
        r.close()
