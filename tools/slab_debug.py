"""Debug helper (GPU): 2 local ranks on one GPU, peer-memory links vs the loopback protocol, step by step."""
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
P = common.product_lib()


def moving_fluid(seed=7):
    pos, vel = common.synthetic_block(14, 12, 40, (24, 22, 96), seed=seed)
    liquid = pos[:, 3] < 2
    rng = np.random.RandomState(seed)
    vel[liquid, 2] = (0.45 + 0.1 * rng.rand(int(liquid.sum()))).astype(np.float32)
    vel[liquid, 0] = (0.05 * (rng.rand(int(liquid.sum())) - 0.5)).astype(np.float32)
    return pos, vel


def build(kind, p, pos, vel, world):
    parts, caps = slab.split_fluid(p, pos, world)
    ranks = [slab.SlabRank(P, p, pos[o], vel[o], o.astype(np.uint32), z0, z1, r, world, **caps) for r, (z0, z1, o) in enumerate(parts)]
    tr = slab.PeerTransport.local(ranks) if kind == "peer" else slab.LoopbackTransport(ranks)
    return ranks, tr


pos, vel = moving_fluid()
p = common.fill_counts(P.default_params(box=(12, 11, 48)), pos)
p.hash_mask = 0xFFFFFFFF
world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
ra, ta = build("loop", p, pos, vel, world)
rb, tb = build("peer", p, pos, vel, world)
for s in range(12):
    slab.slab_step(ta)
    try:
        slab.slab_step(tb)
    except RuntimeError as e:
        print("step", s, "peer failed:", e)
        break
    pa, va = slab.gather_owned(ra, len(pos))
    pb, vb = slab.gather_owned(rb, len(pos))
    same = np.array_equal(pa.view(np.uint32), pb.view(np.uint32))
    print("step", s, "counts loop", [r.counts for r in ra], "peer", [r.counts for r in rb], "info", [r.info() for r in rb], "same", same)
    if not same:
        bad = np.nonzero((pa.view(np.uint32) != pb.view(np.uint32)).any(axis=1))[0]
        print("  differing particles", len(bad), bad[:10], pa[bad[:3]], pb[bad[:3]])
        break
