#!/usr/bin/env python3
"""make_golden.py — generate the committed fixtures under tests/golden/ from the REFERENCE itself.

Run in the build container only (needs /root/reference and oracle/_ref/*.so built by
oracle/build_ref.py).  Nothing here is imported by the product.

  configA.npz   BASELINE config 0/A: /root/reference/configuration/positionPureLiquid.txt +
                velocityPureLiquid.txt parsed to float32 (61 440 x 4 each)
  worm.npz      BASELINE config 1/B: output of the reference's own generateConfiguration
                (oracle/_ref/libowgen.so = /root/reference/src/owHelper.cpp compiled unchanged)
  digests.json  sha256 of every solver buffer (reference layout, raw bytes) after every stage of
                step 0 and of position/velocity/rho/particleIndex after later steps, produced by
                oracle/_ref/libsphref.so (= /root/reference/src/sphFluid.cl compiled as C++),
                plus small literal samples (first rows) for human inspection.
"""
import ctypes as C
import hashlib
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
capi = importlib.import_module("smoothed-particle-hydrodynamics_b200.capi")
HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"

STEP0_BUFFERS = ["position", "velocity", "particleIndex", "particleIndexBack", "sortedPosition", "sortedVelocity",
                 "gridCellIndex", "gridCellIndexFixedUp", "neighborMap", "rho", "pressure", "acceleration"]
LATER_BUFFERS = ["position", "velocity", "rho", "particleIndex"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def muscle_signal(step):
    """Deterministic test signal (NOT main_sim.py): 96 values in [0,1), zero for muscles 96..99."""
    s = np.zeros(100, np.float32)
    k = np.arange(96, dtype=np.float32)
    s[:96] = (0.5 + 0.5 * np.sin(0.37 * k + 0.9 * step)).astype(np.float32)
    return s


def make_config_a():
    pos = np.loadtxt(os.path.join(REF, "configuration/positionPureLiquid.txt"), dtype=np.float32)
    vel = np.loadtxt(os.path.join(REF, "configuration/velocityPureLiquid.txt"), dtype=np.float32)
    np.savez_compressed(os.path.join(HERE, "configA.npz"), position=pos, velocity=vel)
    return pos, vel


def make_worm():
    G = C.CDLL(os.path.join(ROOT, "oracle/_ref/libowgen.so"))
    n, ne, nl, nb, nm = [C.c_int() for _ in range(5)]
    G.owgen_count(C.byref(n), C.byref(ne), C.byref(nl), C.byref(nb), C.byref(nm))
    N, Ne, M = n.value, ne.value, nm.value
    pos = np.zeros((N, 4), np.float32)
    vel = np.zeros((N, 4), np.float32)
    el = np.zeros((Ne * 32, 4), np.float32)
    mem = np.zeros((M, 3), np.int32)
    pml = np.zeros((Ne, 7), np.int32)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    G.owgen_fill(vp(pos), vp(vel), vp(el), vp(mem), vp(pml))
    np.savez_compressed(os.path.join(HERE, "worm.npz"), position=pos, velocity=vel, elasticConnectionsData=el,
                        membraneData=mem, particleMembranesList=pml,
                        counts=np.array([N, Ne, nl.value, nb.value, M], np.int32))
    return pos, vel, el, mem, pml


def params_for(L, pos, box=None, num_membranes=0):
    p = L.default_params(box=box)
    t = pos[:, 3].astype(np.int32)
    p.particle_count = len(pos)
    p.num_elastic = int((t == 2).sum())
    p.num_liquid = int((t == 1).sum())
    p.num_boundary = int((t == 3).sum())
    p.num_membranes = num_membranes
    return p


def digest_run(L, p, pos, vel, el=None, mem=None, pml=None, later_steps=(1, 2, 5), muscles=False):
    ctx = L.create(p, pos, vel, el, mem, pml)
    out = {"step0": {}, "later": {}, "samples": {}}
    if muscles:
        ctx.set_muscles(muscle_signal(0))
    order = capi.STAGES[:10] + capi.STAGES[10:14] * p.max_iteration + capi.STAGES[14:]
    seen = {}
    for st in order:
        ctx.stage(st)
        k = seen.get(st, 0)
        seen[st] = k + 1
        out["step0"]["%s#%d" % (st, k)] = {b: sha(ctx.read(b)) for b in STEP0_BUFFERS}
    out["samples"]["position_after_step0"] = ctx.read("position", 8).tolist()
    out["samples"]["rho_after_step0"] = ctx.read("rho", 8).tolist()
    out["samples"]["neighborMap_row0_after_step0"] = ctx.read("neighborMap", 32).tolist()
    done = 1
    for s in sorted(later_steps):
        while done < s:
            if muscles:
                ctx.set_muscles(muscle_signal(done))
            ctx.step()
            done += 1
        out["later"]["after_%d_steps" % s] = {b: sha(ctx.read(b)) for b in LATER_BUFFERS}
    ctx.close()
    return out


def main():
    R = capi.SphLib(os.path.join(ROOT, "oracle/_ref/libsphref.so"))
    assert R.backend == "reference-cl-as-c++"
    dig = {"generator": "oracle/_ref/libsphref.so (sphFluid.cl compiled as C++), single thread",
           "params_default_hex": R.default_params().as_bytes().hex(),
           "params_boxA_hex": R.default_params(box=(36, 24, 54)).as_bytes().hex()}
    pos, vel = make_config_a()
    dig["configA"] = digest_run(R, params_for(R, pos, box=(36, 24, 54)), pos, vel, later_steps=(1, 2, 5, 10))
    pos, vel, el, mem, pml = make_worm()
    dig["worm"] = digest_run(R, params_for(R, pos, num_membranes=len(mem)), pos, vel, el, mem, pml,
                             later_steps=(1, 2, 3), muscles=True)
    with open(os.path.join(HERE, "digests.json"), "w") as f:
        json.dump(dig, f, indent=1, sort_keys=True)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
