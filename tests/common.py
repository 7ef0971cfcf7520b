"""Shared helpers for the test-suite: library loading, configs, bitwise comparison.

oracle/ is test infrastructure: this module (tests/) is one of the few places allowed to load it.
"""
import ctypes as C
import hashlib
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "smoothed-particle-hydrodynamics_b200"
capi = importlib.import_module(PKG + ".capi")
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libsphref.so")

ALL_STATE = ["position", "velocity", "particleIndex", "particleIndexBack", "sortedPosition", "sortedVelocity",
             "gridCellIndex", "gridCellIndexFixedUp", "neighborMap", "rho", "pressure", "acceleration"]


def oracle_lib():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_oracle
    return capi.SphLib(build_oracle.build())


def ref_lib():
    """The reference's own kernels compiled as C++ (None when oracle/_ref was not built/shipped)."""
    return capi.SphLib(REF_LIB) if os.path.exists(REF_LIB) else None


def product_lib():
    return capi.load_product()


def digests():
    with open(os.path.join(GOLDEN, "digests.json")) as f:
        return json.load(f)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def fill_counts(p, pos, num_membranes=0):
    t = pos[:, 3].astype(np.int32)
    p.particle_count = len(pos)
    p.num_elastic = int((t == 2).sum())
    p.num_liquid = int((t == 1).sum())
    p.num_boundary = int((t == 3).sum())
    p.num_membranes = num_membranes
    return p


def config_a(lib):
    z = np.load(os.path.join(GOLDEN, "configA.npz"))
    pos, vel = z["position"], z["velocity"]
    return dict(params=fill_counts(lib.default_params(box=(36, 24, 54)), pos), position=pos, velocity=vel)


def config_worm(lib):
    z = np.load(os.path.join(GOLDEN, "worm.npz"))
    pos = z["position"]
    return dict(params=fill_counts(lib.default_params(), pos, len(z["membraneData"])), position=pos,
                velocity=z["velocity"], elastic=z["elasticConnectionsData"], membranes=z["membraneData"],
                pml=z["particleMembranesList"])


def create(lib, cfg, device=0):
    return lib.create(cfg["params"], cfg["position"], cfg["velocity"], cfg.get("elastic"), cfg.get("membranes"),
                      cfg.get("pml"), device)


def muscle_signal(step):
    """Same deterministic signal as tests/golden/make_golden.py."""
    s = np.zeros(100, np.float32)
    k = np.arange(96, dtype=np.float32)
    s[:96] = (0.5 + 0.5 * np.sin(0.37 * k + 0.9 * step)).astype(np.float32)
    return s


def stage_order(max_iteration=3):
    return capi.STAGES[:10] + capi.STAGES[10:14] * max_iteration + capi.STAGES[14:]


def first_mismatch(x, y):
    """Bitwise comparison; the only tolerance: any NaN equals any NaN (x86 SSE generates the
    negative default NaN 0xffc00000, CUDA 0x7fffffff — IEEE leaves sign/payload unspecified)."""
    bx, by = bits(x).reshape(len(x), -1), bits(y).reshape(len(y), -1)
    neq = bx != by
    if np.asarray(x).dtype == np.float32 and neq.any():
        fx, fy = np.asarray(x).reshape(len(x), -1), np.asarray(y).reshape(len(y), -1)
        neq &= ~(np.isnan(fx) & np.isnan(fy))
    d = np.argwhere(neq)
    if len(d) == 0:
        return None
    r = int(d[0][0])
    return dict(rows=len(np.unique(d[:, 0])), first_row=r, a=np.asarray(x).reshape(len(x), -1)[r].tolist(),
                b=np.asarray(y).reshape(len(y), -1)[r].tolist())


def canon(name, a):
    """Drop the lanes the reference itself treats as junk: the .w of the two
    acceleration halves (velocity.w / cell-id leftovers, zeroed before use at sphFluid.cl:1721) and the
    .w of the predicted positions sortedPosition[N,2N) (never read). Everything else stays bit-exact,
    including position.w (type tag), velocity.w and sortedPosition[0,N).w (cell id)."""
    a = np.array(a, copy=True)
    if name == "acceleration":
        a[:, 3] = 0
    elif name == "sortedPosition" and len(a) % 2 == 0:
        a[len(a) // 2:, 3] = 0
    return a


def assert_same_state(ca, cb, buffers=ALL_STATE, tag="", no_nan=False):
    """no_nan: also require that neither side holds a NaN anywhere (first_mismatch lets any NaN match any NaN; the
    fixture and bench scenes must never get there)."""
    for b in buffers:
        x, y = canon(b, ca.read(b)), canon(b, cb.read(b))
        if no_nan and x.dtype == np.float32 and b != "neighborMap":
            assert not np.isnan(x).any() and not np.isnan(y).any(), "%s: NaN in buffer %s" % (tag, b)
        mm = first_mismatch(x, y)
        assert mm is None, "%s: buffer %s differs: %s" % (tag, b, mm)


def synthetic_block(nx, ny, nz, box_r0, seed=12345, jitter=0.05, r0=None, h=None, origin_r0=(3, 3, 3)):
    """Liquid block on the reference's r0 lattice + one-layer boundary box with inward normals
    (layout of /root/reference/src/owHelper.cpp:776-928), deterministic counter-based jitter.
    box_r0 = (bx,by,bz) box size in units of r0 (must be even: the box is box_r0/2 * h)."""
    h = np.float32(3.34) if h is None else np.float32(h)
    r0 = np.float32(0.5) * h if r0 is None else np.float32(r0)
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    idx = np.stack([ix.ravel(), iy.ravel(), iz.ravel()], 1).astype(np.float32)
    org = np.asarray(origin_r0, np.float32)
    pos = np.zeros((len(idx), 4), np.float32)
    pos[:, :3] = (idx + org) * r0
    rng = np.random.RandomState(seed)
    pos[:, :3] += (rng.uniform(-jitter, jitter, size=(len(idx), 3)) * float(r0)).astype(np.float32)
    pos[:, 3] = np.float32(1.1)
    vel = np.zeros_like(pos)
    bx, by, bz = box_r0
    bpos, bvel = [], []
    for i in range(bx):
        for j in range(by):
            for k in range(bz):
                fx, fy, fz = (i == 0) - (i == bx - 1), (j == 0) - (j == by - 1), (k == 0) - (k == bz - 1)
                nn = abs(fx) + abs(fy) + abs(fz)
                if nn == 0:
                    continue
                s = np.float32(1.0) / np.sqrt(np.float32(nn))
                bpos.append(((i + 0.5) * r0, (j + 0.5) * r0, (k + 0.5) * r0, 3.0))
                bvel.append((fx * s, fy * s, fz * s, 3.0))
    bpos = np.asarray(bpos, np.float32).reshape(-1, 4)
    bvel = np.asarray(bvel, np.float32).reshape(-1, 4)
    return np.concatenate([pos, bpos]), np.concatenate([vel, bvel])
