"""GPU tests at the sizes BASELINE.json quotes (-m gpu), where the CPU oracle is too slow to be the checker for
more than a step or two:

 * fluid_256k (290 712 particles, the smallest size at which the library picks the tile kernel by itself):
   two fused steps against the oracle, every state buffer bit for bit;
 * fluid_2m: structural invariants of the sort and of the neighbour lists (permutation, monotone cell ranges,
   ids in range / no self / padding after the last neighbour / stored distance within the largest r_thr and equal
   to the distance recomputed from the sorted positions);
 * fluid_8m (the bench configuration, 8 256 352 particles): the two neighbour-search implementations — tile kernel
   and one-thread-per-particle scan — must give bit-identical positions and velocities after two steps.
"""
import importlib
import os

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
scenes = importlib.import_module(common.PKG + ".scenes")


def workload(lib, name):
    nx, ny, nz, fill = scenes.WORKLOADS[name]
    pos, vel, box = scenes.fluid_box(nx, ny, nz, fill_z=fill)
    p = common.fill_counts(lib.default_params(box=box), pos)
    p.hash_mask = 0xFFFFFFFF
    return p, pos, vel


@pytest.fixture(scope="module")
def P():
    return common.product_lib()


def with_path(path):
    old = os.environ.get("SPH_NEIGHBOR_PATH")
    if path is None:
        os.environ.pop("SPH_NEIGHBOR_PATH", None)
    else:
        os.environ["SPH_NEIGHBOR_PATH"] = path
    return old


def test_fluid_256k_two_steps_against_oracle(P):
    O = common.oracle_lib()
    p, pos, vel = workload(P, "fluid_256k")
    old = with_path(None)   # the library's own choice: the list kernel
    try:
        a, b = P.create(p, pos, vel), O.create(p, pos, vel)
        for step in range(2):
            a.step()
            b.step_staged()
            common.assert_same_state(a, b, tag="fluid_256k step %d" % step, no_nan=True)
        a.close()
        b.close()
    finally:
        with_path(old)


def test_fluid_2m_sort_and_neighbour_invariants(P):
    p, pos, vel = workload(P, "fluid_2m")
    n = len(pos)
    ctx = P.create(p, pos, vel)
    ctx.step()
    pi = ctx.read("particleIndex", n)
    assert np.array_equal(np.sort(pi[:, 1]), np.arange(n, dtype=np.uint32)), "particleIndex is not a permutation"
    assert (np.diff(pi[:, 0].astype(np.int64)) >= 0).all(), "cell ids are not sorted"
    back = ctx.read("particleIndexBack", n)
    assert np.array_equal(pi[back, 1], np.arange(n, dtype=np.uint32))
    cells = ctx.read("gridCellIndexFixedUp")
    assert cells[0] == 0 and cells[-1] == n and (np.diff(cells.astype(np.int64)) >= 0).all()
    counts = np.bincount(pi[:, 0], minlength=len(cells) - 1)
    assert np.array_equal(np.diff(cells.astype(np.int64)), counts), "cell ranges do not match the sorted cell ids"
    nm = ctx.read("neighborMap").reshape(n, 32, 2)
    ids, dist = nm[:, :, 0], nm[:, :, 1]
    has = ids >= 0
    assert (ids[has] < n).all() and (ids[has] == np.floor(ids[has])).all()
    assert not (ids == np.arange(n, dtype=np.float32)[:, None]).any(), "a particle lists itself"
    assert not (has[:, 1:] & ~has[:, :-1]).any(), "a neighbour after the padding started"
    assert (dist[~has] == -1).all()
    h_scaled = np.float32(p.h) * np.float32(p.simulation_scale)
    assert (dist[has] <= h_scaled * np.float32(31.0 / 30.0) * np.float32(1.000001)).all()
    sp = ctx.read("sortedPosition", n)
    rows, cols = np.nonzero(has)
    sel = slice(None, None, 97)   # every 97th pair: 700 k distances recomputed on the host
    i, j = rows[sel], ids[rows[sel], cols[sel]].astype(np.int64)
    e = sp[i, :3] - sp[j, :3]
    d2 = (e[:, 0] * e[:, 0] + e[:, 1] * e[:, 1]) + e[:, 2] * e[:, 2]
    want = np.sqrt(d2, dtype=np.float32) * np.float32(p.simulation_scale)
    assert np.array_equal(want, dist[rows[sel], cols[sel]]), "stored distance != distance of the sorted positions"
    rho = ctx.read("rho", n)
    assert np.isfinite(rho).all() and (rho > 0).all()
    ctx.close()


def test_fluid_8m_tile_kernel_equals_plain_scan(P):
    p, pos, vel = workload(P, "fluid_8m")
    n = len(pos)
    out = {}
    old = os.environ.get("SPH_NEIGHBOR_PATH")
    try:
        for path in ("tiles", "generic", "list"):
            with_path(path)
            ctx = P.create(p, pos, vel)
            ctx.step()
            ctx.step()
            out[path] = (ctx.read("position", n), ctx.read("velocity", n))
            ctx.close()
    finally:
        with_path(old)
    for path in ("tiles", "list"):
        assert common.first_mismatch(out[path][0], out["generic"][0]) is None
        assert common.first_mismatch(out[path][1], out["generic"][1]) is None
    assert not np.array_equal(out["tiles"][0], pos), "the fluid did not move"


def test_dambreak_1m_moving_fluid_tile_kernel_equals_plain_scan(P):
    """Config C: a liquid block in a larger box, started imploding. 40 steps move and compress the fluid, so the
    tile kernel meets uneven cells, overflowing tables (deferred list) and empty blocks; the plain scan checks it."""
    p, pos, vel = workload(P, "dambreak_1m")
    n = len(pos)
    liquid = pos[:, 3] < 2
    rng = np.random.RandomState(5)   # an imploding start: ~0.5 grid units per step towards the block's centre
    c = pos[liquid, :3].mean(axis=0)
    dirn = c - pos[liquid, :3]
    dirn /= np.maximum(np.linalg.norm(dirn, axis=1, keepdims=True), 1e-6)
    vel[liquid, :3] = (0.2 * dirn + 0.05 * (rng.rand(int(liquid.sum()), 3) - 0.5)).astype(np.float32)
    out = {}
    old = os.environ.get("SPH_NEIGHBOR_PATH")
    try:
        for path in ("tiles", "generic", "list"):
            with_path(path)
            ctx = P.create(p, pos, vel)
            for _ in range(40):
                ctx.step(sync=False)
            ctx.sync()
            out[path] = (ctx.read("position", n), ctx.read("velocity", n), ctx.read("rho", n))
            ctx.close()
    finally:
        with_path(old)
    for path in ("tiles", "list"):
        for a, b, name in zip(out[path], out["generic"], ("position", "velocity", "rho")):
            assert common.first_mismatch(a, b) is None, (path, name)
    moved = np.abs(out["tiles"][0][liquid, :3] - pos[liquid, :3]).max()
    assert moved > 2.0, "the block did not move (max displacement %g)" % moved


def test_fluid_2m_one_step_against_oracle(P):
    O = common.oracle_lib()
    p, pos, vel = workload(P, "fluid_2m")
    a, b = P.create(p, pos, vel), O.create(p, pos, vel)
    a.step()
    b.step_staged()
    common.assert_same_state(a, b, buffers=("position", "velocity", "rho", "pressure", "neighborMap"), tag="fluid_2m", no_nan=True)
    a.close()
    b.close()


def test_fluid_8m_bench_configuration_one_step_against_oracle(P):
    """The bench workload itself (BASELINE config E, 8 256 352 particles): one fused step, bit for bit against the
    CPU oracle (about half a minute of CPU time)."""
    O = common.oracle_lib()
    p, pos, vel = workload(P, "fluid_8m")
    a, b = P.create(p, pos, vel), O.create(p, pos, vel)
    a.step()
    b.step_staged()
    common.assert_same_state(a, b, buffers=("position", "velocity", "rho", "pressure"), tag="fluid_8m", no_nan=True)
    a.close()
    b.close()
