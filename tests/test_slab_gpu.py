"""GPU tests of the slab decomposition: a fluid cut into 2 / 3 / 4 slabs — all ranks on ONE
GPU, bytes moved by the loopback transport — must give, per global particle id, the positions and velocities
of the single-device run, bit for bit, while particles migrate between slabs."""
import importlib

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
slab = importlib.import_module(common.PKG + ".slab")


@pytest.fixture(scope="module", autouse=True)
def tiled_neighbors():
    """Slab ranks of a bench-sized fluid take the tile kernel; the test scenes are small, so force it."""
    import os
    old = os.environ.get("SPH_NEIGHBOR_PATH")
    os.environ["SPH_NEIGHBOR_PATH"] = "tiles"
    yield
    if old is None:
        os.environ.pop("SPH_NEIGHBOR_PATH", None)
    else:
        os.environ["SPH_NEIGHBOR_PATH"] = old


@pytest.fixture(scope="module")
def P():
    return common.product_lib()


def moving_fluid(seed=7):
    """A liquid column in a tall box, thrown along z so that whole layers of particles change slab."""
    pos, vel = common.synthetic_block(14, 12, 40, (24, 22, 96), seed=seed)
    liquid = pos[:, 3] < 2
    rng = np.random.RandomState(seed)
    vel[liquid, 2] = (0.45 + 0.1 * rng.rand(int(liquid.sum()))).astype(np.float32)   # m/s: ~1.2 grid units per step
    vel[liquid, 0] = (0.05 * (rng.rand(int(liquid.sum())) - 0.5)).astype(np.float32)
    return pos, vel


def make(P, pos, vel):
    p = common.fill_counts(P.default_params(box=(12, 11, 48)), pos)
    p.hash_mask = 0xFFFFFFFF
    return p


def run_single(P, p, pos, vel, steps):
    ctx = P.create(p, pos, vel)
    for _ in range(steps):
        ctx.step()
    out = ctx.read("position", len(pos)), ctx.read("velocity", len(pos))
    ctx.close()
    return out


def run_slabs(P, p, pos, vel, world, steps):
    ranks = []
    slabs, caps = slab.split_fluid(p, pos, world)
    for r, (z0, z1, own) in enumerate(slabs):
        ranks.append(slab.SlabRank(P, p, pos[own], vel[own], own.astype(np.uint32), z0, z1, r, world, **caps))
    tr = slab.LoopbackTransport(ranks)
    moved = 0
    for _ in range(steps):
        slab.slab_step(tr)
        moved += sum(c[0] + c[2] for c in [r.counts for r in ranks])
    for r in ranks:   # the pipelined read of the owned positions delivers what the blocking read does
        want = r.read_owned()[0]
        buf = P.pinned_array((len(want) + 8, 4))
        r.read_owned_begin(buf)
        r.ctx.read_end()
        assert np.array_equal(buf[:len(want)], want)
    out = slab.gather_owned(ranks, len(pos))
    infos = [r.info() for r in ranks]
    for r in ranks:
        r.close()
    return out, moved, infos


@pytest.mark.parametrize("world", [2, 3, 4])
def test_slabs_match_single_device(P, world):
    pos, vel = moving_fluid()
    p = make(P, pos, vel)
    steps = 12
    ref_pos, ref_vel = run_single(P, p, pos, vel, steps)
    (sp, sv), moved, infos = run_slabs(P, p, pos, vel, world, steps)
    assert moved > 0, "the test fluid never crossed a slab boundary"
    assert sum(i[0] for i in infos) == len(pos)
    mm = common.first_mismatch(sp, ref_pos)
    assert mm is None, "position differs from the single-device run: %s" % mm
    mm = common.first_mismatch(sv, ref_vel)
    assert mm is None, "velocity differs from the single-device run: %s" % mm


def test_slab_fluid_at_rest_and_errors(P):
    pos, vel = common.synthetic_block(12, 12, 24, (20, 20, 64), seed=2)
    p = common.fill_counts(P.default_params(box=(10, 10, 32)), pos)
    p.hash_mask = 0xFFFFFFFF
    ref_pos, ref_vel = run_single(P, p, pos, vel, 4)
    (sp, sv), _, _ = run_slabs(P, p, pos, vel, 2, 4)
    assert common.first_mismatch(sp, ref_pos) is None and common.first_mismatch(sv, ref_vel) is None
    # the reference's 16-bit mask aliases distant z layers: slabs refuse it
    p16 = common.fill_counts(P.default_params(box=(10, 10, 32)), pos)
    with pytest.raises(RuntimeError):
        slab.SlabRank(P, p16, pos, vel, np.arange(len(pos), dtype=np.uint32), 0, 8, 0, 1)
