"""GPU tests of the slab decomposition: a fluid cut into 2 / 3 / 4 slabs — all ranks on ONE
GPU, bytes moved by the loopback transport — must give, per global particle id, the positions and velocities
of the single-device run, bit for bit, while particles migrate between slabs."""
import importlib

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
slab = importlib.import_module(common.PKG + ".slab")


@pytest.fixture(scope="module", autouse=True, params=["list", "tiles"])
def neighbor_path(request):
    """Slab ranks run the ordinary kernels: the list kernel (the library's choice) and the tile kernel."""
    import os
    old = os.environ.get("SPH_NEIGHBOR_PATH")
    os.environ["SPH_NEIGHBOR_PATH"] = request.param
    yield
    if old is None:
        os.environ.pop("SPH_NEIGHBOR_PATH", None)
    else:
        os.environ["SPH_NEIGHBOR_PATH"] = old


@pytest.fixture(scope="module")
def P():
    return common.product_lib()


@pytest.fixture(scope="module", autouse=True, params=["loopback", "peer"])
def transport_kind(request):
    """loopback: the library packs / unpacks, the test copies the bytes between the ranks' buffers (round-1 protocol,
    what the NCCL transport also drives). peer: the ranks' kernels write into each other's mailboxes and the whole step
    runs inside the library (sph_slab_step_begin / _end) — the product path of bench.py --gpus N."""
    global KIND
    KIND = request.param
    yield request.param


KIND = "loopback"


def make_transport(ranks):
    return slab.PeerTransport.local(ranks) if KIND == "peer" else slab.LoopbackTransport(ranks)


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
    tr = make_transport(ranks)
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
        xyz = P.pinned_array((3 * (len(want) + 8),))   # and the packed (x, y, z) form
        r.read_owned_xyz_begin(xyz)
        r.ctx.read_end()
        assert np.array_equal(xyz[:3 * len(want)].reshape(-1, 3), want[:, :3])
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


def test_slab_recut_hands_layers_over(P):
    """Load balancing: an uneven initial cut, slab.rebalance() before every step, and one face forced back and forth
    every step. Whole cell layers change owner while the fluid moves; per global id the result stays bit-identical
    to the single-device run."""
    pos, vel = moving_fluid(seed=3)
    p = make(P, pos, vel)
    steps = 10
    ref_pos, ref_vel = run_single(P, p, pos, vel, steps)
    cz = slab.cell_layer(pos[:, 2], p.hash_grid_cell_size_inv)
    top = int(cz.max()) + 1
    cuts = [0, 3, top - 4, int(p.grid_cells_z)]   # most of the fluid starts on the middle rank
    fullest = int(np.bincount(cz).max())
    caps = dict(halo_cap=2 * fullest + 1024, mig_cap=fullest + 1024)
    ranks = []
    for r in range(3):
        own = np.nonzero((cz >= cuts[r]) & (cz < cuts[r + 1]))[0]
        ranks.append(slab.SlabRank(P, p, pos[own], vel[own], own.astype(np.uint32), cuts[r], cuts[r + 1], r, 3,
                                   capacity=len(pos) + 4 * caps["halo_cap"], **caps))
    tr = make_transport(ranks)
    moves, sizes = 0, []
    for s in range(steps):
        if s < 6:
            moves += bool(slab.rebalance(tr))
        else:   # then push the lower face up and down regardless of balance
            d = 1 if s % 2 == 0 else -1
            ranks[0].set_cut(ranks[0].z0, ranks[0].z1 + d)
            ranks[1].set_cut(ranks[1].z0 + d, ranks[1].z1)
        slab.slab_step(tr)
        sizes.append([r.n_owned() for r in ranks])
    sp, sv = slab.gather_owned(ranks, len(pos))
    for r in ranks:
        r.close()
    assert moves >= 2, "rebalance never moved a face: %s" % sizes
    assert common.first_mismatch(sp, ref_pos) is None, sizes
    assert common.first_mismatch(sv, ref_vel) is None
    with pytest.raises(RuntimeError):   # two layers at once are refused
        rr = slab.SlabRank(P, p, pos, vel, np.arange(len(pos), dtype=np.uint32), 0, 8, 0, 2, **caps)
        try:
            rr.set_cut(0, 10)
        finally:
            rr.close()
