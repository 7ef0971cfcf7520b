"""CPU tests of the multi-rank host logic (slab.py): the slab protocol driven by slab.slab_step over
 (a) the loopback transport, 3 ranks in one process, and
 (b) torch.distributed with the gloo backend, world_size 2, one rank per process,
with every rank computed by the CPU oracle (tests/slab_oracle.py). Per global particle id the result must be
bit-identical to the oracle's single-domain run, with particles migrating between slabs."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

import common
import slab_oracle

slab = importlib.import_module(common.PKG + ".slab")
STEPS = 6


def scene():
    pos, vel = common.synthetic_block(6, 5, 14, (12, 12, 40), seed=21)
    liquid = pos[:, 3] < 2
    rng = np.random.RandomState(4)
    vel[liquid, 2] = (0.9 + 0.2 * rng.rand(int(liquid.sum()))).astype(np.float32)   # ~2.5 grid units per step
    return pos, vel


def params_for(lib, pos):
    p = common.fill_counts(lib.default_params(box=(6, 6, 20)), pos)
    p.hash_mask = 0xFFFFFFFF
    return p


def single_domain(lib, p, pos, vel, steps):
    ctx = lib.create(p, pos, vel)
    for _ in range(steps):
        ctx.step_staged()
    out = ctx.read("position", len(pos)), ctx.read("velocity", len(pos))
    ctx.close()
    return out


def host_alloc():
    return slab.DistTransport.allocator("cpu")


def test_cut_layers_balances_and_keeps_two_layers():
    cz = np.repeat(np.arange(10), [5, 5, 50, 50, 5, 5, 5, 5, 5, 5])
    cuts = slab.cut_layers(cz, 3, 12)
    assert cuts[0] == 0 and cuts[-1] == 12 and all(b - a >= 2 for a, b in zip(cuts, cuts[1:]))
    with pytest.raises(ValueError):
        slab.cut_layers(cz, 7, 12)
    z = np.array([0.0, 6.67, 6.69, 13.37], np.float32)
    assert slab.cell_layer(z, 1.0 / 6.68).tolist() == [0, 0, 1, 2]


def test_slab_protocol_loopback_three_ranks_matches_single_domain():
    O = common.oracle_lib()
    pos, vel = scene()
    p = params_for(O, pos)
    ref_pos, ref_vel = single_domain(O, p, pos, vel, STEPS)
    slabs, _ = slab.split_fluid(p, pos, 3)
    ranks = [slab_oracle.OracleSlabRank(O, p, pos[own], vel[own], own, z0, z1, r, 3, host_alloc())
             for r, (z0, z1, own) in enumerate(slabs)]
    tr = slab.LoopbackTransport(ranks)
    moved = 0
    for _ in range(STEPS):
        slab.slab_step(tr)
        moved += sum(r.counts[0] + r.counts[2] for r in ranks)
    sp, sv = slab.gather_owned(ranks, len(pos))
    assert moved > 0, "no particle changed slab"
    assert common.first_mismatch(sp, ref_pos) is None
    assert common.first_mismatch(sv, ref_vel) is None


def slabs_for_cuts(p, pos, cuts):
    cz = slab.cell_layer(pos[:, 2], p.hash_grid_cell_size_inv)
    return [(cuts[r], cuts[r + 1], np.nonzero((cz >= cuts[r]) & (cz < cuts[r + 1]))[0]) for r in range(len(cuts) - 1)]


def test_plan_cuts_moves_one_layer_towards_the_fuller_slab():
    assert slab.plan_cuts([0, 10, 20], [1000, 1000]) == [0, 10, 20]
    assert slab.plan_cuts([0, 10, 20], [3000, 1000]) == [0, 9, 20]
    assert slab.plan_cuts([0, 10, 20], [1000, 3000]) == [0, 11, 20]
    assert slab.plan_cuts([0, 2, 20], [9000, 1000]) == [0, 2, 20]          # never below 2 layers
    assert slab.plan_cuts([0, 4, 8, 12], [4000, 1000, 4000]) == [0, 3, 9, 12]


def test_slab_recut_loopback_hands_layers_over_and_matches_single_domain():
    """Uneven initial cut + rebalance() before every step: whole layers change owner while the fluid moves; the
    result per global id stays the single-domain one."""
    O = common.oracle_lib()
    pos, vel = scene()
    p = params_for(O, pos)
    ref_pos, ref_vel = single_domain(O, p, pos, vel, STEPS)
    gz = int(p.grid_cells_z)
    cuts = [0, 2, 8, gz]                 # the fluid fills layers 0..9: most of it starts on the middle rank
    ranks = [slab_oracle.OracleSlabRank(O, p, pos[own], vel[own], own, z0, z1, r, 3, host_alloc())
             for r, (z0, z1, own) in enumerate(slabs_for_cuts(p, pos, cuts))]
    tr = slab.LoopbackTransport(ranks)
    moved_faces = 0
    history = []
    for _ in range(STEPS):
        moved_faces += bool(slab.rebalance(tr))
        slab.slab_step(tr)
        history.append([r.n_owned() for r in ranks])
    sp, sv = slab.gather_owned(ranks, len(pos))
    assert moved_faces >= 2, "the faces never moved: %s" % history
    assert [r.z0 for r in ranks][1] > 2, "the lower face did not advance into the crowded slab"
    assert common.first_mismatch(sp, ref_pos) is None
    assert common.first_mismatch(sv, ref_vel) is None


def _gloo_worker(rank, world, port, outdir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O = common.oracle_lib()
        pos, vel = scene()
        p = params_for(O, pos)
        # uneven cut (most of the fluid on rank 1) + rebalance() every step: the face walks towards the balance
        # point, whole layers travel as migrants; the counts for the plan go through all_gather on gloo
        z0, z1, own = slabs_for_cuts(p, pos, [0, 2, int(p.grid_cells_z)])[rank]
        r = slab_oracle.OracleSlabRank(O, p, pos[own], vel[own], own, z0, z1, rank, world, host_alloc())
        tr = slab.DistTransport(r)
        moved = 0
        for _ in range(STEPS):
            slab.rebalance(tr)
            slab.slab_step(tr)
            moved += r.counts[0] + r.counts[2]
        assert (r.z1 if rank == 0 else r.z0) > 2, "the face never moved"
        op, ov, og = r.read_owned()
        np.savez(os.path.join(outdir, "rank%d.npz" % rank), pos=op, vel=ov, gid=og, moved=moved)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_slab_protocol_gloo_world_size_2(tmp_path):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    O = common.oracle_lib()
    pos, vel = scene()
    p = params_for(O, pos)
    ref_pos, ref_vel = single_domain(O, p, pos, vel, STEPS)
    got_pos, got_vel = np.zeros_like(pos), np.zeros_like(vel)
    seen, moved = np.zeros(len(pos), int), 0
    for r in range(2):
        z = np.load(os.path.join(str(tmp_path), "rank%d.npz" % r))
        got_pos[z["gid"]], got_vel[z["gid"]] = z["pos"], z["vel"]
        seen[z["gid"]] += 1
        moved += int(z["moved"])
    assert (seen == 1).all() and moved > 0
    assert common.first_mismatch(got_pos, ref_pos) is None
    assert common.first_mismatch(got_vel, ref_vel) is None


if __name__ == "__main__":
    sys.exit(pytest.main([__file__, "-x", "-q"]))
