"""GPU parity tests (-m gpu): libsph_b200.so (CUDA sm_100a) through the C-ABI against the oracle.

Bar: BIT-EXACT on every buffer — integer/index work (hashes, sort order, cell ranges, neighbour
lists) and floating point alike (tolerance 0 ulp; BASELINE.json would allow 1e-5 relative). This
is possible because the kernels keep the reference's operation order and are built without FMA
contraction. Compared against (a) the C oracle run live on the same inputs and (b) the committed
digests produced by the reference's own kernels (tests/golden/digests.json).
"""
import os

import numpy as np
import pytest

import common
from common import capi

pytestmark = pytest.mark.gpu

STAGE_OUTPUTS = {
    "CLEAR_BUFFERS": ["neighborMap"],
    "HASH_PARTICLES": ["particleIndex"],
    "SORT": ["particleIndex"],
    "SORT_POST_PASS": ["sortedVelocity", "particleIndexBack", ("sortedPosition", "first")],
    "INDEXX": ["gridCellIndex"],
    "INDEX_POST_PASS": ["gridCellIndexFixedUp"],
    "FIND_NEIGHBORS": ["neighborMap"],
    "COMPUTE_DENSITY": [("rho", "first")],
    "COMPUTE_FORCES_INIT_PRESSURE": ["acceleration", "pressure"],
    "COMPUTE_ELASTIC_FORCES": ["acceleration"],
    "PREDICT_POSITIONS": ["sortedPosition"],
    "PREDICT_DENSITY": ["rho"],
    "CORRECT_PRESSURE": ["pressure"],
    "COMPUTE_PRESSURE_FORCE_ACCELERATION": ["acceleration"],
    "INTEGRATE": ["position", "velocity"],
    "CLEAR_MEMBRANE_BUFFERS": ["position", "velocity"],
    "MEMBRANE_INTERACTION": ["position"],
    "MEMBRANE_FINALIZE": ["position"],
}


# every parity test runs once per row: neighbour-search path x lanes per particle of the gather kernels x CUDA graph
VARIANTS = {
    "tiles": dict(SPH_NEIGHBOR_PATH="tiles", SPH_LANES="1", SPH_GRAPH="0", SPH_PDL="0"),
    "generic-lanes4-pdl": dict(SPH_NEIGHBOR_PATH="generic", SPH_LANES="4", SPH_GRAPH="0", SPH_PDL="1"),
    "list-histonly-lanes16-graph": dict(SPH_NEIGHBOR_PATH="list", SPH_LANES="16", SPH_GRAPH="1", SPH_PDL="0",
                                        SPH_HIST2_PCT="0"),   # histogram-only first pass from the second step on
    "coop-lanes8-graph-pdl": dict(SPH_NEIGHBOR_PATH="coop", SPH_LANES="8", SPH_GRAPH="1", SPH_PDL="1"),
    "default": dict(),   # what the library picks by size: the product path
}


@pytest.fixture(scope="module", autouse=True, params=list(VARIANTS))
def neighbor_path(request):
    """Force the independent implementations of the step's kernels through their environment switches (read at
    sph_create): the tile / plain / list / cooperative neighbour search, 1 / 4 / 8 / 16 lanes per particle in the
    gather kernels, eager launches or the captured CUDA graph, plain or programmatic dependent launches."""
    keys = ["SPH_NEIGHBOR_PATH", "SPH_LANES", "SPH_GRAPH", "SPH_PDL", "SPH_HIST2_PCT"]
    old = {k: os.environ.get(k) for k in keys}
    for k in keys:
        os.environ.pop(k, None)
    os.environ.update(VARIANTS[request.param])
    yield request.param
    for k in keys:
        os.environ.pop(k, None)
        if old[k] is not None:
            os.environ[k] = old[k]


@pytest.fixture(scope="module")
def P():
    lib = common.product_lib()
    assert lib.backend == "cuda-sm_100a"
    return lib


@pytest.fixture(scope="module")
def O():
    return common.oracle_lib()


@pytest.fixture(scope="module")
def D():
    return common.digests()


def _cmp(a, b, name, tag, half=None):
    x, y = common.canon(name, a.read(name)), common.canon(name, b.read(name))
    if half == "first":
        x, y = x[: len(x) // 2], y[: len(y) // 2]
    mm = common.first_mismatch(x, y)
    assert mm is None, "%s: %s differs (cuda vs oracle): %s" % (tag, name, mm)


def test_constants_bit_equal(P, D):
    assert P.default_params().as_bytes().hex() == D["params_default_hex"]
    assert P.default_params(box=(36, 24, 54)).as_bytes().hex() == D["params_boxA_hex"]


def _staged_vs_oracle(P, O, cfg, steps, muscles=False):
    a, b = common.create(P, cfg), common.create(O, cfg)
    for s in range(steps):
        if muscles:
            a.set_muscles(common.muscle_signal(s))
            b.set_muscles(common.muscle_signal(s))
        for st in common.stage_order(cfg["params"].max_iteration):
            a.stage(st)
            b.stage(st)
            for out in STAGE_OUTPUTS[st]:
                name, half = (out, None) if isinstance(out, str) else out
                _cmp(a, b, name, "step %d stage %s" % (s, st), half)
        common.assert_same_state(a, b, tag="end of staged step %d" % s)
    a.close(); b.close()


def _fused_vs_oracle(P, O, cfg, steps, muscles=False, check_every=1, no_nan=False):
    a, b = common.create(P, cfg), common.create(O, cfg)
    for s in range(steps):
        if muscles:
            a.set_muscles(common.muscle_signal(s))
            b.set_muscles(common.muscle_signal(s))
        a.step()
        b.step()
        if (s + 1) % check_every == 0 or s == steps - 1:
            common.assert_same_state(a, b, tag="after fused step %d" % (s + 1), no_nan=no_nan)
    a.close(); b.close()


def test_config_a_staged_every_stage(P, O):
    _staged_vs_oracle(P, O, common.config_a(P), 2)


def test_config_a_fused_steps(P, O):
    _fused_vs_oracle(P, O, common.config_a(P), 10, no_nan=True)


def test_config_a_fused_matches_reference_digests(P, D):
    cfg = common.config_a(P)
    ctx = common.create(P, cfg)
    done = 0
    for s in (1, 2, 5, 10):
        while done < s:
            ctx.step()
            done += 1
        for b, hx in D["configA"]["later"]["after_%d_steps" % s].items():
            assert common.sha(ctx.read(b)) == hx, "after %d steps: %s differs from the reference kernels" % (s, b)


def test_worm_staged_every_stage(P, O):
    _staged_vs_oracle(P, O, common.config_worm(P), 1, muscles=True)


def test_worm_fused_steps(P, O):
    _fused_vs_oracle(P, O, common.config_worm(P), 4, muscles=True, no_nan=True)


def test_worm_fused_matches_reference_digests(P, D):
    cfg = common.config_worm(P)
    ctx = common.create(P, cfg)
    done = 0
    for s in (1, 2, 3):
        while done < s:
            ctx.set_muscles(common.muscle_signal(done))
            ctx.step()
            done += 1
        for b, hx in D["worm"]["later"]["after_%d_steps" % s].items():
            assert common.sha(ctx.read(b)) == hx, "after %d steps: %s differs from the reference kernels" % (s, b)


@pytest.mark.parametrize("mask", [0xffff, 0xffffffff])
def test_synthetic_block_fused(P, O, mask):
    pos, vel = common.synthetic_block(24, 20, 30, (32, 28, 80), seed=3)
    cfg = dict(params=common.fill_counts(P.default_params(box=(16, 14, 40)), pos), position=pos, velocity=vel)
    cfg["params"].hash_mask = mask
    _fused_vs_oracle(P, O, cfg, 5)


def test_ragged_and_tiny(P, O):
    """N not a multiple of 32/256, a single liquid particle, crowded cells (> 32 candidates within h)."""
    pos, vel = common.synthetic_block(3, 3, 3, (8, 8, 8), seed=5)
    pos, vel = pos[:-3], vel[:-3]  # ragged
    cfg = dict(params=common.fill_counts(P.default_params(box=(4, 4, 4)), pos), position=pos, velocity=vel)
    _fused_vs_oracle(P, O, cfg, 3)
    # crowded: lattice at r0/2 -> ~8x rest density, exercises the >32 histogram branch and r_thr = 0 cases
    pos2, vel2 = common.synthetic_block(10, 10, 10, (10, 10, 10), seed=9, r0=np.float32(0.5 * 3.34))
    pos2[:1000, :3] = (pos2[:1000, :3] - 3 * 1.67) * 0.5 + 3 * 1.67
    pos2[10, :3] = pos2[11, :3]  # two coincident particles (r_ij == 0 paths)
    cfg2 = dict(params=common.fill_counts(P.default_params(box=(5, 5, 5)), pos2), position=pos2, velocity=vel2)
    _fused_vs_oracle(P, O, cfg2, 2)


@pytest.mark.parametrize("per_h3", [8, 24, 48])
def test_random_gas_neighbor_search_and_step(P, O, per_h3):
    """Uniformly random positions (no lattice): distances fall anywhere relative to r_thr and to the cell faces,
    which is what the list kernel's half-precision pre-filter has to be conservative about. 8 particles per h^3 is
    the rest density (~38 hits per particle), 24 fills the shared-memory hit list and spills into the per-thread
    overflow array, 48 exceeds that too (the thread rescans with the plain kernel)."""
    h, r0 = 3.34, 1.67
    box = (32, 28, 40)                                  # in r0: 16 h x 14 h x 20 h
    shell, shell_v = common.synthetic_block(1, 1, 1, box, seed=1)
    shell, shell_v = shell[1:], shell_v[1:]             # the boundary particles only
    rng = np.random.RandomState(100 + per_h3)
    n = int(per_h3 * (box[0] - 2) * (box[1] - 2) * (box[2] - 2) / 8)
    pos = np.zeros((n, 4), np.float32)
    pos[:, :3] = (r0 + rng.uniform(0, 1, size=(n, 3)) * (np.asarray(box, np.float64) - 2) * r0).astype(np.float32)
    pos[:, 3] = np.float32(1.1)
    pos = np.concatenate([pos, shell]); vel = np.concatenate([np.zeros((n, 4), np.float32), shell_v])
    cfg = dict(params=common.fill_counts(P.default_params(box=(box[0] // 2, box[1] // 2, box[2] // 2)), pos),
               position=pos, velocity=vel)
    a, b = common.create(P, cfg), common.create(O, cfg)
    for st in common.stage_order(cfg["params"].max_iteration)[:8]:   # ... FIND_NEIGHBORS, COMPUTE_DENSITY
        a.stage(st); b.stage(st)
    _cmp(a, b, "neighborMap", "random gas %d/h^3" % per_h3)
    _cmp(a, b, "rho", "random gas %d/h^3" % per_h3, "first")
    a.close(); b.close()
    _fused_vs_oracle(P, O, cfg, 2)


def test_read_write_roundtrip_and_errors(P):
    cfg = common.config_a(P)
    ctx = common.create(P, cfg)
    assert np.array_equal(ctx.read("position", ctx.N), cfg["position"])
    newv = np.zeros((ctx.N, 4), np.float32)
    ctx.write("velocity", newv)
    assert np.array_equal(ctx.read("velocity", ctx.N), newv)
    with pytest.raises(capi.SphError):
        ctx.stage(99)
    with pytest.raises(capi.SphError):
        ctx.write("rho", np.zeros(4, np.float32))
    with pytest.raises(capi.SphError):
        P.create(P.default_params(), np.zeros((0, 4), np.float32), np.zeros((0, 4), np.float32))


def test_pipelined_readback_matches_blocking_reads(P):
    """sph_read_begin / sph_read_end: the copy of step k's positions overlaps step k+1 (whose integration kernel
    waits for it) and must deliver exactly what a blocking read after step k sees."""
    pos, vel = common.synthetic_block(24, 20, 30, (32, 28, 80), seed=11)
    cfg = dict(params=common.fill_counts(P.default_params(box=(16, 14, 40)), pos), position=pos, velocity=vel)
    steps = 5
    a = common.create(P, cfg)
    want = []
    for _ in range(steps):
        a.step()
        want.append(a.read("position", a.N))
    a.close()
    b = common.create(P, cfg)
    bufs = [P.pinned_array((b.N, 4)), P.pinned_array((b.N, 4))]
    got = []
    for s in range(steps):
        b.step(sync=False)
        b.read_end()                       # step s-1 has arrived
        if s:
            got.append(bufs[(s - 1) & 1].copy())
        b.read_begin("position", bufs[s & 1])
        with pytest.raises(capi.SphError):
            b.read_begin("position", bufs[s & 1])   # one read in flight
    b.read_end()
    got.append(bufs[(steps - 1) & 1].copy())
    with pytest.raises(capi.SphError):
        b.read_begin("rho", bufs[0])
    b.close()
    for s in range(steps):
        assert common.first_mismatch(got[s], want[s]) is None, "step %d" % s
