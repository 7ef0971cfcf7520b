"""Host-side logic of the drop-in boundary: file formats (owHelper), the Python-3 muscle hook, and —
on a GPU — owPhysicsFluidSimulator::simulationStep through libow_b200.so against the oracle."""
import ctypes as C
import importlib
import math
import os
import sys

import numpy as np
import pytest

import common

HOSTDIR = os.path.join(common.ROOT, common.PKG, "host")


def host_lib():
    b = importlib.import_module(common.PKG + ".build")
    L = C.CDLL(b.build_host())
    L.ow_sim_create.restype = C.c_void_p
    L.ow_sim_create.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    L.ow_sim_create_from_files.restype = C.c_void_p
    L.ow_sim_destroy.argtypes = [C.c_void_p]
    L.ow_sim_step.argtypes = [C.c_void_p, C.c_int]
    L.ow_sim_step.restype = C.c_double
    L.ow_sim_set_fused.argtypes = [C.c_void_p, C.c_int]
    L.ow_sim_set_muscle_hook.argtypes = [C.c_void_p, C.c_int]
    L.ow_sim_position.argtypes = [C.c_void_p]
    L.ow_sim_position.restype = C.POINTER(C.c_float)
    L.ow_sim_density.argtypes = [C.c_void_p]
    L.ow_sim_density.restype = C.POINTER(C.c_float)
    L.ow_sim_particle_index.argtypes = [C.c_void_p]
    L.ow_sim_particle_index.restype = C.POINTER(C.c_uint)
    L.ow_sim_read_velocity.argtypes = [C.c_void_p, C.c_void_p]
    L.ow_muscle_signal.restype = C.POINTER(C.c_float)
    L.ow_set_box.argtypes = [C.c_double] * 3
    L.ow_set_configuration_files.argtypes = [C.c_char_p] * 3
    L.ow_set_buffers_dir.argtypes = [C.c_char_p]
    L.ow_helper_load.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.POINTER(C.c_float))] + [C.POINTER(C.c_int)] * 4
    L.ow_helper_free_floats.argtypes = [C.POINTER(C.c_float)]
    L.ow_helper_write_buffers.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    L.ow_helper_read_frame.argtypes = [C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_float)),
                                       C.POINTER(C.POINTER(C.c_int)), C.c_int]
    L.ow_set_quiet(1)
    return L


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def write_txt(path, arr):
    with open(path, "w") as f:
        for r in arr:
            f.write("%e\t%e\t%e\t%e\n" % tuple(float(v) for v in r))


# ---------------------------------------------------------------- muscle hook
def test_main_sim_known_answers():
    """The Python-3 restatement against known answers produced by the reference's OWN main_sim.py (run under Python 3
    after two mechanical py2 -> py3 rewrites, tests/golden/make_main_sim_kat.py): 96 doubles of calls 0, 1, 2, 500
    and 1000, compared exactly."""
    import json
    sys.path.insert(0, HOSTDIR)
    import main_sim
    kat = json.load(open(os.path.join(common.GOLDEN, "main_sim_kat.json")))["calls"]
    m = main_sim.muscle_simulation()
    for t in range(1001):
        got = m.run()
        assert len(got) == 96
        if str(t) in kat:
            want = [float.fromhex(x) for x in kat[str(t)]]
            assert got == want, "call %d differs from the reference script" % t
    # and two literal values anyone can check by hand: sin(0) = 0 -> 0.5; the antiphase wave mirrors it
    first = main_sim.muscle_simulation().run()
    assert first[0] == 0.5 and first[1] == 0.5 and abs(first[2] + first[26] - 1.0) < 1e-12


@pytest.mark.skipif(not os.path.exists("/root/reference/src/main_sim.py"), reason="reference tree not present")
def test_main_sim_kat_is_current():
    """Regenerate the known answers from /root/reference and compare with the committed fixture."""
    import json
    sys.path.insert(0, common.GOLDEN)
    import make_main_sim_kat as g
    sim = g.reference_module()["muscle_simulation"]()
    kat = json.load(open(os.path.join(common.GOLDEN, "main_sim_kat.json")))["calls"]
    for t in range(3):
        assert [float(x).hex() for x in sim.run()] == kat[str(t)]


# ---------------------------------------------------------------- file formats (no GPU)
def test_configuration_loader_and_buffers_roundtrip(tmp_path):
    L = host_lib()
    z = np.load(os.path.join(common.GOLDEN, "configA.npz"))
    pos, vel = z["position"][:3000].copy(), z["velocity"][:3000].copy()
    pos[10:20, 3] = 2.1  # a few elastic rows
    cfg = tmp_path / "configuration"
    cfg.mkdir()
    write_txt(cfg / "position.txt", pos)
    write_txt(cfg / "velocity.txt", vel)
    con = np.full((10 * 32, 4), 0, np.float32)
    con[:, 0] = -1
    con[0] = (11.1, 3.3e-6, 0, 0)
    con[32] = (10.1, 3.3e-6, 1.1, 0)
    write_txt(cfg / "elasticconnections.txt", con)
    with open(cfg / "position.txt", "a") as f:
        f.write("\n")  # trailing newline-only row must not count (owHelper.cpp:1445)
    L.ow_set_configuration_files(str(cfg).encode(), b"position.txt", b"velocity.txt")
    n = L.ow_helper_preload()
    assert n == 3000
    p2, v2 = np.zeros((n, 4), np.float32), np.zeros((n, 4), np.float32)
    cptr = C.POINTER(C.c_float)()
    nl, ne, nb, nc = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    L.ow_helper_load(vp(p2), vp(v2), C.byref(cptr), C.byref(nl), C.byref(ne), C.byref(nb), C.byref(nc))
    assert np.array_equal(p2, pos) and np.array_equal(v2, vel)  # %e text -> float32 round trip is exact
    t = pos[:, 3].astype(int)
    assert (nl.value, ne.value, nb.value) == ((t == 1).sum(), (t == 2).sum(), (t == 3).sum())
    got = np.ctypeslib.as_array(cptr, shape=(10 * 32, 4)).copy()
    assert np.array_equal(got, con) and nc.value == 2
    L.ow_helper_free_floats(cptr)
    # record (-l_to) then replay (-l_from): header = Ne, Nl; only non-boundary rows; frames appended
    buf = tmp_path / "buffers"
    buf.mkdir()
    L.ow_set_buffers_dir(str(buf).encode())
    L.ow_helper_set_counts(n, ne.value, nl.value, nb.value, 2)
    mem = np.array([[10, 11, 12], [11, 12, 13]], np.int32)
    L.ow_helper_write_buffers(vp(p2), vp(got), vp(mem), 1)
    p3 = p2.copy()
    p3[:, 0] += 1
    L.ow_helper_write_buffers(vp(p3), None, None, 0)
    lines = open(buf / "position_buffer.txt").read().split("\n")
    assert int(lines[0]) == ne.value and int(lines[1]) == nl.value
    keep = t != 3
    assert len([x for x in lines[2:] if x]) == 2 * keep.sum()
    pp, cc, mm = C.POINTER(C.c_float)(), C.POINTER(C.c_float)(), C.POINTER(C.c_int)()
    frames = L.ow_helper_read_frame(C.byref(pp), C.byref(cc), C.byref(mm), 0)
    assert frames == 2
    f0 = np.ctypeslib.as_array(pp, shape=(int(keep.sum()), 4)).copy()
    L.ow_helper_read_frame(C.byref(pp), C.byref(cc), C.byref(mm), 1)
    f1 = np.ctypeslib.as_array(pp, shape=(int(keep.sum()), 4)).copy()
    # ostream << float prints 6 significant digits (as the reference): relative error up to 5e-6
    assert np.allclose(f0, p2[keep], rtol=6e-6) and np.allclose(f1, p3[keep], rtol=6e-6)
    assert np.array_equal(np.ctypeslib.as_array(mm, shape=(2, 3)), mem)


# ---------------------------------------------------------------- simulator on the GPU
@pytest.mark.gpu
def test_simulator_fused_and_staged_match_oracle():
    L = host_lib()
    O = common.oracle_lib()
    cfg = common.config_a(O)
    pos, vel = cfg["position"], cfg["velocity"]
    ref = common.create(O, cfg)
    for fused in (1, 0):
        L.ow_set_box(36.0, 24.0, 54.0)
        sim = L.ow_sim_create(len(pos), vp(pos), vp(vel), 0, None, 0, None, None)
        L.ow_sim_set_fused(sim, fused)
        L.ow_sim_set_muscle_hook(sim, 0)
        o = common.create(O, cfg)
        for s in range(3):
            ms = L.ow_sim_step(sim, 0)
            o.step()
            assert ms > 0
            got = np.ctypeslib.as_array(L.ow_sim_position(sim), shape=(len(pos), 4))
            assert common.first_mismatch(got, o.read("position", len(pos))) is None
        rho = np.ctypeslib.as_array(L.ow_sim_density(sim), shape=(len(pos),))
        assert common.first_mismatch(rho, o.read("rho", len(pos))) is None
        pi = np.ctypeslib.as_array(L.ow_sim_particle_index(sim), shape=(len(pos), 2))
        assert np.array_equal(pi, o.read("particleIndex"))
        v = np.zeros((len(pos), 4), np.float32)
        L.ow_sim_read_velocity(sim, vp(v))
        assert common.first_mismatch(v, o.read("velocity", len(pos))) is None
        L.ow_sim_destroy(sim)
    ref.close()


@pytest.mark.gpu
def test_simulator_worm_with_python_muscle_hook(tmp_path):
    """Full worm config through owPhysicsFluidSimulator with the embedded-Python main_sim hook:
    after each step the 96 signals are fetched and uploaded, as owPhysicsFluidSimulator.cpp:134-141."""
    sys.path.insert(0, HOSTDIR)
    import main_sim
    L = host_lib()
    O = common.oracle_lib()
    cfg = common.config_worm(O)
    pos = cfg["position"]
    L.ow_set_box(30.0, 20.0, 250.0)
    sim = L.ow_sim_create(len(pos), vp(pos), vp(cfg["velocity"]), int(cfg["params"].num_elastic), vp(cfg["elastic"]),
                          len(cfg["membranes"]), vp(cfg["membranes"]), vp(cfg["pml"]))
    o = common.create(O, cfg)
    gen = main_sim.muscle_simulation()
    for s in range(3):
        L.ow_sim_step(sim, 0)
        o.step()
        got = np.ctypeslib.as_array(L.ow_sim_position(sim), shape=(len(pos), 4))
        assert common.first_mismatch(got, o.read("position", len(pos))) is None, "step %d" % s
        sig = np.zeros(100, np.float32)
        sig[:96] = np.asarray(gen.run(), np.float32)
        assert np.array_equal(np.ctypeslib.as_array(L.ow_muscle_signal(), shape=(100,)), sig)
        o.set_muscles(sig)
    L.ow_sim_destroy(sim)


@pytest.mark.gpu
def test_simulator_from_configuration_files_and_l_to(tmp_path):
    L = host_lib()
    O = common.oracle_lib()
    cfg = common.config_a(O)
    d = tmp_path / "configuration"
    d.mkdir()
    write_txt(d / "positionPureLiquid.txt", cfg["position"])
    write_txt(d / "velocityPureLiquid.txt", cfg["velocity"])
    (tmp_path / "buffers").mkdir()
    L.ow_set_configuration_files(str(d).encode(), b"positionPureLiquid.txt", b"velocityPureLiquid.txt")
    L.ow_set_buffers_dir(str(tmp_path / "buffers").encode())
    L.ow_set_box(36.0, 24.0, 54.0)
    sim = L.ow_sim_create_from_files()
    L.ow_sim_set_muscle_hook(sim, 0)
    o = common.create(O, cfg)
    for s in range(11):
        L.ow_sim_step(sim, 1)
        o.step()
    got = np.ctypeslib.as_array(L.ow_sim_position(sim), shape=(len(cfg["position"]), 4))
    assert common.first_mismatch(got, o.read("position", len(got))) is None
    L.ow_flush_recording()   # frames are written by a background thread
    lines = [x for x in open(tmp_path / "buffers" / "position_buffer.txt").read().split("\n") if x]
    nl = int(cfg["params"].num_liquid)
    assert int(lines[0]) == 0 and int(lines[1]) == nl
    assert len(lines) - 2 == 2 * nl  # frames at step 0 and step 10 (iter_step = 10)
    L.ow_sim_destroy(sim)


# ---------------------------------------------------------------- the reference's shipped files, the CLI driver
REF_CFG = "/root/reference/configuration"


@pytest.mark.skipif(not os.path.exists(os.path.join(REF_CFG, "positionPureLiquid.txt")), reason="reference tree not present")
def test_loader_reads_the_shipped_configuration():
    """owHelper::preLoadConfiguration / loadConfiguration pointed at the reference's own
    configuration/positionPureLiquid.txt + velocityPureLiquid.txt (owHelper.cpp:1431-1545) reproduce the committed
    fixture configA.npz (made with numpy.loadtxt from the same files) bit for bit, counts included."""
    L = host_lib()
    z = np.load(os.path.join(common.GOLDEN, "configA.npz"))
    L.ow_set_configuration_files(REF_CFG.encode(), b"positionPureLiquid.txt", b"velocityPureLiquid.txt")
    n = L.ow_helper_preload()
    assert n == len(z["position"]) == 61440
    p, v = np.zeros((n, 4), np.float32), np.zeros((n, 4), np.float32)
    cptr = C.POINTER(C.c_float)()
    nl, ne, nb, nc = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    L.ow_helper_load(vp(p), vp(v), C.byref(cptr), C.byref(nl), C.byref(ne), C.byref(nb), C.byref(nc))
    assert np.array_equal(p.view(np.uint32), z["position"].view(np.uint32))
    assert np.array_equal(v.view(np.uint32), z["velocity"].view(np.uint32))
    t = z["position"][:, 3].astype(int)
    assert (nl.value, ne.value, nb.value) == ((t == 1).sum(), 0, (t == 3).sum())


def test_fixture_scenes_contain_no_nan():
    """common.first_mismatch treats any NaN as equal to any NaN (x86 and CUDA generate different default NaNs); the
    scenes the parity tests run on must therefore never produce one. Inputs here; the GPU tests assert the outputs."""
    for name in ("configA.npz", "worm.npz"):
        z = np.load(os.path.join(common.GOLDEN, name))
        for k in z.files:
            if z[k].dtype.kind == "f":
                assert np.isfinite(z[k]).all(), (name, k)


@pytest.mark.gpu
def test_cli_driver_records_the_oracle_trajectory(tmp_path):
    """The headless driver as a subprocess (reference: `Sibernetic -no_g -l_to`, main.cpp:8-34 +
    owWorldSimulation.cpp:999-1004): loads configuration/*.txt, steps, and records buffers/position_buffer.txt on
    step 0 and every 10th step; the recorded frames are the oracle's trajectory at the 6 significant digits
    `ostream << float` prints."""
    b = importlib.import_module(common.PKG + ".build")
    b.build_host()
    exe = b.HOST_BIN
    O = common.oracle_lib()
    cfg = common.config_a(O)
    d = tmp_path / "configuration"
    d.mkdir()
    write_txt(d / "positionPureLiquid.txt", cfg["position"])
    write_txt(d / "velocityPureLiquid.txt", cfg["velocity"])
    (tmp_path / "buffers").mkdir()
    import subprocess
    env = dict(os.environ, PYTHONPATH=HOSTDIR + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([exe, "-no_g", "-l_to", "--steps", "11", "--box", "36", "24", "54", "--config", str(d),
                        "--position", "positionPureLiquid.txt", "--velocity", "velocityPureLiquid.txt",
                        "--buffers", str(tmp_path / "buffers"), "--quiet"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "particle-steps/s" in r.stdout
    o = common.create(O, cfg)
    frames = []
    for s in range(11):
        o.step()
        if s % 10 == 0:
            frames.append(o.read("position", len(cfg["position"])))
    o.close()
    rows = np.loadtxt(tmp_path / "buffers" / "position_buffer.txt", skiprows=2, dtype=np.float64)
    keep = cfg["position"][:, 3].astype(int) != 3
    want = np.concatenate([f[keep] for f in frames]).astype(np.float64)
    assert rows.shape == want.shape
    assert not np.isnan(want).any()
    assert np.allclose(rows, want, rtol=6e-6, atol=0)
    # replay (-l_from) finds both frames
    r = subprocess.run([exe, "-l_from", "--buffers", str(tmp_path / "buffers")], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and "replayed 2 frame(s)" in r.stdout, r.stdout


@pytest.mark.gpu
def test_simulator_on_several_slabs_matches_single_device():
    """The drop-in class with owSetSlabDevices (the reference is single-device): one fluid cut into z-slabs inside one
    process, linked through peer memory, global position array assembled by read_position_buffer. Two slabs on GPU 0
    here (one GPU is enough: an ordinal may repeat); per particle the positions and velocities of the single-device
    oracle run, bit for bit, while particles migrate between the slabs."""
    L = host_lib()
    L.ow_set_slab_devices.argtypes = [C.POINTER(C.c_int), C.c_int]
    L.ow_set_hash_mask.argtypes = [C.c_uint]
    O = common.oracle_lib()
    pos, vel = common.synthetic_block(14, 12, 40, (24, 22, 96), seed=7)
    liquid = pos[:, 3] < 2
    rng = np.random.RandomState(7)
    vel[liquid, 2] = (0.45 + 0.1 * rng.rand(int(liquid.sum()))).astype(np.float32)   # thrown along z: layers change slab
    p = common.fill_counts(O.default_params(box=(12, 11, 48)), pos)
    p.hash_mask = 0xFFFFFFFF
    o = O.create(p, pos, vel)
    L.ow_set_box(12.0, 11.0, 48.0)
    L.ow_set_hash_mask(0xFFFFFFFF)
    devs = (C.c_int * 2)(0, 0)
    L.ow_set_slab_devices(devs, 2)
    try:
        sim = L.ow_sim_create(len(pos), vp(pos), vp(vel), 0, None, 0, None, None)
        L.ow_sim_set_muscle_hook(sim, 0)
        for s in range(8):
            L.ow_sim_step(sim, 0)
            o.step()
            got = np.ctypeslib.as_array(L.ow_sim_position(sim), shape=(len(pos), 4))
            assert common.first_mismatch(got, o.read("position", len(pos))) is None, "step %d" % s
        v = np.zeros((len(pos), 4), np.float32)
        L.ow_sim_read_velocity(sim, vp(v))
        assert common.first_mismatch(v, o.read("velocity", len(pos))) is None
        L.ow_sim_destroy(sim)
    finally:
        L.ow_set_slab_devices(devs, 0)
        L.ow_set_hash_mask(0xFFFF)
    o.close()


# ---------------------------------------------------------------- config B': the shipped position.txt with regenerated springs
def _bprime():
    sys.path.insert(0, os.path.join(common.ROOT, "tools"))
    import make_connections
    z = np.load(os.path.join(common.GOLDEN, "configBprime.npz"))   # /root/reference/configuration/position.txt + velocity.txt
    pos, vel = z["position"], z["velocity"]
    return pos, vel, make_connections.connections(pos), make_connections


def test_regenerated_connections_follow_the_generator_rule():
    """elasticconnections.txt of the shipped position.txt is a lost blob; tools/make_connections.py rebuilds it with
    the generator's spring rule (owHelper.cpp:972-1001). Properties of that rule: springs only from elastic particles,
    never to liquid ones, candidates in index order, r_ij <= r0*sqrt(2.7), rest length 0.95 r_ij (scaled), elastic-elastic
    springs symmetric."""
    pos, vel, con, mc = _bprime()
    t = pos[:, 3].astype(int)
    el = np.nonzero(t == 2)[0]
    assert con.shape == (32 * len(el), 4) and len(el) == 7327
    rows = con.reshape(len(el), 32, 4)
    jd = rows[:, :, 0]
    used = jd >= 0
    assert used.sum() > 100000 and used.sum(1).max() <= 32
    assert (np.diff(used.astype(int), axis=1) <= 0).all()            # filled from slot 0, no holes
    j = jd[used].astype(int)
    i = np.repeat(el, 32).reshape(len(el), 32)[used]
    assert (t[j] != 1).all() and (j != i).all()
    d = np.linalg.norm(pos[i, :3].astype(np.float64) - pos[j, :3].astype(np.float64), axis=1)
    assert (d <= 0.5 * 3.34 * np.sqrt(2.7) * (1 + 1e-6)).all()
    assert np.allclose(rows[:, :, 1][used], d * float(mc.simulation_scale()) * 0.95, rtol=1e-5)
    jsorted = np.where(used, jd, np.inf)
    assert (np.diff(jsorted, axis=1)[used[:, 1:]] > 0).all()         # index order
    pairs = set(zip(i.tolist(), j.tolist()))
    assert all((b, a) in pairs for a, b in pairs if t[b] == 2)


@pytest.mark.gpu
def test_simulator_config_bprime_from_files(tmp_path):
    """Config B' end to end through the drop-in class: configuration/position.txt, velocity.txt and the regenerated
    elasticconnections.txt are read by owHelper (boundary particles first: elastic offset = numOfBoundaryP,
    owOpenCLSolver.cpp:435), stepped on the GPU, compared with the oracle bit for bit."""
    L = host_lib()
    O = common.oracle_lib()
    pos, vel, con, mc = _bprime()
    d = tmp_path / "configuration"
    d.mkdir()
    write_txt(d / "position.txt", pos)
    write_txt(d / "velocity.txt", vel)
    mc.write_txt(d / "elasticconnections.txt", con)
    L.ow_set_configuration_files(str(d).encode(), b"position.txt", b"velocity.txt")
    L.ow_set_box(36.0, 24.0, 54.5)
    sim = L.ow_sim_create_from_files()
    L.ow_sim_set_muscle_hook(sim, 0)
    p = common.fill_counts(O.default_params(box=(36, 24, 54.5)), pos)
    p.elastic_offset = p.num_boundary
    # the springs as the loader sees them: %e text keeps 7 significant digits of the rest lengths
    con_txt = np.loadtxt(d / "elasticconnections.txt", dtype=np.float32).reshape(-1, 4)
    o = O.create(p, pos, vel, elastic=con_txt)
    for s in range(4):
        L.ow_sim_step(sim, 0)
        o.step()
        got = np.ctypeslib.as_array(L.ow_sim_position(sim), shape=(len(pos), 4))
        want = o.read("position", len(pos))
        assert not np.isnan(want).any()
        assert common.first_mismatch(got, want) is None, "step %d" % s
    v = np.zeros((len(pos), 4), np.float32)
    L.ow_sim_read_velocity(sim, vp(v))
    assert common.first_mismatch(v, o.read("velocity", len(pos))) is None
    L.ow_sim_destroy(sim)
    o.close()


def test_async_recording_writes_in_order_and_the_process_exits(tmp_path):
    """-l_to frames are formatted by a writer thread (owHelper::loadConfigurationToFileAsync). In a process of its own:
    three frames queued back to back arrive in order, flushRecording waits for them, and the process EXITS (a writer
    thread waiting on a condition variable that static destructors tear down would hang it)."""
    import subprocess
    import textwrap
    code = textwrap.dedent('''
        import ctypes as C, numpy as np, sys
        L = C.CDLL(sys.argv[1])
        L.ow_set_buffers_dir.argtypes = [C.c_char_p]
        L.ow_set_buffers_dir(sys.argv[2].encode())
        n = 5000
        L.ow_helper_set_counts(n, 0, n, 0, 0)
        L.ow_helper_write_buffers_async.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        p = np.zeros((n, 4), np.float32); p[:, 3] = 1.1
        for f in range(3):
            p[:, 0] = f + 1          # the frame is copied at the call: changing p afterwards must not matter
            L.ow_helper_write_buffers_async(p.ctypes.data_as(C.c_void_p), None, None, 1 if f == 0 else 0)
        p[:, 0] = 99
        L.ow_flush_recording()
        rows = np.loadtxt(sys.argv[2] + "/position_buffer.txt", skiprows=2)
        assert rows.shape == (3 * n, 4), rows.shape
        assert (rows[:n, 0] == 1).all() and (rows[n:2 * n, 0] == 2).all() and (rows[2 * n:, 0] == 3).all()
        print("frames ok", flush=True)
        ''')
    b = importlib.import_module(common.PKG + ".build")
    lib = b.build_host()
    (tmp_path / "buffers").mkdir()
    r = subprocess.run([sys.executable, "-c", code, lib, str(tmp_path / "buffers")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "frames ok" in r.stdout, r.stdout + r.stderr[-2000:]
