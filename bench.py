#!/usr/bin/env python3
"""bench.py — PCISPH particle-steps/s on N B200s (BASELINE.json metric), one JSON line on rank 0.

  python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference] [--no-configs]

A step = one pass of the whole hot path (owPhysicsFluidSimulator::simulationStep's device part:
hash/sort/neighbour search, PCISPH loop with 3 iterations, integrate, membranes) over the synthetic
fluid named in config.workload (8 M liquid particles per GPU on the
reference's r0 lattice in a boundary box; weak scaling over GPUs).
  value     whole-job particle-steps/s, state resident in HBM, CUDA events on the library's stream,
            max over ranks.
  e2e       the same metric through the drop-in C++ class: owPhysicsFluidSimulator::simulationStep of
            libow_b200.so with HOST buffers — per call the muscle signal goes up (400 B, from the embedded
            Python hook like the reference's), the step runs, the positions come back into position_cpp
            (16 N bytes, pinned) before the call returns (owPhysicsFluidSimulator.cpp:79-149).
  configs   (N = 1) the other BASELINE.json configurations, each with its own device / e2e / CPU-arm numbers
            on the SAME scene: A config_a, B config_b, C dambreak_1m, D fluid_16m (single GPU = the N=1 point
            of the strong-scaling series).
  strong_16m (N > 1) BASELINE config D: fluid_16m cut into N slabs.
  slab_parity (N > 1) a small fluid stepped through the same multi-GPU transport, compared bit for bit with
            a single-device run on rank 0; a mismatch makes the run fail.
  roofline / cpu_baseline: see DESIGN.md "Measurement".
--impl reference times the reference's own CPU kernels (oracle/_ref, sphFluid.cl compiled as C++;
else the C port) with all host threads on the same workload, same steps and warm-up.
"""
import argparse
import ctypes as C
import hashlib
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "smoothed-particle-hydrodynamics_b200"
capi = importlib.import_module(PKG + ".capi")
scenes = importlib.import_module(PKG + ".scenes")

METRIC = "pcisph_particle_steps_per_s"
UNIT = "particle-steps/s"
B_REF = 4164  # algorithmic bytes per particle-step, SURVEY.md §8d / BASELINE.md §4
# algorithmic bytes per particle of the reference stages each fused kernel replaces
KERNEL_BYTES = {
    "hash_count": 24,
    "sort(scan+scatter+rank_reorder)": 16 + 76 + 8,
    "find_neighbors+density": 256 + 272 + 264,
    "forces+predict0+density0": 356 + 92 + 280 + 16,
    "elastic": 0,
    "predict_density+correct_pressure": 280 + 16,
    "pressure_force+predict": 324 + 92,
    "pressure_force+integrate": 324 + 380,
    "membranes": 376,
}
# what ncu says limits each kernel (profiles/README.md); none of the gather kernels is limited by HBM
KERNEL_LIMITER = {
    "find_neighbors+density": "l1-data-pipe+issue",
    "pressure_force+predict": "l1-data-pipe+issue",
    "pressure_force+integrate": "l1-data-pipe+issue",
    "forces+predict0+density0": "l1-data-pipe",
    "predict_density+correct_pressure": "l1-data-pipe+hbm",
    "sort(scan+scatter+rank_reorder)": "hbm",
    "hash_count": "hbm",
}
# profile name -> kernel whose ncu capture (profiles/*_ncu_full_summary.json) gives the measured DRAM traffic per launch
NCU_KERNEL = {"find_neighbors+density": "k_find_neighbors_list<1,2>", "pressure_force+predict": "k_pressure_force<1,0>",
              "pressure_force+integrate": "k_pressure_force<2,0>", "predict_density+correct_pressure": "k_predict_density<1,0>",
              "forces+predict0+density0": "k_forces<1,0>"}


def measured_traffic(kernel, particles):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the newest committed
    `ncu --set full` capture of this same workload (None when there is no capture for this particle count).
    This is a constant read from profiles/, NOT a measurement of the run that prints it."""
    import glob
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_summary.json")), reverse=True):
        try:
            j = json.load(open(f))
        except (OSError, ValueError):
            continue
        k = j.get("kernels", {}).get(NCU_KERNEL.get(kernel, ""))
        if k and j.get("particles") == particles:
            return k["dram_bytes_per_launch"], os.path.basename(f)
    return None, None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, gpu):
        self.rows, self.proc, self.gpu = [], None, gpu

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].startswith("Active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


CONFIG_A = os.path.join(ROOT, "tests", "golden", "configA.npz")  # the reference's shipped pure-liquid scene
CONFIG_B = os.path.join(ROOT, "tests", "golden", "worm.npz")     # the reference generator's worm scene (fixture)
BOXES = {"config_a": (36.0, 24.0, 54.0), "config_b": (30.0, 20.0, 250.0)}
EXTRA = {}   # workload name -> dict(elastic=, membranes=, pml=) for scenes with elastic matter
# BASELINE.json configs[0..4] -> workload names
BASELINE_CONFIGS = [("A", "config_a"), ("B", "config_b"), ("C", "dambreak_1m"), ("D", "fluid_16m"), ("E", "fluid_8m")]


def make_workload(name, lib):
    """(params, position, velocity, box in units of h, hash mask) of a named scene."""
    if name == "config_b":  # BASELINE config 1 of the reference: elastic worm body, membranes, muscles, 16-bit mask
        z = np.load(CONFIG_B)
        pos, vel = z["position"], z["velocity"]
        p = lib.default_params()
        t = pos[:, 3].astype(np.int32)
        p.particle_count, p.num_liquid, p.num_boundary = len(pos), int((t == 1).sum()), int((t == 3).sum())
        p.num_elastic, p.num_membranes = int((t == 2).sum()), int(np.asarray(z["membraneData"]).reshape(-1, 3).shape[0])
        EXTRA[name] = dict(elastic=z["elasticConnectionsData"], membranes=z["membraneData"], pml=z["particleMembranesList"])
        return p, pos, vel, BOXES[name], 0xFFFF
    if name == "config_a":  # BASELINE config 0: positionPureLiquid.txt / velocityPureLiquid.txt, 36h x 24h x 54h, 16-bit mask
        z = np.load(CONFIG_A)
        pos, vel = z["position"], z["velocity"]
        p = lib.default_params(box=BOXES[name])
        t = pos[:, 3].astype(np.int32)
        p.particle_count, p.num_liquid, p.num_boundary = len(pos), int((t == 1).sum()), int((t == 3).sum())
        return p, pos, vel, BOXES[name], 0xFFFF
    nx, ny, nz, fill = scenes.WORKLOADS[name]
    pos, vel, box = scenes.fluid_box(nx, ny, nz, fill_z=fill)
    p = lib.default_params(box=box)
    t = pos[:, 3].astype(np.int32)
    p.particle_count = len(pos)
    p.num_liquid = int((t == 1).sum())
    p.num_boundary = int((t == 3).sum())
    p.hash_mask = 0xFFFFFFFF  # > 65 536 cells: the reference's 16-bit mask would alias
    return p, pos, vel, box, 0xFFFFFFFF


def cpu_oracle():
    """(lib, kind, set_threads) — the reference's own kernels when oracle/_ref was built, else the C port."""
    ref = os.path.join(ROOT, "oracle", "_ref", "libsphref.so")
    if os.path.exists(ref):
        L = capi.SphLib(ref)
        return L, "reference", L.lib.sph_ref_set_threads
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import build_oracle
    L = capi.SphLib(build_oracle.build())
    return L, "port", L.lib.sph_oracle_set_threads


def time_cpu(workload, steps, warmup, scene=None):
    """The CPU arm on a named scene: (cpu_baseline dict, ms per step, particles)."""
    L, kind, set_threads = cpu_oracle()
    cores = os.cpu_count() or 1
    set_threads(cores)
    p, pos, vel = scene if scene is not None else make_workload(workload, L)[:3]
    ctx = L.create(p, pos, vel, **EXTRA.get(workload, {}))
    for _ in range(warmup):
        ctx.step()
    t0 = time.perf_counter()
    for _ in range(steps):
        ctx.step()
    dt = time.perf_counter() - t0
    n = len(pos)
    ctx.close()
    return {"value": n * steps / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%s (%d particles, the whole scene), %d steps after %d warm-up, %.1f s" %
                      (workload, n, steps, warmup, dt)}, dt / steps * 1e3, n


def run_reference(args, rank):
    """--impl reference: the reference's CPU kernels on the GPU arm's workload (N = 1: the same scene, the same
    steps and warm-up; N > 1: one rank's share of the weak-scaling fluid — particle-steps/s does not depend on
    the particle count to within a few per cent, and 66 M particles would take 20 s per step)."""
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    cb, ms, n = time_cpu(args.workload, steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles_per_gpu": n, "particles_total": n,
                       "pcisph_iterations": 3, "hash_mask": "0xffffffff",
                       "note": None if args.gpus == 1 else "one rank's share of the %d-GPU fluid" % args.gpus},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------- the drop-in C++ class (libow_b200.so)
_FACADE = None
TRANSPORT_USED = None
E2E_BYTES = 16
LAST_CHAIN_FLOOR_MS = None   # dependent chain of as many EMPTY kernels as the last timed step had (small scenes)


def facade():
    """ctypes handle on libow_b200.so: owPhysicsFluidSimulator / owOpenCLSolver over the C-ABI (host/ow_capi.cpp)."""
    global _FACADE
    if _FACADE is None:
        path = os.path.join(ROOT, PKG, "libow_b200.so")
        if not os.path.exists(path):
            raise SystemExit("bench.py: %s is missing: run __graft_entry__.build()" % path)
        hostdir = os.path.join(ROOT, PKG, "host")
        if hostdir not in sys.path:
            sys.path.insert(0, hostdir)   # main_sim.py, the muscle-signal script the embedded hook imports
        F = C.CDLL(path)
        F.ow_sim_create.restype = C.c_void_p
        F.ow_sim_create.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        F.ow_sim_destroy.argtypes = [C.c_void_p]
        F.ow_sim_step.argtypes = [C.c_void_p, C.c_int]
        F.ow_sim_step.restype = C.c_double
        for f in (F.ow_sim_set_run_ahead, F.ow_sim_set_keep_velocity, F.ow_sim_set_muscle_hook, F.ow_sim_set_fused):
            f.argtypes = [C.c_void_p, C.c_int]
        F.ow_sim_position.argtypes = [C.c_void_p]
        F.ow_sim_position.restype = C.POINTER(C.c_float)
        F.ow_set_box.argtypes = [C.c_double] * 3
        F.ow_set_hash_mask.argtypes = [C.c_uint]
        F.ow_set_device.argtypes = [C.c_int]
        F.ow_set_quiet(1)
        _FACADE = F
    return _FACADE


def time_facade(name, scene, steps, device=0, run_ahead=True):
    """e2e: wall clock of `steps` calls of owPhysicsFluidSimulator::simulationStep (host buffers in and out every
    call) after 2 untimed calls. Returns (ms per step, checksum of the last positions)."""
    p, pos, vel, box, mask = scene
    F = facade()
    F.ow_set_box(*[float(b) for b in box])
    F.ow_set_hash_mask(int(mask))
    F.ow_set_device(int(device))
    ex = EXTRA.get(name, {})
    vp = lambda a, dt: None if a is None else np.ascontiguousarray(a, dt).ctypes.data_as(C.c_void_p)
    el, mem, pml = ex.get("elastic"), ex.get("membranes"), ex.get("pml")
    keep = [np.ascontiguousarray(pos, np.float32), np.ascontiguousarray(vel, np.float32)]
    sim = F.ow_sim_create(len(pos), vp(keep[0], np.float32), vp(keep[1], np.float32), int(p.num_elastic), vp(el, np.float32),
                          int(p.num_membranes), vp(mem, np.int32), vp(pml, np.int32))
    F.ow_sim_set_keep_velocity(sim, 0)
    F.ow_sim_set_run_ahead(sim, 1 if run_ahead else 0)
    for _ in range(2):
        F.ow_sim_step(sim, 0)
    t0 = time.perf_counter()
    for _ in range(steps):
        F.ow_sim_step(sim, 0)
    ms = (time.perf_counter() - t0) * 1e3 / steps
    got = np.ctypeslib.as_array(F.ow_sim_position(sim), shape=(len(pos), 4))
    ok = bool(np.isfinite(got[:, :3]).all())
    F.ow_sim_destroy(sim)
    return ms, ok


PRESSURE_FRAC_MAIN = None
LAST_PRESSURE_FRAC = None    # fraction of particles with non-zero pressure at the end of the last timed run


def time_device(L, name, scene, steps, warmup, device=0, profile=False):
    """Device-timed fused steps of one scene on one GPU (CUDA events on the library's stream, state resident in
    HBM). profile=True adds the per-kernel events (large scenes only: they disable the CUDA-graph replay)."""
    p, pos, vel = scene[:3]
    ctx = L.create(p, pos, vel, device=device, **EXTRA.get(name, {}))
    ctx.run_steps(warmup)
    if profile:
        ctx.profile(True)
    l0 = ctx.launch_count()
    ms = ctx.run_steps(steps)
    launches = ctx.launch_count() - l0
    prof = None
    if profile:
        prof = ctx.profile_report()
        ctx.profile(False)
    fin = bool(np.isfinite(ctx.read("position", len(pos))[:, :3]).all())
    global LAST_CHAIN_FLOOR_MS, LAST_PRESSURE_FRAC
    LAST_PRESSURE_FRAC = float((ctx.read("pressure") > 0).mean())
    LAST_CHAIN_FLOOR_MS = ctx.empty_chain_ms(max(1, launches // max(1, steps))) if len(pos) <= (1 << 21) else None
    ctx.close()
    return ms / steps, launches, prof, fin


# ---------------------------------------------------------------- multi-GPU: one fluid, z-slabs
def bind_to_gpu_numa_node(local):
    """Pin this rank's process to the CPUs next to its GPU (NVML's ideal affinity) before any pinned host memory is
    allocated: eight ranks reading 131 MB back per step are bound by host memory bandwidth, and pages on the GPU's own
    NUMA node keep the copies off the inter-socket link. Best effort; returns what was done."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",")[local]) if
                                              os.environ.get("CUDA_VISIBLE_DEVICES") else local)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return "cpus %d" % len(os.sched_getaffinity(0))
    except Exception as e:   # noqa: BLE001 — diagnostics only
        return "unbound (%s)" % type(e).__name__


def slab_rank(L, slab, torch, sc, rank, world, local):
    """This process's slab of a scene + the transport that exchanges halos with rank-1 / rank+1."""
    p = L.default_params(box=sc["box"])
    p.hash_mask = 0xFFFFFFFF
    p.particle_count = len(sc["gid"])
    p.num_liquid = sc["n_liquid"]
    p.num_boundary = len(sc["gid"]) - sc["n_liquid"]
    dev = torch.device("cuda", local)
    r = slab.SlabRank(L, p, sc["position"], sc["velocity"], sc["gid"], sc["z0"], sc["z1"], rank, world, device=local,
                      alloc=slab.DistTransport.allocator(dev), halo_cap=sc["halo_cap"], mig_cap=sc["mig_cap"])
    return r, make_transport(slab, torch, r, dev)


def make_transport(slab, torch, r, dev):
    """Default: peer-memory links (the neighbours' kernels write straight into this rank's mailbox over NVLink; the
    whole step runs inside the library). SPH_SLAB_TRANSPORT=nccl selects the round-1 path: pack kernels, NCCL
    point-to-point sends issued from Python, unpack kernels."""
    global TRANSPORT_USED
    tr = None
    if os.environ.get("SPH_SLAB_TRANSPORT", "p2p") != "nccl":
        tr = slab.PeerTransport.distributed(r)
        TRANSPORT_USED = "p2p (peer-memory links over NVLink, no NCCL on the data path)"
    if tr is None:   # asked for, or some rank could not map its neighbour's mailbox
        TRANSPORT_USED = "nccl point-to-point" + ("" if os.environ.get("SPH_SLAB_TRANSPORT") == "nccl" else " (FALLBACK: peer mapping failed)")
        r.set_stream(torch.cuda.current_stream().cuda_stream)  # kernels and NCCL transfers ordered on one stream
        tr = slab.DistTransport(r, dev)
    return tr


def time_slab(slab, torch, dist, L, r, tr, warmup, steps, e2e=True):
    """Device-timed slab steps (CUDA events, barrier on both sides) and, optionally, the e2e variant: muscle signal
    up, step, the owned positions back into pinned host memory every step (pipelined with the next step)."""
    for _ in range(warmup):
        slab.slab_step(tr)
    r.ctx.profile(True)
    dist.barrier()
    torch.cuda.synchronize()
    linked = getattr(tr, "linked", False)
    l0 = r.ctx.launch_count()
    if linked:   # the library's own stream: time it with the library's events (torch's events see torch's streams only)
        t0 = r.ctx.stream_timer_start()
        for _ in range(steps):
            slab.slab_step(tr)
        ms = r.ctx.stream_timer_stop(t0) / steps
    else:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            slab.slab_step(tr)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
    launches = r.ctx.launch_count() - l0 + (0 if linked else steps * slab.EXTRA_LAUNCHES_PER_STEP)
    prof = r.ctx.profile_report()
    r.ctx.profile(False)
    dist.barrier()
    out = dict(ms=ms, launches=launches, prof=prof)
    if not e2e:
        return out
    cap = r.capacity
    host_pos = [L.pinned_array((cap, 4)), L.pinned_array((cap, 4))]
    host_sig = [L.pinned_array((100,)), L.pinned_array((100,))]
    e2e_steps = max(3, min(steps, 10))

    def blocking(s):
        host_sig[0][:] = muscle_signal(s)
        r.ctx.set_muscles(host_sig[0])
        slab.slab_step(tr)
        r._check(r.lib.sph_slab_read_owned(r.ctx.h, host_pos[0].ctypes.data_as(C.c_void_p), None, None), "read_owned")

    marks = []
    # SPH_E2E_XYZ=1: packed (x, y, z), 12 bytes per particle. Off by default: the local order changes with every
    # migration (a few particles cross a cell layer in almost every step even of a fluid at rest), so a consumer needs
    # the ids again each time — measured at 4 GPUs: 3.96 G with 16-byte rows, 2.70 G with packed rows + a blocking id read
    xyz = os.environ.get("SPH_E2E_XYZ", "0") == "1" and hasattr(L.lib, "sph_slab_read_owned_xyz_begin")
    out["e2e_bytes_per_particle"] = 12 if xyz else 16

    def pipelined(s):
        t = [time.perf_counter()]
        host_sig[s & 1][:] = muscle_signal(s)
        r.ctx.set_muscles(host_sig[s & 1])
        t.append(time.perf_counter())
        if linked:
            r.step_begin(); t.append(time.perf_counter())
            r.step_end(); t.append(time.perf_counter())
        else:
            slab.slab_step(tr); t.append(time.perf_counter())
        r.ctx.read_end(); t.append(time.perf_counter())
        if xyz:   # packed (x, y, z): 12 bytes per particle; the ids are fetched again only when particles migrated
            if linked and (r.counts[0] or r.counts[2]):
                r.read_gid()
            r.read_owned_xyz_begin(host_pos[s & 1])
        else:
            r.read_owned_begin(host_pos[s & 1])
        t.append(time.perf_counter())
        marks.append([round((b - a) * 1e3, 2) for a, b in zip(t, t[1:])])

    blocking(0)
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(e2e_steps):
        blocking(s)
    out["e2e_blocking_ms"] = (time.perf_counter() - t0) * 1e3 / e2e_steps
    pipelined(1)
    r.ctx.read_end()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(e2e_steps):
        pipelined(s)
    r.ctx.read_end()
    out["e2e_ms"] = (time.perf_counter() - t0) * 1e3 / e2e_steps
    out["e2e_steps"] = e2e_steps
    if os.environ.get("SPH_BENCH_MARKS"):   # host-side timeline of the pipelined loop: muscles | [begin | end] or step | read_end | read_begin
        print("rank %d marks %s" % (r.index, marks[-4:]), file=sys.stderr, flush=True)
    return out


def muscle_signal(step):
    s = np.zeros(100, np.float32)
    s[:96] = (0.5 + 0.5 * np.sin(0.37 * np.arange(96) + 0.9 * step)).astype(np.float32)
    return s


def parity_scene(L, world):
    """A liquid column in a tall box, thrown along z so that whole layers of particles change slab every few steps
    (the scene of tests/test_slab_gpu.py, scaled with the number of ranks). Identical on every rank."""
    pos, vel, box = scenes.fluid_box(40, 36, 24 * world, fill_z=40 * world, seed=7)
    liquid = pos[:, 3] < 2
    n = int(liquid.sum())
    idx = np.arange(n, dtype=np.uint64)
    vel[liquid, 2] = np.float32(0.5) + scenes.jitter(idx, 0, 99, 0.05)    # m/s: ~1.2 grid units per step
    vel[liquid, 0] = scenes.jitter(idx, 1, 99, 0.025)
    p = L.default_params(box=box)
    t = pos[:, 3].astype(np.int32)
    p.particle_count, p.num_liquid, p.num_boundary = len(pos), int((t == 1).sum()), int((t == 3).sum())
    p.hash_mask = 0xFFFFFFFF
    return p, pos, vel


def slab_parity_check(slab, torch, dist, L, rank, world, local, steps=10):
    """The multi-GPU path that the timed loop uses (real ranks, real transport) against a single-device run of the
    same scene: positions and velocities per global id must be bit-identical. Returns the text for the bench line;
    raises SystemExit on a mismatch."""
    p, pos, vel = parity_scene(L, world)
    parts, caps = slab.split_fluid(p, pos, world)
    z0, z1, own = parts[rank]
    dev = torch.device("cuda", local)
    r = slab.SlabRank(L, p, pos[own], vel[own], own.astype(np.uint32), z0, z1, rank, world, device=local,
                      alloc=slab.DistTransport.allocator(dev), **caps)
    tr = make_transport(slab, torch, r, dev)
    moved = 0
    for _ in range(steps):
        slab.slab_step(tr)
        moved += r.counts[0] + r.counts[2]
    gp, gv, gg = r.read_owned()
    r.close()
    digest = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    mine = dict(gid=gg, pos=gp, vel=gv, moved=moved)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0)
    verdict = [None]
    if rank == 0:
        apos, avel = np.zeros_like(pos), np.zeros_like(vel)
        seen = np.zeros(len(pos), np.int32)
        for g in gathered:
            apos[g["gid"]], avel[g["gid"]] = g["pos"], g["vel"]
            seen[g["gid"]] += 1
        ctx = L.create(p, pos, vel, device=local)
        for _ in range(steps):
            ctx.step()
        spos, svel = ctx.read("position", len(pos)), ctx.read("velocity", len(pos))
        ctx.close()
        same = bool((seen == 1).all()) and digest(apos) == digest(spos) and digest(avel) == digest(svel)
        total_moved = sum(g["moved"] for g in gathered)
        verdict[0] = ("bit-exact vs single device (%d particles, %d steps, %d migrations, %d ranks, sha256 %s)" %
                      (len(pos), steps, total_moved, world, digest(apos)[:16])) if same else \
                     ("MISMATCH vs single device: %d particles differ, ownership %s" %
                      (int((apos.view(np.uint32) != spos.view(np.uint32)).any(axis=1).sum()),
                       "ok" if (seen == 1).all() else "broken"))
    dist.broadcast_object_list(verdict, src=0)
    if verdict[0].startswith("MISMATCH"):
        raise SystemExit("bench.py: slab parity check failed: " + verdict[0])
    return verdict[0]


def time_pressurised(L, name, scene, device=0, chunk=100, max_steps=1500, want_frac=0.5, steps=30):
    """The same scene later, when the fluid carries pressure: the pressure-force kernel drops pairs without pressure
    before the sqrt / divisions, and a fluid at rest (the bench scene during its first steps) has nothing but such pairs.
    Steps in chunks until half of the particles have a non-zero pressure (or max_steps), watching the library's
    out-of-grid / NaN counter so that a diverging run is never stepped on; then times `steps` steps."""
    p, pos, vel = scene[:3]
    ctx = L.create(p, pos, vel, device=device, **EXTRA.get(name, {}))
    done, frac, healthy = 0, 0.0, True
    while done < max_steps:
        ctx.run_steps(chunk)
        done += chunk
        if ctx.out_of_grid_count() != 0:
            healthy = False
            break
        frac = float((ctx.read("pressure") > 0).mean())
        if frac >= want_frac:
            break
    out = {"steps_before": done, "pressure_nonzero_frac": frac, "healthy": healthy}
    if healthy:
        ms = ctx.run_steps(steps) / steps
        healthy = ctx.out_of_grid_count() == 0
        out.update(ms_per_step=ms, value=len(pos) / (ms * 1e-3), healthy=healthy,
                   pressure_nonzero_frac=float((ctx.read("pressure") > 0).mean()))
    ctx.close()
    return out


def roofline_of(prof, n, ms_total_region, value_per_gpu, world):
    """Roofline object of the dominant kernel (per-kernel CUDA events recorded inside the timed region)."""
    peak, peak_src = peaks()
    top = max(prof, key=lambda k: prof[k]["ms"]) if prof else None
    if not top:
        return None
    avg_ms = prof[top]["ms"] / prof[top]["calls"]
    ach = KERNEL_BYTES.get(top, 0) * n / (avg_ms * 1e-3) / 1e9
    traffic, traffic_src = measured_traffic(top, n) if world == 1 else (None, None)
    return {"bound": "hbm", "limiter": KERNEL_LIMITER.get(top, "?"), "kernel": top, "achieved": ach, "peak": peak,
            "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src,
            "traffic_measured_in_run": False,
            "dram_frac": (traffic / (avg_ms * 1e-3) / 1e9 / peak) if traffic else None,
            "note": "achieved/frac use the ALGORITHMIC bytes of the reference stages the kernel replaces; dram_frac = "
                    "ncu-measured DRAM bytes (committed capture) / this run's launch time / peak. The kernel is limited "
                    "by `limiter` (ncu), not by HBM.",
            "peak_source": peak_src, "avg_launch_ms": avg_ms,
            "algorithmic_bytes_per_particle": KERNEL_BYTES.get(top, 0),
            "step_achieved": value_per_gpu * B_REF / 1e9, "step_frac": value_per_gpu * B_REF / 1e9 / peak,
            "kernel_share_of_step": {k: v["ms"] / (ms_total_region if ms_total_region else 1) for k, v in prof.items()}}


def config_block(L, local, args):
    """BASELINE.json configs A-D on one GPU, each on its own scene: device-timed steps, e2e through the drop-in
    class, and the CPU arm on the same scene."""
    out = {}
    plan = {"config_a": (400, 20, 20), "config_b": (200, 10, 10), "dambreak_1m": (100, 10, 5), "fluid_16m": (10, 3, 2)}
    for tag, name in BASELINE_CONFIGS[:4]:
        steps, e2e_steps, cpu_steps = plan[name]
        scene = make_workload(name, L)
        n = len(scene[1])
        if name == "dambreak_1m":
            late = time_pressurised(L, name, scene, device=local)
        ms, launches, _, fin = time_device(L, name, scene, steps, 5, device=local)
        e2e_ms, ok = time_facade(name, scene, e2e_steps, device=local)
        blk_ms, _ = time_facade(name, scene, e2e_steps, device=local, run_ahead=False)
        row = {"workload": name, "particles": n, "ms_per_step": ms, "value": n / (ms * 1e-3),
               "e2e_value": n / (e2e_ms * 1e-3), "e2e_ms_per_step": e2e_ms,
               "e2e_blocking_value": n / (blk_ms * 1e-3), "gpu_launches_per_step": launches / steps,
               "empty_kernel_chain_ms": LAST_CHAIN_FLOOR_MS, "pressure_nonzero_frac": LAST_PRESSURE_FRAC,
               "finite": fin and ok}
        if name == "dambreak_1m":
            row["later_with_pressure"] = late
        if not args.no_cpu_baseline:
            cb, cpu_ms, _ = time_cpu(name, cpu_steps, 1, scene=scene[:3])
            row.update(cpu_value=cb["value"], cpu_ms_per_step=cpu_ms, cpu_kind=cb["kind"], cpu_cores=cb["cores"],
                       cpu_steps=cpu_steps, x_cpu=row["value"] / cb["value"], x_cpu_e2e=row["e2e_value"] / cb["value"])
        out[tag] = row
        del scene
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="fluid_8m", choices=sorted(scenes.WORKLOADS) + ["config_a", "config_b"])
    ap.add_argument("--cpu-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs A-D block / the strong-scaling run")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return 0
    warmup = max(3, args.warmup)
    steps = max(1, args.steps)

    import torch  # device enumeration, torch.distributed plumbing
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU arm)")
    L = capi.load_product()
    extra = {}
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        extra["cpu_binding"] = bind_to_gpu_numa_node(local)
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
        slab = importlib.import_module(PKG + ".slab")
        side = torch.cuda.Stream(device=torch.device("cuda", local))
        with torch.cuda.stream(side):
            extra["slab_parity"] = slab_parity_check(slab, torch, dist, L, rank, world, local)
            sc = scenes.slab_scene(args.workload, world, rank)
            r, tr = slab_rank(L, slab, torch, sc, rank, world, local)
            sampler = ClockSampler(local)
            sampler.start()
            res = time_slab(slab, torch, dist, L, r, tr, warmup, steps)
            clocks = sampler.stop()
            info = r.info()
            r.close()
            n = info[0]
            t = torch.tensor([res["ms"], res["e2e_ms"], res["e2e_blocking_ms"]], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, e2e_ms, e2e_blocking_ms = float(t[0]), float(t[1]), float(t[2])
            cnt = torch.tensor([n], device="cuda", dtype=torch.int64)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
            total = int(cnt[0])
            slab_info = dict(n_local=info[1], halo=info[2:6], z=[sc["z0"], sc["z1"]], n_total=sc["n_total"])
            prof, launches, e2e_steps = res["prof"], res["launches"], res["e2e_steps"]
            global E2E_BYTES
            E2E_BYTES = res["e2e_bytes_per_particle"]
            del sc
            if not args.no_configs:   # BASELINE config D: ONE 16 M fluid divided among the ranks
                sc = scenes.slab_scene("fluid_16m", world, rank, stack=1)
                r, tr = slab_rank(L, slab, torch, sc, rank, world, local)
                rs = time_slab(slab, torch, dist, L, r, tr, 3, 10, e2e=False)
                r.close()
                t = torch.tensor([rs["ms"]], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                n1 = [None]
                if rank == 0:     # the N = 1 point on the same box, for the efficiency
                    scene = make_workload("fluid_16m", L)
                    n1[0] = time_device(L, "fluid_16m", scene, 10, 3, device=local)[0]
                    del scene
                dist.broadcast_object_list(n1, src=0)
                extra["strong_16m"] = {"workload": "fluid_16m", "particles_total": sc["n_total"], "scaling": "strong",
                                       "ms_per_step": float(t[0]), "value": sc["n_total"] / (float(t[0]) * 1e-3),
                                       "n1_ms_per_step": n1[0], "efficiency_vs_n1": n1[0] / (float(t[0]) * world),
                                       "limiter": "halo exchange latency + ghost layers: the per-step exchange cost does "
                                                  "not shrink with the slab while the kernels do"}
        e2e_mode = ("slab ranks: muscle signal up, step, owned positions back into pinned host memory every step, pipelined "
                    "with the next step; %d bytes per particle (%s)" %
                    (E2E_BYTES, "packed x, y, z: the type tag is constant; ids re-read only after a migration" if E2E_BYTES == 12
                     else "x, y, z, type"))
    else:
        dist = None
        scene = make_workload(args.workload, L)
        n = total = len(scene[1])
        sampler = ClockSampler(local)
        sampler.start()
        big = n > (1 << 21)   # per-kernel events cost nothing at 8 M particles; small scenes replay a CUDA graph instead
        ms, launches, prof, fin = time_device(L, args.workload, scene, steps, warmup, device=local, profile=big)
        global PRESSURE_FRAC_MAIN
        PRESSURE_FRAC_MAIN = LAST_PRESSURE_FRAC
        clocks = sampler.stop()
        if not big:
            prof = time_device(L, args.workload, scene, steps, warmup, device=local, profile=True)[2]
        e2e_steps = max(3, min(steps, 10))
        e2e_ms, ok = time_facade(args.workload, scene, e2e_steps, device=local)
        e2e_blocking_ms, _ = time_facade(args.workload, scene, e2e_steps, device=local, run_ahead=False)
        if not (fin and ok):
            raise SystemExit("bench.py: non-finite positions after the timed steps")
        slab_info = None
        e2e_mode = ("owPhysicsFluidSimulator::simulationStep of libow_b200.so (drop-in class): per call muscle signal from "
                    "the Python hook up, step, positions into position_cpp before the call returns; the class queues "
                    "step k+1 before call k returns (run-ahead), so the copy overlaps the next step")
        if not args.no_configs:
            extra["configs"] = config_block(L, local, args)
        del scene
    value = total / (ms * 1e-3)
    roof = roofline_of(prof, n, ms * steps, value / world, world)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles_per_gpu": n,
                       "particles_total": total, "slab": slab_info, "pcisph_iterations": 3,
                       "hash_mask": "0xffffffff", "l2": ("state (%.1f GB) larger than L2, no flush" % (n * 424 / 1e9)) if n * 424 > 2.5e8 else
                       "state (%.0f MB) FITS in L2: not a bench configuration" % (n * 424 / 1e6),
                       "multi_gpu": "none" if world == 1 else
                       "one fluid, %d z-slabs, ghost layers + per-stage halo refresh + migration; transport: %s" %
                       (world, TRANSPORT_USED)},
            "clocks": clocks,
            "e2e": {"value": total / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 400,
                    "d2h_bytes_per_step": E2E_BYTES * n, "ms_per_step": e2e_ms, "steps": e2e_steps, "mode": e2e_mode,
                    "blocking_value": total / (e2e_blocking_ms * 1e-3), "blocking_ms_per_step": e2e_blocking_ms},
            "gpu_launches": launches, "roofline": roof}
    if world == 1:
        line["config"]["pressure_nonzero_frac"] = PRESSURE_FRAC_MAIN
    line.update(extra)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:   # cross-check of the reference arm, same scene, a few steps
            line["cpu_baseline"], _, _ = time_cpu(args.workload, args.cpu_steps, 1)
            line["cpu_baseline"]["note"] = "cross-check only: the driver's --impl reference arm times the same scene over the full step count"
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
