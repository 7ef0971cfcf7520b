#!/usr/bin/env python3
"""bench.py — PCISPH particle-steps/s on N B200s (BASELINE.json metric), one JSON line on rank 0.

  python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference]

A step = one pass of the whole hot path (owPhysicsFluidSimulator::simulationStep's device part:
hash/sort/neighbour search, PCISPH loop with 3 iterations, integrate, membranes) over the synthetic
fluid named in config.workload (8 M liquid particles per GPU on the
reference's r0 lattice in a boundary box; weak scaling).
  value   whole-job particle-steps/s, state resident in HBM, CUDA events on the library's stream,
          max over ranks.
  e2e     the same metric through the reference-facing call sequence with HOST buffers: per step
          upload the muscle signal (pinned), run the step, read position back (pinned) — what
          simulationStep does (owPhysicsFluidSimulator.cpp:115,141).
  roofline / cpu_baseline: see DESIGN.md "Measurement".
--impl reference times the reference's own CPU kernels (oracle/_ref, sphFluid.cl compiled as C++;
else the C port) with all host threads on a bounded sample of the same workload.
"""
import argparse
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
    "forces+predict0": 356 + 92,
    "elastic": 0,
    "predict_density+correct_pressure": 280 + 16,
    "pressure_force+predict": 324 + 92,
    "pressure_force+integrate": 324 + 380,
    "membranes": 376,
}


# profile name -> kernel whose ncu capture (profiles/*_ncu_full_summary.json) gives the measured DRAM traffic per launch
NCU_KERNEL = {"find_neighbors+density": "k_find_neighbors_list<1,2>", "pressure_force+predict": "k_pressure_force<1>",
              "pressure_force+integrate": "k_pressure_force<2>", "predict_density+correct_pressure": "k_predict_density<1>"}


def measured_traffic(kernel, particles):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this same workload (None when the capture is for another particle count)."""
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
EXTRA = {}   # workload name -> dict(elastic=, membranes=, pml=) for scenes with elastic matter


def make_workload(name, lib):
    if name == "config_b":  # BASELINE config 1 of the reference: elastic worm body, membranes, muscles, 16-bit mask
        z = np.load(CONFIG_B)
        pos, vel = z["position"], z["velocity"]
        p = lib.default_params()
        t = pos[:, 3].astype(np.int32)
        p.particle_count, p.num_liquid, p.num_boundary = len(pos), int((t == 1).sum()), int((t == 3).sum())
        p.num_elastic, p.num_membranes = int((t == 2).sum()), int(np.asarray(z["membraneData"]).reshape(-1, 3).shape[0])
        EXTRA[name] = dict(elastic=z["elasticConnectionsData"], membranes=z["membraneData"], pml=z["particleMembranesList"])
        return p, pos, vel
    if name == "config_a":  # BASELINE config 0: positionPureLiquid.txt / velocityPureLiquid.txt, 36h x 24h x 54h, 16-bit mask
        z = np.load(CONFIG_A)
        pos, vel = z["position"], z["velocity"]
        p = lib.default_params(box=(36, 24, 54))
        t = pos[:, 3].astype(np.int32)
        p.particle_count, p.num_liquid, p.num_boundary = len(pos), int((t == 1).sum()), int((t == 3).sum())
        return p, pos, vel
    nx, ny, nz, fill = scenes.WORKLOADS[name]
    pos, vel, box = scenes.fluid_box(nx, ny, nz, fill_z=fill)
    p = lib.default_params(box=box)
    t = pos[:, 3].astype(np.int32)
    p.particle_count = len(pos)
    p.num_liquid = int((t == 1).sum())
    p.num_boundary = int((t == 3).sum())
    p.hash_mask = 0xFFFFFFFF  # > 65 536 cells: the reference's 16-bit mask would alias
    return p, pos, vel


def muscle_signal(step):
    s = np.zeros(100, np.float32)
    s[:96] = (0.5 + 0.5 * np.sin(0.37 * np.arange(96) + 0.9 * step)).astype(np.float32)
    return s


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


def time_cpu(sample, steps, warmup):
    L, kind, set_threads = cpu_oracle()
    cores = os.cpu_count() or 1
    set_threads(cores)
    p, pos, vel = make_workload(sample, L)
    ctx = L.create(p, pos, vel, **EXTRA.get(sample, {}))
    for _ in range(warmup):
        ctx.step()
    t0 = time.perf_counter()
    for _ in range(steps):
        ctx.step()
    dt = time.perf_counter() - t0
    n = len(pos)
    ctx.close()
    return {"value": n * steps / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%s (%d particles of the same lattice fluid), %d steps after %d warm-up, %.1f s" %
                      (sample, n, steps, warmup, dt)}, dt / steps * 1e3, n


def run_reference(args, rank):
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, min(args.warmup, 2))
    steps = min(steps, 20)
    cb, ms, n = time_cpu(args.cpu_sample, steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "sample": args.cpu_sample, "particles_per_step": n,
                       "pcisph_iterations": 3, "hash_mask": "0xffffffff"},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_slab_rank(args, rank, world, local, torch, dist, L, warmup, steps):
    """N > 1: ONE fluid, slab-decomposed along z over the ranks: ghost layers, per-stage halo
    refresh and migration over torch.distributed point-to-point (NCCL / NVLink). Weak scaling: every rank owns
    the named workload's particle count. Returns (ms, e2e_ms, n_owned, profile, launches, info)."""
    slab = importlib.import_module(PKG + ".slab")
    sc = scenes.slab_scene(args.workload, world, rank)
    p = L.default_params(box=sc["box"])
    p.hash_mask = 0xFFFFFFFF
    p.particle_count = len(sc["gid"])
    p.num_liquid = sc["n_liquid"]
    p.num_boundary = len(sc["gid"]) - sc["n_liquid"]
    dev = torch.device("cuda", local)
    # the library's kernels and torch's NCCL transfers must be ordered on ONE stream: a side stream of torch's
    # (a non-zero handle; the context's own stream is non-blocking and would not order with the legacy stream 0)
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        r = slab.SlabRank(L, p, sc["position"], sc["velocity"], sc["gid"], sc["z0"], sc["z1"], rank, world, device=local,
                          alloc=slab.DistTransport.allocator(dev), halo_cap=sc["halo_cap"], mig_cap=sc["mig_cap"])
        r.set_stream(torch.cuda.current_stream().cuda_stream)  # kernels and NCCL transfers ordered on one stream
        tr = slab.DistTransport(r, dev)
        for _ in range(warmup):
            slab.slab_step(tr)
        r.ctx.profile(True)
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = r.ctx.launch_count()
        a.record()
        for _ in range(steps):
            slab.slab_step(tr)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        launches = r.ctx.launch_count() - l0 + steps * (4 + 14)  # + classify/append/iota and the pack/unpack kernels
        prof = r.ctx.profile_report()
        r.ctx.profile(False)
        dist.barrier()
        # e2e: muscle signal up, step, owned positions back to pinned host memory, every step
        cap = r.capacity
        host_pos = L.pinned_array((cap, 4))
        host_sig = L.pinned_array((100,))
        e2e_steps = max(3, min(steps, 10))
        import ctypes as C
        def one(s):
            host_sig[:] = muscle_signal(s)
            r.ctx.set_muscles(host_sig)
            slab.slab_step(tr)
            r._check(r.lib.sph_slab_read_owned(r.ctx.h, host_pos.ctypes.data_as(C.c_void_p), None, None), "read_owned")
        one(0)
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s in range(e2e_steps):
            one(s)
        e2e_blocking_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        # pipelined: the read-back of step s overlaps step s+1 (double-buffered pinned memory), all reads completed
        # inside the timed region
        host_pos2 = [host_pos, L.pinned_array((cap, 4))]
        host_sig2 = [host_sig, L.pinned_array((100,))]
        def two(s):
            host_sig2[s & 1][:] = muscle_signal(s)
            r.ctx.set_muscles(host_sig2[s & 1])
            slab.slab_step(tr)
            r.ctx.read_end()
            r.read_owned_begin(host_pos2[s & 1])
        two(1)
        r.ctx.read_end()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s in range(e2e_steps):
            two(s)
        r.ctx.read_end()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        info = r.info()
        out = (ms, e2e_ms, info[0], prof, launches, dict(n_local=info[1], halo=info[2:6], z=[sc["z0"], sc["z1"]],
                                                           n_total=sc["n_total"], e2e_steps=e2e_steps,
                                                           e2e_blocking_ms=e2e_blocking_ms))
    r.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="fluid_8m", choices=sorted(scenes.WORKLOADS) + ["config_a", "config_b"])
    ap.add_argument("--cpu-sample", default="fluid_2m", choices=sorted(scenes.WORKLOADS) + ["config_a", "config_b"])
    ap.add_argument("--cpu-steps", type=int, default=30)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    L = capi.load_product()
    slab_info = None
    e2e_extra = {"mode": "blocking: every step waits for its read-back"}
    if world > 1:
        sampler = ClockSampler(local)
        sampler.start()
        ms, e2e_ms, n, prof, launches, slab_info = run_slab_rank(args, rank, world, local, torch, dist, L, warmup, steps)
        clocks = sampler.stop()
        e2e_steps = slab_info["e2e_steps"]
        p = L.default_params()
        ctx = None
        t = torch.tensor([ms, e2e_ms, slab_info.pop("e2e_blocking_ms")], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms, e2e_blocking_ms = float(t[0]), float(t[1]), float(t[2])
        cnt = torch.tensor([n], device="cuda", dtype=torch.int64)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        total = int(cnt[0])
        e2e_extra = {"mode": "pipelined: the read-back of step k (sph_slab_read_owned_begin, pinned double buffer) overlaps "
                             "step k+1; every read is completed inside the timed region",
                     "blocking_value": total / (e2e_blocking_ms * 1e-3), "blocking_ms_per_step": e2e_blocking_ms}
    else:
        L = capi.load_product()
        p, pos, vel = make_workload(args.workload, L)
        n = len(pos)
        ctx = L.create(p, pos, vel, device=local, **EXTRA.get(args.workload, {}))
        host_pos = L.pinned_array((n, 4))
        host_sig = L.pinned_array((100,))

        def barrier():
            if dist is not None:
                dist.barrier()
            ctx.sync()

        # ---- warm-up, then the device-timed region (CUDA events on the library's stream) ----
        ctx.run_steps(warmup)
        ctx.profile(True)
        sampler = ClockSampler(local)
        barrier()
        sampler.start()
        launches0 = ctx.launch_count()
        ms = ctx.run_steps(steps)
        launches = ctx.launch_count() - launches0
        clocks = sampler.stop()
        prof = ctx.profile_report()
        ctx.profile(False)
        barrier()
        # ---- e2e: host buffers in the loop, wall clock around a synchronous call sequence ----
        e2e_steps = max(3, min(steps, 10))
        for s in range(2):
            ctx.set_muscles(host_sig)
            ctx.step(sync=False)
            ctx.read_into("position", host_pos)
        barrier()
        t0 = time.perf_counter()
        for s in range(e2e_steps):
            host_sig[:] = muscle_signal(s)
            ctx.set_muscles(host_sig)                 # H2D 400 B (updateMuscleActivityData)
            ctx.step(sync=False)
            ctx.read_into("position", host_pos)       # D2H 16 N B (read_position_buffer), blocking
        e2e_blocking_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        # ---- e2e, pipelined: the same bytes every step, but the read-back of step k (sph_read_begin) overlaps
        # step k+1, whose integration kernel waits for it; two pinned position buffers, two signal buffers.
        # Every read is waited for inside the timed region.
        host_pos2 = [host_pos, L.pinned_array((n, 4))]
        host_sig2 = [host_sig, L.pinned_array((100,))]
        ctx.read_begin("position", host_pos2[1])      # untimed: creates the copy stream, touches the second buffer
        ctx.read_end()
        ctx.sync()
        barrier()
        t0 = time.perf_counter()
        for s in range(e2e_steps):
            host_sig2[s & 1][:] = muscle_signal(s)
            ctx.set_muscles(host_sig2[s & 1])         # H2D 400 B
            ctx.step(sync=False)
            ctx.read_end()                            # positions of step s-1 have arrived in host_pos2[(s-1) & 1]
            ctx.read_begin("position", host_pos2[s & 1])   # D2H 16 N B, starts when step s is done
        ctx.read_end()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        if dist is not None:
            t = torch.tensor([ms, e2e_ms, e2e_blocking_ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, e2e_ms, e2e_blocking_ms = float(t[0]), float(t[1]), float(t[2])
        total = n
        e2e_extra = {"mode": "pipelined: the read-back of step k (sph_read_begin, pinned double buffer) overlaps step "
                             "k+1, whose integration kernel waits for it; every read is completed inside the timed region",
                     "blocking_value": total / (e2e_blocking_ms * 1e-3), "blocking_ms_per_step": e2e_blocking_ms}
    value = total * steps / (ms * 1e-3)
    peak, peak_src = peaks()
    # dominant kernel of the step
    top = max(prof, key=lambda k: prof[k]["ms"]) if prof else None
    roof = None
    if top:
        avg_ms = prof[top]["ms"] / prof[top]["calls"]
        ach = KERNEL_BYTES.get(top, 0) * n / (avg_ms * 1e-3) / 1e9
        traffic, traffic_src = measured_traffic(top, n) if world == 1 else (None, None)
        roof = {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "avg_launch_ms": avg_ms,
                "algorithmic_bytes_per_particle": KERNEL_BYTES.get(top, 0),
                "step_achieved": value / world * B_REF / 1e9, "step_frac": value / world * B_REF / 1e9 / peak,
                "kernel_share_of_step": {k: v["ms"] / (ms if ms else 1) for k, v in prof.items()}}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles_per_gpu": n,
                       "particles_total": total, "slab": slab_info, "pcisph_iterations": 3,
                       "hash_mask": "0xffffffff", "l2": ("state (%.1f GB) larger than L2, no flush" % (n * 424 / 1e9)) if n * 424 > 2.5e8 else
                       "state (%.0f MB) FITS in L2: not a bench configuration, reported for the config-A target only" % (n * 424 / 1e6),
                       "multi_gpu": "none" if world == 1 else
                       "one fluid, %d z-slabs, ghost layers + per-stage halo refresh + migration over NCCL p2p" % world},
            "clocks": clocks,
            "e2e": {"value": total / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 400,
                    "d2h_bytes_per_step": 16 * n, "ms_per_step": e2e_ms, "steps": e2e_steps,
                    **e2e_extra},
            "gpu_launches": launches, "roofline": roof}
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], _, _ = time_cpu(args.cpu_sample, args.cpu_steps, 1)
        print(json.dumps(line), flush=True)
    if ctx is not None:
        ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
