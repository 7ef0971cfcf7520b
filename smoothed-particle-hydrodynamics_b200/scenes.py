"""Synthetic scenes for BASELINE configs C/D/E — host-side tooling, numpy only.

A scene is the reference's own construction re-used at other sizes: liquid particles on the cubic
r0 = h/2 lattice the reference generator uses for its outer liquid
(/root/reference/src/owHelper.cpp:673-689: start 3*r0 from the walls, type tag 1.1, zero velocity)
inside a one-particle-thick boundary box with inward normals stored in the velocity slot
(/root/reference/src/owHelper.cpp:776-928: faces 1, edges 1/sqrt(2), corners 1/sqrt(3), type 3).
A small deterministic jitter breaks the exact distance ties of a perfect lattice. The jitter comes
from a counter-based integer hash of (seed, particle index, axis), so any sub-range of a scene can
be generated independently (multi-GPU slabs) and does not depend on the numpy version.
"""
import numpy as np

H = np.float32(3.34)
R0 = np.float32(0.5) * H


def _hash32(x):
    """lowbias32 integer hash, vectorised on uint32."""
    x = x.astype(np.uint32, copy=True)
    x ^= x >> np.uint32(16)
    x *= np.uint32(0x7FEB352D)
    x ^= x >> np.uint32(15)
    x *= np.uint32(0x846CA68B)
    x ^= x >> np.uint32(16)
    return x


def jitter(index, axis, seed, amplitude):
    """U(-amplitude, +amplitude) float32 for global lattice index `index` (uint64 array)."""
    with np.errstate(over="ignore"):
        k = (index.astype(np.uint64) * np.uint64(3) + np.uint64(axis)) & np.uint64(0xFFFFFFFF)
        hsh = _hash32(k.astype(np.uint32) ^ _hash32(np.full(1, seed, np.uint32)))
    u = hsh.astype(np.float64) / 4294967296.0
    return ((2.0 * u - 1.0) * float(amplitude)).astype(np.float32)


def liquid_block(nx, ny, nz, origin_r0=(3, 3, 3), seed=12345, jitter_frac=0.05, z_range=None, nz_total=None):
    """nx*ny*nz liquid particles, z fastest in memory like the reference generator's loops.
    z_range=(z0,z1) generates only lattice layers z0 <= iz < z1 of a block that is nz_total deep."""
    z0, z1 = (0, nz) if z_range is None else z_range
    nzt = nz if nz_total is None else nz_total
    ix, iy, iz = np.meshgrid(np.arange(nx, dtype=np.int64), np.arange(ny, dtype=np.int64),
                             np.arange(z0, z1, dtype=np.int64), indexing="ij")
    ix, iy, iz = ix.ravel(), iy.ravel(), iz.ravel()
    gidx = ((ix * ny + iy) * nzt + iz).astype(np.uint64)
    pos = np.empty((ix.size, 4), np.float32)
    amp = float(jitter_frac) * float(R0)
    for a, (i, o) in enumerate(zip((ix, iy, iz), origin_r0)):
        pos[:, a] = (i + o).astype(np.float32) * R0 + jitter(gidx, a, seed, amp)
    pos[:, 3] = np.float32(1.1)
    return pos, np.zeros_like(pos)


def boundary_box(bx, by, bz):
    """Boundary particles of a box of bx*by*bz r0-cells (box size = (bx,by,bz)*r0), all six walls."""
    faces = []
    for axis, n in enumerate((bx, by, bz)):
        dims = [bx, by, bz]
        others = [a for a in range(3) if a != axis]
        g = np.meshgrid(*[np.arange(dims[a], dtype=np.int64) for a in others], indexing="ij")
        for side in (0, n - 1):
            idx = np.empty((g[0].size, 3), np.int64)
            idx[:, axis] = side
            idx[:, others[0]] = g[0].ravel()
            idx[:, others[1]] = g[1].ravel()
            faces.append(idx)
    idx = np.unique(np.concatenate(faces), axis=0)
    dims = np.array([bx, by, bz], np.int64)
    f = (idx == 0).astype(np.float32) - (idx == dims - 1).astype(np.float32)
    nn = np.abs(f).sum(1, keepdims=True)
    nrm = f * (np.float32(1.0) / np.sqrt(nn.astype(np.float32)))
    pos = np.empty((len(idx), 4), np.float32)
    pos[:, :3] = (idx.astype(np.float32) + np.float32(0.5)) * R0
    pos[:, 3] = np.float32(3.0)
    vel = np.empty_like(pos)
    vel[:, :3] = nrm
    vel[:, 3] = np.float32(3.0)
    return pos, vel


def fluid_box(nx, ny, nz, fill_z=None, seed=12345, jitter_frac=0.05, margin=3):
    """Liquid block nx*ny*nz (lattice counts) + boundary box leaving `margin` lattice steps to each
    wall. fill_z: box depth in lattice steps if larger than the block (dam-break: block in the low-z
    part). Returns (position, velocity, box multipliers in units of h)."""
    bx, by = nx + 2 * margin, ny + 2 * margin
    bz = (nz if fill_z is None else fill_z) + 2 * margin
    bx, by, bz = bx + (bx & 1), by + (by & 1), bz + (bz & 1)  # box = integer multiples of h
    lp, lv = liquid_block(nx, ny, nz, (margin, margin, margin), seed, jitter_frac)
    bp, bv = boundary_box(bx, by, bz)
    return np.concatenate([lp, bp]), np.concatenate([lv, bv]), (bx / 2.0, by / 2.0, bz / 2.0)


WORKLOADS = {
    # name: (nx, ny, nz, fill_z)  — SURVEY.md §8d
    "dambreak_1m": (100, 100, 100, 312),      # config C: 1 M liquid, block in the low-z third
    "fluid_8m": (200, 160, 250, None),        # config E: 8 M liquid per GPU
    "fluid_16m": (160, 100, 1000, None),      # config D: 16 M liquid
    "fluid_2m": (128, 128, 128, None),
    "fluid_256k": (64, 64, 64, None),
    "fluid_32k": (32, 32, 32, None),
}


def boundary_box_slab(bx, by, bz, k0, k1):
    """The boundary particles of boundary_box(bx, by, bz) whose z index k lies in [k0, k1), with a unique id
    (i*by + j)*bz + k each — what one slab rank needs, without building the whole box."""
    k0, k1 = max(0, int(k0)), min(int(bz), int(k1))
    out = []
    ii, jj = np.meshgrid(np.arange(bx, dtype=np.int64), np.arange(by, dtype=np.int64), indexing="ij")
    ii, jj = ii.ravel(), jj.ravel()
    rim = (ii == 0) | (ii == bx - 1) | (jj == 0) | (jj == by - 1)
    for k in range(k0, k1):
        sel = np.ones(ii.size, bool) if (k == 0 or k == bz - 1) else rim
        idx = np.stack([ii[sel], jj[sel], np.full(int(sel.sum()), k, np.int64)], 1)
        out.append(idx)
    idx = np.concatenate(out) if out else np.zeros((0, 3), np.int64)
    dims = np.array([bx, by, bz], np.int64)
    f = (idx == 0).astype(np.float32) - (idx == dims - 1).astype(np.float32)
    nn = np.maximum(np.abs(f).sum(1, keepdims=True), 1)
    nrm = f * (np.float32(1.0) / np.sqrt(nn.astype(np.float32)))
    pos = np.empty((len(idx), 4), np.float32)
    pos[:, :3] = (idx.astype(np.float32) + np.float32(0.5)) * R0
    pos[:, 3] = np.float32(3.0)
    vel = np.empty_like(pos)
    vel[:, :3] = nrm
    vel[:, 3] = np.float32(3.0)
    ids = (idx[:, 0] * by + idx[:, 1]) * bz + idx[:, 2]
    return pos, vel, ids


def slab_scene(name, world, rank, seed=12345, jitter_frac=0.05, margin=3, stack=None):
    """One rank's part of a slab-decomposed fluid: the named fluid stacked `stack` times along z in one box and cut
    into `world` slabs of whole hash-cell layers. stack = world (default) is the weak-scaling scene (SURVEY.md §8d
    config E: every rank owns one copy of the workload), stack = 1 the strong-scaling one (config D: ONE fluid
    divided among the ranks). Returns dict(position, velocity, gid, z0, z1, box, n_total, halo_cap, mig_cap);
    only this rank's part is generated."""
    nx, ny, nz, fill = WORKLOADS[name]
    stack = world if stack is None else int(stack)
    nzt = nz * stack
    bx, by, bz = nx + 2 * margin, ny + 2 * margin, (nzt if fill is None else fill * stack) + 2 * margin
    bx, by, bz = bx + (bx & 1), by + (by & 1), bz + (bz & 1)
    gz = int(bz / 2.0) + 1  # (int)((zmax - zmin) / h) + 1 with zmax = (bz/2) * h, owOpenCLSolver.cpp:14-17
    cell_inv = np.float32(1.0) / (np.float32(2.0) * H)
    cuts = [0] + [(margin + (r * nzt) // world) // 4 for r in range(1, world)] + [gz]
    z0, z1 = cuts[rank], cuts[rank + 1]
    # lattice / box layers that can fall into cell layers [z0, z1): 4 r0 per cell, one spare layer for the jitter
    lo, hi = 4 * z0 - 1, 4 * z1 + 1
    lz0, lz1 = max(0, lo - margin), min(nzt, hi - margin)
    lp, lv = liquid_block(nx, ny, nzt, (margin, margin, margin), seed, jitter_frac, z_range=(lz0, lz1), nz_total=nzt)
    ix, iy, iz = np.meshgrid(np.arange(nx, dtype=np.int64), np.arange(ny, dtype=np.int64),
                             np.arange(lz0, lz1, dtype=np.int64), indexing="ij")
    lid = ((ix.ravel() * ny + iy.ravel()) * nzt + iz.ravel())
    bp, bv, bid = boundary_box_slab(bx, by, bz, lo, hi)
    n_liquid_total = nx * ny * nzt
    pos, vel = np.concatenate([lp, bp]), np.concatenate([lv, bv])
    gid = np.concatenate([lid, n_liquid_total + bid])
    if gid.size and int(gid.max()) >= 2 ** 32:
        raise ValueError("global ids do not fit 32 bits")
    cz = (pos[:, 2] * cell_inv).astype(np.int32)
    own = (cz >= z0) & (cz < z1)
    n_boundary_total = 2 * bx * by + (bz - 2) * (2 * bx + 2 * by - 4)
    fullest = 4 * nx * ny + 4 * (2 * bx + 2 * by) + bx * by
    return dict(position=pos[own], velocity=vel[own], gid=gid[own].astype(np.uint32), z0=z0, z1=z1,
                box=(bx / 2.0, by / 2.0, bz / 2.0), n_total=n_liquid_total + n_boundary_total,
                n_liquid=int((pos[own, 3] < 2).sum()), halo_cap=2 * fullest + 1024, mig_cap=fullest // 2 + 1024)
