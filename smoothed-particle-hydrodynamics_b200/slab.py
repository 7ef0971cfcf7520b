"""Slab decomposition of one fluid over several GPUs.

The reference is single-device (owOpenCLSolver.cpp:157-182); this module is the host orchestration of
the `sph_slab_*` extension of the C-ABI (include/sph_capi.h, csrc/sph_slab.cu). One rank = one context
= one GPU; the hash grid is cut into slabs of whole cell layers along z (the cell id is z-major,
sphFluid.cl:339-340), every rank keeps ghost copies of its neighbours' boundary layers, and a step is

    classify | counts, migrants + ghost layers -> neighbours | absorb |
    part 0 | rho | part 1 | (p, rho*) | { part 2 | x* | part 3 | (p, rho*) } x (iterations - 1) | part 4

where `| field |` is a halo refresh: pack on the owner, move, unpack into the ghosts. The library does
the device work; a TRANSPORT moves the bytes: `DistTransport` = torch.distributed point-to-point
(NCCL over NVLink for CUDA buffers, gloo for the host counts), `LoopbackTransport` = copies between
ranks living in one process (tests). No collective is involved: a rank talks to rank-1 and rank+1 only.
"""
import ctypes as C

import numpy as np

DOWN, UP = 0, 1
RHO, PRS, XPRED = 0, 1, 2
FIELD_BYTES = {RHO: 4, PRS: 8, XPRED: 16}
EXTRA_LAUNCHES_PER_STEP = 4 + 14   # classify / append / iota x 2 + the 7 x 2 pack / unpack kernels (not in sph_launch_count)


def cell_layer(z, cell_size_inv):
    """z-layer of the hash cell, with hashParticles' float32 arithmetic (sphFluid.cl:361-365)."""
    return (np.asarray(z, np.float32) * np.float32(cell_size_inv)).astype(np.int32)


def cut_layers(cz, world, gz):
    """z-cuts [c0=0, c1, ..., c_world=gz] that balance the particle count; every slab gets >= 2 layers."""
    hist = np.bincount(cz, minlength=gz)[:gz].astype(np.int64)
    cum = np.concatenate([[0], np.cumsum(hist)])
    cuts = [0]
    for r in range(1, world):
        c = int(np.searchsorted(cum, cum[-1] * r / world))
        c = max(c, cuts[-1] + 2)
        c = min(c, gz - 2 * (world - r))
        cuts.append(c)
    cuts.append(gz)
    if any(b - a < 2 for a, b in zip(cuts, cuts[1:])):
        raise ValueError("grid too thin for %d slabs of >= 2 cell layers" % world)
    return cuts


class SlabRank:
    """One slab of the fluid on one GPU: a context created with spare capacity + its exchange buffers."""

    def __init__(self, lib, params, position, velocity, gid, z0, z1, index, world, device=0, alloc=None,
                 capacity=None, mig_cap=None, halo_cap=None):
        self.L, self.lib = lib, lib.lib
        self.index, self.world = index, world
        self.has = [index > 0, index < world - 1]
        self.z0, self.z1 = int(z0), int(z1)
        n = len(position)
        halo_guess = max(1024, int(2.5 * n / max(1, z1 - z0)))      # ~ one cell layer, with head-room
        self.halo_cap = int(halo_cap or halo_guess)
        self.mig_cap = int(mig_cap or max(1024, self.halo_cap // 4))
        self.capacity = int(capacity or (n + 2 * self.halo_cap + 2 * self.mig_cap + n // 8 + 1024))
        p = type(params).from_buffer_copy(params)
        p.particle_count = self.capacity
        pos = np.zeros((self.capacity, 4), np.float32)
        vel = np.zeros((self.capacity, 4), np.float32)
        pos[:n], vel[:n] = position, velocity
        self.ctx = lib.create(p, pos, vel, device=device)
        self.iters = max(1, int(params.max_iteration))
        self._bind()
        g = np.ascontiguousarray(gid, np.uint32)
        self._check(self.lib.sph_slab_begin(self.ctx.h, n, g.ctypes.data_as(C.c_void_p), int(z0), int(z1),
                                            int(self.has[DOWN]), int(self.has[UP]), self.mig_cap, self.halo_cap),
                    "sph_slab_begin")
        rec = self.lib.sph_slab_record_bytes()
        self.rec = rec
        alloc = alloc or self._device_alloc
        self.buf = {}
        for side in (DOWN, UP):
            self.buf["mig_out", side] = alloc(self.mig_cap * rec)
            self.buf["mig_in", side] = alloc(self.mig_cap * rec)
            self.buf["ghost_out", side] = alloc(self.halo_cap * rec)
            self.buf["ghost_in", side] = alloc(self.halo_cap * rec)
            self.buf["halo_out", side] = alloc((self.halo_cap + self.mig_cap) * 16)
            self.buf["halo_in", side] = alloc((self.halo_cap + self.mig_cap) * 16)
        self.counts = [0, 0, 0, 0]
        self.n_send = [0, 0]
        self.n_recv = [0, 0]

    # -- plumbing
    def _bind(self):
        L, V, I = self.lib, C.c_void_p, C.c_int
        L.sph_slab_begin.argtypes = [V, I, V, I, I, I, I, I, I]
        L.sph_slab_classify.argtypes = [V, V, V, V, V, C.POINTER(I)]
        L.sph_slab_absorb.argtypes = [V, V, V, V, V, C.POINTER(I)]
        L.sph_slab_info.argtypes = [V, C.POINTER(I)]
        L.sph_slab_pack.argtypes = [V, I, I, V]
        L.sph_slab_unpack.argtypes = [V, I, I, V]
        L.sph_slab_read_owned.argtypes = [V, V, V, V]
        L.sph_slab_read_owned_begin.argtypes = [V, V]
        if hasattr(L, "sph_slab_read_owned_xyz_begin"):
            L.sph_slab_read_owned_xyz_begin.argtypes = [V, V]
        L.sph_slab_set_cut.argtypes = [V, I, I]
        L.sph_slab_last_error.restype = C.c_char_p
        L.sph_step_part.argtypes = [V, I]
        L.sph_set_stream.argtypes = [V, V]
        L.sph_device_alloc.argtypes = [C.c_size_t]
        L.sph_device_alloc.restype = V
        L.sph_device_free.argtypes = [V]
        L.sph_device_free.restype = None
        L.sph_device_copy.argtypes = [V, V, C.c_size_t]
        if hasattr(L, "sph_slab_link_export"):   # peer-memory exchange (CUDA library only)
            L.sph_slab_link_export.argtypes = [V, V]
            L.sph_slab_link_open.argtypes = [V, I, V]
            L.sph_slab_link_local.argtypes = [V, I, V]
            L.sph_slab_step.argtypes = [V]
            L.sph_slab_step_begin.argtypes = [V]
            L.sph_slab_step_end.argtypes = [V]
            L.sph_slab_counts.argtypes = [V, C.POINTER(I)]

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s failed (%d): %s / %s" % (what, rc, self.lib.sph_slab_last_error().decode(),
                                                            self.L.last_error()))

    def _device_alloc(self, nbytes):
        ptr = self.lib.sph_device_alloc(int(nbytes))
        if not ptr:
            raise MemoryError("sph_device_alloc(%d)" % nbytes)
        return RawBuffer(ptr, int(nbytes))

    def set_stream(self, cuda_stream):
        self._check(self.lib.sph_set_stream(self.ctx.h, C.c_void_p(cuda_stream)), "sph_set_stream")

    # -- the protocol
    def classify(self):
        out = (C.c_int * 4)()
        b = self.buf
        self._check(self.lib.sph_slab_classify(self.ctx.h, b["mig_out", DOWN].ptr, b["ghost_out", DOWN].ptr,
                                               b["mig_out", UP].ptr, b["ghost_out", UP].ptr, out), "sph_slab_classify")
        self.counts = list(out)
        return self.counts

    def absorb(self, incoming):
        inc = (C.c_int * 4)(*[int(x) for x in incoming])
        b = self.buf
        self._check(self.lib.sph_slab_absorb(self.ctx.h, b["mig_in", DOWN].ptr, b["ghost_in", DOWN].ptr,
                                             b["mig_in", UP].ptr, b["ghost_in", UP].ptr, inc), "sph_slab_absorb")
        info = self.info()
        self.n_send, self.n_recv = info[2:4], info[4:6]

    def set_cut(self, z0, z1):
        self._check(self.lib.sph_slab_set_cut(self.ctx.h, int(z0), int(z1)), "sph_slab_set_cut")
        self.z0, self.z1 = int(z0), int(z1)

    def n_owned(self):
        return self.info()[0]

    def info(self):
        out = (C.c_int * 6)()
        self._check(self.lib.sph_slab_info(self.ctx.h, out), "sph_slab_info")
        return list(out)

    def part(self, k):
        self.L.check(self.lib.sph_step_part(self.ctx.h, int(k)), "sph_step_part(%d)" % k)

    def pack(self, field, side):
        self._check(self.lib.sph_slab_pack(self.ctx.h, field, side, self.buf["halo_out", side].ptr), "sph_slab_pack")

    def unpack(self, field, side):
        self._check(self.lib.sph_slab_unpack(self.ctx.h, field, side, self.buf["halo_in", side].ptr), "sph_slab_unpack")

    def sync(self):
        self.ctx.sync()

    def copy(self, dst, src, nbytes):
        if nbytes:
            self._check(self.lib.sph_device_copy(dst.ptr, src.ptr, int(nbytes)), "sph_device_copy")

    def read_owned(self):
        n = self.info()[0]
        pos, vel, gid = np.empty((n, 4), np.float32), np.empty((n, 4), np.float32), np.empty(n, np.uint32)
        self._check(self.lib.sph_slab_read_owned(self.ctx.h, pos.ctypes.data_as(C.c_void_p),
                                                 vel.ctypes.data_as(C.c_void_p), gid.ctypes.data_as(C.c_void_p)),
                    "sph_slab_read_owned")
        return pos, vel, gid

    def read_owned_begin(self, out):
        """Pipelined read of the owned positions into `out` (pinned, >= n_owned rows); finish with ctx.read_end()."""
        self._check(self.lib.sph_slab_read_owned_begin(self.ctx.h, out.ctypes.data_as(C.c_void_p)),
                    "sph_slab_read_owned_begin")

    def read_owned_xyz_begin(self, out):
        """Pipelined read of the owned positions as packed (x, y, z) into `out` (pinned, >= 3 * n_owned floats)."""
        self._check(self.lib.sph_slab_read_owned_xyz_begin(self.ctx.h, out.ctypes.data_as(C.c_void_p)),
                    "sph_slab_read_owned_xyz_begin")

    def read_gid(self):
        n = self.info()[0]
        gid = np.empty(n, np.uint32)
        self._check(self.lib.sph_slab_read_owned(self.ctx.h, None, None, gid.ctypes.data_as(C.c_void_p)), "sph_slab_read_owned")
        return gid

    def close(self):
        self.ctx.close()

    # -- links: the neighbours' kernels write into this rank's mailbox (csrc/sph_slab.cu, "Links")
    def link_export(self):
        """Create the mailbox; returns its CUDA IPC handle (bytes) for neighbours in other processes."""
        n = self.lib.sph_slab_link_handle_bytes()
        buf = (C.c_char * n)()
        self._check(self.lib.sph_slab_link_export(self.ctx.h, buf), "sph_slab_link_export")
        return bytes(buf)

    def link_open(self, side, handle):
        buf = (C.c_char * len(handle)).from_buffer_copy(handle)
        self._check(self.lib.sph_slab_link_open(self.ctx.h, side, buf), "sph_slab_link_open")

    def link_local(self, side, other):
        self._check(self.lib.sph_slab_link_local(self.ctx.h, side, other.ctx.h), "sph_slab_link_local")

    def step_linked(self):
        self.step_begin()
        self.step_end()

    def step_begin(self):
        self._check(self.lib.sph_slab_step_begin(self.ctx.h), "sph_slab_step_begin")

    def step_end(self):
        self._check(self.lib.sph_slab_step_end(self.ctx.h), "sph_slab_step_end")
        out = (C.c_int * 4)()
        self._check(self.lib.sph_slab_counts(self.ctx.h, out), "sph_slab_counts")
        self.counts = list(out)
        info = self.info()
        self.n_send, self.n_recv = info[2:4], info[4:6]


class RawBuffer:
    """A device (or host) allocation known by address only."""

    def __init__(self, ptr, nbytes, tensor=None):
        self.ptr, self.nbytes, self.tensor = ptr, nbytes, tensor


class LoopbackTransport:
    """All ranks live in this process (one GPU or the host): an exchange is a copy between their buffers."""

    def __init__(self, ranks):
        self.ranks = ranks

    def exchange_counts(self, counts):
        """counts[r] = [mig_down, ghost_down, mig_up, ghost_up] -> incoming[r] = [mig_from_down, ghost_from_down,
        mig_from_up, ghost_from_up]."""
        w = len(self.ranks)
        inc = []
        for r in range(w):
            below = counts[r - 1][2:4] if r > 0 else [0, 0]
            above = counts[r + 1][0:2] if r + 1 < w else [0, 0]
            inc.append(list(below) + list(above))
        return inc

    def _move(self, kind_out, kind_in, sizes):
        """sizes[r][side] bytes leave rank r towards `side` and land in the neighbour's opposite-side buffer."""
        for r in self.ranks:
            r.sync()
        for i, r in enumerate(self.ranks):
            if r.has[UP]:
                nb = self.ranks[i + 1]
                nb.copy(nb.buf[kind_in, DOWN], r.buf[kind_out, UP], sizes[i][UP])
            if r.has[DOWN]:
                nb = self.ranks[i - 1]
                nb.copy(nb.buf[kind_in, UP], r.buf[kind_out, DOWN], sizes[i][DOWN])

    def all_gather(self, rows):
        """rows[r] = a few ints of local rank r -> the rows of all ranks, in rank order."""
        return [list(x) for x in rows]

    def exchange_records(self, counts, incoming):
        rec = self.ranks[0].rec
        self._move("mig_out", "mig_in", [[c[0] * rec, c[2] * rec] for c in counts])
        self._move("ghost_out", "ghost_in", [[c[1] * rec, c[3] * rec] for c in counts])

    def exchange_halo(self, field):
        b = FIELD_BYTES[field]
        self._move("halo_out", "halo_in", [[r.n_send[DOWN] * b, r.n_send[UP] * b] for r in self.ranks])


class DistTransport:
    """One rank per process: torch.distributed point-to-point with rank-1 / rank+1. CUDA buffers travel over
    NCCL (NVLink / NVSwitch); the four host-side counts of a step travel over the CPU backend (gloo).
    Needs a process group with both, e.g. init_process_group("cpu:gloo,cuda:nccl"); with gloo only (CPU
    tests) the buffers are host tensors."""

    def __init__(self, rank_obj, device=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.ranks = [rank_obj]
        self.me, self.world = dist.get_rank(), dist.get_world_size()
        self.device = device

    @staticmethod
    def allocator(device):
        """Buffers as torch tensors (what torch.distributed can send), exposed to the library by address."""
        import torch

        def alloc(nbytes):
            t = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
            return RawBuffer(t.data_ptr(), int(nbytes), t)
        return alloc

    def _p2p(self, sends, recvs):
        """sends / recvs: lists of (tensor_slice, peer)."""
        dist = self.dist
        ops = [dist.P2POp(dist.irecv, t, p) for t, p in recvs if t.numel()]
        ops += [dist.P2POp(dist.isend, t, p) for t, p in sends if t.numel()]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def exchange_counts(self, counts):
        torch = self.torch
        r, c = self.ranks[0], counts[0]
        out_dn, out_up = torch.tensor(c[0:2], dtype=torch.int32), torch.tensor(c[2:4], dtype=torch.int32)
        in_dn, in_up = torch.zeros(2, dtype=torch.int32), torch.zeros(2, dtype=torch.int32)
        sends, recvs = [], []
        if r.has[DOWN]:
            sends.append((out_dn, self.me - 1)); recvs.append((in_dn, self.me - 1))
        if r.has[UP]:
            sends.append((out_up, self.me + 1)); recvs.append((in_up, self.me + 1))
        self._p2p(sends, recvs)
        return [in_dn.tolist() + in_up.tolist()]

    def all_gather(self, rows):
        torch = self.torch
        mine = torch.tensor(rows[0], dtype=torch.int64)
        out = [torch.zeros_like(mine) for _ in range(self.world)]
        self.dist.all_gather(out, mine)   # CPU tensors: the gloo side of the process group
        return [t.tolist() for t in out]

    def _move(self, kind_out, kind_in, out_bytes, in_bytes):
        r = self.ranks[0]
        sends, recvs = [], []
        for side, peer in ((DOWN, self.me - 1), (UP, self.me + 1)):
            if not r.has[side]:
                continue
            sends.append((r.buf[kind_out, side].tensor[:out_bytes[side]], peer))
            recvs.append((r.buf[kind_in, side].tensor[:in_bytes[side]], peer))
        self._p2p(sends, recvs)

    def exchange_records(self, counts, incoming):
        rec, c, i = self.ranks[0].rec, counts[0], incoming[0]
        self._move("mig_out", "mig_in", [c[0] * rec, c[2] * rec], [i[0] * rec, i[2] * rec])
        self._move("ghost_out", "ghost_in", [c[1] * rec, c[3] * rec], [i[1] * rec, i[3] * rec])

    def exchange_halo(self, field):
        r, b = self.ranks[0], FIELD_BYTES[field]
        self._move("halo_out", "halo_in", [r.n_send[DOWN] * b, r.n_send[UP] * b], [r.n_recv[DOWN] * b, r.n_recv[UP] * b])


class PeerTransport:
    """Exchange through peer memory (NVLink / NVSwitch P2P, or plain device memory when the ranks share a GPU): the
    whole step, halo refreshes included, runs inside the library (sph_slab_step_begin / _end); nothing travels through
    Python, NCCL or the host. Two set-ups:
      PeerTransport.local(ranks)        all ranks in this process (tests on one GPU, a C++-style multi-GPU host);
      PeerTransport.distributed(rank)   one rank per process: the mailbox handles (CUDA IPC) are exchanged once through
                                        torch.distributed's object all-gather, then torch is out of the picture."""
    linked = True

    def __init__(self, ranks):
        self.ranks = ranks

    @classmethod
    def local(cls, ranks):
        for r in ranks:
            r.link_export()
        for i, r in enumerate(ranks):
            if r.has[DOWN]:
                r.link_local(DOWN, ranks[i - 1])
            if r.has[UP]:
                r.link_local(UP, ranks[i + 1])
        return cls(ranks)

    @classmethod
    def distributed(cls, rank_obj):
        """Returns the transport, or None (on EVERY rank) when some rank could not map a neighbour's mailbox — e.g. a
        process that sees only its own GPU: the caller then falls back to the NCCL protocol."""
        import torch.distributed as dist
        me, world = dist.get_rank(), dist.get_world_size()
        handles = [None] * world
        ok = True
        try:
            mine = rank_obj.link_export()
        except RuntimeError:
            mine, ok = None, False
        dist.all_gather_object(handles, mine)
        if ok and all(h is not None for h in handles):
            try:
                if rank_obj.has[DOWN]:
                    rank_obj.link_open(DOWN, handles[me - 1])
                if rank_obj.has[UP]:
                    rank_obj.link_open(UP, handles[me + 1])
            except RuntimeError:
                ok = False
        else:
            ok = False
        flags = [None] * world
        dist.all_gather_object(flags, ok)
        if not all(flags):
            return None
        return cls([rank_obj])

    def all_gather(self, rows):
        if len(self.ranks) > 1:
            return [list(x) for x in rows]
        import torch
        import torch.distributed as dist
        mine = torch.tensor(rows[0], dtype=torch.int64)
        out = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
        dist.all_gather(out, mine)
        return [t.tolist() for t in out]

    def step(self):
        for r in self.ranks:      # every rank delivers before any rank waits for its neighbours
            r.step_begin()
        for r in self.ranks:
            r.step_end()


def slab_step(transport):
    """One PCISPH step of every rank the transport drives (all of them for loopback, this process's for dist)."""
    ranks = transport.ranks
    if getattr(transport, "linked", False):
        transport.step()
        return

    def halo(field):
        for r in ranks:
            for side in (DOWN, UP):
                if r.has[side]:
                    r.pack(field, side)
        transport.exchange_halo(field)
        for r in ranks:
            for side in (DOWN, UP):
                if r.has[side]:
                    r.unpack(field, side)

    counts = [r.classify() for r in ranks]
    incoming = transport.exchange_counts(counts)
    transport.exchange_records(counts, incoming)
    for r, inc in zip(ranks, incoming):
        r.absorb(inc)
    for r in ranks:
        r.part(0)
    halo(RHO)
    for r in ranks:
        r.part(1)
    halo(PRS)
    for _ in range(1, ranks[0].iters):
        for r in ranks:
            r.part(2)
        halo(XPRED)
        for r in ranks:
            r.part(3)
        halo(PRS)
    for r in ranks:
        r.part(4)


def plan_cuts(cuts, owned):
    """New face positions from the current ones and the owned counts of all ranks: a face moves one layer towards
    the fuller of its two slabs when they differ by more than two average layers, and never leaves a slab with
    fewer than 2 layers. Pure function of its arguments: every rank computes the same plan."""
    new, n = list(cuts), [float(x) for x in owned]
    for b in range(1, len(cuts) - 1):
        per_layer = (n[b - 1] + n[b]) / max(1, new[b + 1] - new[b - 1])
        if n[b - 1] - n[b] > 2 * per_layer and new[b] - new[b - 1] > 2:
            new[b] -= 1
            n[b - 1] -= per_layer
            n[b] += per_layer
        elif n[b] - n[b - 1] > 2 * per_layer and new[b + 1] - new[b] > 2:
            new[b] += 1
            n[b] -= per_layer
            n[b - 1] += per_layer
    return new


def rebalance(transport):
    """Re-cut the slabs by particle count (call between steps, on every rank, at the same step). Returns True
    when a face moved; the layer that changed hands migrates in the next slab_step."""
    ranks = transport.ranks
    table = transport.all_gather([[r.z0, r.z1, r.n_owned()] for r in ranks])
    cuts = [row[0] for row in table] + [table[-1][1]]
    new = plan_cuts(cuts, [row[2] for row in table])
    for r in ranks:
        r.set_cut(new[r.index], new[r.index + 1])
    return new != cuts


def split_fluid(params, position, world):
    """Cut a fluid into `world` slabs, balanced by particle count: ([(z0, z1, indices of the owned particles)],
    caps) where caps = dict(halo_cap, mig_cap) are buffer sizes every rank must share (a neighbour's records
    land in this rank's buffers): twice the fullest cell layer."""
    cz = cell_layer(position[:, 2], params.hash_grid_cell_size_inv)
    gz = int(params.grid_cells_z)
    cuts = cut_layers(cz, world, gz)
    fullest = int(np.bincount(cz, minlength=gz).max())
    caps = dict(halo_cap=2 * fullest + 1024, mig_cap=fullest + 1024)
    return [(cuts[r], cuts[r + 1], np.nonzero((cz >= cuts[r]) & (cz < cuts[r + 1]))[0]) for r in range(world)], caps


def gather_owned(ranks, n_total):
    """Positions / velocities of all ranks, ordered by global id (loopback tests)."""
    pos, vel = np.zeros((n_total, 4), np.float32), np.zeros((n_total, 4), np.float32)
    seen = np.zeros(n_total, np.int32)
    for r in ranks:
        p, v, g = r.read_owned()
        pos[g], vel[g] = p, v
        seen[g] += 1
    if not (seen == 1).all():
        raise RuntimeError("ownership is not a partition: %d particles lost, %d duplicated" %
                           (int((seen == 0).sum()), int((seen > 1).sum())))
    return pos, vel
