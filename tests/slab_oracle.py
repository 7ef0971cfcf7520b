"""TEST INFRASTRUCTURE: one slab rank run by the CPU oracle through the plain stage API (sph_stage / sph_read /
sph_write), with the slab protocol of csrc/sph_slab.cu restated in numpy. It offers the interface of
slab.SlabRank, so slab.slab_step and slab.DistTransport (gloo) drive it unchanged: this is how the multi-rank
host logic is tested on CPU with world_size 2. The oracle sorts within a cell by local index, so the local
arrays are kept ordered by global id (what the CUDA library gets from ranking by global id)."""
import importlib

import numpy as np

import common

slab = importlib.import_module(common.PKG + ".slab")
DOWN, UP = slab.DOWN, slab.UP
REC = 48  # position, velocity, (gid, 0, 0, 0)

PART_STAGES = {
    0: [0, 1, 2, 3, 4, 5, 6, 7],
    1: [8, 9, 10, 11, 12],
    2: [13, 10],
    3: [11, 12],
    4: [13, 14, 15, 16, 17],
}


def _records(pos, vel, gid):
    r = np.zeros((len(gid), 12), np.float32)
    r[:, 0:4], r[:, 4:8] = pos, vel
    r[:, 8] = np.asarray(gid, np.uint32).view(np.float32)
    return r


def _unrecords(raw, n):
    r = np.frombuffer(raw[:n * REC].tobytes(), np.float32).reshape(n, 12)
    return r[:, 0:4].copy(), r[:, 4:8].copy(), r[:, 8].copy().view(np.uint32)


class OracleSlabRank:
    def __init__(self, lib, params, position, velocity, gid, z0, z1, index, world, alloc, cap=1 << 16):
        self.L, self.params = lib, params
        self.index, self.world = index, world
        self.has = [index > 0, index < world - 1]
        self.z0, self.z1 = z0, z1
        self.pos, self.vel, self.gid = position.copy(), velocity.copy(), np.asarray(gid, np.uint32).copy()
        self.iters = max(1, int(params.max_iteration))
        self.rec = REC
        self.buf = {}
        for side in (DOWN, UP):
            for kind in ("mig_out", "mig_in", "ghost_out", "ghost_in", "halo_out", "halo_in"):
                self.buf[kind, side] = alloc(cap * REC)
        self.counts, self.n_send, self.n_recv = [0] * 4, [0, 0], [0, 0]
        self.ctx = None

    def _np(self, kind, side):
        return self.buf[kind, side].tensor.numpy()

    def _put(self, kind, side, rec):
        raw = rec.view(np.uint8).reshape(-1)
        self._np(kind, side)[:len(raw)] = raw

    def set_cut(self, z0, z1):
        self.z0, self.z1 = int(z0), int(z1)

    def n_owned(self):
        return len(self.gid)

    def classify(self):
        cz = slab.cell_layer(self.pos[:, 2], self.params.hash_grid_cell_size_inv)
        down = (cz < self.z0) if self.has[DOWN] else np.zeros(len(cz), bool)
        up = (cz >= self.z1) if self.has[UP] else np.zeros(len(cz), bool)
        keep = ~(down | up)
        gd = keep & (cz == self.z0) if self.has[DOWN] else np.zeros(len(cz), bool)
        gu = keep & (cz == self.z1 - 1) if self.has[UP] else np.zeros(len(cz), bool)
        self._sel = dict(keep=keep, down=down, up=up, gd=gd, gu=gu)
        for kind, side, m in (("mig_out", DOWN, down), ("mig_out", UP, up), ("ghost_out", DOWN, gd), ("ghost_out", UP, gu)):
            self._put(kind, side, _records(self.pos[m], self.vel[m], self.gid[m]))
        self.counts = [int(down.sum()), int(gd.sum()), int(up.sum()), int(gu.sum())]
        return self.counts

    def absorb(self, incoming):
        s = self._sel
        parts = [(self.pos[s["keep"]], self.vel[s["keep"]], self.gid[s["keep"]])]
        parts.append(_unrecords(self._np("mig_in", DOWN), incoming[0]))
        parts.append(_unrecords(self._np("mig_in", UP), incoming[2]))
        n_own = sum(len(p[2]) for p in parts)
        parts.append(_unrecords(self._np("ghost_in", DOWN), incoming[1]))
        parts.append((self.pos[s["down"]], self.vel[s["down"]], self.gid[s["down"]]))
        parts.append(_unrecords(self._np("ghost_in", UP), incoming[3]))
        parts.append((self.pos[s["up"]], self.vel[s["up"]], self.gid[s["up"]]))
        pos = np.concatenate([p[0] for p in parts])
        vel = np.concatenate([p[1] for p in parts])
        gid = np.concatenate([p[2] for p in parts])
        off = np.cumsum([0] + [len(p[2]) for p in parts])
        # what the neighbours hold as ghosts, in the agreed order: kept boundary layer, then the migrants they sent
        kept_gid = parts[0][2]
        send_gid = [np.concatenate([self.gid[s["gd"]], parts[1][2]]), np.concatenate([self.gid[s["gu"]], parts[2][2]])]
        recv_gid = [gid[off[3]:off[5]], gid[off[5]:off[7]]]
        order = np.argsort(gid, kind="stable")                      # local arrays ordered by global id
        where = np.empty(len(gid), np.int64)
        where[order] = np.arange(len(gid))
        lookup = dict(zip(gid.tolist(), where.tolist()))
        self.send_idx = [np.array([lookup[g] for g in sg.tolist()], np.int64) for sg in send_gid]
        self.recv_idx = [np.array([lookup[g] for g in rg.tolist()], np.int64) for rg in recv_gid]
        self.n_send, self.n_recv = [len(x) for x in self.send_idx], [len(x) for x in self.recv_idx]
        self.lpos, self.lvel, self.lgid = pos[order], vel[order], gid[order]
        self.own_local = where[:n_own]
        self.n_own = n_own
        assert len(kept_gid) + incoming[0] + incoming[2] == n_own
        p = type(self.params).from_buffer_copy(self.params)
        p.particle_count = len(gid)
        if self.ctx is not None:
            self.ctx.close()
        self.ctx = self.L.create(p, self.lpos, self.lvel)
        self.N = len(gid)

    def part(self, k):
        for st in PART_STAGES[k]:
            self.ctx.stage(st)
        if k == 4:
            pos, vel = self.ctx.read("position", self.N), self.ctx.read("velocity", self.N)
            self.pos, self.vel, self.gid = pos[self.own_local], vel[self.own_local], self.lgid[self.own_local]

    def _field(self, field):
        back = self.ctx.read("particleIndexBack").astype(np.int64)
        if field == slab.RHO:
            return back, "rho", None
        if field == slab.PRS:
            return back, "pressure", "rho"
        return back, "sortedPosition", None

    def pack(self, field, side):
        back, a, b = self._field(field)
        sid = back[self.send_idx[side]]
        if field == slab.RHO:
            out = self.ctx.read("rho")[sid]
        elif field == slab.PRS:
            out = np.stack([self.ctx.read("pressure")[sid], self.ctx.read("rho")[self.N + sid]], 1)
        else:
            out = self.ctx.read("sortedPosition")[self.N + sid]
        self._put("halo_out", side, np.ascontiguousarray(out, np.float32))

    def unpack(self, field, side):
        back, a, b = self._field(field)
        sid = back[self.recv_idx[side]]
        n = len(sid)
        raw = self._np("halo_in", side)
        width = slab.FIELD_BYTES[field] // 4
        vals = np.frombuffer(raw[:n * width * 4].tobytes(), np.float32).reshape(n, width)
        if field == slab.RHO:
            rho = self.ctx.read("rho"); rho[sid] = vals[:, 0]; self.ctx.write("rho", rho)
        elif field == slab.PRS:
            pr = self.ctx.read("pressure"); pr[sid] = vals[:, 0]; self.ctx.write("pressure", pr)
            rho = self.ctx.read("rho"); rho[self.N + sid] = vals[:, 1]; self.ctx.write("rho", rho)
        else:
            sp = self.ctx.read("sortedPosition"); sp[self.N + sid] = vals; self.ctx.write("sortedPosition", sp)

    def sync(self):
        pass

    def copy(self, dst, src, nbytes):
        dst.tensor[:nbytes] = src.tensor[:nbytes]

    def read_owned(self):
        return self.pos, self.vel, self.gid

    def info(self):
        return [len(self.gid), getattr(self, "N", 0)] + self.n_send + self.n_recv

    def close(self):
        if self.ctx is not None:
            self.ctx.close()
            self.ctx = None
