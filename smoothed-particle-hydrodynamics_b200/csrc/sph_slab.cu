// sph_slab.cu — slab decomposition of one fluid over several devices. The reference is
// single-device (owOpenCLSolver.cpp:157-182: one context, one queue, one device); this file is an addition.
//
// A rank owns the particles whose hash-cell z-layer cz = (int)(z * hashGridCellSizeInv) (sphFluid.cl:361-365)
// lies in [z0, z1), plus GHOST copies of the neighbours' boundary layers z0-1 and z1. Every rank runs the
// ordinary kernels over owned + ghost particles on the GLOBAL hash grid (same cell ids, same searchCell wrap),
// and values a ghost cannot compute itself are refreshed from its owner at the three points of the step where
// an owned particle gathers them: rho(t) before the forces, (p, rho*) before each pressure force, x* before each
// density prediction after the first. Within a cell particles are ranked by their GLOBAL id, so sorted order,
// neighbour order and the keep-first-32 truncation are those of the single-device run: results are
// bit-identical per global id (tests/test_slab_gpu.py).
//
// The library packs and unpacks; moving bytes between ranks is the caller's business (torch.distributed /
// NCCL in slab.py, device-to-device copies in the single-process tests).
#include <stdlib.h>
#include <string.h>

#include "sph_ctx.h"

struct SlabRec { float4 pos, vel, tag; };  // tag.x = global id (bit pattern)

struct SlabState {
  int z0, z1, hasDown, hasUp;
  int cap, migCap, haloCap;
  int nOwn;
  unsigned *gid, *gidAlt;
  float4 *posAlt, *velAlt;
  int *counters, *hostCounters;   // 8 ints: migD, ghostD, migU, ghostU, keep, flags
  int *hostCountersDev;           // device-side address of the pinned hostCounters
  int *sendList[2];               // local original indices of the particles a neighbour holds as ghosts
  int migOut[2], ghostOut[2];
  const SlabRec *myMig[2];
  int nSend[2], nRecv[2], recvStart[2];
  int recut;                      // the faces moved since the last classify (sph_slab_set_cut)
  // ---- peer-memory exchange (sph_slab_link_* / sph_slab_step): see the "links" section below
  char *mail;                     // this rank's mailbox (device memory; neighbours write into it)
  char *peer[2];                  // the neighbours' mailboxes as mapped here: [0] = rank-1, [1] = rank+1
  bool peerIpc[2];                // opened with cudaIpcOpenMemHandle (to be closed)
  SlabRec *migLocal[2];           // local copies of the outgoing migrants (they stay here as ghosts)
  unsigned epoch;                 // step counter = value of the delivery flags of this step
  int xchg;                       // halo exchanges done in the current step
  int *doneCounter;               // last-block detection of the send kernels
  bool stepOpen;                  // between sph_slab_step_begin and _end
  float *xyzStage;                // packed (x, y, z) of the owned particles for sph_slab_read_owned_xyz_begin
};

struct SlabGeom { int z0, z1, hasDown, hasUp, nOwn, migCap, haloCap, slack; };

__global__ void __launch_bounds__(256) k_slab_classify(DevParams d, SlabGeom s, const float4 *__restrict__ pos,
                                                       const float4 *__restrict__ vel, const unsigned *__restrict__ gid,
                                                       float4 *__restrict__ posN, float4 *__restrict__ velN,
                                                       unsigned *__restrict__ gidN, SlabRec *__restrict__ migD,
                                                       SlabRec *__restrict__ ghD, SlabRec *__restrict__ migU,
                                                       SlabRec *__restrict__ ghU, int *__restrict__ listD,
                                                       int *__restrict__ listU, int *__restrict__ counters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.nOwn) return;
  const float4 p = pos[i], v = vel[i];
  const unsigned g = gid[i];
  const int cz = (int)(p.z * d.cellSizeInv);  // the z factor of hashParticles' cell id, sphFluid.cl:361-365
  SlabRec r;
  r.pos = p; r.vel = v; r.tag = make_float4(__uint_as_float(g), 0.f, 0.f, 0.f);
  // moved more than one layer (two right after a re-cut, which hands a whole layer to the neighbour)
  if ((s.hasDown && cz < s.z0 - 1 - s.slack) || (s.hasUp && cz > s.z1 + s.slack)) atomicOr(&counters[5], 1);
  if (s.hasDown && cz < s.z0) {
    const int slot = atomicAdd(&counters[0], 1);
    if (slot < s.migCap) migD[slot] = r; else atomicOr(&counters[5], 2);
  } else if (s.hasUp && cz >= s.z1) {
    const int slot = atomicAdd(&counters[2], 1);
    if (slot < s.migCap) migU[slot] = r; else atomicOr(&counters[5], 2);
  } else {
    const int ni = atomicAdd(&counters[4], 1);
    posN[ni] = p; velN[ni] = v; gidN[ni] = g;
    if (s.hasDown && cz == s.z0) {
      const int slot = atomicAdd(&counters[1], 1);
      if (slot < s.haloCap) { ghD[slot] = r; listD[slot] = ni; } else atomicOr(&counters[5], 4);
    }
    if (s.hasUp && cz == s.z1 - 1) {
      const int slot = atomicAdd(&counters[3], 1);
      if (slot < s.haloCap) { ghU[slot] = r; listU[slot] = ni; } else atomicOr(&counters[5], 4);
    }
  }
}

// Append up to 6 record arrays to the local particle arrays, back to back from `base`.
struct SlabAppend { const SlabRec *src[6]; int n[6]; int base; };

__global__ void __launch_bounds__(256) k_slab_append(SlabAppend a, float4 *__restrict__ pos, float4 *__restrict__ vel,
                                                     unsigned *__restrict__ gid) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int off = 0;
#pragma unroll
  for (int q = 0; q < 6; q++) {
    if (i >= off && i < off + a.n[q]) {
      const SlabRec r = a.src[q][i - off];
      pos[a.base + i] = r.pos; vel[a.base + i] = r.vel; gid[a.base + i] = __float_as_uint(r.tag.x);
    }
    off += a.n[q];
  }
}

__global__ void k_slab_publish(const int *__restrict__ counters, volatile int *hostCounters) {
  if (threadIdx.x < 8) hostCounters[threadIdx.x] = counters[threadIdx.x];
  __threadfence_system();
}

__global__ void __launch_bounds__(256) k_slab_iota(int *__restrict__ dst, int n, int first) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = first + i;
}

// field: 0 = rho(t) (float), 1 = (pressure, rho*) (float2), 2 = predicted position (float4)
template <int FIELD, bool PACK>
__global__ void __launch_bounds__(256) k_slab_halo(int n, const int *__restrict__ list, int first,
                                                   const unsigned *__restrict__ back, float *__restrict__ rho,
                                                   float4 *__restrict__ recA, float4 *__restrict__ recB,
                                                   float4 *__restrict__ spPred, float *__restrict__ rhoPred,
                                                   void *__restrict__ buf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned sid = back[list ? list[i] : first + i];
  if (FIELD == 0) {  // rho(t): the array and the rho slot of the gather record
    float *b = (float *)buf;
    if (PACK) b[i] = rho[sid];
    else { rho[sid] = b[i]; reinterpret_cast<float *>(recA)[8 * (size_t)sid + 7] = b[i]; }
  }
  if (FIELD == 1) {  // (p, rho*): second half of the pressure-force record
    float2 *b = (float2 *)buf;
    float *q = reinterpret_cast<float *>(recB + sid) + 3;   // pressure slot of the 16-byte record; rho* is rho[N + sid]
    if (PACK) b[i] = make_float2(*q, rhoPred[sid]); else { *q = b[i].x; rhoPred[sid] = b[i].y; }
  }
  if (FIELD == 2) { float4 *b = (float4 *)buf; if (PACK) b[i] = spPred[sid]; else spPred[sid] = b[i]; }
}

static thread_local char g_slab_err[256];
static int slab_fail(const char *m) { strncpy(g_slab_err, m, sizeof g_slab_err - 1); return SPH_E_INVALID; }
extern "C" const char *sph_slab_last_error(void) { return g_slab_err; }

#define SCU(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { slab_fail(cudaGetErrorString(_e)); return SPH_E_DEVICE; } } while (0)

static inline int nblk(int n) { return n > 0 ? (n + 255) / 256 : 1; }

extern "C" {

int sph_slab_record_bytes(void) { return (int)sizeof(SlabRec); }

// Turn a context created with particle_count = CAPACITY into a slab rank. The first n_owned particles given to
// sph_create are the rank's own; gid[n_owned] are their global ids. Cell layers [z0, z1) belong to this rank.
int sph_slab_begin(sph_ctx *c, int n_owned, const uint32_t *gid, int z0, int z1, int has_down, int has_up,
                   int mig_cap, int halo_cap) {
  if (!c || c->slab) return slab_fail("sph_slab_begin: bad context");
  if (n_owned < 0 || n_owned > c->p.particle_count || z1 <= z0 || !gid) return slab_fail("sph_slab_begin: bad arguments");
  if (c->d.mask != 0xffffffffu) return slab_fail("sph_slab_begin: slabs need the wide hash mask (SURVEY 8e)");
  SCU(cudaSetDevice(c->device));
  SlabState *s = new SlabState();
  memset(s, 0, sizeof *s);
  s->z0 = z0; s->z1 = z1; s->hasDown = has_down; s->hasUp = has_up;
  s->cap = c->p.particle_count; s->migCap = mig_cap; s->haloCap = halo_cap; s->nOwn = n_owned;
  const size_t cap = (size_t)s->cap;
  SCU(cudaMalloc(&s->gid, 4 * cap)); SCU(cudaMalloc(&s->gidAlt, 4 * cap));
  SCU(cudaMalloc(&s->posAlt, 32 * cap)); SCU(cudaMalloc(&s->velAlt, 32 * cap));
  SCU(cudaMemset(s->posAlt, 0, 32 * cap)); SCU(cudaMemset(s->velAlt, 0, 32 * cap));
  SCU(cudaMalloc(&c->tmpRankKey, 4 * cap));
  SCU(cudaMalloc(&s->counters, 32)); SCU(cudaMallocHost(&s->hostCounters, 64));
  memset(s->hostCounters, 0, 64);
  SCU(cudaHostGetDevicePointer((void **)&s->hostCountersDev, s->hostCounters, 0));
  SCU(cudaMalloc(&s->sendList[0], 4 * (size_t)(halo_cap + mig_cap)));
  SCU(cudaMalloc(&s->sendList[1], 4 * (size_t)(halo_cap + mig_cap)));
  SCU(cudaMemcpy(s->gid, gid, 4 * (size_t)n_owned, cudaMemcpyHostToDevice));
  c->slab = s;
  c->rankKey = s->gid;
  return 0;
}

// Step phase 1: sort the owned particles into keep / migrate-down / migrate-up, and copy the kept particles of
// the two boundary layers (the neighbours' ghosts). Records go to caller-provided device buffers (mig_cap /
// halo_cap records each). counts = {mig_down, ghost_down, mig_up, ghost_up}. Synchronises (counts are needed on the host).
int sph_slab_classify(sph_ctx *c, void *mig_down, void *ghost_down, void *mig_up, void *ghost_up, int counts[4]) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  SlabState *s = c->slab;
  SCU(cudaSetDevice(c->device));
  SCU(cudaMemsetAsync(s->counters, 0, 32, c->stream));
  if (c->readPending && c->readSrc == (const void *)s->posAlt) read_fence(c);  // a pipelined read still copies from the buffer rebuilt now
  SlabGeom g = {s->z0, s->z1, s->hasDown, s->hasUp, s->nOwn, s->migCap, s->haloCap, s->recut};
  s->recut = 0;
  k_slab_classify<<<nblk(s->nOwn), 256, 0, c->stream>>>(c->d, g, c->position, c->velocity, s->gid, s->posAlt, s->velAlt,
                                                        s->gidAlt, (SlabRec *)mig_down, (SlabRec *)ghost_down,
                                                        (SlabRec *)mig_up, (SlabRec *)ghost_up, s->sendList[0],
                                                        s->sendList[1], s->counters);
  // the counts go to pinned host memory through a store, not a copy: a device-to-host memcpy would queue on the DMA
  // engine behind a pipelined read-back of 130 MB (sph_slab_read_owned_begin) and stall every step by its duration
  k_slab_publish<<<1, 32, 0, c->stream>>>(s->counters, s->hostCountersDev);
  SCU(cudaStreamSynchronize(c->stream));
  const int *h = s->hostCounters;
  if (h[5]) return slab_fail(h[5] & 1 ? "a particle crossed more than one cell layer in one step"
                                       : "migration / halo buffer capacity exceeded");
  s->migOut[0] = h[0]; s->ghostOut[0] = h[1]; s->migOut[1] = h[2]; s->ghostOut[1] = h[3];
  s->myMig[0] = (const SlabRec *)mig_down; s->myMig[1] = (const SlabRec *)mig_up;
  for (int q = 0; q < 4; q++) counts[q] = h[q];
  return 0;
}

// Step phase 2: incoming = {mig_from_down, ghost_from_down, mig_from_up, ghost_from_up} record counts in the four
// received buffers. New owned set = kept + incoming migrants; ghosts = received ghosts + the rank's own outgoing
// migrants (one layer per step: they landed in the neighbour's boundary layer). Asynchronous.
int sph_slab_absorb(sph_ctx *c, const void *mig_from_down, const void *ghost_from_down, const void *mig_from_up,
                    const void *ghost_from_up, const int incoming[4]) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  SlabState *s = c->slab;
  SCU(cudaSetDevice(c->device));
  const int keep = s->hostCounters[4];
  const int inD = incoming[0], ghD = incoming[1], inU = incoming[2], ghU = incoming[3];
  const int nOwn = keep + inD + inU;
  const int nGhost = ghD + s->migOut[0] + ghU + s->migOut[1];
  if (nOwn + nGhost > s->cap) return slab_fail("slab capacity exceeded");
  if (s->ghostOut[0] + inD > s->haloCap + s->migCap || s->ghostOut[1] + inU > s->haloCap + s->migCap)
    return slab_fail("halo list capacity exceeded");
  SlabAppend a;
  a.src[0] = (const SlabRec *)mig_from_down; a.n[0] = inD;
  a.src[1] = (const SlabRec *)mig_from_up; a.n[1] = inU;
  a.src[2] = (const SlabRec *)ghost_from_down; a.n[2] = ghD;
  a.src[3] = s->myMig[0]; a.n[3] = s->migOut[0];
  a.src[4] = (const SlabRec *)ghost_from_up; a.n[4] = ghU;
  a.src[5] = s->myMig[1]; a.n[5] = s->migOut[1];
  a.base = keep;
  const int total = inD + inU + nGhost;
  if (total > 0) k_slab_append<<<nblk(total), 256, 0, c->stream>>>(a, s->posAlt, s->velAlt, s->gidAlt);
  if (inD > 0) k_slab_iota<<<nblk(inD), 256, 0, c->stream>>>(s->sendList[0] + s->ghostOut[0], inD, keep);
  if (inU > 0) k_slab_iota<<<nblk(inU), 256, 0, c->stream>>>(s->sendList[1] + s->ghostOut[1], inU, keep + inD);
  // the compacted arrays become the live ones
  float4 *t4; unsigned *tu;
  t4 = c->position; c->position = s->posAlt; s->posAlt = t4;
  t4 = c->velocity; c->velocity = s->velAlt; s->velAlt = t4;
  tu = s->gid; s->gid = s->gidAlt; s->gidAlt = tu;
  c->rankKey = s->gid;
  s->nOwn = nOwn;
  s->nSend[0] = s->ghostOut[0] + inD; s->nSend[1] = s->ghostOut[1] + inU;
  s->recvStart[0] = nOwn; s->nRecv[0] = ghD + s->migOut[0];
  s->recvStart[1] = nOwn + s->nRecv[0]; s->nRecv[1] = ghU + s->migOut[1];
  c->d.N = nOwn + nGhost;
  c->d.Npad = ((c->d.N + 31) / 32) * 32;
  SCU(cudaGetLastError());
  return 0;
}

// out = {n_owned, n_local, n_send_down, n_send_up, n_recv_down, n_recv_up}
int sph_slab_info(const sph_ctx *c, int out[6]) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  const SlabState *s = c->slab;
  out[0] = s->nOwn; out[1] = c->d.N; out[2] = s->nSend[0]; out[3] = s->nSend[1]; out[4] = s->nRecv[0]; out[5] = s->nRecv[1];
  return 0;
}

static int halo(sph_ctx *c, int field, int side, void *buf, bool pack) {
  if (!c || !c->slab || side < 0 || side > 1 || field < 0 || field > 2) return slab_fail("sph_slab halo: bad arguments");
  SlabState *s = c->slab;
  SCU(cudaSetDevice(c->device));
  const int n = pack ? s->nSend[side] : s->nRecv[side];
  if (n <= 0) return 0;
  const int *list = pack ? s->sendList[side] : nullptr;
  const int first = s->recvStart[side];
  float4 *pred = c->sortedPosition + c->d.N;
#define H(F, P) k_slab_halo<F, P><<<nblk(n), 256, 0, c->stream>>>(n, list, first, c->particleIndexBack, c->rho, c->recA, c->recB, pred, c->rho + c->d.N, buf)
  if (pack) { if (field == 0) H(0, true); else if (field == 1) H(1, true); else H(2, true); }
  else { if (field == 0) H(0, false); else if (field == 1) H(1, false); else H(2, false); }
#undef H
  SCU(cudaGetLastError());
  return 0;
}

// Halo refresh: gather `field` of the particles the neighbour on `side` (0 = down, 1 = up) holds as ghosts into
// buf (n_send[side] elements of 4 / 8 / 16 bytes), or scatter a received buffer into the local ghosts. Asynchronous.
int sph_slab_pack(sph_ctx *c, int field, int side, void *buf) { return halo(c, field, side, buf, true); }
int sph_slab_unpack(sph_ctx *c, int field, int side, const void *buf) { return halo(c, field, side, (void *)buf, false); }

// The owned particles (any order) with their global ids; synchronises.
int sph_slab_read_owned(sph_ctx *c, float *pos4, float *vel4, uint32_t *gid) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  SlabState *s = c->slab;
  SCU(cudaSetDevice(c->device));
  SCU(cudaStreamSynchronize(c->stream));
  const size_t n = (size_t)s->nOwn;
  if (pos4) SCU(cudaMemcpy(pos4, c->position, 16 * n, cudaMemcpyDeviceToHost));
  if (vel4) SCU(cudaMemcpy(vel4, c->velocity, 16 * n, cudaMemcpyDeviceToHost));
  if (gid) SCU(cudaMemcpy(gid, s->gid, 4 * n, cudaMemcpyDeviceToHost));
  return 0;
}

// Move the slab's faces (load balancing): the next classify hands the layers that fell outside to the neighbours as
// ordinary migrants. One layer per face and call; both ranks that share a face must move it before the same step.
int sph_slab_set_cut(sph_ctx *c, int z0, int z1) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  SlabState *s = c->slab;
  if (z1 - z0 < 2) return slab_fail("sph_slab_set_cut: a slab keeps at least 2 cell layers");
  if (abs(z0 - s->z0) > 1 || abs(z1 - s->z1) > 1) return slab_fail("sph_slab_set_cut: a face moves one layer at a time");
  if ((!s->hasDown && z0 != s->z0) || (!s->hasUp && z1 != s->z1)) return slab_fail("sph_slab_set_cut: outer faces are fixed");
  if (z0 != s->z0 || z1 != s->z1) s->recut = 1;
  s->z0 = z0;
  s->z1 = z1;
  return 0;
}

// Pipelined variant for the positions: the copy of this step's owned particles overlaps the next step (it reads the
// buffer that the next absorb only reads and the one after rebuilds). Finish with sph_read_end.
int sph_slab_read_owned_begin(sph_ctx *c, float *pos4) {
  if (!c || !c->slab || !pos4) return slab_fail("sph_slab_read_owned_begin: not a slab context / null");
  if (c->readPending) return slab_fail("sph_slab_read_owned_begin: a read is already in flight");
  SCU(cudaSetDevice(c->device));
  if (!c->copyStream) {
    SCU(cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking));
    SCU(cudaEventCreateWithFlags(&c->evReadReady, cudaEventDisableTiming));
    SCU(cudaEventCreateWithFlags(&c->evReadDone, cudaEventDisableTiming));
  }
  SCU(cudaEventRecord(c->evReadReady, c->stream));
  SCU(cudaStreamWaitEvent(c->copyStream, c->evReadReady, 0));
  SCU(cudaMemcpyAsync(pos4, c->position, 16 * (size_t)c->slab->nOwn, cudaMemcpyDeviceToHost, c->copyStream));
  SCU(cudaEventRecord(c->evReadDone, c->copyStream));
  c->readPending = 1;
  c->readSrc = c->position;
  return 0;
}

}  // extern "C"

// (x, y, z) of the owned particles, packed: the type tag in .w never changes and a quarter of the read-back is saved
__global__ void __launch_bounds__(256) k_slab_pack_xyz(int n, const float4 *__restrict__ pos, float *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = pos[i];
  out[3 * (size_t)i] = p.x; out[3 * (size_t)i + 1] = p.y; out[3 * (size_t)i + 2] = p.z;
}

extern "C" {
// Pipelined read-back of the owned POSITIONS as packed (x, y, z) — 12 bytes per particle instead of 16: with eight
// ranks reading back every step the host side is the bottleneck (8 x 131 MB per step), and the type tag is constant.
// The order is the rank's local order; it changes only in steps that migrate particles (sph_slab_counts), and
// sph_slab_read_owned(ctx, NULL, NULL, gid) gives the ids when it did. Finish with sph_read_end.
int sph_slab_read_owned_xyz_begin(sph_ctx *c, float *xyz3) {
  if (!c || !c->slab || !xyz3) return slab_fail("sph_slab_read_owned_xyz_begin: not a slab context / null");
  if (c->readPending) return slab_fail("sph_slab_read_owned_xyz_begin: a read is already in flight");
  SlabState *s = c->slab;
  SCU(cudaSetDevice(c->device));
  if (!c->copyStream) {
    SCU(cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking));
    SCU(cudaEventCreateWithFlags(&c->evReadReady, cudaEventDisableTiming));
    SCU(cudaEventCreateWithFlags(&c->evReadDone, cudaEventDisableTiming));
  }
  if (!s->xyzStage) SCU(cudaMalloc((void **)&s->xyzStage, 12 * (size_t)s->cap));
  // the staging buffer is free again: the previous read was completed by sph_read_end (readPending == 0)
  k_slab_pack_xyz<<<nblk(s->nOwn), 256, 0, c->stream>>>(s->nOwn, c->position, s->xyzStage);
  SCU(cudaEventRecord(c->evReadReady, c->stream));
  SCU(cudaStreamWaitEvent(c->copyStream, c->evReadReady, 0));
  SCU(cudaMemcpyAsync(xyz3, s->xyzStage, 12 * (size_t)s->nOwn, cudaMemcpyDeviceToHost, c->copyStream));
  SCU(cudaEventRecord(c->evReadDone, c->copyStream));
  c->readPending = 1;
  c->readSrc = s->xyzStage;   // not position: the next step may rewrite the particle arrays while the copy runs
  c->readStaged = true;
  return 0;
}
}  // extern "C"

void slab_destroy(sph_ctx *c) {
  SlabState *s = c->slab;
  if (!s) return;
  // position / velocity may have been swapped with the Alt buffers: both pairs are freed exactly once
  // (sph_destroy frees c->position / c->velocity, this frees the Alt side).
  cudaFree(s->gid); cudaFree(s->gidAlt); cudaFree(s->posAlt); cudaFree(s->velAlt);
  cudaFree(s->counters); cudaFreeHost(s->hostCounters); cudaFree(s->sendList[0]); cudaFree(s->sendList[1]);
  for (int q = 0; q < 2; q++) {
    if (s->peer[q] && s->peerIpc[q]) cudaIpcCloseMemHandle(s->peer[q]);
    if (s->migLocal[q]) cudaFree(s->migLocal[q]);
  }
  if (s->mail) cudaFree(s->mail);
  if (s->xyzStage) cudaFree(s->xyzStage);
  if (s->doneCounter) cudaFree(s->doneCounter);
  delete s;
  c->slab = nullptr;
}

// =====================================================================================================
// Links: halo exchange through PEER MEMORY (NVLink / NVSwitch P2P), no NCCL call and no host round trip on the
// data path. Every rank owns a MAILBOX in device memory; its two neighbours map it (CUDA IPC between processes,
// peer access inside one process) and their kernels store records, halo values and delivery flags straight into
// it. A consumer kernel spins on the flag of the delivery it needs. One host synchronisation per step remains:
// the particle counts after migration size the rank's launches.
//
// Mailbox = header | per origin side f (0 = written by rank-1, 1 = written by rank+1):
//           migrant records [migCap] | ghost records [haloCap] | halo values [2 parities][(haloCap+migCap) x 16 B]
// Deliveries of a step: slot 0 = records + counts, slot 1 + e = halo exchange e (rho | (p,rho*) | { x* | (p,rho*) }).
// flag[f][slot] = epoch of the last complete delivery; epochs only grow, nothing is ever reset.
// Buffer reuse is safe without a handshake: a rank writes delivery e + 2 (same parity as e) only after it has
// received its neighbour's delivery e + 1, which the neighbour sent after consuming delivery e (stream order);
// the record buffers of step k + 1 are written after the neighbour's first delivery of step k + 1 is awaited,
// i.e. after the neighbour consumed everything of step k.
#define SLAB_SLOTS 16
struct SlabMailHeader {
  unsigned flag[2][SLAB_SLOTS];   // [origin side][1 + e] = epoch of the last complete halo delivery e
  // [origin side][0 = migrants, 1 = ghosts] record delivery of a step: epoch << 32 | count in ONE 64-bit word each,
  // so that a count and its flag arrive together and no fence is needed between them
  unsigned long long records[2][2];
  unsigned pad[24];
};
static_assert(sizeof(SlabMailHeader) == 256, "mailbox header");

struct MailLayout { size_t mig[2], ghost[2], halo[2][2], total; };
static MailLayout mail_layout(int migCap, int haloCap) {
  MailLayout m;
  size_t o = sizeof(SlabMailHeader);
  auto take = [&](size_t bytes) { size_t at = o; o += (bytes + 255) & ~(size_t)255; return at; };
  for (int f = 0; f < 2; f++) {
    m.mig[f] = take(sizeof(SlabRec) * (size_t)migCap);
    m.ghost[f] = take(sizeof(SlabRec) * (size_t)haloCap);
    for (int q = 0; q < 2; q++) m.halo[f][q] = take(16 * (size_t)(haloCap + migCap));
  }
  m.total = o;
  return m;
}

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Wait until *flag has reached `epoch` (wrap-safe), with acquire semantics at system scope: the data the neighbour
// wrote before it raised the flag is visible to the loads that follow. No __threadfence_system() anywhere on this
// path: a system-scope fence waits for the device's outstanding PCIe traffic too, and with a 130 MB device-to-host
// read-back in flight (the pipelined e2e loop) every fence cost ~70 us — 2.1 ms per step when each of the classify
// kernel's 32 k blocks fenced (tools/fence_probe.py). Gives up after 10 s and raises the sticky error word *err: a
// lost neighbour must not hang the GPU.
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void wait_delivery(const unsigned *flag, unsigned epoch, int *err) {
  const unsigned long long t0 = global_ns();
  while ((int)(ld_acquire_sys(flag) - epoch) < 0) {
    __nanosleep(200);
    if (global_ns() - t0 > 10000000000ull) { atomicOr(err, 256); break; }
  }
}

// classify of the linked step: k_slab_classify writing into the neighbours' mailboxes. The delivery is announced by
// the NEXT kernel of the stream (k_slab_records): a kernel boundary orders all of this kernel's writes — peer memory
// included — before anything the next kernel does, so no fence is needed here.
struct SlabLinkOut {
  SlabRec *mig[2], *ghost[2], *migLocal[2];  // [0] = towards rank-1, [1] = towards rank+1
  SlabMailHeader *hdr[2];
  unsigned epoch;
};

__global__ void __launch_bounds__(256) k_slab_classify_linked(DevParams d, SlabGeom s, SlabLinkOut o,
                                                              const float4 *__restrict__ pos, const float4 *__restrict__ vel,
                                                              const unsigned *__restrict__ gid, float4 *__restrict__ posN,
                                                              float4 *__restrict__ velN, unsigned *__restrict__ gidN,
                                                              int *__restrict__ listD, int *__restrict__ listU,
                                                              int *counters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  // Kept particles are compacted in BLOCK ORDER (one atomic per block, ranks inside the block by ballot): the arrays
  // keep the locality of the previous step's order, which the sort's gathers and histogram atomics live on (a
  // per-thread atomic scrambled it: sort 0.48 -> 0.62 ms, neighbour search + 0.05 ms at 8 M particles).
  __shared__ int warpKept[8], blockBase;
  float4 p = f4(0.f, 0.f, 0.f, 0.f), v = p;
  unsigned g = 0u;
  int cz = 0;
  bool keep = false;
  if (i < s.nOwn) {
    p = pos[i]; v = vel[i]; g = gid[i];
    cz = (int)(p.z * d.cellSizeInv);
    keep = !((s.hasDown && cz < s.z0) || (s.hasUp && cz >= s.z1));
  }
  const unsigned kmask = __ballot_sync(0xffffffffu, keep);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warpKept[warp] = __popc(kmask);
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int w = 0; w < 8; w++) { const int c = warpKept[w]; warpKept[w] = tot; tot += c; }
    blockBase = tot ? atomicAdd(&counters[4], tot) : 0;
  }
  __syncthreads();
  if (i < s.nOwn) {
    SlabRec r;
    r.pos = p; r.vel = v; r.tag = make_float4(__uint_as_float(g), 0.f, 0.f, 0.f);
    if ((s.hasDown && cz < s.z0 - 1 - s.slack) || (s.hasUp && cz > s.z1 + s.slack)) atomicOr(&counters[5], 1);
    if (s.hasDown && cz < s.z0) {
      const int slot = atomicAdd(&counters[0], 1);
      if (slot < s.migCap) { o.mig[0][slot] = r; o.migLocal[0][slot] = r; } else atomicOr(&counters[5], 2);
    } else if (s.hasUp && cz >= s.z1) {
      const int slot = atomicAdd(&counters[2], 1);
      if (slot < s.migCap) { o.mig[1][slot] = r; o.migLocal[1][slot] = r; } else atomicOr(&counters[5], 2);
    } else {
      const int ni = blockBase + warpKept[warp] + __popc(kmask & ((1u << lane) - 1u));
      posN[ni] = p; velN[ni] = v; gidN[ni] = g;
      if (s.hasDown && cz == s.z0) {
        const int slot = atomicAdd(&counters[1], 1);
        if (slot < s.haloCap) { o.ghost[0][slot] = r; listD[slot] = ni; } else atomicOr(&counters[5], 4);
      }
      if (s.hasUp && cz == s.z1 - 1) {
        const int slot = atomicAdd(&counters[3], 1);
        if (slot < s.haloCap) { o.ghost[1][slot] = r; listU[slot] = ni; } else atomicOr(&counters[5], 4);
      }
    }
  }
}

// After k_slab_classify_linked: announce this rank's record deliveries (counts + epoch in one word per neighbour),
// wait for the neighbours' announcements, and hand all counts to the host (pinned memory: [0..7] own counters,
// [8..11] incoming {mig from below, ghosts from below, mig from above, ghosts from above}, [5] error bits, [6] sticky).
__global__ void k_slab_records(SlabMailHeader *mine, SlabMailHeader *down, SlabMailHeader *up, unsigned epoch, int migCap,
                               int haloCap, const int *counters, volatile int *hostCounters, int *sticky) {
  const int t = threadIdx.x;
  if (t < 4) {   // t = 2 * side + kind: towards rank-1 (it arrives there from above = origin side 1) / rank+1
    SlabMailHeader *h = (t >> 1) == 0 ? down : up;
    if (h) {
      const unsigned n = (unsigned)min(counters[t], (t & 1) ? haloCap : migCap);
      st_relaxed_sys64(&h->records[1 - (t >> 1)][t & 1], ((unsigned long long)epoch << 32) | n);
    }
  }
  if (t < 4) {   // t = 2 * origin side + kind
    const bool has = (t >> 1) == 0 ? down != nullptr : up != nullptr;
    int n = 0;
    if (has) {
      const unsigned long long t0 = global_ns();
      unsigned long long w;
      while ((int)((unsigned)((w = ld_acquire_sys64(&mine->records[t >> 1][t & 1])) >> 32) - epoch) < 0) {
        __nanosleep(200);
        if (global_ns() - t0 > 10000000000ull) { atomicOr(sticky, 256); break; }
      }
      n = (int)(unsigned)w;
    }
    hostCounters[8 + t] = n;
  }
  if (t < 8) hostCounters[t] = counters[t];
  __syncwarp();
  if (t == 0) hostCounters[6] = *(volatile int *)sticky;   // a delivery of this or an earlier step timed out
}

// Halo refresh, sending half: gather `FIELD` of the particles the two neighbours hold as ghosts and store the values
// into the neighbours' mailboxes. The delivery flags are raised by the kernel that follows (k_slab_halo_recv).
struct SlabHaloSend {
  int n[2];                 // values towards rank-1, rank+1
  const int *list[2];       // local original indices
  void *dst[2];             // halo buffer (this exchange's parity) in the neighbour's mailbox
};

template <int FIELD>
__global__ void __launch_bounds__(256) k_slab_halo_send(SlabHaloSend a, const unsigned *__restrict__ back,
                                                        const float *__restrict__ rho, const float4 *__restrict__ recB,
                                                        const float4 *__restrict__ spPred, const float *__restrict__ rhoPred) {
  const int total = a.n[0] + a.n[1];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int side = i < a.n[0] ? 0 : 1;
    const int k = side ? i - a.n[0] : i;
    const unsigned sid = back[a.list[side][k]];
    if (FIELD == 0) ((float *)a.dst[side])[k] = rho[sid];
    if (FIELD == 1) ((float2 *)a.dst[side])[k] = make_float2(recB[sid].w, rhoPred[sid]);
    if (FIELD == 2) ((float4 *)a.dst[side])[k] = spPred[sid];
  }
}

// Receiving half: wait for the deliveries, scatter the values into the local ghosts.
struct SlabHaloRecv {
  int n[2], first[2];             // ghosts received from rank-1, rank+1: counts and first local original index
  const void *src[2];             // halo buffer (this exchange's parity) in this rank's mailbox
  const unsigned *flag[2];        // this rank's delivery flags of the exchange
  unsigned *raise[2];             // the neighbours' flags for what k_slab_halo_send (the previous kernel) stored there
  unsigned epoch;
};

template <int FIELD>
__global__ void __launch_bounds__(256) k_slab_halo_recv(SlabHaloRecv a, const unsigned *__restrict__ back,
                                                        float *__restrict__ rho, float4 *__restrict__ recA,
                                                        float4 *__restrict__ recB, float4 *__restrict__ spPred,
                                                        float *__restrict__ rhoPred, int *sticky) {
  // the send kernel has completed (stream order): its stores are out, announce them; then wait for the neighbours'
  if (blockIdx.x == 0 && threadIdx.x < 2 && a.raise[threadIdx.x]) st_relaxed_sys(a.raise[threadIdx.x], a.epoch);
  if (threadIdx.x < 2 && a.flag[threadIdx.x]) wait_delivery(a.flag[threadIdx.x], a.epoch, sticky);
  __syncthreads();
  const int total = a.n[0] + a.n[1];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int side = i < a.n[0] ? 0 : 1;
    const int k = side ? i - a.n[0] : i;
    const unsigned sid = back[a.first[side] + k];
    if (FIELD == 0) {
      const float v = __ldcg((const float *)a.src[side] + k);
      rho[sid] = v;
      reinterpret_cast<float *>(recA)[8 * (size_t)sid + 7] = v;
    }
    if (FIELD == 1) {
      const float2 v = __ldcg((const float2 *)a.src[side] + k);
      reinterpret_cast<float *>(recB + sid)[3] = v.x;
      rhoPred[sid] = v.y;
    }
    if (FIELD == 2) spPred[sid] = __ldcg((const float4 *)a.src[side] + k);
  }
}

void preload_sort_kernels();
void preload_neighbor_kernels();
void preload_pcisph_kernels();
void preload_membrane_kernels();

// see preload() in sph_sort.cu: everything a linked step launches is loaded before the first flag is waited for
static void preload_step_kernels() {
  static bool done = false;
  if (done) return;
  done = true;
  preload_sort_kernels();
  preload_neighbor_kernels();
  preload_pcisph_kernels();
  preload_membrane_kernels();
  const void *const f[] = {(const void *)k_slab_classify_linked, (const void *)k_slab_records, (const void *)k_slab_append,
                           (const void *)k_slab_iota, (const void *)k_slab_halo_send<0>, (const void *)k_slab_halo_send<1>,
                           (const void *)k_slab_halo_send<2>, (const void *)k_slab_halo_recv<0>, (const void *)k_slab_halo_recv<1>,
                           (const void *)k_slab_halo_recv<2>};
  cudaFuncAttributes a;
  for (const void *k : f) cudaFuncGetAttributes(&a, k);
  cudaGetLastError();
}

extern "C" {

int sph_slab_link_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

static int link_alloc(sph_ctx *c) {
  SlabState *s = c->slab;
  if (s->mail) return 0;
  preload_step_kernels();
  const MailLayout m = mail_layout(s->migCap, s->haloCap);
  SCU(cudaSetDevice(c->device));
  SCU(cudaMalloc((void **)&s->mail, m.total));
  SCU(cudaMemset(s->mail, 0, m.total));
  for (int q = 0; q < 2; q++) SCU(cudaMalloc((void **)&s->migLocal[q], sizeof(SlabRec) * (size_t)s->migCap));
  SCU(cudaMalloc((void **)&s->doneCounter, 2 * sizeof(int)));   // [0] last-block counter, [1] sticky error word (delivery timeouts)
  SCU(cudaMemset(s->doneCounter, 0, 2 * sizeof(int)));
  return 0;
}

// Create this rank's mailbox and return its CUDA IPC handle (sph_slab_link_handle_bytes() bytes) for the
// neighbours in OTHER processes. handle may be NULL when all ranks live in this process.
int sph_slab_link_export(sph_ctx *c, void *handle) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  int rc = link_alloc(c);
  if (rc) return rc;
  if (handle) SCU(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)handle, c->slab->mail));
  return 0;
}

// Map the mailbox of the neighbour on `side` (0 = rank-1, 1 = rank+1) exported by another process.
int sph_slab_link_open(sph_ctx *c, int side, const void *handle) {
  if (!c || !c->slab || side < 0 || side > 1 || !handle) return slab_fail("sph_slab_link_open: bad arguments");
  SlabState *s = c->slab;
  SCU(cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  void *p = nullptr;
  SCU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  s->peer[side] = (char *)p;
  s->peerIpc[side] = true;
  return 0;
}

// The neighbour on `side` lives in this process: use its mailbox directly (peer access between the two devices).
int sph_slab_link_local(sph_ctx *c, int side, sph_ctx *peer) {
  if (!c || !c->slab || !peer || !peer->slab || side < 0 || side > 1) return slab_fail("sph_slab_link_local: bad arguments");
  int rc = link_alloc(peer);
  if (rc) return rc;
  if (peer->device != c->device) {
    int can = 0;
    SCU(cudaDeviceCanAccessPeer(&can, c->device, peer->device));
    if (!can) return slab_fail("sph_slab_link_local: no peer access between the two devices");
    SCU(cudaSetDevice(c->device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { slab_fail(cudaGetErrorString(e)); return SPH_E_DEVICE; }
    cudaGetLastError();
  }
  c->slab->peer[side] = peer->slab->mail;
  c->slab->peerIpc[side] = false;
  return 0;
}

// First half of a linked step: classify, deliver migrants + ghost layers to the neighbours, wait for theirs.
// Asynchronous; every rank of a process must have begun before any of them ends (the end waits for the neighbours).
int sph_slab_step_begin(sph_ctx *c) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  SlabState *s = c->slab;
  if (!s->mail || (s->hasDown && !s->peer[0]) || (s->hasUp && !s->peer[1])) return slab_fail("sph_slab_step: links are not set up");
  if (s->stepOpen) return slab_fail("sph_slab_step_begin: the previous step was not ended");
  SCU(cudaSetDevice(c->device));
  const MailLayout m = mail_layout(s->migCap, s->haloCap);
  s->epoch++;
  s->xchg = 0;
  SCU(cudaMemsetAsync(s->counters, 0, 32, c->stream));
  if (c->readPending && c->readSrc == (const void *)s->posAlt) read_fence(c);
  SlabGeom g = {s->z0, s->z1, s->hasDown, s->hasUp, s->nOwn, s->migCap, s->haloCap, s->recut};
  s->recut = 0;
  SlabLinkOut o;
  for (int side = 0; side < 2; side++) {
    const bool has = side == 0 ? s->hasDown : s->hasUp;
    const int from = 1 - side;  // what goes to rank-1 arrives there "from above" and vice versa
    o.mig[side] = has ? (SlabRec *)(s->peer[side] + m.mig[from]) : nullptr;
    o.ghost[side] = has ? (SlabRec *)(s->peer[side] + m.ghost[from]) : nullptr;
    o.hdr[side] = has ? (SlabMailHeader *)s->peer[side] : nullptr;
    o.migLocal[side] = s->migLocal[side];
  }
  o.epoch = s->epoch;
  ProfMark pm = prof_mark_begin(c, PK_SLAB_RECORDS);
  k_slab_classify_linked<<<nblk(s->nOwn), 256, 0, c->stream>>>(c->d, g, o, c->position, c->velocity, s->gid, s->posAlt,
                                                               s->velAlt, s->gidAlt, s->sendList[0], s->sendList[1],
                                                               s->counters);
  k_slab_records<<<1, 32, 0, c->stream>>>((SlabMailHeader *)s->mail, o.hdr[0], o.hdr[1], s->epoch, s->migCap, s->haloCap,
                                          s->counters, s->hostCountersDev, s->doneCounter + 1);
  prof_mark_end(c, pm);
  SCU(cudaGetLastError());
  s->stepOpen = true;
  return 0;
}

static int linked_halo(sph_ctx *c, int field) {
  SlabState *s = c->slab;
  const MailLayout m = mail_layout(s->migCap, s->haloCap);
  const int e = s->xchg++;
  if (1 + e >= SLAB_SLOTS) return slab_fail("sph_slab_step: too many halo exchanges per step (max_iteration too large)");
  const int par = e & 1;
  SlabHaloSend a;
  SlabHaloRecv b;
  for (int side = 0; side < 2; side++) {
    const bool has = side == 0 ? s->hasDown : s->hasUp;
    const int from = 1 - side;
    a.n[side] = has ? s->nSend[side] : 0;
    a.list[side] = s->sendList[side];
    a.dst[side] = has ? (void *)(s->peer[side] + m.halo[from][par]) : nullptr;
    b.raise[side] = has ? &((SlabMailHeader *)s->peer[side])->flag[from][1 + e] : nullptr;
    b.n[side] = has ? s->nRecv[side] : 0;
    b.first[side] = s->recvStart[side];
    b.src[side] = s->mail + m.halo[side][par];
    b.flag[side] = has ? &((SlabMailHeader *)s->mail)->flag[side][1 + e] : nullptr;
  }
  b.epoch = s->epoch;
  float4 *pred = c->sortedPosition + c->d.N;
  const int gs = min(nblk(a.n[0] + a.n[1]), 4 * c->numSMs), gr = min(nblk(b.n[0] + b.n[1]), 4 * c->numSMs);
#define SEND(F) k_slab_halo_send<F><<<gs, 256, 0, c->stream>>>(a, c->particleIndexBack, c->rho, c->recB, pred, c->rho + c->d.N)
#define RECV(F) k_slab_halo_recv<F><<<gr, 256, 0, c->stream>>>(b, c->particleIndexBack, c->rho, c->recA, c->recB, pred, c->rho + c->d.N, s->doneCounter + 1)
  ProfMark pm = prof_mark_begin(c, PK_SLAB_HALO);
  if (field == 0) { SEND(0); RECV(0); } else if (field == 1) { SEND(1); RECV(1); } else { SEND(2); RECV(2); }
  prof_mark_end(c, pm);
#undef SEND
#undef RECV
  c->launches += 2;
  return 0;
}

int sph_step_part(sph_ctx *c, int part);

// Second half: the one host synchronisation of the step (the counts size the launches), then absorb, and the five
// parts of the fused step with the halo refreshes between them, all asynchronous.
int sph_slab_step_end(sph_ctx *c) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  SlabState *s = c->slab;
  if (!s->stepOpen) return slab_fail("sph_slab_step_end without sph_slab_step_begin");
  s->stepOpen = false;
  SCU(cudaSetDevice(c->device));
  SCU(cudaStreamSynchronize(c->stream));
  const int *h = s->hostCounters;
  if (h[6]) return slab_fail("timed out waiting for a neighbour's delivery (the neighbour failed, or never stepped)");
  if (h[5]) return slab_fail(h[5] & 1 ? "a particle crossed more than one cell layer in one step"
                                       : "migration / halo buffer capacity exceeded");
  const MailLayout m = mail_layout(s->migCap, s->haloCap);
  s->migOut[0] = h[0]; s->ghostOut[0] = h[1]; s->migOut[1] = h[2]; s->ghostOut[1] = h[3];
  s->myMig[0] = s->migLocal[0]; s->myMig[1] = s->migLocal[1];
  const int incoming[4] = {h[8], h[9], h[10], h[11]};
  ProfMark pm = prof_mark_begin(c, PK_SLAB_ABSORB);
  int rc = sph_slab_absorb(c, s->mail + m.mig[0], s->mail + m.ghost[0], s->mail + m.mig[1], s->mail + m.ghost[1], incoming);
  prof_mark_end(c, pm);
  if (rc) return rc;
  c->launches += 5;
  const int iters = c->p.max_iteration < 1 ? 1 : c->p.max_iteration;
  if ((rc = sph_step_part(c, 0))) return rc;
  if ((rc = linked_halo(c, 0))) return rc;
  if ((rc = sph_step_part(c, 1))) return rc;
  if ((rc = linked_halo(c, 1))) return rc;
  for (int it = 1; it < iters; it++) {
    if ((rc = sph_step_part(c, 2))) return rc;
    if ((rc = linked_halo(c, 2))) return rc;
    if ((rc = sph_step_part(c, 3))) return rc;
    if ((rc = linked_halo(c, 1))) return rc;
  }
  return sph_step_part(c, 4);
}

// {migrants down, ghosts down, migrants up, ghosts up} this rank sent in its last step.
int sph_slab_counts(const sph_ctx *c, int out[4]) {
  if (!c || !c->slab) return slab_fail("not a slab context");
  const SlabState *s = c->slab;
  out[0] = s->migOut[0]; out[1] = s->ghostOut[0]; out[2] = s->migOut[1]; out[3] = s->ghostOut[1];
  return 0;
}

// One linked step of a rank that has its process to itself (one process per GPU).
int sph_slab_step(sph_ctx *c) {
  int rc = sph_slab_step_begin(c);
  return rc ? rc : sph_slab_step_end(c);
}

}  // extern "C"
