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
                                                   float4 *__restrict__ spPred,
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
    float4 *q = recB + 2 * (size_t)sid + 1;
    if (PACK) b[i] = make_float2(q->x, q->y); else *q = make_float4(b[i].x, b[i].y, 0.f, 0.f);
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
  SCU(cudaMalloc(&s->counters, 32)); SCU(cudaMallocHost(&s->hostCounters, 32));
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
#define H(F, P) k_slab_halo<F, P><<<nblk(n), 256, 0, c->stream>>>(n, list, first, c->particleIndexBack, c->rho, c->recA, c->recB, pred, buf)
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

void slab_destroy(sph_ctx *c) {
  SlabState *s = c->slab;
  if (!s) return;
  // position / velocity may have been swapped with the Alt buffers: both pairs are freed exactly once
  // (sph_destroy frees c->position / c->velocity, this frees the Alt side).
  cudaFree(s->gid); cudaFree(s->gidAlt); cudaFree(s->posAlt); cudaFree(s->velAlt);
  cudaFree(s->counters); cudaFreeHost(s->hostCounters); cudaFree(s->sendList[0]); cudaFree(s->sendList[1]);
  delete s;
  c->slab = nullptr;
}
