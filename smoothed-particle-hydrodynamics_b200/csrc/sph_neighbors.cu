// sph_neighbors.cu — fixed-32 neighbour search.
//
// Replaces, with bit-identical neighbour lists, the reference's
//   clearBuffers        sphFluid.cl:64-92
//   findNeighbors       sphFluid.cl:207-329 (+ searchCell :94-112, searchForNeighbors :118-184)
// and fuses pcisph_computeDensity (sphFluid.cl:472-518) into the selection pass, because
// rho(t) depends on nothing but the distances of the selected neighbours.
//
// Internal list = int32 ids only (128 B/particle, 32x32 tiles, see nb_index); the reference's
// float2 (id, distance) layout is materialised on demand: distance = sqrtf(dx²+dy²+dz²)*scale
// recomputed from sortedPosition is bit-identical to the value findNeighbors stored
// (same operands, same operation order, sphFluid.cl:154-171).
#include "sph_ctx.h"

__global__ void __launch_bounds__(256) k_clear_neighbors(int *__restrict__ nb, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) nb[i] = SPH_NO_ID;
}

// searchCell, sphFluid.cl:94-112: single wrap-around correction in each direction.
__device__ __forceinline__ int search_cell(int cell, int off, int G) {
  int n = cell + off;
  n = (n < 0) ? n + G : n;
  n = (n >= G) ? n - G : n;
  return n;
}

// The 8 cells visited by a particle, in the reference's order (sphFluid.cl:248-308):
// own, +x, +y, +z, +xy, +xz, +yz, +xyz with the per-axis direction chosen by which half of
// its 2h cell the particle sits in (:253-271).
__device__ __forceinline__ void neighbor_cells(const DevParams &d, float4 pos, int myCell, int cells[8]) {
  float cfx = (float)(int)(pos.x * d.cellSizeInv) * d.cellSize;
  float cfy = (float)(int)(pos.y * d.cellSizeInv) * d.cellSize;
  float cfz = (float)(int)(pos.z * d.cellSizeInv) * d.cellSize;
  int dx = ((pos.x - d.xmin) - cfx) < d.h ? -1 : 1;
  int dy = (((pos.y - d.ymin) - cfy) < d.h ? -1 : 1) * d.gx;
  int dz = (((pos.z - d.zmin) - cfz) < d.h ? -1 : 1) * d.gx * d.gy;
  cells[0] = myCell;
  cells[1] = search_cell(myCell, dx, d.G);
  cells[2] = search_cell(myCell, dy, d.G);
  cells[3] = search_cell(myCell, dz, d.G);
  cells[4] = search_cell(myCell, dx + dy, d.G);
  cells[5] = search_cell(myCell, dx + dz, d.G);
  cells[6] = search_cell(myCell, dy + dz, d.G);
  cells[7] = search_cell(myCell, dx + dy + dz, d.G);
}

// Generic path: one thread per particle, the reference's two passes. Used for cells the tiled kernel
// cannot take (more candidate chunks than its mask table holds) and by the stage API.
template <bool WITH_DENSITY>
__device__ __forceinline__ void find_neighbors_one(const DevParams &d, int id, const float4 *__restrict__ sp,
                                                   const unsigned *__restrict__ cellStart, int *__restrict__ nb,
                                                   float *__restrict__ rho) {
  const float4 pos = sp[id];
  const int myCell = (int)pos.w & d.mask;
  int cells[8];
  neighbor_cells(d, pos, myCell, cells);
  int cb[8], ce[8];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    cb[k] = (int)cellStart[cells[k]];
    ce[k] = (int)cellStart[cells[k] + 1];
  }
  // pass 0: radial histogram of everything within h (sphFluid.cl:157-161)
  int hist[SPH_RADIUS_SEGMENTS];
#pragma unroll
  for (int j = 0; j < SPH_RADIUS_SEGMENTS; j++) hist[j] = 0;
  const float h2 = d.h * d.h;
#pragma unroll 1
  for (int k = 0; k < 8; k++) {
    for (int j = cb[k]; j < ce[k]; ++j) {
      if (j == id) continue;
      float4 q = ldg4(sp + j);
      float ex = pos.x - q.x, ey = pos.y - q.y, ez = pos.z - q.z;
      float d2 = ex * ex + ey * ey + ez * ez;
      if (d2 <= h2) {
        int b = (int)(sqrtf(d2) * SPH_RADIUS_SEGMENTS / d.h);
        if (b < SPH_RADIUS_SEGMENTS) hist[b]++;
      }
    }
  }
  // radius threshold (sphFluid.cl:310-324)
  int j = 0, sum = 0;
  while (j < SPH_RADIUS_SEGMENTS) {
    sum += hist[j];
    if (sum == SPH_NB) break;
    if (sum > SPH_NB) { j--; break; }
    j++;
  }
  const float r_thr = (float)(j + 1) * d.h / SPH_RADIUS_SEGMENTS;
  const float r2 = r_thr * r_thr;
  // pass 1: first 32 candidates within r_thr, in scan order (:166-174)
  int found = 0;
  double density = 0.0;
#pragma unroll 1
  for (int k = 0; k < 8 && found < SPH_NB; k++) {
    for (int jj = cb[k]; jj < ce[k]; ++jj) {
      if (jj == id) continue;
      float4 q = ldg4(sp + jj);
      float ex = pos.x - q.x, ey = pos.y - q.y, ez = pos.z - q.z;
      float d2 = ex * ex + ey * ey + ez * ez;
      if (d2 <= r2) {
        if (found >= SPH_NB) break;
        nb[nb_index(id, found)] = jj;
        if (WITH_DENSITY) {  // pcisph_computeDensity term, sphFluid.cl:505-507
          float r = sqrtf(d2) * d.simScale;
          float rr = r * r;
          density += (double)((d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr));
        }
        found++;
      }
    }
  }
  for (int k = found; k < SPH_NB; k++) nb[nb_index(id, k)] = SPH_NO_ID;
  if (WITH_DENSITY) {  // sphFluid.cl:514-517
    if (density < (double)d.hScaled6) density = (double)d.hScaled6;
    density *= d.massWpoly6;
    rho[id] = (float)density;
  }
}

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(128) k_find_neighbors_generic(DevParams d, const float4 *__restrict__ sp,
                                                                const unsigned *__restrict__ cellStart,
                                                                int *__restrict__ nb, float *__restrict__ rho) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho);
}

// ---------------------------------------------------------------- tiled path (the product path)
// One CTA per non-empty hash cell ("home cell"). All particles of a cell that sit in the same
// octant of it visit the same 8 cells in the same order, so the home particles are grouped by
// octant and every 32-candidate chunk of those cells is tested against the whole group at once:
// lanes hold candidates (one coalesced float4 load per chunk), home positions are broadcast from
// shared memory two at a time and the distance test runs on Blackwell's packed fp32x2 pipe
// (add.rn.f32x2 / mul.rn.f32x2: per-lane IEEE results, identical to the scalar reference order
// d.x*d.x + d.y*d.y + d.z*d.z). Each (home particle, chunk) ballot is kept as a 32-bit hit mask
// in shared memory — hits within the largest radius the reference can choose, r = 31h/30.
// A second phase (one thread per home particle) walks its masks in the reference's scan order,
// re-derives the exact distance of the ~35 hits, builds the 30-bin radial histogram, picks r_thr
// (sphFluid.cl:310-324) and keeps the first 32 hits within it — the two passes of the reference
// collapse into one sweep over the candidates. rho(t) is accumulated on the selected neighbours.
#define FN_THREADS 64
#define FN_WARPS (FN_THREADS / 32)
#define FN_HB 64                 // home particles per batch
#define FN_MAXSLOTS 48           // 32-candidate chunks per octant group (8 cells)
#define FN_ROWS (FN_HB + 8)      // rows = home particles in group order, groups padded to even size
#define FN_CAP 54                // hits kept per home particle (within 31h/30); more -> generic path
#define FN_PAIRS 4               // home-particle pairs held in registers per sweep over the chunks

struct FnShared {
  float2 px[FN_ROWS / 2 + FN_PAIRS], py[FN_ROWS / 2 + FN_PAIRS], pz[FN_ROWS / 2 + FN_PAIRS];
  union {
    unsigned mask[FN_MAXSLOTS][FN_ROWS];  // phase 1: hit bits [slot][row]
    float d2[FN_CAP][FN_HB];              // phase 2: exact squared distance of each kept hit
  } u;
  int hits[FN_CAP][FN_HB];                // phase 2: candidate ids in scan order; compacted in place to the result
  int slotBase[8][FN_MAXSLOTS];
  int slotCnt[8][FN_MAXSLOTS];
  unsigned char hist[SPH_RADIUS_SEGMENTS + 2][FN_HB];
  int cellB[27], cellN[27];
  int nslots[8], grpStart[9], grpCount[8];
  int fallback, nextCell;
};
static_assert(sizeof(float) * FN_CAP * FN_HB <= sizeof(unsigned) * FN_MAXSLOTS * FN_ROWS, "d2 must fit in the mask area");

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(FN_THREADS) k_find_neighbors_tiled(DevParams d, const float4 *__restrict__ sp,
                                                                     const unsigned *__restrict__ cellStart,
                                                                     const unsigned *__restrict__ cellList,
                                                                     const unsigned *__restrict__ cellListCount,
                                                                     unsigned *__restrict__ cellCursor,
                                                                     int *__restrict__ nb, float *__restrict__ rho) {
  __shared__ FnShared S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nCells = (int)*cellListCount;
  const float rmaxf = (float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS;  // largest r_thr, j == 30
  // Phase 1 is only a conservative filter (phase 2 re-derives every distance exactly), so it may use
  // FMA and a slightly inflated radius; nothing inexact reaches the output.
  const float rfilt2 = rmaxf * rmaxf * 1.0001f;
  const float h2 = d.h * d.h;
  const float FAR = 1.0e30f;
  while (true) {
    __syncthreads();
    if (tid == 0) S.nextCell = (int)atomicAdd(cellCursor, 1u);
    __syncthreads();
    const int ci = S.nextCell;
    if (ci >= nCells) break;
    const int cell = (int)cellList[ci];
    const int cbeg = (int)cellStart[cell], cend = (int)cellStart[cell + 1];
    if (tid < 27) {  // ranges of the 27 surrounding cells (searchCell wrap, sphFluid.cl:108-110)
      int ox = tid % 3 - 1, oy = (tid / 3) % 3 - 1, oz = tid / 9 - 1;
      int nc = search_cell(cell, ox + oy * d.gx + oz * d.gx * d.gy, d.G);
      int b = (int)cellStart[nc];
      S.cellB[tid] = b;
      S.cellN[tid] = (int)cellStart[nc + 1] - b;
    }
    for (int b0 = cbeg; b0 < cend; b0 += FN_HB) {
      const int hb = min(FN_HB, cend - b0);
      const int id = b0 + tid;
      __syncthreads();
      if (tid < 8) S.grpCount[tid] = 0;
      if (tid == 0) S.fallback = 0;
      __syncthreads();
      // ---- group the home particles by octant (which half of the 2h cell, per axis: :253-271)
      float4 pos = f4(0.f, 0.f, 0.f, 0.f);
      int g = 0, rankInG = 0;
      if (tid < hb) {
        pos = sp[id];
        float cfx = (float)(int)(pos.x * d.cellSizeInv) * d.cellSize;
        float cfy = (float)(int)(pos.y * d.cellSizeInv) * d.cellSize;
        float cfz = (float)(int)(pos.z * d.cellSizeInv) * d.cellSize;
        g = ((((pos.x - d.xmin) - cfx) < d.h) ? 0 : 1) | ((((pos.y - d.ymin) - cfy) < d.h) ? 0 : 2) |
            ((((pos.z - d.zmin) - cfz) < d.h) ? 0 : 4);
        rankInG = atomicAdd(&S.grpCount[g], 1);
      }
      for (int r = tid; r < FN_ROWS / 2 + FN_PAIRS; r += FN_THREADS) {  // unused rows: far-away dummies
        S.px[r] = make_float2(FAR, FAR); S.py[r] = make_float2(FAR, FAR); S.pz[r] = make_float2(FAR, FAR);
      }
      __syncthreads();
      if (tid < 8) {  // slot table of group `tid`: its 8 cells in the reference's visiting order (:248-308)
        int gs = 0;
        for (int q = 0; q < tid; q++) gs += (S.grpCount[q] + 1) & ~1;
        S.grpStart[tid] = gs;
        int ns = 0;
        if (S.grpCount[tid] > 0) {
          const int sx = (tid & 1) ? 1 : -1, sy = (tid & 2) ? 1 : -1, sz = (tid & 4) ? 1 : -1;
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const int kx = (0xB2 >> k) & 1, ky = (0xD4 >> k) & 1, kz = (0xE8 >> k) & 1;  // own,x,y,z,xy,xz,yz,xyz
            const int i27 = (kx * sx + 1) + (ky * sy + 1) * 3 + (kz * sz + 1) * 9;
            const int base = S.cellB[i27], n = S.cellN[i27];
            for (int j = 0; j < n; j += 32) {
              if (ns < FN_MAXSLOTS) { S.slotBase[tid][ns] = base + j; S.slotCnt[tid][ns] = min(32, n - j); }
              ns++;
            }
          }
          if (ns > FN_MAXSLOTS) S.fallback = 1;
        }
        S.nslots[tid] = ns;
      }
      __syncthreads();
      if (S.fallback) {  // a crowded neighbourhood: the generic two-pass scan for this batch
        if (tid < hb) find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho);
        continue;
      }
      const int row = S.grpStart[g] + rankInG;
      if (tid < hb) { ((float *)S.px)[row] = pos.x; ((float *)S.py)[row] = pos.y; ((float *)S.pz)[row] = pos.z; }
      __syncthreads();
      // ---- phase 1: hit masks. Home pairs live in registers; lanes hold candidates.
#pragma unroll 1
      for (int gg = 0; gg < 8; gg++) {
        const int cnt = S.grpCount[gg];
        if (!cnt) continue;
        const int pairs = (cnt + 1) >> 1, rowBase = S.grpStart[gg], ns = S.nslots[gg];
#pragma unroll 1
        for (int q0 = 0; q0 < pairs; q0 += FN_PAIRS) {
          float2 hx[FN_PAIRS], hy[FN_PAIRS], hz[FN_PAIRS];
#pragma unroll
          for (int q = 0; q < FN_PAIRS; q++) {
            const int pr = (rowBase >> 1) + q0 + q;
            hx[q] = S.px[pr]; hy[q] = S.py[pr]; hz[q] = S.pz[pr];
          }
#pragma unroll 1
          for (int s = warp; s < ns; s += FN_WARPS) {
            const int n = S.slotCnt[gg][s];
            float4 c = f4(FAR, FAR, FAR, 0.f);
            if (lane < n) c = ldg4(sp + S.slotBase[gg][s] + lane);
            const float2 ncx = make_float2(-c.x, -c.x), ncy = make_float2(-c.y, -c.y), ncz = make_float2(-c.z, -c.z);
            unsigned m[2 * FN_PAIRS];
#pragma unroll
            for (int q = 0; q < FN_PAIRS; q++) {
              float2 ex = __fadd2_rn(hx[q], ncx), ey = __fadd2_rn(hy[q], ncy), ez = __fadd2_rn(hz[q], ncz);
              float2 dd = __ffma2_rn(ez, ez, __ffma2_rn(ey, ey, __fmul2_rn(ex, ex)));
              m[2 * q] = __ballot_sync(0xffffffffu, dd.x <= rfilt2);
              m[2 * q + 1] = __ballot_sync(0xffffffffu, dd.y <= rfilt2);
            }
            if (lane == 0) {
              unsigned *dst = &S.u.mask[s][rowBase + 2 * q0];
#pragma unroll
              for (int q = 0; q < FN_PAIRS; q++)
                if (q0 + q < pairs) *reinterpret_cast<uint2 *>(dst + 2 * q) = make_uint2(m[2 * q], m[2 * q + 1]);
            }
          }
        }
      }
      __syncthreads();
      // ---- phase 2a: hit masks -> per-particle candidate list in the reference's scan order
      int nh = 0;
      if (tid < hb) {
        const int ns = S.nslots[g];
        for (int s = 0; s < ns; s++) {
          unsigned m = S.u.mask[s][row];
          const int base = S.slotBase[g][s];
          while (m) {
            const int cidx = base + __ffs(m) - 1;
            m &= m - 1;
            if (cidx != id) {
              if (nh < FN_CAP) S.hits[nh][tid] = cidx;
              nh++;
            }
          }
        }
#pragma unroll
        for (int b = 0; b < SPH_RADIUS_SEGMENTS; b++) S.hist[b][tid] = 0;
      }
      __syncthreads();  // masks are dead from here on: their storage becomes d2[][]
      if (tid < hb) {
        if (nh > FN_CAP) {  // more hits than the list holds: generic path for this particle
          find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho);
        } else {
          // ---- phase 2b: exact distances + 30-bin radial histogram of everything within h (:157-161)
          float4 q = (nh > 0) ? ldg4(sp + S.hits[0][tid]) : pos;
          for (int t = 0; t < nh; t++) {
            const float4 cur = q;
            if (t + 1 < nh) q = ldg4(sp + S.hits[t + 1][tid]);
            float ex = pos.x - cur.x, ey = pos.y - cur.y, ez = pos.z - cur.z;
            float dd = ex * ex + ey * ey + ez * ez;
            S.u.d2[t][tid] = dd;
            if (dd <= h2) {
              int b = (int)(sqrtf(dd) * SPH_RADIUS_SEGMENTS / d.h);
              if (b < SPH_RADIUS_SEGMENTS) {
                unsigned v = S.hist[b][tid];
                if (v < 255u) S.hist[b][tid] = (unsigned char)(v + 1u);  // saturation keeps the >32 logic exact
              }
            }
          }
          // ---- radius threshold (:310-324)
          int j = 0, sum = 0;
          while (j < SPH_RADIUS_SEGMENTS) {
            sum += S.hist[j][tid];
            if (sum == SPH_NB) break;
            if (sum > SPH_NB) { j--; break; }
            j++;
          }
          const float r_thr = (float)(j + 1) * d.h / SPH_RADIUS_SEGMENTS;
          const float r2 = r_thr * r_thr;
          // ---- phase 2c: first 32 hits within r_thr, in scan order (:166-174) + rho(t)
          int found = 0;
          double density = 0.0;
          for (int t = 0; t < nh; t++) {
            const float dd = S.u.d2[t][tid];
            if (dd <= r2) {
              S.hits[found][tid] = S.hits[t][tid];
              if (WITH_DENSITY) {
                float r = sqrtf(dd) * d.simScale;
                float rr = r * r;
                density += (double)((d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr));
              }
              if (++found == SPH_NB) break;
            }
          }
#pragma unroll 8
          for (int k = 0; k < SPH_NB; k++) nb[nb_index(id, k)] = (k < found) ? S.hits[k][tid] : SPH_NO_ID;
          if (WITH_DENSITY) {
            if (density < (double)d.hScaled6) density = (double)d.hScaled6;
            density *= d.massWpoly6;
            rho[id] = (float)density;
          }
        }
      }
    }
  }
}

// neighborMap in the reference layout (32N float2: id as float, distance in metres; -1,-1 padding)
__global__ void __launch_bounds__(256) k_materialize_neighbor_map(DevParams d, const float4 *__restrict__ sp,
                                                                  const int *__restrict__ nb, float2 *__restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)d.N * SPH_NB) return;
  int id = (int)(i / SPH_NB), k = (int)(i % SPH_NB);
  int jd = nb[nb_index(id, k)];
  float2 r = make_float2(-1.f, -1.f);
  if (jd != SPH_NO_ID) {
    float4 a = sp[id], b = sp[jd];
    float ex = a.x - b.x, ey = a.y - b.y, ez = a.z - b.z;
    float d2 = ex * ex + ey * ey + ez * ez;
    r = make_float2((float)jd, sqrtf(d2) * d.simScale);
  }
  out[i] = r;
}

void launch_clear_neighbors(sph_ctx *c) {
  size_t n = (size_t)c->d.Npad * SPH_NB;
  k_clear_neighbors<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->nbIds, n);
}

void launch_find_neighbors(sph_ctx *c, bool with_density) {
  if (c->force_generic_neighbors) {
    int blocks = (c->d.N + 127) / 128;
    if (with_density)
      k_find_neighbors_generic<true><<<blocks, 128, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->nbIds, c->rho);
    else
      k_find_neighbors_generic<false><<<blocks, 128, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->nbIds, c->rho);
    return;
  }
  // persistent CTAs (one resident wave), cells handed out through an atomic cursor
  static int occ = 0;
  if (!occ) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_find_neighbors_tiled<true>, FN_THREADS, 0);
    if (occ < 1) occ = 1;
  }
  const int grid = c->numSMs * occ;
  cudaMemsetAsync(c->cellCursor, 0, sizeof(unsigned), c->stream);
  if (with_density)
    k_find_neighbors_tiled<true><<<grid, FN_THREADS, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->cellList,
                                                                   c->cellListCount, c->cellCursor, c->nbIds, c->rho);
  else
    k_find_neighbors_tiled<false><<<grid, FN_THREADS, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->cellList,
                                                                    c->cellListCount, c->cellCursor, c->nbIds, c->rho);
}

void launch_materialize_neighbor_map(sph_ctx *c, float2 *out) {
  size_t n = (size_t)c->d.N * SPH_NB;
  k_materialize_neighbor_map<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d, c->sortedPosition, c->nbIds, out);
}
