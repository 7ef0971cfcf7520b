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
// Two kernels.
//
// P1  k_neighbor_masks — one CTA per non-empty hash cell ("home cell"). All particles of a cell that
//     sit in the same octant of it visit the same 8 cells in the same order (sphFluid.cl:248-308), so
//     the home particles are grouped by octant and the group's candidates are the concatenation of
//     those 8 cells ("virtual candidate list", 32-candidate chunks). A warp takes one PAIR of home
//     particles, keeps the pair's coordinates in registers as fp32x2 operands and sweeps the chunks:
//     lanes hold candidates (one float4 load each), the squared distances of both home particles come
//     out of Blackwell's packed fp32x2 pipe (add/mul/fma.rn.f32x2), two ballots give the 32-bit hit
//     masks of the chunk, lane c keeps chunk c's masks, and at the end the warp writes each particle's
//     mask row with one coalesced store. The test is a conservative FILTER (radius 31h/30 — the largest
//     r_thr the reference can choose — slightly inflated, FMA allowed): nothing inexact reaches the output.
// P2  k_neighbor_select — one thread per particle walks its mask row in the reference's scan order,
//     re-derives the exact distance of the ~35 hits with the reference's operation order, builds the
//     30-bin radial histogram, picks r_thr (sphFluid.cl:310-324), keeps the first 32 hits within it
//     (:166-174) and accumulates rho(t) (pcisph_computeDensity, :472-518) on them. The two scan passes
//     of the reference become one sweep over the candidates plus two passes over the hits.
// Particles whose neighbourhood does not fit the tables (more than FN_LMAX candidates or FN_CAP hits)
// take find_neighbors_one, the generic two-pass scan.
#define FN_THREADS 128
#define FN_WARPS (FN_THREADS / 32)
#define FN_HB 64                 // home particles per batch
#define FN_ROWS (FN_HB + 8)      // rows = home particles in group order, groups padded to even size
#define FN_LMAX 768              // candidates per octant group (8 cells) held in the index table
#define FN_ROWW 32               // mask words per particle row (128 B); FN_LMAX / 32 <= FN_ROWW
#define FN_CAP 52                // hits kept per particle in P2

struct FnShared {
  float2 px[FN_ROWS / 2], py[FN_ROWS / 2], pz[FN_ROWS / 2];
  int rowId[FN_ROWS];
  int cidx[8][FN_LMAX];
  int cellB[27], cellN[27];
  int cum[8][9], cbase[8][8];
  int grpStart[9], grpCount[8], pairStart[9], bad[8];
  int nextCell;
};

template <int DUMMY>
__global__ void __launch_bounds__(FN_THREADS) k_neighbor_masks(DevParams d, const float4 *__restrict__ sp,
                                                               const unsigned *__restrict__ cellStart,
                                                               const unsigned *__restrict__ cellList,
                                                               const unsigned *__restrict__ cellListCount,
                                                               unsigned *__restrict__ cellCursor,
                                                               unsigned *__restrict__ masks,
                                                               unsigned char *__restrict__ flags) {
  __shared__ FnShared S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nCells = (int)*cellListCount;
  const float rmaxf = (float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS;  // largest r_thr (j == 30)
  const float rfilt2 = rmaxf * rmaxf * 1.0001f;
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
      if (tid < FN_ROWS) S.rowId[tid] = -1;
      if (tid < FN_ROWS / 2) { S.px[tid] = make_float2(FAR, FAR); S.py[tid] = make_float2(FAR, FAR); S.pz[tid] = make_float2(FAR, FAR); }
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
      __syncthreads();
      if (tid < 8) {  // tables of group `tid`: its 8 cells in the reference's visiting order (:248-308)
        int gs = 0, ps = 0;
        for (int q = 0; q < tid; q++) { gs += (S.grpCount[q] + 1) & ~1; ps += (S.grpCount[q] + 1) >> 1; }
        S.grpStart[tid] = gs;
        S.pairStart[tid] = ps;
        if (tid == 7) S.pairStart[8] = ps + ((S.grpCount[7] + 1) >> 1);
        const int sx = (tid & 1) ? 1 : -1, sy = (tid & 2) ? 1 : -1, sz = (tid & 4) ? 1 : -1;
        int L = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
          const int kx = (0xB2 >> k) & 1, ky = (0xD4 >> k) & 1, kz = (0xE8 >> k) & 1;  // own,x,y,z,xy,xz,yz,xyz
          const int i27 = (kx * sx + 1) + (ky * sy + 1) * 3 + (kz * sz + 1) * 9;
          S.cum[tid][k] = L;
          S.cbase[tid][k] = S.cellB[i27];
          L += S.cellN[i27];
        }
        S.cum[tid][8] = L;
        S.bad[tid] = (L > FN_LMAX) ? 1 : 0;
      }
      __syncthreads();
      if (tid < hb) {
        const int row = S.grpStart[g] + rankInG;
        ((float *)S.px)[row] = pos.x; ((float *)S.py)[row] = pos.y; ((float *)S.pz)[row] = pos.z;
        S.rowId[row] = id;
        flags[id] = (unsigned char)S.bad[g];
      }
      // ---- virtual candidate list -> sorted index, per group
#pragma unroll 1
      for (int gg = 0; gg < 8; gg++) {
        if (!S.grpCount[gg] || S.bad[gg]) continue;
        const int L = S.cum[gg][8];
        for (int v = tid; v < L; v += FN_THREADS) {
          int k = 0;
#pragma unroll
          for (int q = 1; q < 8; q++) k += (v >= S.cum[gg][q]) ? 1 : 0;
          S.cidx[gg][v] = S.cbase[gg][k] + v - S.cum[gg][k];
        }
      }
      __syncthreads();
      // ---- one warp per home pair: sweep the chunks
      const int totalPairs = S.pairStart[8];
#pragma unroll 1
      for (int t = warp; t < totalPairs; t += FN_WARPS) {
        int gg = 0;
#pragma unroll
        for (int q = 1; q < 8; q++) gg += (t >= S.pairStart[q]) ? 1 : 0;
        if (S.bad[gg]) continue;
        const int pr = (S.grpStart[gg] >> 1) + (t - S.pairStart[gg]);
        const float2 hx = S.px[pr], hy = S.py[pr], hz = S.pz[pr];
        const int L = S.cum[gg][8];
        const int nch = (L + 31) >> 5;
        unsigned keep0 = 0, keep1 = 0;
#pragma unroll 2
        for (int c = 0; c < nch; c++) {
          const int v = (c << 5) + lane;
          float4 cd = f4(FAR, FAR, FAR, 0.f);
          if (v < L) cd = ldg4(sp + S.cidx[gg][v]);
          const float2 ncx = make_float2(-cd.x, -cd.x), ncy = make_float2(-cd.y, -cd.y), ncz = make_float2(-cd.z, -cd.z);
          float2 ex = __fadd2_rn(hx, ncx), ey = __fadd2_rn(hy, ncy), ez = __fadd2_rn(hz, ncz);
          float2 dd = __ffma2_rn(ez, ez, __ffma2_rn(ey, ey, __fmul2_rn(ex, ex)));
          const unsigned m0 = __ballot_sync(0xffffffffu, dd.x <= rfilt2);
          const unsigned m1 = __ballot_sync(0xffffffffu, dd.y <= rfilt2);
          if (lane == c) { keep0 = m0; keep1 = m1; }
        }
        const int id0 = S.rowId[2 * pr], id1 = S.rowId[2 * pr + 1];
        if (lane < nch) {
          masks[(size_t)id0 * FN_ROWW + lane] = keep0;
          if (id1 >= 0) masks[(size_t)id1 * FN_ROWW + lane] = keep1;
        }
      }
    }
  }
}

struct FnSelShared {
  int hits[FN_CAP][FN_THREADS];
  int cumS[9][FN_THREADS];
  int adjS[8][FN_THREADS];
  unsigned char hist[SPH_RADIUS_SEGMENTS + 2][FN_THREADS];
};

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(FN_THREADS) k_neighbor_select(DevParams d, const float4 *__restrict__ sp,
                                                                const unsigned *__restrict__ cellStart,
                                                                const unsigned *__restrict__ masks,
                                                                const unsigned char *__restrict__ flags,
                                                                int *__restrict__ nb, float *__restrict__ rho) {
  __shared__ FnSelShared S;
  const int tid = threadIdx.x;
  const int id = blockIdx.x * FN_THREADS + tid;
  if (id >= d.N) return;
  if (flags[id]) { find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho); return; }
  const float4 pos = sp[id];
  const float h2 = d.h * d.h;
  {  // the 8 cells of this particle and the prefix of their sizes (virtual candidate index space)
    int cells[8];
    neighbor_cells(d, pos, (int)pos.w & d.mask, cells);
    int L = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int b = (int)cellStart[cells[k]], n = (int)cellStart[cells[k] + 1] - b;
      S.cumS[k][tid] = L;
      S.adjS[k][tid] = b - L;
      L += n;
    }
    S.cumS[8][tid] = L;
  }
  const int L = S.cumS[8][tid];
  const int nch = (L + 31) >> 5;
  // ---- mask row -> candidate list in scan order (each thread walks words and bits at its own pace)
  const unsigned *row = masks + (size_t)id * FN_ROWW;
  int nh = 0;
  {
    int c = 0, k = 0, kend = S.cumS[1][tid], adj = S.adjS[0][tid];
    unsigned m = nch > 0 ? __ldg(row) : 0u;
    while (true) {
      if (m == 0u) {
        if (++c >= nch) break;
        m = __ldg(row + c);
        continue;
      }
      const int v = (c << 5) + __ffs(m) - 1;
      m &= m - 1;
      while (v >= kend) { ++k; kend = S.cumS[k + 1][tid]; adj = S.adjS[k][tid]; }
      const int cidx = v + adj;
      if (cidx != id) {
        if (nh < FN_CAP) S.hits[nh][tid] = cidx;
        nh++;
      }
    }
  }
  if (nh > FN_CAP) { find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho); return; }
#pragma unroll
  for (int b = 0; b < SPH_RADIUS_SEGMENTS; b++) S.hist[b][tid] = 0;
  // ---- exact distances + 30-bin radial histogram of everything within h (:157-161)
  {
    float4 q = (nh > 0) ? ldg4(sp + S.hits[0][tid]) : pos;
    for (int t = 0; t < nh; t++) {
      const float4 cur = q;
      if (t + 1 < nh) q = ldg4(sp + S.hits[t + 1][tid]);
      float ex = pos.x - cur.x, ey = pos.y - cur.y, ez = pos.z - cur.z;
      float dd = ex * ex + ey * ey + ez * ez;
      if (dd <= h2) {
        int b = (int)(sqrtf(dd) * SPH_RADIUS_SEGMENTS / d.h);
        if (b < SPH_RADIUS_SEGMENTS) {
          unsigned v = S.hist[b][tid];
          if (v < 255u) S.hist[b][tid] = (unsigned char)(v + 1u);  // saturation keeps the >32 logic exact
        }
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
  // ---- first 32 hits within r_thr, in scan order (:166-174) + rho(t)
  int found = 0;
  double density = 0.0;
  {
    float4 q = (nh > 0) ? ldg4(sp + S.hits[0][tid]) : pos;
    for (int t = 0; t < nh; t++) {
      const float4 cur = q;
      if (t + 1 < nh) q = ldg4(sp + S.hits[t + 1][tid]);
      float ex = pos.x - cur.x, ey = pos.y - cur.y, ez = pos.z - cur.z;
      float dd = ex * ex + ey * ey + ez * ez;
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
  }
#pragma unroll 8
  for (int k = 0; k < SPH_NB; k++) nb[nb_index(id, k)] = (k < found) ? S.hits[k][tid] : SPH_NO_ID;
  if (WITH_DENSITY) {
    if (density < (double)d.hScaled6) density = (double)d.hScaled6;
    density *= d.massWpoly6;
    rho[id] = (float)density;
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
  // P1: persistent CTAs (one resident wave), cells handed out through an atomic cursor; P2: one thread per particle
  static int occ = 0;
  if (!occ) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_neighbor_masks<0>, FN_THREADS, 0);
    if (occ < 1) occ = 1;
  }
  cudaMemsetAsync(c->cellCursor, 0, sizeof(unsigned), c->stream);
  k_neighbor_masks<0><<<c->numSMs * occ, FN_THREADS, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->cellList,
                                                                   c->cellListCount, c->cellCursor, c->fnMasks,
                                                                   c->fnFlags);
  const int blocks = (c->d.N + FN_THREADS - 1) / FN_THREADS;
  if (with_density)
    k_neighbor_select<true><<<blocks, FN_THREADS, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->fnMasks,
                                                                 c->fnFlags, c->nbIds, c->rho);
  else
    k_neighbor_select<false><<<blocks, FN_THREADS, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->fnMasks,
                                                                  c->fnFlags, c->nbIds, c->rho);
}

size_t sph_neighbor_mask_words(size_t n) { return n * FN_ROWW; }

void launch_materialize_neighbor_map(sph_ctx *c, float2 *out) {
  size_t n = (size_t)c->d.N * SPH_NB;
  k_materialize_neighbor_map<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d, c->sortedPosition, c->nbIds, out);
}
