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

// Generic path: one thread per particle, the reference's two passes.
template <bool WITH_DENSITY>
__global__ void __launch_bounds__(128) k_find_neighbors_generic(DevParams d, const float4 *__restrict__ sp,
                                                                const unsigned *__restrict__ cellStart,
                                                                int *__restrict__ nb, float *__restrict__ rho) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
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
  int blocks = (c->d.N + 127) / 128;
  if (with_density)
    k_find_neighbors_generic<true><<<blocks, 128, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->nbIds, c->rho);
  else
    k_find_neighbors_generic<false><<<blocks, 128, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->nbIds, c->rho);
}

void launch_materialize_neighbor_map(sph_ctx *c, float2 *out) {
  size_t n = (size_t)c->d.N * SPH_NB;
  k_materialize_neighbor_map<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d, c->sortedPosition, c->nbIds, out);
}
