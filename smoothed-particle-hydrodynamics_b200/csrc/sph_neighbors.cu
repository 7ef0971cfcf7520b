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
#include <cuda_fp16.h>

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

// Cell id of a sorted particle (sortedPosition.w, sphFluid.cl:229). A particle that left the hash grid (NaN or a
// position outside the box: cannot happen while integrate clamps positions) was binned into
// the last cell by k_hash_count; the reference would index past gridCellIndex here — keep the loads in range instead.
__device__ __forceinline__ int home_cell(const DevParams &d, float w) {
  const int c = (int)w & d.mask;
  return min(max(c, 0), d.G - 1);
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

#ifndef SPH_SCAN_MAX_CELL
#define SPH_SCAN_MAX_CELL 16384
#endif
// Generic path: one thread per particle, the reference's two passes. Used for cells the tiled kernel
// cannot take (more candidate chunks than its mask table holds) and by the stage API.
template <bool WITH_DENSITY>
__device__ __forceinline__ void find_neighbors_one(const DevParams &d, int id, const float4 *__restrict__ sp,
                                                   const unsigned *__restrict__ cellStart, int *__restrict__ nb,
                                                   float *__restrict__ rho, float4 *__restrict__ recA = nullptr) {
  const float4 pos = sp[id];
  const int myCell = home_cell(d, pos.w);
  int cells[8];
  neighbor_cells(d, pos, myCell, cells);
  int cb[8], ce[8];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    cb[k] = (int)cellStart[cells[k]];
    // a cell beyond SPH_SCAN_MAX_CELL particles exists only in a diverged simulation (see k_rank_reorder): its scan
    // is cut so that a step stays finite
    ce[k] = min((int)cellStart[cells[k] + 1], cb[k] + SPH_SCAN_MAX_CELL);
  }
  // pass 0: radial histogram of everything within h (sphFluid.cl:157-161). 30 bins x 6 bits in three 64-bit
  // registers, saturating at 63: the threshold rule below only asks whether a running sum is < 32, == 32 or > 32,
  // so a saturated bin gives the same answer as the true count. Bin = (int)(sqrtf(d2)*30/h), obtained exactly
  // from the host-built thresholds d.binT (see DevParams) around an approximate first guess.
  unsigned long long hist0 = 0ull, hist1 = 0ull, hist2 = 0ull;
  const float h2 = d.h * d.h;
  auto hist_add = [&](float dd) {
    if (dd <= h2) {
      float sq;
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(dd));
      int e = min((int)(sq * d.binScale), SPH_RADIUS_SEGMENTS);
      const int b = e + ((dd >= d.binT[e + 1]) ? 1 : 0) - ((dd < d.binT[e]) ? 1 : 0);
      if (b < SPH_RADIUS_SEGMENTS) {
        const int w = (b >= 10 ? 1 : 0) + (b >= 20 ? 1 : 0);
        const int sh = (b - 10 * w) * 6;
        unsigned long long &hw = (w == 0) ? hist0 : (w == 1) ? hist1 : hist2;
        if (((hw >> sh) & 63ull) != 63ull) hw += 1ull << sh;
      }
    }
  };
  auto dist2 = [&](const float4 &q) {
    const float ex = pos.x - q.x, ey = pos.y - q.y, ez = pos.z - q.z;
    return ex * ex + ey * ey + ez * ez;
  };
#pragma unroll 1
  for (int k = 0; k < 8; k++) {
    int j = cb[k];
    for (; j + 3 < ce[k]; j += 4) {  // four candidates per pass: independent loads and distance chains
      const float4 q0 = ldg4(sp + j), q1 = ldg4(sp + j + 1), q2 = ldg4(sp + j + 2), q3 = ldg4(sp + j + 3);
      const float d0 = dist2(q0), d1 = dist2(q1), d2 = dist2(q2), d3 = dist2(q3);
      if (j != id) hist_add(d0);
      if (j + 1 != id) hist_add(d1);
      if (j + 2 != id) hist_add(d2);
      if (j + 3 != id) hist_add(d3);
    }
    for (; j < ce[k]; ++j)
      if (j != id) hist_add(dist2(ldg4(sp + j)));
  }
  // radius threshold (sphFluid.cl:310-324)
  int j = SPH_RADIUS_SEGMENTS;
  {
    int sum = 0;
    bool open = true;
#pragma unroll
    for (int jj = 0; jj < SPH_RADIUS_SEGMENTS; jj++) {
      const unsigned long long hw = jj < 10 ? hist0 : (jj < 20 ? hist1 : hist2);
      sum += (int)((hw >> ((jj % 10) * 6)) & 63ull);
      if (open && sum >= SPH_NB) { j = (sum == SPH_NB) ? jj : jj - 1; open = false; }
    }
  }
  const float r_thr = (float)(j + 1) * d.h / SPH_RADIUS_SEGMENTS;
  const float r2 = r_thr * r_thr;
  // pass 1: first 32 candidates within r_thr, in scan order (:166-174)
  int found = 0;
  double density = 0.0;
  auto take = [&](int jj, float dd) {
    if (jj != id && dd <= r2 && found < SPH_NB) {
      nb[nb_index(id, found)] = jj;
      if (WITH_DENSITY) {  // pcisph_computeDensity term, sphFluid.cl:505-507
        float r = sqrtf(dd) * d.simScale;
        float rr = r * r;
        density += (double)((d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr));
      }
      found++;
    }
  };
#pragma unroll 1
  for (int k = 0; k < 8 && found < SPH_NB; k++) {
    int jj = cb[k];
    for (; jj + 3 < ce[k] && found < SPH_NB; jj += 4) {
      const float4 q0 = ldg4(sp + jj), q1 = ldg4(sp + jj + 1), q2 = ldg4(sp + jj + 2), q3 = ldg4(sp + jj + 3);
      const float d0 = dist2(q0), d1 = dist2(q1), d2 = dist2(q2), d3 = dist2(q3);
      take(jj, d0);
      take(jj + 1, d1);
      take(jj + 2, d2);
      take(jj + 3, d3);
    }
    for (; jj < ce[k] && found < SPH_NB; ++jj) take(jj, dist2(ldg4(sp + jj)));
  }
  for (int k = found; k < SPH_NB; k++) nb[nb_index(id, k)] = SPH_NO_ID;
  if (WITH_DENSITY) {  // sphFluid.cl:514-517
    if (density < (double)d.hScaled6) density = (double)d.hScaled6;
    density *= d.massWpoly6;
    rho[id] = (float)density;
    if (recA) reinterpret_cast<float *>(recA)[8 * (size_t)id + 7] = (float)density;  // rho slot of the gather record
  }
}

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(128) k_find_neighbors_generic(DevParams d, const float4 *__restrict__ sp,
                                                                const unsigned *__restrict__ cellStart,
                                                                int *__restrict__ nb, float *__restrict__ rho,
                                                                float4 *__restrict__ recA) {
  pdl_entry();
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho, recA);
}

// ---------------------------------------------------------------- list path (the product path)
// One thread per particle, like the generic path, but the 8-cell scan is done ONCE:
//   group   the CTA's 128 consecutive particles are dealt to the threads grouped by (cell, octant), so that the
//           lanes of a warp walk the same cells and load the same addresses (broadcast);
//   sweep   16 candidates per iteration from a half-precision, cell-local copy of the sorted positions (three
//           256-bit loads, k_soa_positions_half), packed HADD2/HFMA2 distances against a conservative filter radius
//           (the largest r_thr the reference can choose, 31h/30, plus the fp16 error bound), the sign bits gathered
//           into a 16-bit mask, and the (rare) hits appended as (cell k, rank in cell) to the thread's hit list in
//           shared memory — in scan order, because the thread walks its cells in the reference's visiting order.
//           Grids where a cell id could stand for two nearby cells (tiny or aliased grids, see launch_find_neighbors)
//           sweep the fp32 copy instead (VAR 1; VAR 0 is the first, branchy version, kept for A/B runs);
//   select  the ~38 hits are re-evaluated with the reference's operation order (exact distances from the float4
//           sorted positions): histogram, r_thr, first 32 within r_thr, rho(t).
// Nothing inexact reaches the output. Hits beyond the shared-memory list spill to a per-thread array in local
// memory; a thread whose neighbourhood exceeds that as well (or a cell with more than 8191 particles) takes
// find_neighbors_one. The kernel is bound by the L1 data pipe (81 %) and instruction issue (83 %), see profiles/.
#ifndef LP_THREADS
#define LP_THREADS 128
#endif
// hit-list capacity and resident CTAs per SM the register allocation is tuned for: the kernel is bound by load
// latency and the L1 data pipe; 10 CTAs (40 warps, 47 registers, 16 KB of shared memory each) measured fastest
#ifndef LP_CAP
#define LP_CAP 56
#endif
#ifndef LP_MINB
#define LP_MINB 10
#endif
// hits beyond the shared-memory list go to a per-thread overflow array in local memory (denser-than-lattice
// neighbourhoods, e.g. the settled config A); only beyond LP_CAP + LP_OVF does a thread rescan with the plain kernel
#ifndef LP_OVF
#define LP_OVF 72
#endif
// VAR 2 keeps one 32-bit ENTRY per non-empty group of 16 candidates instead of one 16-bit code per hit:
// (cell number k << 13 | rank of the group's first candidate in the cell + 16) << 16 | 16-bit hit mask. ~25 entries per particle on the lattice fluid.
// Shared memory is what limits the kernel to 10 CTAs per SM (20 624 bytes each: one more row of entries and only 9
// fit, measured 2.69 -> 2.75 ms): 28 rows of entries + the 8 rows of cell bases (cbS) measured faster than 36 rows of
// entries with the bases in local memory (2.75 -> 2.69 ms; 26 / 27 / 28 rows: 2.71 / 2.69 / 2.69 ms).
#ifndef LP_ECAP
#define LP_ECAP 28
#endif
#ifndef LP_EOVF
#define LP_EOVF 36
#endif
#ifndef SPH_LP_CBS
#define SPH_LP_CBS 1
#endif
#ifndef SPH_LP_PRMT
#define SPH_LP_PRMT 1
#endif
#ifndef SPH_LP_WADDR
#define SPH_LP_WADDR 1
#endif
#ifndef SPH_LP_BMSK
#define SPH_LP_BMSK 1
#endif


// Sweep copy of the sorted positions: groups of 8 particles as (x[8] | y[8] | z[8]) = 96 bytes, so that the three
// 256-bit loads of a sweep pass share one address computation.
__global__ void __launch_bounds__(256) k_soa_positions(int n, const float4 *__restrict__ sp, float *__restrict__ S) {
  pdl_entry();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float *o = S + 24 * (size_t)(i >> 3) + (i & 7);
  if (i < n) {
    const float4 p = sp[i];
    o[0] = p.x; o[8] = p.y; o[16] = p.z;
  } else if (i < n + 16) {  // the sweep reads whole groups of 8: finite padding behind the last particle
    o[0] = 1.0e30f; o[8] = 1.0e30f; o[16] = 1.0e30f;
  }
}

// Half-precision sweep copy (list kernel, VAR 2): cell-LOCAL coordinates in units of h, u = (p - cell corner) / h in
// [0, 2], where the cell corner is the one the hash uses (cx = (int)(p.x * cellSizeInv), sphFluid.cl:197-199). Groups
// of 16 particles as x[16] | y[16] | z[16] = 96 bytes; inside a 16-vector, half2 register r holds particles (r, r + 8)
// so that the sign bits of the packed distances land in ascending particle order (bits 0-7 and 16-23).
#define LP_HPAD 100.0f   // padding coordinate: far from everything, its square still finite in half precision
__device__ __forceinline__ float3 cell_local_h(const DevParams &d, const float4 p, const float hInv) {
  const float lx = p.x - (float)(int)(p.x * d.cellSizeInv) * d.cellSize;
  const float ly = p.y - (float)(int)(p.y * d.cellSizeInv) * d.cellSize;
  const float lz = p.z - (float)(int)(p.z * d.cellSizeInv) * d.cellSize;
  return make_float3(lx * hInv, ly * hInv, lz * hInv);
}
__global__ void __launch_bounds__(256) k_soa_positions_half(DevParams d, const float4 *__restrict__ sp,
                                                            __half *__restrict__ S) {
  pdl_entry();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int s = i & 15;
  __half *o = S + 48 * (size_t)(i >> 4) + 2 * (s & 7) + (s >> 3);
  if (i < d.N) {
    const float3 u = cell_local_h(d, sp[i], 1.0f / d.h);
    o[0] = __float2half_rn(u.x); o[16] = __float2half_rn(u.y); o[32] = __float2half_rn(u.z);
  } else if (i < d.N + 32) {  // the sweep reads whole groups of 16
    o[0] = __float2half_rn(LP_HPAD); o[16] = __float2half_rn(LP_HPAD); o[32] = __float2half_rn(LP_HPAD);
  }
}

__device__ __forceinline__ void ld256u(const void *p, uint4 &a, uint4 &b) {
  asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
               : "l"(p));
}

struct LpYes { static constexpr bool value = true; };
struct LpNo { static constexpr bool value = false; };

template <bool WITH_DENSITY, int VAR>
__global__ void __launch_bounds__(LP_THREADS, LP_MINB) k_find_neighbors_list(DevParams d, const float4 *__restrict__ sp,
                                                                    const float *__restrict__ S,
                                                                    const unsigned *__restrict__ cellStart,
                                                                    int *__restrict__ nb, float *__restrict__ rho,
                                                                    float4 *__restrict__ recA, unsigned *nsStats) {
  __shared__ unsigned short hits[VAR == 2 ? 1 : LP_CAP + 1][LP_THREADS];
  __shared__ unsigned ent[VAR == 2 ? LP_ECAP + 1 : 1][LP_THREADS];
  __shared__ int binCount[64], binBase[64];
  __shared__ int binTot0;
  __shared__ unsigned char srcOf[LP_THREADS];
#if SPH_LP_CBS
  // first index of the 8 visited cells, one column per thread: the select phase looks it up by a run-time cell number
  // for every entry — from a per-thread local array that was 7 L1 sectors per warp access (lanes differ in the cell
  // they are at: 16 of the kernel's 59 L1 sectors per particle, ncu source view), from here it is one wavefront
  __shared__ int cbS[8][LP_THREADS];
#endif
  pdl_entry();
  const int tid = threadIdx.x;
  // Lanes that share (cell, octant) walk the same cells in lockstep and load the same addresses. The CTA's 128
  // consecutive particles (about two cells) are therefore dealt to the threads grouped by (cell, octant): a counting
  // sort over 8 cells x 8 octants in shared memory (cells beyond the eighth share the last bins; the order inside
  // a bin is arbitrary — grouping only affects speed, every thread still handles exactly one particle).
  int id;
  {
    const int first = blockIdx.x * LP_THREADS, mine = first + tid;
    int bin = 63;  // beyond N
    if (mine < d.N) {
      const float4 p = sp[mine];
      const float cfx = (float)(int)(p.x * d.cellSizeInv) * d.cellSize;
      const float cfy = (float)(int)(p.y * d.cellSizeInv) * d.cellSize;
      const float cfz = (float)(int)(p.z * d.cellSizeInv) * d.cellSize;
      const int oct = ((((p.x - d.xmin) - cfx) < d.h) ? 0 : 1) | ((((p.y - d.ymin) - cfy) < d.h) ? 0 : 2) |
                      ((((p.z - d.zmin) - cfz) < d.h) ? 0 : 4);
      const int c0 = (int)sp[first].w & d.mask;  // sorted order: cell ids ascend inside the CTA
      bin = (min(max(((int)p.w & d.mask) - c0, 0), 7) << 3) | oct;
    }
    if (tid < 64) binCount[tid] = 0;
    __syncthreads();
    const int slot = atomicAdd(&binCount[bin], 1);
    __syncthreads();
    if (tid < 64) {
      const int c = binCount[tid];
      int inc = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc, o);
        if ((tid & 31) >= o) inc += u;
      }
      binBase[tid] = inc - c;
      if (tid == 31) binTot0 = inc;
    }
    __syncthreads();
    srcOf[binBase[bin] + (bin >= 32 ? binTot0 : 0) + slot] = (unsigned char)tid;
    __syncthreads();
    id = first + (int)srcOf[tid];
  }
  if (id >= d.N) return;
  const float4 pos = sp[id];
  const int myCell = home_cell(d, pos.w);
  int cells[8];
  neighbor_cells(d, pos, myCell, cells);
  int cb[8], ce[8];
  bool fits = true;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    cb[k] = (int)cellStart[cells[k]];
    ce[k] = (int)cellStart[cells[k] + 1];
    fits = fits && (ce[k] - cb[k] < 8176);  // 13-bit ranks, biased by 16 in the entries of VAR 2
#if SPH_LP_CBS
    cbS[k][tid] = cb[k];
#endif
  }
  const float rmaxf = (float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS;  // largest r_thr (j == 30)
  const float rfilt2 = rmaxf * rmaxf * 1.0001f;
  const float2 npx = make_float2(-pos.x, -pos.x), npy = make_float2(-pos.y, -pos.y), npz = make_float2(-pos.z, -pos.z);
  unsigned short ovf[VAR == 2 ? 1 : LP_OVF];
  unsigned ovfE[VAR == 2 ? LP_EOVF : 1];
  int n = 0, ne = 0;
  if (fits && VAR == 2) {
    // half-precision variant: 16 candidates per pass from the cell-local fp16 copy (three 256-bit loads = half the
    // bytes per candidate through the L1 data pipe, which bounds the fp32 sweep). The home particle is expressed in
    // the visited cell's frame: u_p - 2 (kx sx, ky sy, kz sz). Everything is in units of h; the filter constant
    // covers the largest r_thr, (31/30)^2 * 1.0001 = 1.0679, plus the worst-case fp16 error of a true hit
    // (coordinates 0.002 per axis -> 0.007 on d2, three fma roundings 0.0015): T = 1 + 92/1024 = 1.0898.
    const float3 up = cell_local_h(d, pos, 1.0f / d.h);
    const float sgx = (((pos.x - d.xmin) - (float)(int)(pos.x * d.cellSizeInv) * d.cellSize) < d.h) ? -2.0f : 2.0f;
    const float sgy = (((pos.y - d.ymin) - (float)(int)(pos.y * d.cellSizeInv) * d.cellSize) < d.h) ? -2.0f : 2.0f;
    const float sgz = (((pos.z - d.zmin) - (float)(int)(pos.z * d.cellSizeInv) * d.cellSize) < d.h) ? -2.0f : 2.0f;
    const __half2 nT = __float2half2_rn(-1.08984375f);
    const __half *Sh = reinterpret_cast<const __half *>(S);
    int nh = 0;
#if SPH_LP_WADDR
    // shared-memory address of this thread's next entry slot, kept in a register: recomputing it from threadIdx and
    // the shared window base cost five instructions per iteration (the compiler rematerialises both at 48 registers)
    const unsigned wAddr = (unsigned)__cvta_generic_to_shared(&ent[0][tid]);
#endif
#pragma unroll 1
    for (int k = 0; k < 8; k++) {
      const int b = cb[k], e = ce[k];
      const int enc = (k << 13) - b + 16;  // entry code of group j = (k << 13) | (j - b + 16): the aligned group may start up to 15 before the cell
      const __half2 px = __float2half2_rn(up.x - (((0xB2 >> k) & 1) ? sgx : 0.0f));
      const __half2 py = __float2half2_rn(up.y - (((0xD4 >> k) & 1) ? sgy : 0.0f));
      const __half2 pz = __float2half2_rn(up.z - (((0xE8 >> k) & 1) ? sgz : 0.0f));
#pragma unroll 1
      for (int j = b & ~15; j < e; j += 16) {
        uint4 xa, xb, ya, yb, za, zb;
        const __half *g = Sh + 3 * (size_t)j;
        ld256u(g, xa, xb);
        ld256u(g + 16, ya, yb);
        ld256u(g + 32, za, zb);
        unsigned m = 0u;
        auto dist2T = [&](unsigned x, unsigned y, unsigned z) {  // (d2 - T) of candidates r (low half) and 8 + r (high)
          const __half2 ex = __hsub2(*reinterpret_cast<const __half2 *>(&x), px);
          const __half2 ey = __hsub2(*reinterpret_cast<const __half2 *>(&y), py);
          const __half2 ez = __hsub2(*reinterpret_cast<const __half2 *>(&z), pz);
          const __half2 a = __hfma2(ez, ez, __hfma2(ey, ey, __hfma2(ex, ex, nT)));  // negative = hit
          return *reinterpret_cast<const unsigned *>(&a);
        };
#if SPH_LP_PRMT
        // The four sign bits of two such words (candidates r, 8 + r, r + 4, 12 + r) are turned into four byte masks
        // by ONE byte permute with sign replication (selector nibbles 9, B, D, F: the sign of bytes 1 and 3 of either
        // operand) and placed by ONE and-or: 2 instructions per 4 candidates instead of 4 (shift + and-or per pair).
        auto quad = [&](unsigned a, unsigned b, int r) {
          unsigned w;
          asm("prmt.b32 %0, %1, %2, 0xFDB9;" : "=r"(w) : "r"(a), "r"(b));
          m |= w & (0x10100101u << r);  // bits r | 8 + r | 20 + r | 28 + r
        };
        quad(dist2T(xa.x, ya.x, za.x), dist2T(xb.x, yb.x, zb.x), 0);
        quad(dist2T(xa.y, ya.y, za.y), dist2T(xb.y, yb.y, zb.y), 1);
        quad(dist2T(xa.z, ya.z, za.z), dist2T(xb.z, yb.z, zb.z), 2);
        quad(dist2T(xa.w, ya.w, za.w), dist2T(xb.w, yb.w, zb.w), 3);
        // fold to 16 ascending bits (bit i = candidate j + i); the group is aligned to 16: its ends may reach into the
        // neighbouring cells
        m = (m | (m >> 16)) & 0xffffu;
#else
        // bit r = candidate j + r, bit 16 + r = candidate j + 8 + r (r < 8): inside the filter
        auto lane2 = [&](unsigned x, unsigned y, unsigned z, int r) {
          m |= (dist2T(x, y, z) >> (15 - r)) & (0x00010001u << r);
        };
        lane2(xa.x, ya.x, za.x, 0); lane2(xa.y, ya.y, za.y, 1); lane2(xa.z, ya.z, za.z, 2); lane2(xa.w, ya.w, za.w, 3);
        lane2(xb.x, yb.x, zb.x, 4); lane2(xb.y, yb.y, zb.y, 5); lane2(xb.z, yb.z, zb.z, 6); lane2(xb.w, yb.w, zb.w, 7);
        // fold to 16 ascending bits; the group is aligned to 16: its ends may reach into the neighbouring cells
        m = (m & 0xffu) | ((m >> 8) & 0xff00u);
#endif
#if SPH_LP_BMSK
        {  // candidates [max(b - j, 0), min(e - j, 16)) of the group belong to the cell: one bit-mask instruction
          const int lo = max(b - j, 0);
          unsigned vm;
          asm("bmsk.clamp.b32 %0, %1, %2;" : "=r"(vm) : "r"(lo), "r"(min(e - j, 16) - lo));
          m &= vm;
        }
#else
        m &= (0xffffu << max(b - j, 0)) & (0xffffu >> max(j + 16 - e, 0));
#endif
        // One entry per non-empty group; the hits are unpacked by the select phase, whose loop runs at nearly full
        // lane occupancy. (Appending hit by hit here cost 27 % of the kernel's instructions at 6 active lanes of 32:
        // most lanes of a warp have no hit in a given group.)
        // Branch-free: every lane stores, a lane without a hit does not advance and overwrites the slot next time
        // (as a predicated block this was 22 instructions at 9 active lanes in 85 % of the iterations).
        const unsigned entry = ((unsigned)(enc + j) << 16) | m;
#if SPH_LP_WADDR
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(wAddr + ((unsigned)min(ne, LP_ECAP) << 9)), "r"(entry) : "memory");  // row LP_ECAP is never read
        static_assert(LP_THREADS * 4 == 512, "row stride of ent");
#else
        ent[min(ne, LP_ECAP)][tid] = entry;  // row LP_ECAP is never read
#endif
        if (__builtin_expect(m != 0u && ne >= LP_ECAP, 0)) ovfE[min(ne - LP_ECAP, LP_EOVF - 1)] = entry;
        asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p add.s32 %0, %0, 1;\n\t}" : "+r"(ne) : "r"(m));  // ne += (m != 0)
        nh += __popc(m);
      }
    }
    n = nh;
  } else if (fits && VAR == 1) {
    // branch-free variant: the sign bits of (d2 - filter radius^2) are funnel-shifted into an 8-bit mask, the slots
    // of the aligned group that lie outside the cell are masked off, and the (rare) hits are appended by a loop over
    // the set bits
    const float2 nr2 = make_float2(-rfilt2, -rfilt2);
    int nh = 0;
#pragma unroll 1
    for (int k = 0; k < 8; k++) {
      const int b = cb[k], e = ce[k];
      const int enc = (k << 13) - b - 1;  // hit = enc + j + ffs = (k << 13) | rank in cell
#pragma unroll 1
      for (int j = b & ~7; j < e; j += 8) {  // 8 candidates per pass: three 256-bit loads
        float4 xa, xb, ya, yb, za, zb;
        const float4 *g = reinterpret_cast<const float4 *>(S + 3 * (size_t)j);
        ld256(g, xa, xb);
        ld256(g + 2, ya, yb);
        ld256(g + 4, za, zb);
        auto pair = [&](float x0, float x1, float y0, float y1, float z0, float z1) {
          const float2 ex = __fadd2_rn(make_float2(x0, x1), npx), ey = __fadd2_rn(make_float2(y0, y1), npy),
                       ez = __fadd2_rn(make_float2(z0, z1), npz);
          return __fadd2_rn(__ffma2_rn(ez, ez, __ffma2_rn(ey, ey, __fmul2_rn(ex, ex))), nr2);  // < 0: inside the filter
        };
        const float2 d0 = pair(xa.x, xa.y, ya.x, ya.y, za.x, za.y), d1 = pair(xa.z, xa.w, ya.z, ya.w, za.z, za.w);
        const float2 d2 = pair(xb.x, xb.y, yb.x, yb.y, zb.x, zb.y), d3 = pair(xb.z, xb.w, yb.z, yb.w, zb.z, zb.w);
        unsigned m = 0u;  // bit i = candidate j + i inside the filter radius
        m = __funnelshift_l(__float_as_uint(d3.y), m, 1); m = __funnelshift_l(__float_as_uint(d3.x), m, 1);
        m = __funnelshift_l(__float_as_uint(d2.y), m, 1); m = __funnelshift_l(__float_as_uint(d2.x), m, 1);
        m = __funnelshift_l(__float_as_uint(d1.y), m, 1); m = __funnelshift_l(__float_as_uint(d1.x), m, 1);
        m = __funnelshift_l(__float_as_uint(d0.y), m, 1); m = __funnelshift_l(__float_as_uint(d0.x), m, 1);
        // the group is aligned to 8: its ends may reach into the neighbouring cells
        m &= (0xffu << max(b - j, 0)) & (0xffu >> max(j + 8 - e, 0));
        const int v0 = enc + j;
        while (m) {
          const int i = __ffs(m);
          m &= m - 1u;
          const unsigned short hv = (unsigned short)(v0 + i);
          hits[min(nh, LP_CAP)][tid] = hv;  // row LP_CAP is never read
          if (nh >= LP_CAP) ovf[min(nh - LP_CAP, LP_OVF - 1)] = hv;
          nh++;
        }
      }
    }
    n = nh;
  } else if (fits) {
#pragma unroll 1
    for (int k = 0; k < 8; k++) {
      const int b = cb[k];
      const unsigned len = (unsigned)(ce[k] - b);
      const int enc = (k << 13) - b;  // hit = enc + j = (k << 13) | rank in cell
#pragma unroll 1
      for (int j = b & ~7; j < ce[k]; j += 8) {  // 8 candidates per pass: three 256-bit loads
        float4 xa, xb, ya, yb, za, zb;
        const float4 *g = reinterpret_cast<const float4 *>(S + 3 * (size_t)j);
        ld256(g, xa, xb);
        ld256(g + 2, ya, yb);
        ld256(g + 4, za, zb);
        auto pair = [&](float x0, float x1, float y0, float y1, float z0, float z1) {
          const float2 ex = __fadd2_rn(make_float2(x0, x1), npx), ey = __fadd2_rn(make_float2(y0, y1), npy),
                       ez = __fadd2_rn(make_float2(z0, z1), npz);
          return __ffma2_rn(ez, ez, __ffma2_rn(ey, ey, __fmul2_rn(ex, ex)));
        };
        const float2 d0 = pair(xa.x, xa.y, ya.x, ya.y, za.x, za.y), d1 = pair(xa.z, xa.w, ya.z, ya.w, za.z, za.w);
        const float2 d2 = pair(xb.x, xb.y, yb.x, yb.y, zb.x, zb.y), d3 = pair(xb.z, xb.w, yb.z, yb.w, zb.z, zb.w);
        // the group is aligned to 8: its ends may reach into the neighbouring cells, masked by the rank test
        auto keep = [&](float dd, int i) {
          if (dd <= rfilt2 && (unsigned)(j + i - b) < len) {
            if (n < LP_CAP) hits[n][tid] = (unsigned short)(enc + j + i);
            else ovf[min(n - LP_CAP, LP_OVF - 1)] = (unsigned short)(enc + j + i);
            n++;
          }
        };
        keep(d0.x, 0); keep(d0.y, 1); keep(d1.x, 2); keep(d1.y, 3);
        keep(d2.x, 4); keep(d2.y, 5); keep(d3.x, 6); keep(d3.y, 7);
      }
    }
  }
  if (!fits || n > LP_CAP + LP_OVF || ne > LP_ECAP + LP_EOVF) {
    if (nsStats) atomicAdd(&nsStats[3], 1u);   // particles that took the plain two-pass scan so far (sph_debug_counters)
    find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho, recA);
    return;
  }
  // exact distances of the hits (the particle itself is among them: skipped by index)
#if SPH_LP_CBS
  auto hit_index = [&](unsigned hv) { return cbS[hv >> 13][tid] + (int)(hv & 8191u); };
#else
  auto hit_index = [&](unsigned hv) { return cb[hv >> 13] + (int)(hv & 8191u); };
#endif
  auto dist2 = [&](int j) {
    const float4 q = ldg4(sp + j);
    const float ex = pos.x - q.x, ey = pos.y - q.y, ez = pos.z - q.z;
    return ex * ex + ey * ey + ez * ez;
  };
  // 30 bins x 7 bits, 8 bins per word: at most LP_CAP + LP_OVF - 1 = 127 other particles, no bin can overflow
  static_assert(LP_CAP + LP_OVF <= 128, "7-bit histogram bins");
  unsigned long long hist0 = 0ull, hist1 = 0ull, hist2 = 0ull, hist3 = 0ull;
  const float h2 = d.h * d.h;
  int found = 0;
  double density = 0.0;
  auto take = [&](int j, float dd) {  // :166-174 + the pcisph_computeDensity term, sphFluid.cl:505-507
    nb[nb_index(id, found)] = j;
    if (WITH_DENSITY) {
      const float r = sqrtf(dd) * d.simScale;
      const float rr = r * r;
      density += (double)((d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr));
    }
    found++;
  };
  // One pass builds the histogram AND selects as if the histogram left r_thr at its largest value (fewer than 32
  // candidates within h: the usual case at rest density). Only the other particles repeat the selection with their
  // smaller r_thr: the exact distances are gathered once for most lanes.
  const float r2open = ((float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS) *
                       ((float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS);
  auto hist_add = [&](float dd) {
    float sq;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(dd));
    int e = min((int)(sq * d.binScale), SPH_RADIUS_SEGMENTS);
    const int bin = e + ((dd >= d.binT[e + 1]) ? 1 : 0) - ((dd < d.binT[e]) ? 1 : 0);
    if (bin < SPH_RADIUS_SEGMENTS) {
      const int w = bin >> 3;
      const unsigned long long inc = 1ull << ((bin & 7) * 7);
      hist0 += (w == 0) ? inc : 0ull;
      hist1 += (w == 1) ? inc : 0ull;
      hist2 += (w == 2) ? inc : 0ull;
      hist3 += (w == 3) ? inc : 0ull;
    }
  };
  // VAR 2 can leave the radial histogram out of the first pass: with fewer than 32 candidates within h its running sum
  // never reaches 32 and r_thr keeps its largest value whatever the bins are (sphFluid.cl:310-324); a particle that
  // turns out to have 32 or more builds it in a pass of its own. That pays in a fluid near rest density (fluid_8m:
  // 2.93 -> 2.74 ms) and costs in dense matter, where most particles need the extra pass (worm scene 0.141 -> 0.155 ms).
  // The choice is made per STEP for all threads alike (a per-thread choice diverges inside the warps and costs both):
  // nsStats[1], set by k_hash_count from the number of particles that needed the histogram in the previous step.
  const bool histFirst = VAR != 2 || nsStats == nullptr || nsStats[1] != 0u;
  // third mode (nsStats[1] == 2, more than SPH_HIST2_PCT % needy): the first pass only builds the histogram and EVERY
  // particle selects in the second pass — the speculative selection (store, sqrt, fp64 add per hit) would be thrown
  // away by most of them
  const bool histOnly = VAR == 2 && nsStats != nullptr && nsStats[1] == 2u;
  int within = 0;
  // HIST is a compile-time constant per call site: the loop without the histogram must not keep its four 64-bit
  // accumulators alive (a run-time flag inside one loop cost 0.3 ms at 8 M particles)
  auto first_pass = [&](int j, auto HIST) {
    if (j == id) return;
    const float dd = dist2(j);
    if (dd <= h2) {
      within++;
      if (decltype(HIST)::value) hist_add(dd);
    }
    if (dd <= r2open && found < SPH_NB) take(j, dd);
  };
  // VAR 2: walk the hits of the entry list in scan order — a flat loop over the n hits (one per iteration for every
  // lane), fetching the next entry whenever the current mask is used up (entries are never empty)
  // The sorted index of candidate i of an entry is (index of the group's first candidate) + i, decoded once per entry.
  unsigned curMask = 0u;
  int curBase = 0, curEntry = 0;
  auto next_hit = [&]() -> int {
    if (curMask == 0u) {
      const unsigned e = (curEntry < LP_ECAP) ? ent[curEntry][tid] : ovfE[min(curEntry - LP_ECAP, LP_EOVF - 1)];
      curEntry++;
      curMask = e & 0xffffu;
#if SPH_LP_CBS
      curBase = cbS[e >> 29][tid] + (int)((e >> 16) & 8191u) - 16;  // sorted index of the group's first candidate
#else
      curBase = cb[e >> 29] + (int)((e >> 16) & 8191u) - 16;
#endif
    }
    const int i = __ffs(curMask) - 1;
    curMask &= curMask - 1u;
    return curBase + i;
  };
  if (VAR == 2) {
    if (histOnly) {  // most particles will select again with a smaller r_thr: no speculative selection at all
      for (int t = 0; t < n; t++) {
        const int j = next_hit();
        if (j == id) continue;
        const float dd = dist2(j);
        if (dd <= h2) { within++; hist_add(dd); }
      }
    } else if (histFirst) { for (int t = 0; t < n; t++) first_pass(next_hit(), LpYes()); }
    else { for (int t = 0; t < n; t++) first_pass(next_hit(), LpNo()); }
    if (!histFirst && within >= SPH_NB) {   // the histogram decides r_thr after all: build it
      curMask = 0u;
      curEntry = 0;
      for (int t = 0; t < n; t++) {
        const int j = next_hit();
        if (j == id) continue;
        const float dd = dist2(j);
        if (dd <= h2) hist_add(dd);
      }
    }
  } else {
    const int nList = min(n, LP_CAP);
    for (int t = 0; t < nList; t++) first_pass(hit_index(hits[t][tid]), LpYes());
    for (int t = LP_CAP; t < n; t++) first_pass(hit_index(ovf[t - LP_CAP]), LpYes());
  }
  int jthr = SPH_RADIUS_SEGMENTS;  // radius threshold (sphFluid.cl:310-324)
  if (within >= SPH_NB) {  // the bins hold at most `within` candidates: below 32 the running sum cannot reach 32
    int sum = 0;
    bool open = true;
#pragma unroll
    for (int jj = 0; jj < SPH_RADIUS_SEGMENTS; jj++) {
      const unsigned long long hw = jj < 8 ? hist0 : (jj < 16 ? hist1 : (jj < 24 ? hist2 : hist3));
      sum += (int)((hw >> ((jj & 7) * 7)) & 127ull);
      if (open && sum >= SPH_NB) { jthr = (sum == SPH_NB) ? jj : jj - 1; open = false; }
    }
  }
  if (jthr != SPH_RADIUS_SEGMENTS || histOnly) {  // a smaller r_thr: select again (histOnly: select at all;
                                                  // jthr == SEGMENTS gives r2 == r2open, the same expression)
    const float r_thr = (float)(jthr + 1) * d.h / SPH_RADIUS_SEGMENTS;
    const float r2 = r_thr * r_thr;
    found = 0;
    density = 0.0;
    curMask = 0u;
    curEntry = 0;
    for (int t = 0; t < n && found < SPH_NB; t++) {
      const int j = VAR == 2 ? next_hit() : hit_index((unsigned)((t < LP_CAP) ? hits[t][tid] : ovf[t - LP_CAP]));
      if (j == id) continue;
      const float dd = dist2(j);
      if (dd <= r2) take(j, dd);
    }
  }
  for (int k = found; k < SPH_NB; k++) nb[nb_index(id, k)] = SPH_NO_ID;
  if (WITH_DENSITY) {  // sphFluid.cl:514-517
    if (density < (double)d.hScaled6) density = (double)d.hScaled6;
    density *= d.massWpoly6;
    rho[id] = (float)density;
    if (recA) reinterpret_cast<float *>(recA)[8 * (size_t)id + 7] = (float)density;
  }
  if (VAR == 2 && nsStats) {  // statistic for the next step's choice: one atomic per warp
    const unsigned act = __activemask();
    const unsigned m = __ballot_sync(act, within >= SPH_NB);
    if ((threadIdx.x & 31) == __ffs(act) - 1 && m) atomicAdd(&nsStats[0], (unsigned)__popc(m));
  }
}

// ---------------------------------------------------------------- cooperative path (small scenes)
// EIGHT LANES PER PARTICLE, for scenes that do not fill the GPU with one thread per particle (the reference's own
// 61 k / 233 k particle scenes): there the list kernel lasts as long as ONE thread's walk over ~500-800 candidates
// and ~40-60 dependent gathers. Here the candidate sequence of a particle — its 8 cells in the reference's visiting
// order, cut into aligned groups of 16 — is dealt round-robin to the 8 lanes:
//   sweep   round r: lane l filters group 8r + l with the fp16 test of the list kernel; a prefix sum of the hit
//           counts over the 8 lanes (groups of a round are in scan order) places every hit at its scan-order
//           position in the particle's list in shared memory: one pass, ordered output;
//   select  lane l re-evaluates hits l, l + 8, ... exactly (fp32, reference operation order), the 8 partial radial
//           histograms are added with shuffles, every lane derives r_thr, and the "first 32 within r_thr" of
//           sphFluid.cl:166-174 is a ballot + prefix popcount over the hits in list order;
//   rho(t)  the terms of the selected neighbours go to shared memory by slot, lane 0 adds them in slot order (fp64).
// A particle with more than CP_CAP hits or a cell of more than 8191 particles takes find_neighbors_one (lane 0).
#ifndef SPH_COOP_MAX_N
#define SPH_COOP_MAX_N (1 << 17)   // measured: config A (61 k) 0.088 -> 0.059 ms, config B (233 k) 0.136 -> 0.179 ms
#endif
#ifndef CP_MINB
#define CP_MINB 3
#endif
#define CP_L 8
#define CP_THREADS 256
#define CP_P (CP_THREADS / CP_L)
#define CP_CAP 128

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(CP_THREADS, CP_MINB) k_find_neighbors_coop(DevParams d, const float4 *__restrict__ sp,
                                                                    const float *__restrict__ S,
                                                                    const unsigned *__restrict__ cellStart,
                                                                    int *__restrict__ nb, float *__restrict__ rho,
                                                                    float4 *__restrict__ recA) {
  __shared__ int hitJ[CP_P][CP_CAP + 1];      // sorted index of hit t (odd row stride)
  __shared__ float hitD[CP_P][CP_CAP + 1];    // its exact squared distance
  __shared__ float rhoT[CP_P][SPH_NB + 1];    // rho(t) term of the neighbour in slot s
  pdl_entry();
  const int t = threadIdx.x, pl = t / CP_L, l = t % CP_L;
  const int id = blockIdx.x * CP_P + pl;
  const bool live = id < d.N;
  const int gshift = (t & 31) - l;            // first lane of this particle's group inside the warp
  const float4 pos = sp[live ? id : 0];
  const int myCell = home_cell(d, pos.w);
  int cells[8];
  neighbor_cells(d, pos, myCell, cells);
  int cb[8], ce[8], ng[8];                    // first / end index, number of aligned 16-groups per visited cell
  int total = 0;
  bool fits = true;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int b = (int)cellStart[cells[k]], e = (int)cellStart[cells[k] + 1];
    cb[k] = b;
    ce[k] = e;
    ng[k] = e > b ? ((e + 15) >> 4) - (b >> 4) : 0;
    total += ng[k];
    fits = fits && (e - b < 8192);
  }
  // sweep
  const float3 up = cell_local_h(d, pos, 1.0f / d.h);
  const float sgx = (((pos.x - d.xmin) - (float)(int)(pos.x * d.cellSizeInv) * d.cellSize) < d.h) ? -2.0f : 2.0f;
  const float sgy = (((pos.y - d.ymin) - (float)(int)(pos.y * d.cellSizeInv) * d.cellSize) < d.h) ? -2.0f : 2.0f;
  const float sgz = (((pos.z - d.zmin) - (float)(int)(pos.z * d.cellSizeInv) * d.cellSize) < d.h) ? -2.0f : 2.0f;
  const __half2 nT = __float2half2_rn(-1.08984375f);  // see k_find_neighbors_list
  const __half *Sh = reinterpret_cast<const __half *>(S);
  int n = 0;
  const int rounds = (live && fits) ? (total + CP_L - 1) / CP_L : 0;
  // every lane of the warp runs the same number of rounds (shuffles inside): the maximum over the 4 particles
  int wrounds = rounds;
#pragma unroll
  for (int o = CP_L; o < 32; o <<= 1) wrounds = max(wrounds, __shfl_xor_sync(0xffffffffu, wrounds, o));
#pragma unroll 1
  for (int r = 0; r < wrounds; r++) {
    const int g = r * CP_L + l;
    unsigned m = 0u;
    int j0 = 0;
    if (r < rounds && g < total) {
      int k = 0, gl = g;                     // visited cell of group g and the group's index inside it
#pragma unroll
      for (int q = 0; q < 7; q++) {
        const bool next = (k == q) && gl >= ng[q];
        gl -= next ? ng[q] : 0;
        k += next ? 1 : 0;
      }
      int b = cb[0], e = ce[0];
#pragma unroll
      for (int q = 1; q < 8; q++) { b = (k == q) ? cb[q] : b; e = (k == q) ? ce[q] : e; }
      j0 = (b & ~15) + 16 * gl;
      const __half2 px = __float2half2_rn(up.x - (((0xB2 >> k) & 1) ? sgx : 0.0f));
      const __half2 py = __float2half2_rn(up.y - (((0xD4 >> k) & 1) ? sgy : 0.0f));
      const __half2 pz = __float2half2_rn(up.z - (((0xE8 >> k) & 1) ? sgz : 0.0f));
      uint4 xa, xb, ya, yb, za, zb;
      const __half *gp = Sh + 3 * (size_t)j0;
      ld256u(gp, xa, xb);
      ld256u(gp + 16, ya, yb);
      ld256u(gp + 32, za, zb);
      auto lane2 = [&](unsigned x, unsigned y, unsigned z, int rr) {
        const __half2 ex = __hsub2(*reinterpret_cast<const __half2 *>(&x), px);
        const __half2 ey = __hsub2(*reinterpret_cast<const __half2 *>(&y), py);
        const __half2 ez = __hsub2(*reinterpret_cast<const __half2 *>(&z), pz);
        const __half2 a = __hfma2(ez, ez, __hfma2(ey, ey, __hfma2(ex, ex, nT)));  // d2 - T: negative = hit
        m |= ((*reinterpret_cast<const unsigned *>(&a)) >> (15 - rr)) & (0x00010001u << rr);
      };
      lane2(xa.x, ya.x, za.x, 0); lane2(xa.y, ya.y, za.y, 1); lane2(xa.z, ya.z, za.z, 2); lane2(xa.w, ya.w, za.w, 3);
      lane2(xb.x, yb.x, zb.x, 4); lane2(xb.y, yb.y, zb.y, 5); lane2(xb.z, yb.z, zb.z, 6); lane2(xb.w, yb.w, zb.w, 7);
      // bits 0-7 = candidates j0..j0+7, bits 16-23 = candidates j0+8..j0+15: fold into 16 ascending bits
      m = (m & 0xffu) | ((m >> 8) & 0xff00u);
      // the group is aligned to 16: its ends may reach into the neighbouring cells
      m &= (0xffffu << max(b - j0, 0)) & (0xffffu >> max(j0 + 16 - e, 0));
    }
    // scan-order position of this lane's hits: hits of the lanes before it in this round come first
    const int cnt = __popc(m);
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < CP_L; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, inc, o);
      if (l >= o) inc += u;
    }
    int at = n + inc - cnt;
    n += __shfl_sync(0xffffffffu, inc, gshift + CP_L - 1);
    while (m) {
      const int i = __ffs(m) - 1;
      m &= m - 1u;
      if (at < CP_CAP) hitJ[pl][at] = j0 + i;
      at++;
    }
  }
  __syncwarp();
  const bool generic = live && (!fits || n > CP_CAP);
  if (generic) {
    if (l == 0) find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho, recA);
    n = 0;
  }
  const bool work = live && !generic;
  // exact distances and the radial histogram (30 bins x 7 bits, as in the list kernel)
  unsigned long long hist0 = 0ull, hist1 = 0ull, hist2 = 0ull, hist3 = 0ull;
  const float h2 = d.h * d.h;
#pragma unroll 1
  for (int q = l; q < n; q += CP_L) {
    const int j = hitJ[pl][q];
    const float4 c = ldg4(sp + j);
    const float ex = pos.x - c.x, ey = pos.y - c.y, ez = pos.z - c.z;
    const float dd = ex * ex + ey * ey + ez * ez;
    hitD[pl][q] = dd;
    if (j != id && dd <= h2) {
      float sq;
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(dd));
      const int e = min((int)(sq * d.binScale), SPH_RADIUS_SEGMENTS);
      const int bin = e + ((dd >= d.binT[e + 1]) ? 1 : 0) - ((dd < d.binT[e]) ? 1 : 0);
      if (bin < SPH_RADIUS_SEGMENTS) {
        const int w = bin >> 3;
        const unsigned long long inc = 1ull << ((bin & 7) * 7);
        hist0 += (w == 0) ? inc : 0ull;
        hist1 += (w == 1) ? inc : 0ull;
        hist2 += (w == 2) ? inc : 0ull;
        hist3 += (w == 3) ? inc : 0ull;
      }
    }
  }
#pragma unroll
  for (int o = 1; o < CP_L; o <<= 1) {  // at most CP_CAP - 1 = 127 other particles: no 7-bit bin can overflow
    hist0 += __shfl_xor_sync(0xffffffffu, hist0, o);
    hist1 += __shfl_xor_sync(0xffffffffu, hist1, o);
    hist2 += __shfl_xor_sync(0xffffffffu, hist2, o);
    hist3 += __shfl_xor_sync(0xffffffffu, hist3, o);
  }
  int jthr = SPH_RADIUS_SEGMENTS;  // radius threshold (sphFluid.cl:310-324)
  {
    int sum = 0;
    bool open = true;
#pragma unroll
    for (int jj = 0; jj < SPH_RADIUS_SEGMENTS; jj++) {
      const unsigned long long hw = jj < 8 ? hist0 : (jj < 16 ? hist1 : (jj < 24 ? hist2 : hist3));
      sum += (int)((hw >> ((jj & 7) * 7)) & 127ull);
      if (open && sum >= SPH_NB) { jthr = (sum == SPH_NB) ? jj : jj - 1; open = false; }
    }
  }
  const float r_thr = (float)(jthr + 1) * d.h / SPH_RADIUS_SEGMENTS;
  const float r2 = r_thr * r_thr;
  // first 32 within r_thr, in list (= scan) order
  int found = 0;
  int wn = n;
#pragma unroll
  for (int o = CP_L; o < 32; o <<= 1) wn = max(wn, __shfl_xor_sync(0xffffffffu, wn, o));
#pragma unroll 1
  for (int q0 = 0; q0 < wn; q0 += CP_L) {
    const int q = q0 + l;
    bool take = false;
    int j = 0;
    float dd = 0.f;
    if (q < n) {
      j = hitJ[pl][q];
      dd = hitD[pl][q];
      take = (j != id) && (dd <= r2);
    }
    const unsigned bits = (__ballot_sync(0xffffffffu, take) >> gshift) & ((1u << CP_L) - 1u);
    const int slot = found + __popc(bits & ((1u << l) - 1u));
    if (take && slot < SPH_NB) {
      nb[nb_index(id, slot)] = j;
      if (WITH_DENSITY) {  // pcisph_computeDensity term, sphFluid.cl:505-507
        const float rr0 = sqrtf(dd) * d.simScale;
        const float rr = rr0 * rr0;
        rhoT[pl][slot] = (d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr);
      }
    }
    found += __popc(bits);
  }
  __syncwarp();
  if (!work) return;
  found = min(found, SPH_NB);
  for (int k = found + l; k < SPH_NB; k += CP_L) nb[nb_index(id, k)] = SPH_NO_ID;
  if (WITH_DENSITY && l == 0) {  // sphFluid.cl:514-517
    double density = 0.0;
    for (int k = 0; k < found; k++) density += (double)rhoT[pl][k];
    if (density < (double)d.hScaled6) density = (double)d.hScaled6;
    density *= d.massWpoly6;
    rho[id] = (float)density;
    if (recA) reinterpret_cast<float *>(recA)[8 * (size_t)id + 7] = (float)density;
  }
}

// ---------------------------------------------------------------- tiled path (round-1 design, superseded by the list
// path above; kept as an independent implementation for the parity tests and A/B runs: SPH_NEIGHBOR_PATH=tiles|auto)
// One persistent kernel, k_neighbor_tiles. A CTA takes a HOME BLOCK = the two hash cells (2b, 2b+1),
// whose particles are one contiguous range of the sorted arrays, and
//   stage   copies the positions of the 4x3x3 = 36 surrounding cells (offsets -1..2 in id, -1..1 in
//           y-row and z-slab, searchCell wrap included) into shared memory as SoA x/y/z: every
//           candidate any home particle of the block can visit, read from HBM/L2 once, coalesced;
//   group   sorts the home particles by (home cell, octant): all particles of a cell that sit in the
//           same octant of it visit the same 8 cells in the same order (sphFluid.cl:248-308). Those 8
//           cells are 4 RUNS of two id-adjacent cells, and a run is contiguous in the staging area;
//   sweep   (P1) one warp per QUAD of home particles of a group: the quad's coordinates live in
//           registers as two sets of fp32x2 operands, lanes hold 32 consecutive staged candidates of a run, the four
//           squared distances come out of Blackwell's packed fp32x2 pipe, four ballots give the
//           32-bit hit masks of the chunk, lane w keeps mask word w, and the row of words goes to
//           shared memory. The test is a conservative FILTER (radius 31h/30 — the largest r_thr the
//           reference can choose — slightly inflated, FMA allowed): nothing inexact reaches the output;
//   select  (P2) one thread per home particle walks its row through the group's SEGMENT table (one
//           segment = the bits of one mask word that belong to one cell, listed in the reference's
//           visiting order), re-derives the exact distance of the ~35 hits with the reference's
//           operation order (from the staged copies), builds the 30-bin radial histogram, picks r_thr
//           (sphFluid.cl:310-324), keeps the first 32 hits within it (:166-174), accumulates rho(t) on
//           them (pcisph_computeDensity, :472-518) and writes the ids in the 32x32 tile layout.
// The two scan passes of the reference (2 x ~512 candidates per particle) become one packed sweep plus
// two passes over the hits; the mask rows never leave the SM. While a block is swept, warp 0 fetches the
// next block's cell table, so the cursor atomic and the cellStart loads are off the critical path.
// Particles whose neighbourhood does not fit the tables (staging area, FT_NCH mask words, FT_NSEG
// segments, FT_CAP hits) take find_neighbors_one, the generic scan.
#define FT_THREADS 160
#define FT_WARPS (FT_THREADS / 32)
#define FT_ROWS FT_THREADS        // home particles per batch (thread t of P2 owns particle b0 + t)
#define FT_GROUPS 16              // (home cell 0/1) x octant
#define FT_ROWPAD (FT_ROWS + 3 * FT_GROUPS)  // rows in group order, every group padded to a multiple of 4
#define FT_NCELL 36               // staged cells: id offset hx in [-1,2], y-row and z-slab offsets in [-1,1]
#define FT_CMAX 2560              // staged candidates (36 cells x 71); staged indices fit 12 bits
#define FT_NCH 24                 // mask words per row (<= 32: lane w of the sweeping warp keeps word w)
#define FT_MSTRIDE (FT_NCH + 1)   // odd row stride: P2's per-thread row walks are bank-conflict free
#define FT_NSEG 32                // segments per group
#define FT_CAP 48                 // hits kept per particle in P2

struct FtHalo {                   // cell table of one home block
  int gStart[FT_NCELL], cnt[FT_NCELL], sStart[FT_NCELL + 1];
  int hBeg, hMid, hEnd, bad, valid;
};

struct FtShared {
  float X[FT_CMAX + 32], Y[FT_CMAX + 32], Z[FT_CMAX + 32];  // +32: a run's last chunk may read past its end
  unsigned masks[FT_THREADS * FT_MSTRIDE];   // mask row of P2 thread t (particle b0 + t)
  unsigned short hits[FT_CAP][FT_THREADS];   // (k << 12) | staged index, scan order
  __align__(16) float px[FT_ROWPAD], py[FT_ROWPAD], pz[FT_ROWPAD];  // home coordinates, row order (FAR = no particle)
  __align__(4) unsigned char rowTid[FT_ROWPAD];                     // row -> owning P2 thread (255 = none)
  unsigned short wordOff[FT_GROUPS][FT_NCH];  // mask word w of a group: byte offset of its bit-0 candidate in X/Y/Z
  unsigned wordValid[FT_GROUPS][FT_NCH]; // ... and which of its 32 bits are real candidates
  unsigned segMask[FT_GROUPS][FT_NSEG];  // segment: bits of mask word (info >> 16) that belong to cell k
  unsigned segInfo[FT_GROUPS][FT_NSEG];  // (staged index of bit 0 + 32) | k << 13 | word << 16
  int dG[FT_GROUPS][8];                  // staged index + dG = sorted particle id, per visited cell k
  int nWords[FT_GROUPS], nSeg[FT_GROUPS], bad[FT_GROUPS];
  int bitLo0[FT_GROUPS];                 // row bit of the first particle of the group's own cell (k = 0)
  float binT[SPH_RADIUS_SEGMENTS + 2];
  FtHalo halo[2];
  int grpCount[FT_GROUPS], quadStart[FT_GROUPS + 1];
  unsigned char quadGroup[FT_ROWPAD / 4];
  int pairCursor;
};

// warp-level: take the next home block from the work list and build its cell table
__device__ __forceinline__ void ft_fetch_block(const DevParams &d, FtHalo &H, const unsigned *__restrict__ cellStart,
                                               const unsigned *__restrict__ cellList, int nBlocks,
                                               unsigned *__restrict__ cellCursor, int lane) {
  int ci = 0;
  if (lane == 0) ci = (int)atomicAdd(cellCursor, 1u);
  ci = __shfl_sync(0xffffffffu, ci, 0);
  if (ci >= nBlocks) {
    if (lane == 0) H.valid = 0;
    return;
  }
  const int base = (int)cellList[ci] & ~1;
  int c0 = 0, c1 = 0;
  {
    const int t = lane;
    const int nc = search_cell(base, (t & 3) - 1 + ((t >> 2) % 3 - 1) * d.gx + (t / 12 - 1) * d.gx * d.gy, d.G);
    const int b = (int)cellStart[nc];
    c0 = (int)cellStart[nc + 1] - b;
    H.gStart[t] = b;
    H.cnt[t] = c0;
  }
  if (lane < FT_NCELL - 32) {
    const int t = 32 + lane;
    const int nc = search_cell(base, (t & 3) - 1 + ((t >> 2) % 3 - 1) * d.gx + (t / 12 - 1) * d.gx * d.gy, d.G);
    const int b = (int)cellStart[nc];
    c1 = (int)cellStart[nc + 1] - b;
    H.gStart[t] = b;
    H.cnt[t] = c1;
  }
  int inc = c0;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += u;
  }
  const int tot32 = __shfl_sync(0xffffffffu, inc, 31);
  H.sStart[lane] = inc - c0;
  int inc1 = c1;
#pragma unroll
  for (int o = 1; o < 4; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, inc1, o);
    if (lane >= o) inc1 += u;
  }
  if (lane < FT_NCELL - 32) H.sStart[32 + lane] = tot32 + inc1 - c1;
  if (lane == FT_NCELL - 32 - 1) {
    H.sStart[FT_NCELL] = tot32 + inc1;
    H.bad = (tot32 + inc1 > FT_CMAX) ? 1 : 0;
  }
  if (lane == 0) {
    H.hBeg = (int)cellStart[base];
    H.hMid = (int)cellStart[base + 1];
    H.hEnd = (int)cellStart[min(base + 2, d.G)];
    H.valid = 1;
  }
}

// geometry of group g = (home cell hc, octant): its 4 runs, ordered (0,0), (sy,0), (0,sz), (sy,sz)
struct FtGroupGeom {
  int sx, sy, sz, hc, hxlo;
  int wOff[5];  // first mask word of each run
  __device__ __forceinline__ void init(int g, const FtHalo &H) {
    sx = (g & 1) ? 1 : -1; sy = (g & 2) ? 1 : -1; sz = (g & 4) ? 1 : -1; hc = g >> 3;
    hxlo = hc + ((sx < 0) ? -1 : 0);  // the group's two id-adjacent cells are hxlo, hxlo + 1
    int w = 0;
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int tlo = run_slot(r);
      wOff[r] = w;
      w += (H.cnt[tlo] + H.cnt[tlo + 1] + 31) >> 5;
    }
    wOff[4] = w;
  }
  __device__ __forceinline__ int run_slot(int r) const {
    const int oy = (r & 1) ? sy : 0, oz = (r & 2) ? sz : 0;
    return (hxlo + 1) + (oy + 1) * 4 + (oz + 1) * 12;
  }
  // cell k in the reference's visiting order (:248-308): own, x, y, z, xy, xz, yz, xyz -> staged slot, first row bit
  __device__ __forceinline__ void cell(int k, const FtHalo &H, int &t, int &lo) const {
    const int kx = (0xB2 >> k) & 1, ky = (0xD4 >> k) & 1, kz = (0xE8 >> k) & 1;
    const int hx = hc + kx * sx;
    t = (hx + 1) + (ky * sy + 1) * 4 + (kz * sz + 1) * 12;
    lo = wOff[ky + 2 * kz] * 32 + ((hx == hxlo) ? 0 : H.cnt[t - 1]);
  }
};

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(FT_THREADS, 3) k_neighbor_tiles(DevParams d, const float4 *__restrict__ sp,
                                                                  const unsigned *__restrict__ cellStart,
                                                                  const unsigned *__restrict__ cellList,
                                                                  const unsigned *__restrict__ cellListCount,
                                                                  unsigned *__restrict__ cellCursor,
                                                                  int *__restrict__ nb, float *__restrict__ rho,
                                                                  float4 *__restrict__ recA, int *__restrict__ deferred) {
  extern __shared__ __align__(16) unsigned char ft_smem[];
  FtShared &S = *reinterpret_cast<FtShared *>(ft_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nBlocks = (int)*cellListCount;
  const float rmaxf = (float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS;  // largest r_thr (j == 30)
  const float rfilt2 = rmaxf * rmaxf * 1.0001f;
  const float FAR = 1.0e30f;
  const float h2 = d.h * d.h;
  if (tid < SPH_RADIUS_SEGMENTS + 2) S.binT[tid] = d.binT[tid];
  if (tid < 32) { S.X[FT_CMAX + tid] = FAR; S.Y[FT_CMAX + tid] = FAR; S.Z[FT_CMAX + tid] = FAR; }
  if (tid < FT_GROUPS) S.grpCount[tid] = 0;
  if (tid == 0) S.pairCursor = 0;
  for (int r = tid; r < FT_ROWPAD; r += FT_THREADS) {
    S.px[r] = FAR; S.py[r] = FAR; S.pz[r] = FAR;
    S.rowTid[r] = 255;
  }
  if (warp == 0) ft_fetch_block(d, S.halo[0], cellStart, cellList, nBlocks, cellCursor, lane);
  int cur = 0;
  while (true) {
    __syncthreads();  // (A) halo[cur] is ready; the previous block's P2 is over
    const FtHalo &H = S.halo[cur];
    if (!H.valid) break;
    const int hBeg = H.hBeg, hMid = H.hMid, hEnd = H.hEnd, blockBad = H.bad;
    // the first batch's home particle of this thread: requested now, needed after the tables and the staging
    const float4 posFirst = (hBeg + tid < hEnd) ? ldg4(sp + hBeg + tid) : f4(0.f, 0.f, 0.f, 0.f);
    // ---- group tables of the block (depend on the cell table only)
    if (tid < FT_GROUPS * 4) {  // thread (g, r): the mask words of run r
      const int g = tid >> 2, r = tid & 3;
      FtGroupGeom q;
      q.init(g, H);
      const int tlo = q.run_slot(r);
      const int s0 = H.sStart[tlo], len = H.cnt[tlo] + H.cnt[tlo + 1];
      int w = q.wOff[r];
      for (int c = 0; c * 32 < len; c++, w++) {
        if (w < FT_NCH) {
          S.wordOff[g][w] = (unsigned short)((s0 + c * 32) * 4);
          S.wordValid[g][w] = 0xffffffffu >> (32 - min(32, len - c * 32));
        }
      }
      if (r == 3) S.nWords[g] = q.wOff[4];
    }
    if (tid < FT_GROUPS * 8) {  // thread (g, k): the segments of visited cell k
      const int g = tid >> 3, k = tid & 7;
      FtGroupGeom q;
      q.init(g, H);
      int sg = 0;
      for (int kk = 0; kk < k; kk++) {
        int t, lo;
        q.cell(kk, H, t, lo);
        const int n = H.cnt[t];
        if (n) sg += ((lo + n - 1) >> 5) - (lo >> 5) + 1;
      }
      int t, lo;
      q.cell(k, H, t, lo);
      const int n = H.cnt[t], hi = lo + n;
      S.dG[g][k] = H.gStart[t] - H.sStart[t];
      if (k == 0) S.bitLo0[g] = lo;
      if (n) {
        for (int w = lo >> 5; w <= (hi - 1) >> 5; w++, sg++) {
          const int a = max(lo, w * 32) - w * 32, b = min(hi, w * 32 + 32) - w * 32;  // bits [a, b) of word w
          if (sg < FT_NSEG) {
            S.segMask[g][sg] = (0xffffffffu << a) & (0xffffffffu >> (32 - b));
            S.segInfo[g][sg] = (unsigned)(H.sStart[t] - lo + w * 32 + 32) | ((unsigned)k << 13) | ((unsigned)w << 16);
          }
        }
      }
      if (k == 7) {
        S.nSeg[g] = sg;
        S.bad[g] = (blockBad || q.wOff[4] > FT_NCH || sg > FT_NSEG) ? 1 : 0;
      }
    }
    // ---- stage the positions of the 36 cells: two cells per pass, six loads in flight per lane
    if (!blockBad) {
      auto put = [&](int s, const float4 &q) { S.X[s] = q.x; S.Y[s] = q.y; S.Z[s] = q.z; };
#pragma unroll 1
      for (int t = warp; t < FT_NCELL; t += 2 * FT_WARPS) {
        const int t2 = t + FT_WARPS;
        const int g0 = H.gStart[t] + lane, n0 = H.cnt[t] - lane, s0 = H.sStart[t] + lane;
        int g1 = 0, n1 = 0, s1 = 0;
        if (t2 < FT_NCELL) { g1 = H.gStart[t2] + lane; n1 = H.cnt[t2] - lane; s1 = H.sStart[t2] + lane; }
        float4 q[6];
#pragma unroll
        for (int c = 0; c < 3; c++) {
          if (32 * c < n0) q[c] = ldg4(sp + g0 + 32 * c);
          if (32 * c < n1) q[3 + c] = ldg4(sp + g1 + 32 * c);
        }
#pragma unroll
        for (int c = 0; c < 3; c++) {
          if (32 * c < n0) put(s0 + 32 * c, q[c]);
          if (32 * c < n1) put(s1 + 32 * c, q[3 + c]);
        }
        for (int i = 96; i < n0; i += 32) put(s0 + i, ldg4(sp + g0 + i));
        for (int i = 96; i < n1; i += 32) put(s1 + i, ldg4(sp + g1 + i));
      }
    }
    for (int b0 = hBeg; b0 < hEnd; b0 += FT_ROWS) {
      const int hb = min(FT_ROWS, hEnd - b0);
      const int id = b0 + tid;
      if (b0 != hBeg) __syncthreads();  // (A') the previous batch's P2 is over
      // ---- group the home particles by (home cell, octant); octant = which half of the 2h cell, per axis (:253-271)
      float4 pos = f4(0.f, 0.f, 0.f, 0.f);
      int g = 0, rankInG = 0;
      if (tid < hb) {
        pos = (b0 == hBeg) ? posFirst : sp[id];
        float cfx = (float)(int)(pos.x * d.cellSizeInv) * d.cellSize;
        float cfy = (float)(int)(pos.y * d.cellSizeInv) * d.cellSize;
        float cfz = (float)(int)(pos.z * d.cellSizeInv) * d.cellSize;
        g = ((((pos.x - d.xmin) - cfx) < d.h) ? 0 : 1) | ((((pos.y - d.ymin) - cfy) < d.h) ? 0 : 2) |
            ((((pos.z - d.zmin) - cfz) < d.h) ? 0 : 4) | ((id >= hMid) ? 8 : 0);
        rankInG = atomicAdd(&S.grpCount[g], 1);
      }
      __syncthreads();  // (C) group sizes are final; tables and staging are visible
      if (tid < hb) {   // rows in group order, every group padded to a multiple of 4 (one QUAD = one sweep)
        int gs = 0;
#pragma unroll
        for (int q = 0; q < FT_GROUPS; q++) gs += (q < g) ? ((S.grpCount[q] + 3) & ~3) : 0;
        const int row = gs + rankInG;
        S.px[row] = pos.x; S.py[row] = pos.y; S.pz[row] = pos.z;
        S.rowTid[row] = (unsigned char)tid;
      }
      if (tid < FT_GROUPS) {
        int qs = 0;
        for (int q = 0; q < tid; q++) qs += (S.grpCount[q] + 3) >> 2;
        const int nq = (S.grpCount[tid] + 3) >> 2;
        S.quadStart[tid] = qs;
        if (tid == FT_GROUPS - 1) S.quadStart[FT_GROUPS] = qs + nq;
        for (int q = 0; q < nq; q++) S.quadGroup[qs + q] = (unsigned char)tid;
      }
      __syncthreads();  // (E) rows and quad tables are ready
      if (tid < FT_GROUPS) S.grpCount[tid] = 0;  // consumed; clean for the next batch
      // ---- while the block is swept, warp 0 fetches the next block's cell table
      if (warp == 0 && b0 == hBeg) ft_fetch_block(d, S.halo[cur ^ 1], cellStart, cellList, nBlocks, cellCursor, lane);
      // ---- P1 sweep: one warp per quad of home particles, quads handed out through a shared cursor
      const int totalQuads = S.quadStart[FT_GROUPS];
      const unsigned char *xb = reinterpret_cast<const unsigned char *>(S.X) + lane * 4;
      const unsigned char *yb = reinterpret_cast<const unsigned char *>(S.Y) + lane * 4;
      const unsigned char *zb = reinterpret_cast<const unsigned char *>(S.Z) + lane * 4;
#pragma unroll 1
      while (true) {
        int t = 0;
        if (lane == 0) t = atomicAdd(&S.pairCursor, 1);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= totalQuads) break;
        const int gg = S.quadGroup[t];
        if (S.bad[gg]) continue;
        const float2 hxa = reinterpret_cast<const float2 *>(S.px)[2 * t], hxb = reinterpret_cast<const float2 *>(S.px)[2 * t + 1];
        const float2 hya = reinterpret_cast<const float2 *>(S.py)[2 * t], hyb = reinterpret_cast<const float2 *>(S.py)[2 * t + 1];
        const float2 hza = reinterpret_cast<const float2 *>(S.pz)[2 * t], hzb = reinterpret_cast<const float2 *>(S.pz)[2 * t + 1];
        const unsigned short *wtab = S.wordOff[gg];
        const int nch = S.nWords[gg];
        unsigned keep0 = 0, keep1 = 0, keep2 = 0, keep3 = 0;
#pragma unroll 4
        for (int c = 0; c < nch; c++) {
          const unsigned off = wtab[c];
          const float cx = *reinterpret_cast<const float *>(xb + off);
          const float cy = *reinterpret_cast<const float *>(yb + off);
          const float cz = *reinterpret_cast<const float *>(zb + off);
          const float2 ncx = make_float2(-cx, -cx), ncy = make_float2(-cy, -cy), ncz = make_float2(-cz, -cz);
          const float2 exa = __fadd2_rn(hxa, ncx), eya = __fadd2_rn(hya, ncy), eza = __fadd2_rn(hza, ncz);
          const float2 exb = __fadd2_rn(hxb, ncx), eyb = __fadd2_rn(hyb, ncy), ezb = __fadd2_rn(hzb, ncz);
          const float2 da = __ffma2_rn(eza, eza, __ffma2_rn(eya, eya, __fmul2_rn(exa, exa)));
          const float2 db = __ffma2_rn(ezb, ezb, __ffma2_rn(eyb, eyb, __fmul2_rn(exb, exb)));
          const unsigned m0 = __ballot_sync(0xffffffffu, da.x <= rfilt2);
          const unsigned m1 = __ballot_sync(0xffffffffu, da.y <= rfilt2);
          const unsigned m2 = __ballot_sync(0xffffffffu, db.x <= rfilt2);
          const unsigned m3 = __ballot_sync(0xffffffffu, db.y <= rfilt2);
          if (lane == c) { keep0 = m0; keep1 = m1; keep2 = m2; keep3 = m3; }
        }
        if (lane < nch) {  // mask rows are indexed by the owning P2 thread; padding rows have none
          const unsigned valid = S.wordValid[gg][lane];
          const uchar4 own = reinterpret_cast<const uchar4 *>(S.rowTid)[t];
          if (own.x != 255) S.masks[own.x * FT_MSTRIDE + lane] = keep0 & valid;
          if (own.y != 255) S.masks[own.y * FT_MSTRIDE + lane] = keep1 & valid;
          if (own.z != 255) S.masks[own.z * FT_MSTRIDE + lane] = keep2 & valid;
          if (own.w != 255) S.masks[own.w * FT_MSTRIDE + lane] = keep3 & valid;
        }
      }
      __syncthreads();  // (F) all mask rows are written
      // ---- P2 select: one thread per home particle
      if (tid == 0) S.pairCursor = 0;
      for (int r = tid; r < FT_ROWPAD; r += FT_THREADS) {  // rows back to "no particle" for the next batch
        S.px[r] = FAR; S.py[r] = FAR; S.pz[r] = FAR;
        S.rowTid[r] = 255;
      }
      if (tid >= hb) continue;
      // neighbourhoods the tables cannot hold go to the deferred list: k_neighbors_deferred scans them with
      // one thread each, spread over the whole GPU, instead of stalling this CTA (cellCursor[1] = list length)
      if (S.bad[g]) { deferred[atomicAdd(cellCursor + 1, 1u)] = id; continue; }
      unsigned *mrow = S.masks + tid * FT_MSTRIDE;
      {  // j == id (:148): the particle's own bit sits in its own cell, the first one visited
        const int selfBit = S.bitLo0[g] + id - ((g & 8) ? hMid : hBeg);
        mrow[selfBit >> 5] &= ~(1u << (selfBit & 31));
      }
      // walk the row through the segment table -> hit list in scan order
      int nh = 0;
      {
        const unsigned *segM = S.segMask[g], *segI = S.segInfo[g];
        const int nSeg = S.nSeg[g];
#pragma unroll 1
        for (int sg = 0; sg < nSeg; sg++) {
          const unsigned info = segI[sg];
          unsigned m = mrow[info >> 16] & segM[sg];
          const unsigned hbase = (((info >> 13) & 7u) << 12) + (info & 0x1fffu) - 32u;  // (k << 12) | staged index of bit 0
          while (m) {
            const unsigned hv = hbase + (unsigned)__ffs(m) - 1u;
            m &= m - 1;
            if (nh < FT_CAP) S.hits[nh][tid] = (unsigned short)hv;
            nh++;
          }
        }
      }
      if (nh > FT_CAP) { deferred[atomicAdd(cellCursor + 1, 1u)] = id; continue; }
      // exact distances + 30-bin radial histogram of everything within h (:157-161), in registers:
      // 3 words x 10 bins x 6 bits (a particle keeps <= FT_CAP < 64 hits)
      unsigned long long hist0 = 0ull, hist1 = 0ull, hist2 = 0ull;
      auto hist_add = [&](float dd) {
        if (dd <= h2) {
          // bin (int)(sqrtf(dd)*30/h): approximate guess, made exact with the host-built thresholds
          float sq;
          asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(dd));
          int e = (int)(sq * d.binScale);
          e = min(e, SPH_RADIUS_SEGMENTS);
          const int b = e + ((dd >= S.binT[e + 1]) ? 1 : 0) - ((dd < S.binT[e]) ? 1 : 0);
          if (b < SPH_RADIUS_SEGMENTS) {
            const int w = (b >= 10 ? 1 : 0) + (b >= 20 ? 1 : 0);
            const unsigned long long inc = 1ull << ((b - 10 * w) * 6);
            hist0 += (w == 0) ? inc : 0ull;
            hist1 += (w == 1) ? inc : 0ull;
            hist2 += (w == 2) ? inc : 0ull;
          }
        }
      };
      {  // two hits per pass: their loads and distance chains are independent
        int t = 0;
        for (; t + 1 < nh; t += 2) {
          const int s0 = S.hits[t][tid] & 0xfff, s1 = S.hits[t + 1][tid] & 0xfff;
          const float ex0 = pos.x - S.X[s0], ey0 = pos.y - S.Y[s0], ez0 = pos.z - S.Z[s0];
          const float ex1 = pos.x - S.X[s1], ey1 = pos.y - S.Y[s1], ez1 = pos.z - S.Z[s1];
          const float dd0 = ex0 * ex0 + ey0 * ey0 + ez0 * ez0;
          const float dd1 = ex1 * ex1 + ey1 * ey1 + ez1 * ez1;
          hist_add(dd0);
          hist_add(dd1);
        }
        if (t < nh) {
          const int s0 = S.hits[t][tid] & 0xfff;
          const float ex0 = pos.x - S.X[s0], ey0 = pos.y - S.Y[s0], ez0 = pos.z - S.Z[s0];
          hist_add(ex0 * ex0 + ey0 * ey0 + ez0 * ez0);
        }
      }
      // radius threshold (:310-324)
      int j = SPH_RADIUS_SEGMENTS;
      {
        int sum = 0;
        bool open = true;
#pragma unroll
        for (int jj = 0; jj < SPH_RADIUS_SEGMENTS; jj++) {
          const unsigned long long hw = jj < 10 ? hist0 : (jj < 20 ? hist1 : hist2);
          sum += (int)((hw >> ((jj % 10) * 6)) & 63ull);
          if (open && sum >= SPH_NB) { j = (sum == SPH_NB) ? jj : jj - 1; open = false; }
        }
      }
      const float r_thr = (float)(j + 1) * d.h / SPH_RADIUS_SEGMENTS;
      const float r2 = r_thr * r_thr;
      // first 32 hits within r_thr, in scan order (:166-174) + rho(t)
      int found = 0;
      double density = 0.0;
      auto take = [&](unsigned hv, float dd) {
        if (found < SPH_NB && dd <= r2) {
          S.hits[found][tid] = (unsigned short)hv;  // in-place compaction: found <= index of hv
          if (WITH_DENSITY) {
            float r = sqrtf(dd) * d.simScale;
            float rr = r * r;
            density += (double)((d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr));
          }
          found++;
        }
      };
      {  // two hits per pass (independent loads and distance chains), taken in scan order
        int t = 0;
        for (; t + 1 < nh && found < SPH_NB; t += 2) {
          const unsigned hv0 = S.hits[t][tid], hv1 = S.hits[t + 1][tid];
          const int s0 = hv0 & 0xfff, s1 = hv1 & 0xfff;
          const float ex0 = pos.x - S.X[s0], ey0 = pos.y - S.Y[s0], ez0 = pos.z - S.Z[s0];
          const float ex1 = pos.x - S.X[s1], ey1 = pos.y - S.Y[s1], ez1 = pos.z - S.Z[s1];
          const float dd0 = ex0 * ex0 + ey0 * ey0 + ez0 * ez0;
          const float dd1 = ex1 * ex1 + ey1 * ey1 + ez1 * ez1;
          take(hv0, dd0);
          take(hv1, dd1);
        }
        if (t < nh && found < SPH_NB) {
          const unsigned hv0 = S.hits[t][tid];
          const int s0 = hv0 & 0xfff;
          const float ex0 = pos.x - S.X[s0], ey0 = pos.y - S.Y[s0], ez0 = pos.z - S.Z[s0];
          take(hv0, ex0 * ex0 + ey0 * ey0 + ez0 * ez0);
        }
      }
#pragma unroll 8
      for (int q = 0; q < SPH_NB; q++) {
        const unsigned hv = S.hits[q][tid];
        nb[nb_index(id, q)] = (q < found) ? (int)(hv & 0xfffu) + S.dG[g][hv >> 12] : SPH_NO_ID;
      }
      if (WITH_DENSITY) {
        if (density < (double)d.hScaled6) density = (double)d.hScaled6;
        density *= d.massWpoly6;
        rho[id] = (float)density;
        reinterpret_cast<float *>(recA)[8 * (size_t)id + 7] = (float)density;  // rho slot of the gather record
      }
    }
    cur ^= 1;
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

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(128) k_neighbors_deferred(DevParams d, const float4 *__restrict__ sp,
                                                            const unsigned *__restrict__ cellStart,
                                                            const unsigned *__restrict__ cellCursor,
                                                            const int *__restrict__ deferred, int *__restrict__ nb,
                                                            float *__restrict__ rho, float4 *__restrict__ recA) {
  const int n = (int)cellCursor[1];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    find_neighbors_one<WITH_DENSITY>(d, deferred[i], sp, cellStart, nb, rho, recA);
}

static int ft_ctas_per_sm = 0;
template <bool WD>
static void launch_tiles(sph_ctx *c) {
  if (!ft_ctas_per_sm) {
    cudaFuncSetAttribute(k_neighbor_tiles<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FtShared));
    cudaFuncSetAttribute(k_neighbor_tiles<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FtShared));
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_neighbor_tiles<true>, FT_THREADS, sizeof(FtShared));
    ft_ctas_per_sm = occ < 1 ? 1 : occ;
  }
  cudaMemsetAsync(c->cellCursor, 0, 2 * sizeof(unsigned), c->stream);  // [0] work cursor, [1] deferred count
  k_neighbor_tiles<WD><<<c->numSMs * ft_ctas_per_sm, FT_THREADS, sizeof(FtShared), c->stream>>>(
      c->d, c->sortedPosition, c->cellStart, c->cellList, c->cellListCount, c->cellCursor, c->nbIds, c->rho, c->recA, c->fnDeferred);
  k_neighbors_deferred<WD><<<c->numSMs * 8, 128, 0, c->stream>>>(c->d, c->sortedPosition, c->cellStart, c->cellCursor,
                                                                 c->fnDeferred, c->nbIds, c->rho, c->recA);
}

void launch_find_neighbors(sph_ctx *c, bool with_density) {
  // Default: the list kernel (below). SPH_NEIGHBOR_PATH=tiles|generic force the earlier implementations, =auto the
  // round-1 policy (tiles for fluids of >= 131 072 particles, watched by their own statistics).
  // the tiled kernel assumes the 36 cells around a home block are distinct: true for grids >= 4 x 3 x 3.
  // Below ~1000 home blocks (two per resident CTA) its persistent CTAs do not balance and the plain scan wins
  // (config A, 61 k particles: 0.17 ms against 0.32 ms).
  const bool tilesFit = c->d.gx >= 4 && c->d.gy >= 3 && c->d.gz >= 3;
  bool useTiles = tilesFit && !c->force_generic_neighbors && !c->force_list_neighbors && (c->force_tiled_neighbors || (c->auto_tiles && c->d.N >= 131072));
  const bool adaptive = useTiles && !c->force_tiled_neighbors;
  bool wantStats = false, syncStats = false;
  if (adaptive) {
    // The tile kernel pays a fixed price per home block (cell table, staging of 36 cells) and a high one per
    // deferred particle. Scenes with many thinly populated blocks (the worm scene: long box, wall-only cells; 0.25 ms
    // with the plain scan against 0.46 ms) or many overflowing neighbourhoods are faster with the plain scan. Both
    // paths give identical lists, so switching has no side effect. The statistics of a tile launch come back
    // without a sync (except once, after the very first launch) and are acted upon when they arrive.
    if (c->deferredPending && cudaEventQuery(c->evDeferred) == cudaSuccess) {
      c->deferredPending = false;
      const double deferred = (double)c->hostDeferred[0], blocks = (double)c->hostDeferred[1], n = (double)c->d.N;
      c->nbMode = (deferred > 0.10 * n || n < 64.0 * blocks) ? 2 : 1;
      if (c->nbMode == 2) c->genericHold = 64;
    }
    if (c->nbMode == 0) {            // first search of this context: tiles, then look at the statistics at once
      wantStats = syncStats = true;
    } else if (c->nbMode == 1) {     // tiles suit the fluid: keep watching
      wantStats = !c->deferredPending;
    } else {                         // plain scan; every 64 steps one probing tile launch
      useTiles = false;
      if (!c->deferredPending && --c->genericHold <= 0) { useTiles = wantStats = true; c->genericHold = 64; }
    }
  }
  if (!useTiles && !c->force_generic_neighbors) {
    const int n = c->d.N;
    // the half-precision sweep assumes that a true neighbour found in visited cell k sits in the cell at spatial
    // offset (kx sx, ky sy, kz sz): true when no other offset in {-2..2}^3 maps to the same (masked, wrapped) id
    const long long span = 2 + 2 * (long long)c->d.gx + 2 * (long long)c->d.gx * c->d.gy;
    const bool halfOk = c->d.gx >= 4 && c->d.gy >= 4 && c->d.gz >= 4 && span < (long long)c->d.G &&
                        span < (long long)c->d.mask + 1;
    const int var = (c->list_variant == 2 && !halfOk) ? 1 : c->list_variant;
    if (var == 2)
      launch_k(c, k_soa_positions_half, dim3((n + 32 + 255) / 256), dim3(256), c->d, c->sortedPosition, reinterpret_cast<__half *>(c->soa));
    else
      launch_k(c, k_soa_positions, dim3((n + 16 + 255) / 256), dim3(256), n, c->sortedPosition, c->soa);
    // small scenes: eight lanes per particle (SPH_NEIGHBOR_PATH=coop forces it at any size, =list forbids it)
    const bool coop = var == 2 && (c->force_coop_neighbors || (!c->force_list_neighbors && n <= SPH_COOP_MAX_N));
    if (coop) {
      const int cb = (n + CP_P - 1) / CP_P;
      if (with_density)
        launch_k(c, k_find_neighbors_coop<true>, dim3(cb), dim3(CP_THREADS), c->d, c->sortedPosition, c->soa, c->cellStart, c->nbIds, c->rho, c->recA);
      else
        launch_k(c, k_find_neighbors_coop<false>, dim3(cb), dim3(CP_THREADS), c->d, c->sortedPosition, c->soa, c->cellStart, c->nbIds, c->rho, (float4 *)nullptr);
      return;
    }
    const int blocks = (n + LP_THREADS - 1) / LP_THREADS;
#define LP_LAUNCH(WD, V, REC)                                                                                      \
  launch_k(c, k_find_neighbors_list<WD, V>, dim3(blocks), dim3(LP_THREADS), c->d, c->sortedPosition, c->soa, c->cellStart, \
           c->nbIds, c->rho, (float4 *)(REC), c->nsStats)
    if (var == 0) { if (with_density) LP_LAUNCH(true, 0, c->recA); else LP_LAUNCH(false, 0, nullptr); }
    else if (var == 1) { if (with_density) LP_LAUNCH(true, 1, c->recA); else LP_LAUNCH(false, 1, nullptr); }
    else { if (with_density) LP_LAUNCH(true, 2, c->recA); else LP_LAUNCH(false, 2, nullptr); }
#undef LP_LAUNCH
    return;
  }
  if (!useTiles) {
    int blocks = (c->d.N + 127) / 128;
    if (with_density)
      launch_k(c, k_find_neighbors_generic<true>, dim3(blocks), dim3(128), c->d, c->sortedPosition, c->cellStart, c->nbIds, c->rho, c->recA);
    else
      launch_k(c, k_find_neighbors_generic<false>, dim3(blocks), dim3(128), c->d, c->sortedPosition, c->cellStart, c->nbIds, c->rho, (float4 *)nullptr);
    return;
  }
  // persistent CTAs (one resident wave); home blocks handed out through an atomic cursor over cellList
  if (with_density) launch_tiles<true>(c); else launch_tiles<false>(c);
  if (wantStats) {
    if (!c->hostDeferred) {
      if (cudaMallocHost((void **)&c->hostDeferred, 2 * sizeof(unsigned)) != cudaSuccess) { c->hostDeferred = nullptr; return; }
      cudaEventCreateWithFlags(&c->evDeferred, cudaEventDisableTiming);
    }
    cudaMemcpyAsync(c->hostDeferred, c->cellCursor + 1, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream);
    cudaMemcpyAsync(c->hostDeferred + 1, c->cellListCount, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream);
    cudaEventRecord(c->evDeferred, c->stream);
    c->deferredPending = true;
    if (syncStats) {
      cudaEventSynchronize(c->evDeferred);
      c->nbMode = 1;  // evaluated by the query at the next call
    }
  }
}

void launch_materialize_neighbor_map(sph_ctx *c, float2 *out) {
  size_t n = (size_t)c->d.N * SPH_NB;
  k_materialize_neighbor_map<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d, c->sortedPosition, c->nbIds, out);
}

// Force the (lazily loaded, CUDA >= 12.2) code of this file's step kernels onto the device. A kernel that spins on a
// neighbour's delivery flag (sph_slab.cu) must never be resident while the host loads another kernel for the first
// time — module loading synchronises the context and would wait for the spinning kernel forever.
static void preload(const void *const *fns, int n) {
  cudaFuncAttributes a;
  for (int i = 0; i < n; i++) cudaFuncGetAttributes(&a, fns[i]);
  cudaGetLastError();
}
void preload_neighbor_kernels() {
  const void *const f[] = {(const void *)k_soa_positions, (const void *)k_soa_positions_half,
                           (const void *)k_find_neighbors_list<true, 0>, (const void *)k_find_neighbors_list<true, 1>,
                           (const void *)k_find_neighbors_list<true, 2>, (const void *)k_find_neighbors_coop<true>,
                           (const void *)k_find_neighbors_generic<true>, (const void *)k_neighbor_tiles<true>,
                           (const void *)k_neighbors_deferred<true>};
  preload(f, sizeof f / sizeof f[0]);
}
