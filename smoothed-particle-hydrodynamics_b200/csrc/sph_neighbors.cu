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
// One kernel, k_neighbor_tiles. A CTA takes a HOME BLOCK = the two hash cells (2b, 2b+1), whose
// particles are one contiguous range of the sorted arrays, and
//   stage   copies the positions of the 4x3x3 = 36 surrounding cells (offsets -1..2 in id, -1..1 in
//           y-row and z-slab, searchCell wrap included) into shared memory as SoA x/y/z: every
//           candidate any home particle of the block can visit, read from HBM/L2 once, coalesced;
//   group   sorts the home particles by (home cell, octant): all particles of a cell that sit in the
//           same octant of it visit the same 8 cells in the same order (sphFluid.cl:248-308). Those 8
//           cells are 4 RUNS of two id-adjacent cells, and a run is contiguous in the staging area;
//   sweep   (P1) one warp per PAIR of home particles of a group: the pair's coordinates live in
//           registers as fp32x2 operands, lanes hold 32 consecutive staged candidates of a run, both
//           squared distances come out of Blackwell's packed fp32x2 pipe, two ballots give the
//           32-bit hit masks of the chunk, lane w keeps mask word w, and the row of words goes to
//           shared memory. The test is a conservative FILTER (radius 31h/30 — the largest r_thr the
//           reference can choose — slightly inflated, FMA allowed): nothing inexact reaches the output;
//   select  (P2) one thread per home particle walks the bit ranges of its 8 cells in the reference's
//           visiting order, re-derives the exact distance of the ~35 hits with the reference's
//           operation order (from the staged copies), builds the 30-bin radial histogram, picks r_thr
//           (sphFluid.cl:310-324), keeps the first 32 hits within it (:166-174), accumulates rho(t) on
//           them (pcisph_computeDensity, :472-518) and writes the ids in the 32x32 tile layout.
// The two scan passes of the reference (2 x ~512 candidates per particle) become one packed sweep plus
// two passes over the hits; the mask rows never leave the SM. Particles whose neighbourhood does not fit
// the tables (staging area, FT_NCH mask words, FT_CAP hits) take find_neighbors_one, the generic scan.
#define FT_THREADS 128
#define FT_WARPS (FT_THREADS / 32)
#define FT_ROWS FT_THREADS        // home particles per batch (thread t of P2 owns particle b0 + t)
#define FT_GROUPS 16              // (home cell 0/1) x octant
#define FT_ROWPAD (FT_ROWS + FT_GROUPS)  // rows in group order, every group padded to an even size
#define FT_NCELL 36               // staged cells: id offset hx in [-1,2], y-row and z-slab offsets in [-1,1]
#define FT_CMAX 2816              // staged candidates (36 cells x 78); staged indices fit 12 bits
#define FT_NCH 32                 // mask words per row: one per lane of the sweeping warp
#define FT_MSTRIDE (FT_NCH + 1)   // odd row stride: P2's per-thread row walks are bank-conflict free; a row
                                  // is reused for the 32 selected ids once its masks have been consumed
#define FT_CAP 48                 // hits kept per particle in P2

struct FtShared {
  float X[FT_CMAX + 32], Y[FT_CMAX + 32], Z[FT_CMAX + 32];  // +32: a run's last chunk may read past its end
  unsigned masks[FT_ROWPAD * FT_MSTRIDE];
  unsigned short hits[FT_CAP][FT_THREADS];   // (k << 12) | staged index, scan order
  unsigned char hist[SPH_RADIUS_SEGMENTS + 2][FT_THREADS];
  float2 px[FT_ROWPAD / 2], py[FT_ROWPAD / 2], pz[FT_ROWPAD / 2];
  unsigned word[FT_GROUPS][FT_NCH];  // mask word w of a group: staged index of its bit 0 | (valid bits - 1) << 16
  int bitLo[FT_GROUPS][8];           // cell k (reference visiting order) = bits [bitLo, bitLo + cntK) of the row
  int cntK[FT_GROUPS][8];
  int sB[FT_GROUPS][8];              // staged index of the cell's first particle
  int dG[FT_GROUPS][8];              // staged index + dG = sorted particle id
  int nWords[FT_GROUPS];
  float binT[SPH_RADIUS_SEGMENTS + 2];
  int gStart[FT_NCELL], cnt[FT_NCELL], sStart[FT_NCELL + 1];
  int grpCount[FT_GROUPS], grpStart[FT_GROUPS], pairStart[FT_GROUPS + 1], bad[FT_GROUPS];
  unsigned char pairGroup[FT_ROWPAD / 2];
  int nextCell, blockBad;
};

template <bool WITH_DENSITY>
__global__ void __launch_bounds__(FT_THREADS, 3) k_neighbor_tiles(DevParams d, const float4 *__restrict__ sp,
                                                                  const unsigned *__restrict__ cellStart,
                                                                  const unsigned *__restrict__ cellList,
                                                                  const unsigned *__restrict__ cellListCount,
                                                                  unsigned *__restrict__ cellCursor,
                                                                  int *__restrict__ nb, float *__restrict__ rho) {
  extern __shared__ __align__(16) unsigned char ft_smem[];
  FtShared &S = *reinterpret_cast<FtShared *>(ft_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nCells = (int)*cellListCount;
  const float rmaxf = (float)(SPH_RADIUS_SEGMENTS + 1) * d.h / SPH_RADIUS_SEGMENTS;  // largest r_thr (j == 30)
  const float rfilt2 = rmaxf * rmaxf * 1.0001f;
  const float FAR = 1.0e30f;
  const float h2 = d.h * d.h;
  if (tid < SPH_RADIUS_SEGMENTS + 2) S.binT[tid] = d.binT[tid];
  for (int i = tid; i < 32; i += FT_THREADS) { S.X[FT_CMAX + i] = FAR; S.Y[FT_CMAX + i] = FAR; S.Z[FT_CMAX + i] = FAR; }
  while (true) {
    __syncthreads();
    if (tid == 0) S.nextCell = (int)atomicAdd(cellCursor, 1u);
    __syncthreads();
    const int ci = S.nextCell;
    if (ci >= nCells) break;
    const int cell = (int)cellList[ci];
    const int base = cell & ~1;
    // every non-empty cell appears once in cellList; the block is run by its first non-empty cell
    if (cell != base && cellStart[base + 1] != cellStart[base]) continue;
    const int hBeg = (int)cellStart[base];
    const int hMid = (int)cellStart[base + 1];
    const int hEnd = (int)cellStart[min(base + 2, d.G)];
    // ---- stage: ranges of the 36 surrounding cells, prefix of their sizes, then the positions
    if (warp == 0) {
      int c0 = 0, c1 = 0;
      {
        const int t = lane;
        const int nc = search_cell(base, (t & 3) - 1 + ((t >> 2) % 3 - 1) * d.gx + (t / 12 - 1) * d.gx * d.gy, d.G);
        const int b = (int)cellStart[nc];
        c0 = (int)cellStart[nc + 1] - b;
        S.gStart[t] = b;
        S.cnt[t] = c0;
      }
      if (lane < FT_NCELL - 32) {
        const int t = 32 + lane;
        const int nc = search_cell(base, (t & 3) - 1 + ((t >> 2) % 3 - 1) * d.gx + (t / 12 - 1) * d.gx * d.gy, d.G);
        const int b = (int)cellStart[nc];
        c1 = (int)cellStart[nc + 1] - b;
        S.gStart[t] = b;
        S.cnt[t] = c1;
      }
      int inc = c0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
      }
      const int tot32 = __shfl_sync(0xffffffffu, inc, 31);
      S.sStart[lane] = inc - c0;
      int inc1 = c1;
#pragma unroll
      for (int o = 1; o < 4; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc1, o);
        if (lane >= o) inc1 += u;
      }
      if (lane < FT_NCELL - 32) S.sStart[32 + lane] = tot32 + inc1 - c1;
      if (lane == FT_NCELL - 32 - 1) {
        S.sStart[FT_NCELL] = tot32 + inc1;
        S.blockBad = (tot32 + inc1 > FT_CMAX) ? 1 : 0;
      }
    }
    __syncthreads();
    const int blockBad = S.blockBad;
    if (!blockBad) {
#pragma unroll 1
      for (int t = warp; t < FT_NCELL; t += FT_WARPS) {
        const int g0 = S.gStart[t], n = S.cnt[t], s0 = S.sStart[t];
        for (int i = lane; i < n; i += 32) {
          const float4 q = ldg4(sp + g0 + i);
          S.X[s0 + i] = q.x; S.Y[s0 + i] = q.y; S.Z[s0 + i] = q.z;
        }
      }
    }
    for (int b0 = hBeg; b0 < hEnd; b0 += FT_ROWS) {
      const int hb = min(FT_ROWS, hEnd - b0);
      const int id = b0 + tid;
      __syncthreads();
      if (tid < FT_GROUPS) S.grpCount[tid] = 0;
      for (int r = tid; r < FT_ROWPAD / 2; r += FT_THREADS) {
        S.px[r] = make_float2(FAR, FAR); S.py[r] = make_float2(FAR, FAR); S.pz[r] = make_float2(FAR, FAR);
      }
      __syncthreads();
      // ---- group the home particles by (home cell, octant); octant = which half of the 2h cell, per axis (:253-271)
      float4 pos = f4(0.f, 0.f, 0.f, 0.f);
      int g = 0, rankInG = 0;
      if (tid < hb) {
        pos = sp[id];
        float cfx = (float)(int)(pos.x * d.cellSizeInv) * d.cellSize;
        float cfy = (float)(int)(pos.y * d.cellSizeInv) * d.cellSize;
        float cfz = (float)(int)(pos.z * d.cellSizeInv) * d.cellSize;
        g = ((((pos.x - d.xmin) - cfx) < d.h) ? 0 : 1) | ((((pos.y - d.ymin) - cfy) < d.h) ? 0 : 2) |
            ((((pos.z - d.zmin) - cfz) < d.h) ? 0 : 4) | ((id >= hMid) ? 8 : 0);
        rankInG = atomicAdd(&S.grpCount[g], 1);
      }
      __syncthreads();
      if (tid < FT_GROUPS) {  // tables of group `tid`
        int gs = 0, ps = 0;
        for (int q = 0; q < tid; q++) { gs += (S.grpCount[q] + 1) & ~1; ps += (S.grpCount[q] + 1) >> 1; }
        const int np = (S.grpCount[tid] + 1) >> 1;
        S.grpStart[tid] = gs;
        S.pairStart[tid] = ps;
        if (tid == FT_GROUPS - 1) S.pairStart[FT_GROUPS] = ps + np;
        for (int q = 0; q < np; q++) S.pairGroup[ps + q] = (unsigned char)tid;
        const int sx = (tid & 1) ? 1 : -1, sy = (tid & 2) ? 1 : -1, sz = (tid & 4) ? 1 : -1, hc = tid >> 3;
        const int hxlo = hc + ((sx < 0) ? -1 : 0);  // the group's two id-adjacent cells are hxlo, hxlo + 1
        // the 4 runs, ordered (0,0), (sy,0), (0,sz), (sy,sz): mask words of the row
        int wOff[5];
        int w = 0;
#pragma unroll
        for (int r = 0; r < 4; r++) {
          const int oy = (r & 1) ? sy : 0, oz = (r & 2) ? sz : 0;
          const int tlo = (hxlo + 1) + (oy + 1) * 4 + (oz + 1) * 12;
          const int s0 = S.sStart[tlo], len = S.cnt[tlo] + S.cnt[tlo + 1];
          wOff[r] = w;
          for (int c = 0; c * 32 < len; c++) {
            const int nv = min(32, len - c * 32);
            if (w < FT_NCH) S.word[tid][w] = (unsigned)(s0 + c * 32) | ((unsigned)(nv - 1) << 16);
            w++;
          }
        }
        wOff[4] = w;
        S.nWords[tid] = w;
        S.bad[tid] = (blockBad || w > FT_NCH) ? 1 : 0;
        // the 8 cells in the reference's visiting order (:248-308): own, x, y, z, xy, xz, yz, xyz
#pragma unroll
        for (int k = 0; k < 8; k++) {
          const int kx = (0xB2 >> k) & 1, ky = (0xD4 >> k) & 1, kz = (0xE8 >> k) & 1;
          const int hx = hc + kx * sx;
          const int t = (hx + 1) + (ky * sy + 1) * 4 + (kz * sz + 1) * 12;
          const int r = ky + 2 * kz;
          S.bitLo[tid][k] = wOff[r] * 32 + ((hx == hxlo) ? 0 : S.cnt[t - 1]);
          S.cntK[tid][k] = S.cnt[t];
          S.sB[tid][k] = S.sStart[t];
          S.dG[tid][k] = S.gStart[t] - S.sStart[t];
        }
      }
      __syncthreads();
      int row = 0;
      if (tid < hb) {
        row = S.grpStart[g] + rankInG;
        ((float *)S.px)[row] = pos.x; ((float *)S.py)[row] = pos.y; ((float *)S.pz)[row] = pos.z;
      }
      __syncthreads();
      // ---- P1 sweep: one warp per home pair
      const int totalPairs = S.pairStart[FT_GROUPS];
#pragma unroll 1
      for (int t = warp; t < totalPairs; t += FT_WARPS) {
        const int gg = S.pairGroup[t];
        if (S.bad[gg]) continue;
        const int pr = (S.grpStart[gg] >> 1) + (t - S.pairStart[gg]);
        const float2 hx = S.px[pr], hy = S.py[pr], hz = S.pz[pr];
        const unsigned *wtab = S.word[gg];
        const int nch = S.nWords[gg];
        unsigned keep0 = 0, keep1 = 0;
#pragma unroll 4
        for (int c = 0; c < nch; c++) {
          const unsigned wd = wtab[c];
          const int s = (int)(wd & 0xffffu) + lane;
          const unsigned valid = 0xffffffffu >> (31u - (wd >> 16));
          const float cx = S.X[s], cy = S.Y[s], cz = S.Z[s];
          const float2 ncx = make_float2(-cx, -cx), ncy = make_float2(-cy, -cy), ncz = make_float2(-cz, -cz);
          float2 ex = __fadd2_rn(hx, ncx), ey = __fadd2_rn(hy, ncy), ez = __fadd2_rn(hz, ncz);
          float2 dd = __ffma2_rn(ez, ez, __ffma2_rn(ey, ey, __fmul2_rn(ex, ex)));
          const unsigned m0 = __ballot_sync(0xffffffffu, dd.x <= rfilt2) & valid;
          const unsigned m1 = __ballot_sync(0xffffffffu, dd.y <= rfilt2) & valid;
          if (lane == c) { keep0 = m0; keep1 = m1; }
        }
        if (lane < nch) {
          S.masks[(2 * pr) * FT_MSTRIDE + lane] = keep0;
          S.masks[(2 * pr + 1) * FT_MSTRIDE + lane] = keep1;
        }
      }
      __syncthreads();
      // ---- P2 select: one thread per home particle
      if (tid >= hb) continue;
      if (S.bad[g]) { find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho); continue; }
      unsigned *mrow = S.masks + row * FT_MSTRIDE;
      // bit ranges of the 8 cells, in visiting order -> hit list (scan order)
      int nh = 0;
#pragma unroll 1
      for (int k = 0; k < 8; k++) {
        const int n = S.cntK[g][k];
        if (n == 0) continue;
        const int lo = S.bitLo[g][k], hi = lo + n;  // [lo, hi)
        const int sOff = S.sB[g][k] - lo;           // bit position + sOff = staged index
        const int self = id - S.dG[g][k] - sOff;    // bit position of this particle, if it lies in this cell
        const unsigned tag = (unsigned)k << 12;
        int c = lo >> 5;
        const int cLast = (hi - 1) >> 5;
        unsigned m = mrow[c] & (0xffffffffu << (lo & 31));
        while (true) {
          if (c == cLast) m &= 0xffffffffu >> (31 - ((hi - 1) & 31));
          while (m) {
            const int bit = (c << 5) + __ffs(m) - 1;
            m &= m - 1;
            if (bit != self) {
              if (nh < FT_CAP) S.hits[nh][tid] = (unsigned short)(tag | (unsigned)(bit + sOff));
              nh++;
            }
          }
          if (c == cLast) break;
          m = mrow[++c];
        }
      }
      if (nh > FT_CAP) { find_neighbors_one<WITH_DENSITY>(d, id, sp, cellStart, nb, rho); continue; }
#pragma unroll
      for (int b = 0; b < SPH_RADIUS_SEGMENTS; b++) S.hist[b][tid] = 0;
      // exact distances + 30-bin radial histogram of everything within h (:157-161)
      for (int t = 0; t < nh; t++) {
        const int s = S.hits[t][tid] & 0xfff;
        float ex = pos.x - S.X[s], ey = pos.y - S.Y[s], ez = pos.z - S.Z[s];
        float dd = ex * ex + ey * ey + ez * ez;
        if (dd <= h2) {
          // bin (int)(sqrtf(dd)*30/h): approximate guess, made exact with the host-built thresholds
          float sq;
          asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(dd));
          int e = (int)(sq * d.binScale);
          e = min(e, SPH_RADIUS_SEGMENTS);
          const int b = e + ((dd >= S.binT[e + 1]) ? 1 : 0) - ((dd < S.binT[e]) ? 1 : 0);
          if (b < SPH_RADIUS_SEGMENTS) {
            unsigned hv = S.hist[b][tid];
            if (hv < 255u) S.hist[b][tid] = (unsigned char)(hv + 1u);  // saturation keeps the >32 logic exact
          }
        }
      }
      // radius threshold (:310-324)
      int j = 0, sum = 0;
      while (j < SPH_RADIUS_SEGMENTS) {
        sum += S.hist[j][tid];
        if (sum == SPH_NB) break;
        if (sum > SPH_NB) { j--; break; }
        j++;
      }
      const float r_thr = (float)(j + 1) * d.h / SPH_RADIUS_SEGMENTS;
      const float r2 = r_thr * r_thr;
      // first 32 hits within r_thr, in scan order (:166-174) + rho(t)
      int found = 0;
      double density = 0.0;
      for (int t = 0; t < nh; t++) {
        const unsigned hv = S.hits[t][tid];
        const int s = hv & 0xfff;
        float ex = pos.x - S.X[s], ey = pos.y - S.Y[s], ez = pos.z - S.Z[s];
        float dd = ex * ex + ey * ey + ez * ez;
        if (dd <= r2) {
          mrow[found] = (unsigned)(s + S.dG[g][hv >> 12]);  // own row: its masks are consumed
          if (WITH_DENSITY) {
            float r = sqrtf(dd) * d.simScale;
            float rr = r * r;
            density += (double)((d.hScaled2 - rr) * (d.hScaled2 - rr) * (d.hScaled2 - rr));
          }
          if (++found == SPH_NB) break;
        }
      }
#pragma unroll 8
      for (int q = 0; q < SPH_NB; q++) nb[nb_index(id, q)] = (q < found) ? (int)mrow[q] : SPH_NO_ID;
      if (WITH_DENSITY) {
        if (density < (double)d.hScaled6) density = (double)d.hScaled6;
        density *= d.massWpoly6;
        rho[id] = (float)density;
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
  cudaMemsetAsync(c->cellCursor, 0, sizeof(unsigned), c->stream);
  k_neighbor_tiles<WD><<<c->numSMs * ft_ctas_per_sm, FT_THREADS, sizeof(FtShared), c->stream>>>(
      c->d, c->sortedPosition, c->cellStart, c->cellList, c->cellListCount, c->cellCursor, c->nbIds, c->rho);
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
  // persistent CTAs (one resident wave); home blocks handed out through an atomic cursor over cellList
  if (with_density) launch_tiles<true>(c); else launch_tiles<false>(c);
}

void launch_materialize_neighbor_map(sph_ctx *c, float2 *out) {
  size_t n = (size_t)c->d.N * SPH_NB;
  k_materialize_neighbor_map<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->d, c->sortedPosition, c->nbIds, out);
}
