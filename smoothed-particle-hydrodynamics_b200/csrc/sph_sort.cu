// sph_sort.cu — particle-to-cell hash and a single-digit device radix (counting) sort.
//
// Replaces, with identical results, the reference's
//   hashParticles        sphFluid.cl:346-383  (+ cellFactors :187-201, cellId :332-342)
//   _runSort             owOpenCLSolver.cpp:255-261 (D2H, host qsort on the cell id, H2D)
//   sortPostPass         sphFluid.cl:441-466
//   indexx               sphFluid.cl:385-439   (per-cell binary search)
//   _runIndexPostPass    owOpenCLSolver.cpp:305-319 (D2H, host backward scan, H2D)
//
// B200 design: the radix IS the cell id. One pass builds the cell histogram with
// global atomics, a single-pass decoupled-look-back scan turns it into
// gridCellIndexFixedUp directly (the closed form of indexx + the host fix-up is
// "number of particles with a smaller cell id"), an atomic scatter bins the
// particles, and a rank pass orders each cell by serial id — the canonical order a
// stable sort (glibc qsort = merge sort) gives the reference — while gathering
// position/velocity into sorted order. Nothing crosses PCIe.
#include "sph_ctx.h"

#ifndef SPH_RANK_MAX_CELL
#define SPH_RANK_MAX_CELL 4096u
#endif
#define SCAN_THREADS 512
#define SCAN_ITEMS 4
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

// ---------------------------------------------------------------- hash + histogram
template <bool WRITE_PAIRS>
__global__ void __launch_bounds__(256) k_hash_count(DevParams d, const float4 *__restrict__ position,
                                                    unsigned *__restrict__ key, unsigned *__restrict__ cellCount,
                                                    uint2 *__restrict__ particleIndex,
                                                    unsigned long long *__restrict__ scanState,
                                                    unsigned *__restrict__ scanTicket, int scanTiles,
                                                    unsigned *__restrict__ nsStats, unsigned histPct) {
  pdl_entry();
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  // neighbour search of this step: histogram in the first pass if more than histPct % of the particles needed it in
  // the previous step (see k_find_neighbors_list); then restart the count
  // (histPct bits 0-7); above the percentage in bits 8-15 the first pass builds the histogram only (mode 2)
  if (id == 0) {
    const unsigned long long needy = 100ull * nsStats[0], n = (unsigned long long)d.N;
    nsStats[1] = (needy > (unsigned long long)(histPct >> 8) * n) ? 2u : (needy > (unsigned long long)(histPct & 255u) * n) ? 1u : 0u;
    nsStats[0] = 0u;
  }
  // reset the look-back scan of this step's sort (k_scan_cells runs after this kernel): the step's launch
  // parameters are then the same every step, which is what lets sph_step replay a captured CUDA graph
  for (int t = id; t < scanTiles; t += gridDim.x * blockDim.x) scanState[t] = 0ull;
  if (id == 0) *scanTicket = 0u;
  if (id >= d.N) return;
  float4 p = position[id];
  // cellFactors: no xmin subtraction (sphFluid.cl:197-199); cellId (:339-340); low-bit mask (:377)
  int cx = (int)(p.x * d.cellSizeInv);
  int cy = (int)(p.y * d.cellSizeInv);
  int cz = (int)(p.z * d.cellSizeInv);
  unsigned cell = (unsigned)(cx + cy * d.gx + cz * d.gx * d.gy) & d.mask;
  if (WRITE_PAIRS) particleIndex[id] = make_uint2(cell, (unsigned)id);
  // The reference would index past gridCellIndex for a cell id >= G (cannot happen while
  // integrate clamps positions into the box); keep memory safe.
  unsigned bin = cell < (unsigned)d.G ? cell : (unsigned)d.G - 1u;
  // particles outside the hash grid, or with a NaN coordinate ((int)NaN = 0: they land in cell 0), so far
  // (sph_out_of_grid_count): the sign that a simulation has diverged
  if (bin != cell || !(p.x == p.x) || !(p.y == p.y) || !(p.z == p.z)) atomicAdd(&nsStats[2], 1u);
  key[id] = cell;
  atomicAdd(&cellCount[bin], 1u);
}

// ---------------------------------------------------------------- exclusive scan (decoupled look-back)
// state word: [63:34] epoch, [33:32] flag (1 = aggregate, 2 = inclusive prefix), [31:0] value
__device__ __forceinline__ unsigned long long pack_state(unsigned epoch, unsigned flag, unsigned v) {
  return ((unsigned long long)(epoch & 0x3fffffffu) << 34) | ((unsigned long long)flag << 32) | v;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_cells(unsigned *__restrict__ count, unsigned *__restrict__ start,
                                                             int n, volatile unsigned long long *state,
                                                             unsigned *ticket, unsigned ticketBase, unsigned epoch,
                                                             unsigned *cellListCount) {
  __shared__ unsigned s_tile;
  __shared__ unsigned s_warp[SCAN_THREADS / 32];
  __shared__ unsigned s_prefix;
  pdl_entry();
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u) - ticketBase;
  __syncthreads();
  const unsigned tile = s_tile;
  if (tile == 0 && threadIdx.x == 0) *cellListCount = 0u;  // refilled by the rank pass below
  const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  unsigned v[SCAN_ITEMS];
  unsigned sum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) {
    v[i] = (base + i < n) ? count[base + i] : 0u;
    sum += v[i];
  }
  // warp inclusive scan of per-thread sums
  unsigned inc = sum;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    unsigned w = lane < SCAN_THREADS / 32 ? s_warp[lane] : 0u;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    if (lane < SCAN_THREADS / 32) s_warp[lane] = w;  // inclusive warp totals
  }
  __syncthreads();
  const unsigned tileTotal = s_warp[SCAN_THREADS / 32 - 1];
  if (threadIdx.x == 0) {
    unsigned prefix = 0;
    if (tile == 0) {
      state[0] = pack_state(epoch, 2u, tileTotal);
    } else {
      state[tile] = pack_state(epoch, 1u, tileTotal);
      int t = (int)tile - 1;
      while (true) {
        unsigned long long s = state[t];
        if ((unsigned)(s >> 34) != (epoch & 0x3fffffffu)) continue;  // not published this step yet
        unsigned flag = (unsigned)(s >> 32) & 3u;
        if (flag == 0u) continue;
        prefix += (unsigned)s;
        if (flag == 2u) break;
        --t;
      }
      state[tile] = pack_state(epoch, 2u, prefix + tileTotal);
    }
    s_prefix = prefix;
  }
  __syncthreads();
  unsigned excl = s_prefix + (warp ? s_warp[warp - 1] : 0u) + (inc - sum);
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) {
    if (base + i < n) {
      start[base + i] = excl;
      count[base + i] = 0u;  // becomes the scatter cursor
    }
    excl += v[i];
  }
}

// ---------------------------------------------------------------- scatter into cells (order inside a cell arbitrary)
__global__ void __launch_bounds__(256) k_scatter(DevParams d, const unsigned *__restrict__ key,
                                                 const unsigned *__restrict__ cellStart, unsigned *__restrict__ cursor,
                                                 unsigned *__restrict__ tmpSerial,
                                                 const unsigned *__restrict__ rankKey, unsigned *__restrict__ tmpRankKey) {
  pdl_entry();
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  unsigned cell = key[id];
  unsigned bin = cell < (unsigned)d.G ? cell : (unsigned)d.G - 1u;
  unsigned slot = cellStart[bin] + atomicAdd(&cursor[bin], 1u);
  tmpSerial[slot] = (unsigned)id;
  if (rankKey) tmpRankKey[slot] = rankKey[id];  // slab ranks: the ranking key travels with the serial id
}

// ---------------------------------------------------------------- rank inside the cell (+ sortPostPass gather)
// Final slot = cellStart + #(cell-mates with a smaller serial id): (cell, serial) ascending.
template <bool REORDER>
__global__ void __launch_bounds__(256) k_rank_reorder(DevParams d, const unsigned *__restrict__ key,
                                                      const unsigned *__restrict__ cellStart,
                                                      const unsigned *__restrict__ tmpSerial,
                                                      uint2 *__restrict__ particleIndex,
                                                      unsigned *__restrict__ particleIndexBack,
                                                      const float4 *__restrict__ position,
                                                      const float4 *__restrict__ velocity,
                                                      float4 *__restrict__ sortedPosition,
                                                      float4 *__restrict__ sortedVelocity,
                                                      float *__restrict__ sortedType, float4 *__restrict__ spT, unsigned *__restrict__ cellList,
                                                      unsigned *__restrict__ cellListCount,
                                                      const unsigned *__restrict__ rankKey,
                                                      float4 *__restrict__ recA, float4 *__restrict__ recB,
                                                      unsigned *__restrict__ cellCount) {
  pdl_entry();
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  // the scatter cursors (the cell histogram, consumed by k_scatter) go back to zero for the next step's hash
  for (int t = s; t <= d.G; t += gridDim.x * blockDim.x) cellCount[t] = 0u;
  if (s >= d.N) return;
  unsigned serial = tmpSerial[s];
  unsigned cell = key[serial];
  unsigned bin = cell < (unsigned)d.G ? cell : (unsigned)d.G - 1u;
  unsigned b = cellStart[bin], e = cellStart[bin + 1];
  unsigned rank = 0;
  // order within a cell = ascending original index (glibc's stable merge sort in _runSort); a slab rank
  // ranks by GLOBAL id instead, which is the original index of the single-device run.
  // A cell with more than SPH_RANK_MAX_CELL particles (64x the rest density: only a diverged simulation, whose NaN
  // and out-of-grid particles collapse into one cell, gets there) keeps the scatter order: the quadratic ranking of a
  // million cell-mates would look like a hang, and there is no meaningful order left to preserve.
  if (e - b > SPH_RANK_MAX_CELL) {
    rank = (unsigned)s - b;
  } else if (rankKey) {  // rankKey = the keys in scatter order (tmpRankKey), contiguous per cell
    const unsigned mine = rankKey[s];
    for (unsigned t = b; t < e; ++t) rank += (rankKey[t] < mine) ? 1u : 0u;
  } else {
    for (unsigned t = b; t < e; ++t) rank += (tmpSerial[t] < serial) ? 1u : 0u;
  }
  unsigned dst = b + rank;
  // work list of the neighbour search: one entry per HOME BLOCK (cells 2b, 2b+1), posted by the first
  // particle of the block's first non-empty cell
  if (rank == 0 && ((bin & 1u) == 0u || cellStart[bin - 1] == b)) cellList[atomicAdd(cellListCount, 1u)] = bin;
  particleIndex[dst] = make_uint2(cell, serial);
  if (REORDER) {  // sortPostPass, sphFluid.cl:457-465
    float4 p = position[serial];
    sortedType[dst] = p.w;
    spT[dst] = p;
    const float4 v = velocity[serial];
    recA[2 * (size_t)dst] = p;
    recA[2 * (size_t)dst + 1] = f4(v.x, v.y, v.z, 0.f);  // .w = rho(t), set by the neighbour search
    recB[dst] = f4(p.x, p.y, p.z, 0.f);   // (x, y, z, pressure = 0): computeForcesAndInitPressure's pressure[id] = 0 (:697)
    p.w = (float)(int)cell;
    sortedVelocity[dst] = v;
    sortedPosition[dst] = p;
    particleIndexBack[serial] = dst;
  }
}

__global__ void __launch_bounds__(256) k_reorder(DevParams d, const uint2 *__restrict__ particleIndex,
                                                 unsigned *__restrict__ particleIndexBack,
                                                 const float4 *__restrict__ position, const float4 *__restrict__ velocity,
                                                 float4 *__restrict__ sortedPosition, float4 *__restrict__ sortedVelocity,
                                                 float *__restrict__ sortedType, float4 *__restrict__ spT) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  uint2 spi = particleIndex[id];
  float4 p = position[spi.y];
  sortedType[id] = p.w;
  spT[id] = p;
  p.w = (float)(int)spi.x;
  sortedVelocity[id] = velocity[spi.y];
  sortedPosition[id] = p;
  particleIndexBack[spi.y] = (unsigned)id;
}

// indexx output for API parity (sphFluid.cl:385-439): first index per cell, -1 when empty, [0]=0, [G]=N
__global__ void __launch_bounds__(256) k_indexx_compat(DevParams d, const unsigned *__restrict__ cellStart,
                                                       unsigned *__restrict__ gridCellIndex) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > d.G) return;
  unsigned v;
  if (c == d.G) v = (unsigned)d.N;
  else if (c == 0) v = 0u;
  else v = (cellStart[c + 1] > cellStart[c]) ? cellStart[c] : 0xffffffffu;
  gridCellIndex[c] = v;
}

static inline int blocks_for(int n, int t) { return (n + t - 1) / t; }

void launch_hash_count(sph_ctx *c, bool write_pairs) {
  const int N = c->d.N;
  // k_rank_reorder leaves the histogram zeroed; only a histogram somebody else dirtied (a HASH stage without its SORT,
  // a changed particle count) needs the memset — which would also break the chain of programmatic launches
  if (!c->cellCountClean) cudaMemsetAsync(c->cellCount, 0, sizeof(unsigned) * ((size_t)c->d.G + 1), c->stream);
  c->cellCountClean = false;
  if (write_pairs)
    launch_k(c, k_hash_count<true>, dim3(blocks_for(N, 256)), dim3(256), c->d, c->position, c->key, c->cellCount,
             c->particleIndex, c->scanState, c->scanTicket, c->scanTiles, c->nsStats, c->hist_pct);
  else
    launch_k(c, k_hash_count<false>, dim3(blocks_for(N, 256)), dim3(256), c->d, c->position, c->key, c->cellCount,
             c->particleIndex, c->scanState, c->scanTicket, c->scanTiles, c->nsStats, c->hist_pct);
}

void launch_sort(sph_ctx *c, bool fuse_reorder) {
  const int N = c->d.N, n = c->d.G + 1;
  // scan state and ticket were zeroed by k_hash_count: constant epoch, no per-step launch parameter
  launch_k(c, k_scan_cells, dim3(c->scanTiles), dim3(SCAN_THREADS), c->cellCount, c->cellStart, n, c->scanState, c->scanTicket,
           0u, 1u, c->cellListCount);
  launch_k(c, k_scatter, dim3(blocks_for(N, 256)), dim3(256), c->d, c->key, c->cellStart, c->cellCount, c->tmpSerial, c->rankKey,
           c->tmpRankKey);
  const unsigned *rk = c->rankKey ? c->tmpRankKey : nullptr;
  // enough threads to zero the G + 1 histogram words too (sparse scenes have more cells than particles)
  const int blocks = blocks_for(N, 256);
  if (fuse_reorder)
    launch_k(c, k_rank_reorder<true>, dim3(blocks), dim3(256), c->d, c->key, c->cellStart, c->tmpSerial, c->particleIndex,
             c->particleIndexBack, c->position, c->velocity, c->sortedPosition, c->sortedVelocity, c->sortedType, c->spT,
             c->cellList, c->cellListCount, rk, c->recA, c->recB, c->cellCount);
  else
    launch_k(c, k_rank_reorder<false>, dim3(blocks), dim3(256), c->d, c->key, c->cellStart, c->tmpSerial, c->particleIndex,
             c->particleIndexBack, c->position, c->velocity, c->sortedPosition, c->sortedVelocity, c->sortedType, c->spT,
             c->cellList, c->cellListCount, rk, c->recA, c->recB, c->cellCount);
  c->cellCountClean = true;
}

void launch_reorder(sph_ctx *c) {
  k_reorder<<<blocks_for(c->d.N, 256), 256, 0, c->stream>>>(c->d, c->particleIndex, c->particleIndexBack, c->position,
                                                            c->velocity, c->sortedPosition, c->sortedVelocity,
                                                            c->sortedType, c->spT);
}

void launch_indexx_compat(sph_ctx *c) {
  k_indexx_compat<<<blocks_for(c->d.G + 1, 256), 256, 0, c->stream>>>(c->d, c->cellStart, c->gridCellIndex);
}

int sph_scan_tiles_for(int n) { return (n + SCAN_TILE - 1) / SCAN_TILE; }

// Force the (lazily loaded, CUDA >= 12.2) code of this file's step kernels onto the device. A kernel that spins on a
// neighbour's delivery flag (sph_slab.cu) must never be resident while the host loads another kernel for the first
// time — module loading synchronises the context and would wait for the spinning kernel forever.
static void preload(const void *const *fns, int n) {
  cudaFuncAttributes a;
  for (int i = 0; i < n; i++) cudaFuncGetAttributes(&a, fns[i]);
  cudaGetLastError();
}
void preload_sort_kernels() {
  const void *const f[] = {(const void *)k_hash_count<false>, (const void *)k_scan_cells, (const void *)k_scatter,
                           (const void *)k_rank_reorder<true>};
  preload(f, sizeof f / sizeof f[0]);
}
