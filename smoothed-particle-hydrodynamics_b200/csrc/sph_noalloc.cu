// sph_noalloc.cu — sph_cell_start_noalloc: the particle-to-cell histogram + exclusive scan (hashParticles
// sphFluid.cl:346-383 + the closed form of indexx / _runIndexPostPass, see sph_sort.cu) WITHOUT any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The scratch the sort
// needs — the cell histogram — lives in a file-scope __device__ array of a fixed maximum size; every GPU that loads
// the library holds it (4 MiB) until exit. All calls share the array and it keeps the last call's contents: each call
// clears the part it uses first, and calls that overlap on the GPU (different streams) race. Nothing in this file
// calls cudaMalloc / cudaFree or any allocating routine; no existing code path uses it.
#include <string.h>

#include "sph_ctx.h"

#define NA_MAX_CELLS SPH_NOALLOC_MAX_CELLS          // histogram words (G + 1 <= this)
#define NA_MAX_DEVICES 64

static __device__ __align__(256) unsigned g_na_hist[NA_MAX_CELLS];

// In host code `g_na_hist` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static unsigned *g_na_ptr[NA_MAX_DEVICES];

static thread_local char g_na_err[256];
static int na_fail(int code, const char *m) { strncpy(g_na_err, m, sizeof g_na_err - 1); return code; }

struct NaGrid { int n, G, gx, gy; unsigned mask; float cellSizeInv; };

__global__ void __launch_bounds__(256) k_na_clear(unsigned *__restrict__ hist, int words) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) hist[i] = 0u;
}

// same cell id and clamping as k_hash_count (sph_sort.cu)
__global__ void __launch_bounds__(256) k_na_hash_count(NaGrid g, const float4 *__restrict__ position,
                                                       unsigned *__restrict__ hist) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= g.n) return;
  const float4 p = position[id];
  const int cx = (int)(p.x * g.cellSizeInv), cy = (int)(p.y * g.cellSizeInv), cz = (int)(p.z * g.cellSizeInv);
  const unsigned cell = (unsigned)(cx + cy * g.gx + cz * g.gx * g.gy) & g.mask;
  atomicAdd(&hist[cell < (unsigned)g.G ? cell : (unsigned)g.G - 1u], 1u);
}

// exclusive scan of hist[0, words) into out: one CTA, 1024 elements per pass, running carry
__global__ void __launch_bounds__(1024) k_na_scan(const unsigned *__restrict__ hist, unsigned *__restrict__ out, int words) {
  __shared__ unsigned warpSum[32];
  __shared__ unsigned carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0u;
  __syncthreads();
  for (int base = 0; base < words; base += 1024) {
    const int i = base + threadIdx.x;
    const unsigned v = i < words ? hist[i] : 0u;
    unsigned inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) warpSum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      unsigned w = warpSum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warpSum[lane] = w;  // inclusive totals of the warps
    }
    __syncthreads();
    const unsigned excl = carry + (warp ? warpSum[warp - 1] : 0u) + (inc - v);
    if (i < words) out[i] = excl;
    __syncthreads();
    if (threadIdx.x == 0) carry += warpSum[31];
    __syncthreads();
  }
}

extern "C" {

const char *sph_noalloc_last_error(void) { return g_na_err; }

int sph_noalloc_max_cells(void) { return NA_MAX_CELLS; }

int sph_cell_start_noalloc(const sph_params *p, const float *pos4_device, int n, uint32_t *cell_start_device,
                           void *cuda_stream) {
  if (!p || !pos4_device || !cell_start_device || n <= 0 || p->grid_cell_count <= 0 || p->grid_cells_x <= 0 ||
      p->grid_cells_y <= 0)
    return na_fail(SPH_E_INVALID, "sph_cell_start_noalloc: bad arguments");
  if ((long long)p->grid_cell_count + 1 > (long long)NA_MAX_CELLS)
    return na_fail(SPH_E_SIZE, "sph_cell_start_noalloc: grid_cell_count + 1 exceeds the static scratch (SPH_NOALLOC_MAX_CELLS = 1048576 words)");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(SPH_E_DEVICE, "sph_cell_start_noalloc: no current CUDA device");
  if (!g_na_ptr[dev]) {  // set-up, once per device, with that device current
    unsigned *ptr = nullptr;
    e = cudaGetSymbolAddress((void **)&ptr, g_na_hist);
    if (e != cudaSuccess || !ptr) return na_fail(SPH_E_DEVICE, cudaGetErrorString(e));
    g_na_ptr[dev] = ptr;
  }
  unsigned *hist = g_na_ptr[dev];
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const int words = p->grid_cell_count + 1;
  NaGrid g = {n, p->grid_cell_count, p->grid_cells_x, p->grid_cells_y, p->hash_mask, p->hash_grid_cell_size_inv};
  k_na_clear<<<(words + 255) / 256 < 1184 ? (words + 255) / 256 : 1184, 256, 0, st>>>(hist, words);  // the array keeps the last call's counts
  k_na_hash_count<<<(n + 255) / 256, 256, 0, st>>>(g, (const float4 *)pos4_device, hist);
  k_na_scan<<<1, 1024, 0, st>>>(hist, cell_start_device, words);
  e = cudaGetLastError();
  if (e != cudaSuccess) return na_fail(SPH_E_DEVICE, cudaGetErrorString(e));
  return 0;
}

}  // extern "C"
