// sph_device.cuh — shared device-side definitions of the B200 PCISPH step.
//
// Arithmetic contract (DESIGN.md "Numerics"): every float expression is evaluated
// with IEEE round-to-nearest single operations in the reference's association
// order (sphFluid.cl is built with -cl-opt-disable: no FMA contraction), so the
// library is compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
// and results are bit-identical to the reference kernels compiled as C++.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "sph_capi.h"

#define SPH_NB 32                 // MAX_NEIGHBOR_COUNT, owOpenCLConstant.h:4
#define SPH_NO_ID (-1)            // NO_PARTICLE_ID
#define SPH_RADIUS_SEGMENTS 30    // sphFluid.cl:116

// Kernel-visible parameters: sph_params plus sub-expressions hoisted to the host
// (computed there with the reference's exact expressions, see sph_capi.cu).
struct DevParams {
  int N;        // particles
  int Npad;     // N rounded up to 32 (one neighbour tile = 32 particles)
  int G;        // hash grid cells
  int gx, gy, gz;
  unsigned mask;
  int numElastic, elasticOffset, muscleCount;
  float h, cellSize, cellSizeInv, simScale, simScaleInv, r0;
  float xmin, xmax, ymin, ymax, zmin, zmax;
  float rho0, dt, delta, gravx, gravy, gravz;
  float hScaled, hScaled2, hScaled6;  // sphFluid.cl:495-497
  float surfCoeff;                    // sphFluid.cl:662
  float massMu;                       // sphFluid.cl:688
  float posTimeStep;                  // timeStep*simulationScaleInv, sphFluid.cl:928,1738
  float rho0delta;                    // rho0*delta, sphFluid.cl:1168
  float xmaxClamp, ymaxClamp, zmaxClamp;  // xmax-0.000001f ..., sphFluid.cl:1753-1755
  double massWpoly6;                  // sphFluid.cl:516,1057
  double del2W;                       // del2WviscosityCoefficient
  double massGradW;                   // sphFluid.cl:1194
  double closeThr;                    // 0.5*(hScaled/2), sphFluid.cl:1166
  // binT[j] = smallest float d2 whose radial-histogram bin (int)(sqrtf(d2)*30/h) (sphFluid.cl:157-160) is >= j,
  // found on the host by bisection over the float bit patterns with the same IEEE operations; binT[0] = 0,
  // binT[31] = +inf. The bin of d2 is then max{j : binT[j] <= d2}, exactly, without sqrt/div per candidate.
  float binT[SPH_RADIUS_SEGMENTS + 2];
  float binScale;                     // 30/h, for the approximate first guess of the bin
  // r_ij = sqrtf(d2) * simulationScale is a monotone function of d2, so the two range tests of the pressure force
  // (sphFluid.cl:1150, :1166) are tests on d2 against host-built thresholds (bisection over float bit patterns with
  // the same IEEE operations): r_ij < hScaled <=> d2 < d2H, r_ij < closeThr <=> d2 < d2Close. No sqrt needed to decide.
  float d2H, d2Close;
};

// Programmatic dependent launch (sm_90+): first statement of every kernel of the fused step. wait blocks until the
// PREVIOUS kernel of the stream has completed and its writes are visible; launch_dependents then lets the NEXT kernel
// be scheduled as soon as all CTAs of this one have got here, so its launch latency and ramp-up overlap this kernel's
// execution. Wait comes FIRST: at most two grids of the chain are then in flight. (Triggering before waiting let a
// third grid start while the first still ran, and its wait did not cover the first: measured as stale reads in the
// worm scene, tools/diag_worm.py.) Both are no-ops for launches without the programmatic attribute.
__device__ __forceinline__ void pdl_entry() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ---- float4 lane-wise helpers (OpenCL vector semantics, one IEEE op per lane) ----
__device__ __forceinline__ float4 f4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return f4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 operator-(float4 a, float4 b) { return f4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
__device__ __forceinline__ float4 operator-(float4 a) { return f4(-a.x, -a.y, -a.z, -a.w); }
__device__ __forceinline__ float4 operator*(float4 a, float s) { return f4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4 operator*(float s, float4 a) { return f4(s * a.x, s * a.y, s * a.z, s * a.w); }
__device__ __forceinline__ float4 operator/(float4 a, float s) { return f4(a.x / s, a.y / s, a.z / s, a.w / s); }
__device__ __forceinline__ float dot4(float4 a, float4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
__device__ __forceinline__ float max0(float x) { return (0.f < x) ? x : 0.f; }  // max(0.f,x), sphFluid.cl:859,1582

__device__ __forceinline__ float4 ldg4(const float4 *p) { return __ldg(p); }

// Gather RECORDS of the fused path: one 32-byte, 32-byte-aligned record per sorted particle, fetched with a
// single 256-bit load (LDG.E.256, one L1/L2 sector) instead of two or three 16-byte gathers from separate arrays.
//   recA[i] = (x, y, z, type | vx, vy, vz, rho(t))   gathered by the viscosity / surface-tension kernel
//   recB[i] = (x, y, z, p)  — 16 bytes, ONE 128-bit gather — for the pressure-force kernel: positions at time t and
//             the current pressure. rho*_j lives in rho[N + j] and is fetched only for the pairs that carry pressure
//             (most pairs of a fluid near rest do not, see pressure_term); the neighbour's TYPE, which round 1 kept in
//             this record for the integration's boundary handling, is now a bit mask per particle (bMask / elasticMask)
//             that the forces kernel derives from recA once per step. A 16-byte gather costs ~12 L1 wavefronts per warp
//             request against ~22 for a 32-byte one, and the kernel is bound by exactly that (L1 data pipe 98 %).
__device__ __forceinline__ void ld256(const float4 *rec, float4 &a, float4 &b) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(rec));
}

// Neighbour ids are kept as int32 in 32x32 tiles: tile t holds particles [32t, 32t+32),
// element (k, lane) at nb[(t*32 + k)*32 + lane] — a warp reading slot k of its 32
// particles touches one 128-byte line.
__device__ __forceinline__ size_t nb_index(int id, int k) {
  return ((size_t)(id >> 5) * SPH_NB + (size_t)k) * 32 + (size_t)(id & 31);
}

// Ids of slots [k0, k0 + 8) of particle id: eight independent, coalesced loads issued together. The hot loops
// fetch the next eight while they work on the current eight, so no iteration waits for its id to arrive.
template <int B>
__device__ __forceinline__ void nb_load8(const int *__restrict__ nb, int id, int k0, int (&v)[B]) {
  const int *p = nb + nb_index(id, k0);
#pragma unroll
  for (int j = 0; j < B; j++) v[j] = __ldg(p + 32 * j);
}
