// sph_lanes.cu — the gather kernels of the fused step with SEVERAL LANES PER PARTICLE, for scenes that do not
// fill the GPU with one thread per particle (the reference's own scenes: 61 k and 233 k particles).
//
// With one thread per particle such a launch is a single partial wave of CTAs and lasts as long as one thread's
// chain of 32 dependent gathers. Here L lanes share a particle: lane l evaluates the neighbour slots l, l + L, ...
// (32 / L independent gathers, all requested at once), the per-slot terms go to shared memory, and the sums are
// then formed SEQUENTIALLY IN SLOT ORDER 0..31 by one lane per component — the reference's do{}while(++nc<32)
// order, so the results stay bit-identical (terms the reference skips are added as +0, see sph_terms.cuh).
// Same per-neighbour device functions as sph_pcisph.cu.
#include "sph_ctx.h"
#include "sph_terms.cuh"

#define LN_THREADS 256
#ifndef LN_T4
#define LN_T4 128
#endif

// shared-memory layout of the term table: element (component c, particle pl, slot k) at c * CS + pl * S + k with
// S = 32 + L and CS = P * S + 1: the L x (32 / L) lanes of a warp write distinct banks, and so do the summing lanes.
template <int L> struct LaneGeom {
  static constexpr int T = L <= 4 ? LN_T4 : LN_THREADS;  // threads per CTA (the L = 4 term table of the forces kernel with 256 threads exceeds 48 KB)
  static constexpr int P = T / L;            // particles per CTA
  static constexpr int S = 32 + L;
  static constexpr int CS = P * S + 1;
  static constexpr int SLOTS = SPH_NB / L;    // slots per lane
};

// forces + predictPositions #0 + the first predictDensity / correctPressure (see k_forces<true, .>)
template <int L>
__global__ void __launch_bounds__(LaneGeom<L>::T) k_forces_lanes(DevParams d, const float4 *__restrict__ sp,
                                                             const float4 *__restrict__ sv,
                                                             const float *__restrict__ sortedType,
                                                             const int *__restrict__ nb, float *__restrict__ rho,
                                                             float4 *__restrict__ acc, float *__restrict__ pressure,
                                                             float4 *__restrict__ spPred, const float4 *__restrict__ recA,
                                                             unsigned *__restrict__ bMask, unsigned *__restrict__ elasticMask,
                                                             float4 *__restrict__ recB) {
  using G = LaneGeom<L>;
  static_assert(L >= 4, "lane 3 sums the density terms");
  __shared__ float term[7 * G::CS];   // 3 viscosity + 3 surface tension + 1 predicted-density term per slot
  pdl_entry();
  const int t = threadIdx.x, pl = t / L, l = t % L;
  const int id = blockIdx.x * G::P + pl;
  const bool live = id < d.N;
  const float4 zero = f4(0.f, 0.f, 0.f, 0.f);
  float4 pi = zero, vi = zero, piP = zero;
  bool active = false;   // not a boundary particle: takes part in the forces
  unsigned bmask = 0u, emask = 0u;
  if (live) {
    pi = sp[id];
    active = !is_boundary(sortedType[id]);
    if (active) vi = sv[id];
    piP = active ? predict_position(d, pi, vi, zero) : pi;
    int jd[G::SLOTS];
    float4 Pj[G::SLOTS], Vj[G::SLOTS];
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) jd[i] = __ldg(nb + nb_index(id, l + L * i));
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) ld256(recA + 2 * (size_t)(jd[i] != SPH_NO_ID ? jd[i] : id), Pj[i], Vj[i]);
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) {
      float a[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (jd[i] != SPH_NO_ID) {
        const bool jBoundary = is_boundary(Pj[i].w);
        // type masks of the neighbour list for the integration / membrane kernels (see k_forces)
        if (jBoundary) bmask |= 1u << (l + L * i);
        if ((int)Pj[i].w == SPH_ELASTIC_PARTICLE) emask |= 1u << (l + L * i);
        if (!density_term(d, piP, jBoundary ? Pj[i] : predict_position(d, Pj[i], Vj[i], zero), a[6])) a[6] = 0.f;
        if (!active || !forces_term(d, pi, vi, Pj[i], Vj[i], a[0], a[1], a[2], a[3], a[4], a[5]))
          a[0] = a[1] = a[2] = a[3] = a[4] = a[5] = 0.f;
      }
      float *o = term + pl * G::S + l + L * i;
#pragma unroll
      for (int c = 0; c < 7; c++) o[c * G::CS] = a[c];
    }
  }
  __syncthreads();
  // lane c < 3 sums component c of the viscosity sum and of the surface-tension sum, lane 3 the density terms (fp64),
  // all over the slots in order
  float s = 0.f, u = 0.f;
  if (active && l < 3) {
    const float *a = term + l * G::CS + pl * G::S, *b = a + 3 * G::CS;
#pragma unroll 8
    for (int k = 0; k < SPH_NB; k++) { s += a[k]; u += b[k]; }
  }
  if (live && l == 3) {
    double density = 0.0;
    const float *a = term + 6 * G::CS + pl * G::S;
#pragma unroll 8
    for (int k = 0; k < SPH_NB; k++) density += (double)a[k];
    // predictDensity tail (:1054-1058) + correctPressure (:1093-1096) on pressure = 0 (:697)
    if (density < (double)d.hScaled6) density = (double)d.hScaled6;
    density *= d.massWpoly6;
    const float rhoP = (float)density;
    rho[d.N + id] = rhoP;
    float p_corr = (rhoP - d.rho0) * d.delta;
    if (p_corr < 0.f) p_corr = 0.f;
    const float p = 0.f + p_corr;
    pressure[id] = p;
    recB[id] = f4(pi.x, pi.y, pi.z, p);
    spPred[id] = piP;
  }
  const int base = (t & 31) - l;
  const float sx = __shfl_sync(0xffffffffu, s, base), sy = __shfl_sync(0xffffffffu, s, base + 1),
              sz = __shfl_sync(0xffffffffu, s, base + 2);
  const float tx = __shfl_sync(0xffffffffu, u, base), ty = __shfl_sync(0xffffffffu, u, base + 1),
              tz = __shfl_sync(0xffffffffu, u, base + 2);
#pragma unroll
  for (int o = 1; o < L; o <<= 1) {
    bmask |= __shfl_xor_sync(0xffffffffu, bmask, o);
    emask |= __shfl_xor_sync(0xffffffffu, emask, o);
  }
  if (!live || l != 0) return;
  if (!active) {  // boundary particle, as in k_forces
    acc[id] = zero;
    acc[d.N + id] = zero;
    return;
  }
  const float cv = d.massMu * (float)(d.del2W / (double)rho[id]);         // :688
  float4 a = f4(sx * cv, sy * cv, sz * cv, 0.f);
  a = a + f4(d.gravx, d.gravy, d.gravz, 0.f);
  a = a + f4(tx, ty, tz, 0.f);
  acc[id] = a;
  acc[d.N + id] = zero;
  bMask[id] = bmask;
  if (elasticMask) elasticMask[id] = emask;
}

// predictDensity + correctPressure (k_predict_density<true, .>)
template <int L>
__global__ void __launch_bounds__(LaneGeom<L>::T) k_predict_density_lanes(DevParams d, const float4 *__restrict__ spPred,
                                                                      const int *__restrict__ nb, float *__restrict__ rho,
                                                                      float *__restrict__ pressure,
                                                                      float4 *__restrict__ recB) {
  using G = LaneGeom<L>;
  __shared__ float term[G::CS];
  pdl_entry();
  const int t = threadIdx.x, pl = t / L, l = t % L;
  const int id = blockIdx.x * G::P + pl;
  const bool live = id < d.N;
  if (live) {
    const float4 pi = spPred[id];
    int jd[G::SLOTS];
    float4 Pj[G::SLOTS];
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) jd[i] = __ldg(nb + nb_index(id, l + L * i));
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) Pj[i] = ldg4(spPred + (jd[i] != SPH_NO_ID ? jd[i] : id));
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) {
      float w = 0.f;
      if (jd[i] != SPH_NO_ID && !density_term(d, pi, Pj[i], w)) w = 0.f;
      term[pl * G::S + l + L * i] = w;
    }
  }
  __syncthreads();
  if (!live || l != 0) return;
  double density = 0.0;
  const float *a = term + pl * G::S;
#pragma unroll 8
  for (int k = 0; k < SPH_NB; k++) density += (double)a[k];
  if (density < (double)d.hScaled6) density = (double)d.hScaled6;
  density *= d.massWpoly6;
  const float rhoP = (float)density;
  rho[d.N + id] = rhoP;
  float p_corr = (rhoP - d.rho0) * d.delta;  // :1093-1096
  if (p_corr < 0.f) p_corr = 0.f;
  float4 rec = recB[id];
  rec.w += p_corr;
  pressure[id] = rec.w;
  recB[id] = rec;
}

// computePressureForceAcceleration + predictPositions (MODE 1) or + integrate (MODE 2): k_pressure_force<MODE, .>
template <int MODE, int L>
__global__ void __launch_bounds__(LaneGeom<L>::T) k_pressure_force_lanes(DevParams d, const float4 *__restrict__ spT,
                                                                     const float4 *__restrict__ sv,
                                                                     const int *__restrict__ nb,
                                                                     const float4 *__restrict__ recB,
                                                                     float4 *__restrict__ acc, const float4 *__restrict__ sp,
                                                                     float4 *__restrict__ spPred,
                                                                     const uint2 *__restrict__ particleIndex,
                                                                     float4 *__restrict__ position,
                                                                     float4 *__restrict__ velocity,
                                                                     const unsigned *__restrict__ elasticMask,
                                                                     int *__restrict__ memList, int *__restrict__ memCount,
                                                                     const unsigned *__restrict__ bMask,
                                                                     const float *__restrict__ rhoP) {
  using G = LaneGeom<L>;
  __shared__ float term[3 * G::CS];
  pdl_entry();
  const int t = threadIdx.x, pl = t / L, l = t % L;
  const int id = blockIdx.x * G::P + pl;
  const bool live = id < d.N;
  float4 pi = f4(0.f, 0.f, 0.f, 0.f);
  bool active = false;
  if (live) {
    pi = spT[id];
    active = !is_boundary(pi.w);
  }
  float rho_i = 1.f;
  if (active) {
    const float p_i = recB[id].w;
    rho_i = rhoP[id];
    const float closeThr = (float)d.closeThr;
    const float hq = d.hScaled * 0.25f;
    int jd[G::SLOTS];
    float4 Pj[G::SLOTS];
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) jd[i] = __ldg(nb + nb_index(id, l + L * i));
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) Pj[i] = ldg4(recB + (jd[i] != SPH_NO_ID ? jd[i] : id));
#pragma unroll
    for (int i = 0; i < G::SLOTS; i++) {
      const int k = l + L * i;
      float ax = 0.f, ay = 0.f, az = 0.f;
      if (jd[i] != SPH_NO_ID && !pressure_term(d, pi, p_i, Pj[i], Pj[i].w, rhoP, jd[i], closeThr, hq, ax, ay, az)) ax = ay = az = 0.f;
      float *o = term + pl * G::S + k;
      o[0] = ax; o[G::CS] = ay; o[2 * G::CS] = az;
    }
  }
  __syncthreads();
  float s = 0.f;
  if (active && l < 3) {
    const float *a = term + l * G::CS + pl * G::S;
#pragma unroll 8
    for (int k = 0; k < SPH_NB; k++) s += a[k];
  }
  const int base = (t & 31) - l;
  const float rx = __shfl_sync(0xffffffffu, s, base), ry = __shfl_sync(0xffffffffu, s, base + 1),
              rz = __shfl_sync(0xffffffffu, s, base + 2);
  if (!live || l != 0) return;
  if (!active) {
    acc[d.N + id] = f4(0.f, 0.f, 0.f, 0.f);  // :1126-1129
    return;
  }
  const float cf = (float)(d.massGradW / (double)rho_i);  // :1194
  const float4 result = f4(rx * cf, ry * cf, rz * cf, 0.f);
  acc[d.N + id] = result;
  if (MODE == 1) spPred[id] = predict_position(d, sp[id], sv[id], result);
  if (MODE == 2) {
    const unsigned serial = particleIndex[id].y;
    integrate_particle(d, id, sp[id], sv[id], acc[id] + result, pi.w, serial, bMask[id], spT, sv, nb, position, velocity,
                       /*foldFinalize=*/elasticMask == nullptr);
    if (elasticMask) membrane_prologue(d, id, serial, elasticMask[id], pi.w, memList, memCount, position, velocity);
  }
}

// ---------------------------------------------------------------- launchers
template <int L> static inline int lanes_grid(int n) { return (n + LaneGeom<L>::P - 1) / LaneGeom<L>::P; }

bool launch_forces_lanes(sph_ctx *c, int L) {
  float4 *pred = c->sortedPosition + c->d.N;
#define F_ARGS c->d, c->sortedPosition, c->sortedVelocity, c->sortedType, c->nbIds, c->rho, c->acceleration, c->pressure, pred, c->recA, c->bMask, c->elasticMask, c->recB
  if (L == 4) launch_k(c, k_forces_lanes<4>, dim3(lanes_grid<4>(c->d.N)), dim3(LaneGeom<4>::T), F_ARGS);
  else if (L == 8) launch_k(c, k_forces_lanes<8>, dim3(lanes_grid<8>(c->d.N)), dim3(LaneGeom<8>::T), F_ARGS);
  else if (L == 16) launch_k(c, k_forces_lanes<16>, dim3(lanes_grid<16>(c->d.N)), dim3(LaneGeom<16>::T), F_ARGS);
  else return false;
#undef F_ARGS
  return true;
}

bool launch_predict_density_lanes(sph_ctx *c, int L) {
  const float4 *pred = c->sortedPosition + c->d.N;
#define D_ARGS c->d, pred, c->nbIds, c->rho, c->pressure, c->recB
  if (L == 4) launch_k(c, k_predict_density_lanes<4>, dim3(lanes_grid<4>(c->d.N)), dim3(LaneGeom<4>::T), D_ARGS);
  else if (L == 8) launch_k(c, k_predict_density_lanes<8>, dim3(lanes_grid<8>(c->d.N)), dim3(LaneGeom<8>::T), D_ARGS);
  else if (L == 16) launch_k(c, k_predict_density_lanes<16>, dim3(lanes_grid<16>(c->d.N)), dim3(LaneGeom<16>::T), D_ARGS);
  else return false;
#undef D_ARGS
  return true;
}

bool launch_pressure_force_lanes(sph_ctx *c, int mode, int L) {
  float4 *pred = c->sortedPosition + c->d.N;
#define P_ARGS c->d, c->spT, c->sortedVelocity, c->nbIds, c->recB, c->acceleration, c->sortedPosition, pred, c->particleIndex, \
               c->position, c->velocity, c->elasticMask, c->memList, c->memCount, c->bMask, (const float *)(c->rho + c->d.N)
#define P_LAUNCH(M, LL) launch_k(c, k_pressure_force_lanes<M, LL>, dim3(lanes_grid<LL>(c->d.N)), dim3(LaneGeom<LL>::T), P_ARGS)
  if (mode != 1 && mode != 2) return false;
  if (L == 4) { if (mode == 1) P_LAUNCH(1, 4); else P_LAUNCH(2, 4); }
  else if (L == 8) { if (mode == 1) P_LAUNCH(1, 8); else P_LAUNCH(2, 8); }
  else if (L == 16) { if (mode == 1) P_LAUNCH(1, 16); else P_LAUNCH(2, 16); }
  else return false;
#undef P_LAUNCH
#undef P_ARGS
  return true;
}

// Force the (lazily loaded, CUDA >= 12.2) code of this file's step kernels onto the device. A kernel that spins on a
// neighbour's delivery flag (sph_slab.cu) must never be resident while the host loads another kernel for the first
// time — module loading synchronises the context and would wait for the spinning kernel forever.
static void preload(const void *const *fns, int n) {
  cudaFuncAttributes a;
  for (int i = 0; i < n; i++) cudaFuncGetAttributes(&a, fns[i]);
  cudaGetLastError();
}
void preload_lanes_kernels() {
  const void *const f[] = {(const void *)k_forces_lanes<4>, (const void *)k_forces_lanes<8>, (const void *)k_forces_lanes<16>,
                           (const void *)k_predict_density_lanes<4>, (const void *)k_predict_density_lanes<8>,
                           (const void *)k_predict_density_lanes<16>, (const void *)k_pressure_force_lanes<1, 4>,
                           (const void *)k_pressure_force_lanes<1, 8>, (const void *)k_pressure_force_lanes<1, 16>,
                           (const void *)k_pressure_force_lanes<2, 4>, (const void *)k_pressure_force_lanes<2, 8>,
                           (const void *)k_pressure_force_lanes<2, 16>};
  preload(f, sizeof f / sizeof f[0]);
}
