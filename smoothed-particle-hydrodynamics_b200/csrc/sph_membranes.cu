// sph_membranes.cu — liquid <-> membrane repulsion.
//
// Replaces, with identical results, the reference's
//   clearMembraneBuffers                       sphFluid.cl:1214-1227
//   computeInteractionWithMembranes            sphFluid.cl:1369-1650
//     (+ calculateProjectionOfPointToPlane :1250-1308, calcDeterminant3x3 :1229-1247)
//   computeInteractionWithMembranes_finalize   sphFluid.cl:1652-1682
//
// The reference collects, per liquid particle, up to 32 (normal, distance) pairs in private
// arrays (32 x float4 + 32 x float + 32 x int) and reduces them in a second loop in the same
// order; here the Ihmsen sums are accumulated on the fly in registers (same order, same
// operations) and discarded if the reference would have returned early, so no local memory.
#include "sph_ctx.h"

__device__ __forceinline__ float det3(float4 c1, float4 c2, float4 c3) {  // sphFluid.cl:1244-1245
  return c1.x * c2.y * c3.z + c1.y * c2.z * c3.x + c1.z * c2.x * c3.y - c1.z * c2.y * c3.x - c1.x * c2.z * c3.y -
         c1.y * c2.x * c3.z;
}

// sphFluid.cl:1250-1308. The reference passes b = (0, b_1, b_2, b_3) to a determinant that reads
// .x,.y,.z, so the right-hand side it actually uses is (0, b_1, b_2); reproduced as is.
__device__ __forceinline__ float4 project_point_to_plane(float4 ps, float4 pa, float4 pb, float4 pc) {
  float4 pm = f4(0.f, 0.f, 0.f, 0.f);
  float b_1 = pa.x * ((pb.y - pa.y) * (pc.z - pa.z) - (pb.z - pa.z) * (pc.y - pa.y)) +
              pa.y * ((pb.z - pa.z) * (pc.x - pa.x) - (pb.x - pa.x) * (pc.z - pa.z)) +
              pa.z * ((pb.x - pa.x) * (pc.y - pa.y) - (pb.y - pa.y) * (pc.x - pa.x));
  float b_2 = ps.x * (pb.x - pa.x) + ps.y * (pb.y - pa.y) + ps.z * (pb.z - pa.z);
  float4 a_1 = f4((pb.y - pa.y) * (pc.z - pa.z) - (pb.z - pa.z) * (pc.y - pa.y), pb.x - pa.x, pc.x - pa.x, 0.f);
  float4 a_2 = f4((pb.z - pa.z) * (pc.x - pa.x) - (pb.x - pa.x) * (pc.z - pa.z), pb.y - pa.y, pc.y - pa.y, 0.f);
  float4 a_3 = f4((pb.x - pa.x) * (pc.y - pa.y) - (pb.y - pa.y) * (pc.x - pa.x), pb.z - pa.z, pc.z - pa.z, 0.f);
  float4 b = f4(0.f, b_1, b_2, 0.f);
  float den = det3(a_1, a_2, a_3);
  if (den != 0.f) {
    pm.x = det3(b, a_2, a_3) / den;
    pm.y = det3(a_1, b, a_3) / den;
    pm.z = det3(a_1, a_2, b) / den;
  } else {
    pm.w = -1.f;
  }
  return pm;
}

__global__ void __launch_bounds__(256) k_clear_membrane_buffers(DevParams d, float4 *__restrict__ position,
                                                                float4 *__restrict__ velocity) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  position[d.N + id] = f4(0.f, 0.f, 0.f, 0.f);
  velocity[d.N + id] = f4(0.f, 0.f, 0.f, 0.f);
}

// One elastic neighbour of a liquid particle (the body of the reference's neighbour loop, sphFluid.cl:1425-1550):
// the averaged normal of the membranes the neighbour belongs to, weighted as in :1580-1586.
//   returns 0: the neighbour contributes nothing (no membrane, or not elastic)
//           1: contribution in (nw = normal * w, w, w2 = w * (r0 - dist))
//           2: the reference aborts the whole particle here (:1468-1472, :1501-1505)
__device__ __forceinline__ int membrane_neighbour(const DevParams &d, const float4 ps, const int jd,
                                                  const float4 *__restrict__ position,
                                                  const uint2 *__restrict__ particleIndex,
                                                  const float *__restrict__ sortedType, const int *__restrict__ pml,
                                                  const int *__restrict__ membraneData, float4 &nw, float &w, float &w2) {
  if ((int)__ldg(sortedType + jd) != SPH_ELASTIC_PARTICLE) return 0;
  const int jsrc = (int)particleIndex[jd].y;
  float4 vij = ps - position[jsrc];
  vij.z = 0.f;                               // :1437 — z dropped, but the 4-lane dot keeps the type-tag lane
  const float dist = sqrtf(dot4(vij, vij));  // :1438
  float4 normal = f4(0.f, 0.f, 0.f, 0.f);
  int ijk = 0;
  for (int mli = 0; mli < SPH_MAX_MEMBRANES_INCLUDING_SAME_PARTICLE; mli++) {
    int mdi = __ldg(pml + jsrc * SPH_MAX_MEMBRANES_INCLUDING_SAME_PARTICLE + mli);
    if (!(mdi > -1)) break;
    float4 pi = position[__ldg(membraneData + mdi * 3 + 0)];
    float4 pj = position[__ldg(membraneData + mdi * 3 + 1)];
    float4 pk = position[__ldg(membraneData + mdi * 3 + 2)];
    float4 pp = project_point_to_plane(ps, pi, pj, pk);
    if (pp.w == -1.f) return 2;  // :1468-1472
    float4 nrm = ps - pp;
    float len = sqrtf(nrm.x * nrm.x + nrm.y * nrm.y + nrm.z * nrm.z);
    if (len > 0.f) {
      nrm = nrm / len;
      normal = normal + nrm;
      ijk++;
    } else {
      return 2;  // :1501-1505
    }
  }
  if (ijk == 0) return 0;
  normal = normal / (float)ijk;  // :1544-1550 then :1580-1586 for this entry
  w = max0((d.r0 - dist) / d.r0);
  nw = normal * w;
  w2 = w * (d.r0 - dist);
  return 1;
}

// :1590-1603: the displacement from the accumulated sums
__device__ __forceinline__ bool membrane_displacement(float4 n_c_i, float w_sum, float w_second, float4 &disp) {
  n_c_i.w = 0.f;
  float len = dot4(n_c_i, n_c_i);
  if (len == 0.f) return false;
  len = sqrtf(len);
  float4 dp = (1.0f * ((n_c_i / len) * w_second)) / w_sum;  // :1600
  disp.x = 0.f + dp.x;  // position[N+src] was cleared to 0, then += (:1601-1603)
  disp.y = 0.f + dp.y;
  disp.z = 0.f + dp.z;
  return true;
}

// Stage API: one thread per sorted particle; clearMembraneBuffers zeroed the displacement slots just before.
__global__ void __launch_bounds__(128) k_membrane_interaction(DevParams d, float4 *__restrict__ position,
                                                              const uint2 *__restrict__ particleIndex,
                                                              const float *__restrict__ sortedType,
                                                              const int *__restrict__ nb, const int *__restrict__ pml,
                                                              const int *__restrict__ membraneData) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  if ((int)sortedType[id] != SPH_LIQUID_PARTICLE || pml == nullptr) return;
  const int src = (int)particleIndex[id].y;
  const float4 ps = position[src];
  float4 n_c_i = f4(0.f, 0.f, 0.f, 0.f);
  float w_sum = 0.f, w_second = 0.f;
  int jd_counter = 0;
  for (int nc = 0; nc < SPH_NB; nc++) {
    const int jd = nb[nb_index(id, nc)];
    if (jd == SPH_NO_ID) break;  // :1557
    float4 nw;
    float w, w2;
    const int r = membrane_neighbour(d, ps, jd, position, particleIndex, sortedType, pml, membraneData, nw, w, w2);
    if (r == 2) return;
    if (r == 1) { n_c_i = n_c_i + nw; w_sum += w; w_second += w2; jd_counter++; }
  }
  float4 disp = f4(0.f, 0.f, 0.f, 0.f);
  if (jd_counter > 0 && membrane_displacement(n_c_i, w_sum, w_second, disp)) position[d.N + src] = disp;
}

// Fused path: EIGHT LANES per liquid particle that has elastic neighbours (the list the integration kernel wrote).
// Lane l evaluates the elastic neighbour slots l, l + 8, ... (the expensive part: up to 7 plane projections each), the
// contributions go to shared memory in slot order, lane 0 adds them in that order (the reference's order) and writes
// the displacement. No position is written here, so reading the membrane vertices races with nothing; the finalize
// kernel below applies the displacements. One thread per particle took 0.10 ms of the worm scene's 0.46 ms step,
// almost all of it waiting for the few lanes of a warp that had work.
#define ML_L 8
#define ML_THREADS 256
#define ML_P (ML_THREADS / ML_L)
__global__ void __launch_bounds__(ML_THREADS) k_membrane_lanes(DevParams d, float4 *__restrict__ position,
                                                               const uint2 *__restrict__ particleIndex,
                                                               const float *__restrict__ sortedType,
                                                               const int *__restrict__ nb, const int *__restrict__ pml,
                                                               const int *__restrict__ membraneData,
                                                               const unsigned *__restrict__ elasticMask,
                                                               const int *__restrict__ memList,
                                                               const int *__restrict__ memCount) {
  __shared__ float4 sNw[ML_P][SPH_NB];
  __shared__ float sW[ML_P][SPH_NB + 1], sW2[ML_P][SPH_NB + 1];
  __shared__ unsigned char sKind[ML_P][SPH_NB + 4];
  pdl_entry();
  const int t = threadIdx.x, pl = t / ML_L, l = t % ML_L;
  const int item = blockIdx.x * ML_P + pl;
  const bool live = item < *memCount;
  int id = 0, src = 0, ne = 0;
  if (live) {
    id = memList[item];
    src = (int)particleIndex[id].y;
    unsigned m = elasticMask[id];
    ne = __popc(m);
    const float4 ps = position[src];
    // q-th set bit of the mask for q = l, l + 8, ...
    for (int q = 0; q < l; q++) m &= m - 1u;
    for (int q = l; q < ne; q += ML_L) {
      const int nc = __ffs(m) - 1;
      const int jd = nb[nb_index(id, nc)];
      float4 nw = f4(0.f, 0.f, 0.f, 0.f);
      float w = 0.f, w2 = 0.f;
      const int r = membrane_neighbour(d, ps, jd, position, particleIndex, sortedType, pml, membraneData, nw, w, w2);
      sNw[pl][q] = nw; sW[pl][q] = w; sW2[pl][q] = w2; sKind[pl][q] = (unsigned char)r;
      for (int k = 0; k < ML_L && m; k++) m &= m - 1u;
    }
  }
  __syncthreads();
  if (!live || l != 0) return;
  float4 n_c_i = f4(0.f, 0.f, 0.f, 0.f);
  float w_sum = 0.f, w_second = 0.f;
  int jd_counter = 0;
  bool aborted = false;
  for (int q = 0; q < ne; q++) {
    const int r = sKind[pl][q];
    if (r == 2) { aborted = true; break; }
    if (r == 1) { n_c_i = n_c_i + sNw[pl][q]; w_sum += sW[pl][q]; w_second += sW2[pl][q]; jd_counter++; }
  }
  float4 disp = f4(0.f, 0.f, 0.f, 0.f);
  if (!aborted && jd_counter > 0 && membrane_displacement(n_c_i, w_sum, w_second, disp)) position[d.N + src] = disp;
}

// Fused path, after k_membrane_lanes: computeInteractionWithMembranes_finalize (:1652-1682) for every non-boundary
// particle (the displacement slots were zeroed by the integration kernel, sphFluid.cl:1214-1227), and the list
// counter is reset for the next step.
__global__ void __launch_bounds__(256) k_membrane_finalize_fused(DevParams d, float4 *__restrict__ position,
                                                                 const uint2 *__restrict__ particleIndex,
                                                                 const float *__restrict__ sortedType, int *memCount) {
  pdl_entry();
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id == 0) *memCount = 0;
  if (id >= d.N) return;
  if ((int)sortedType[id] == SPH_BOUNDARY_PARTICLE) return;
  const int src = (int)particleIndex[id].y;
  const float4 p = position[src];
  const float4 q = p + position[d.N + src];
  if (__float_as_uint(q.x) != __float_as_uint(p.x) || __float_as_uint(q.y) != __float_as_uint(p.y) ||
      __float_as_uint(q.z) != __float_as_uint(p.z) || __float_as_uint(q.w) != __float_as_uint(p.w))
    position[src] = q;
}

__global__ void __launch_bounds__(256) k_membrane_finalize(DevParams d, float4 *__restrict__ position,
                                                           const uint2 *__restrict__ particleIndex,
                                                           const float *__restrict__ sortedType) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  if ((int)sortedType[id] == SPH_BOUNDARY_PARTICLE) return;
  const int src = (int)particleIndex[id].y;
  position[src] = position[src] + position[d.N + src];
}

void launch_clear_membrane_buffers(sph_ctx *c) {
  k_clear_membrane_buffers<<<(c->d.N + 255) / 256, 256, 0, c->stream>>>(c->d, c->position, c->velocity);
}

void launch_membrane_interaction(sph_ctx *c, bool fuse_finalize) {
  const int *pml = c->has_pml ? c->pml : nullptr;
  if (fuse_finalize && pml) {
    // the reference's memset race (k_membrane_finalize_fused resets *memCount after the lanes kernel has read it) is
    // ordered by the stream: lanes kernel, then finalize kernel
    launch_k(c, k_membrane_lanes, dim3((c->d.N + ML_P - 1) / ML_P), dim3(ML_THREADS), c->d, c->position, c->particleIndex,
             c->sortedType, c->nbIds, pml, c->membraneData, c->elasticMask, c->memList, c->memCount);
    launch_k(c, k_membrane_finalize_fused, dim3((c->d.N + 255) / 256), dim3(256), c->d, c->position, c->particleIndex,
             c->sortedType, c->memCount);
  } else {
    k_membrane_interaction<<<(c->d.N + 127) / 128, 128, 0, c->stream>>>(c->d, c->position, c->particleIndex, c->sortedType,
                                                                        c->nbIds, pml, c->membraneData);
  }
}

void launch_membrane_finalize(sph_ctx *c) {
  k_membrane_finalize<<<(c->d.N + 255) / 256, 256, 0, c->stream>>>(c->d, c->position, c->particleIndex, c->sortedType);
}

// Force the (lazily loaded, CUDA >= 12.2) code of this file's step kernels onto the device. A kernel that spins on a
// neighbour's delivery flag (sph_slab.cu) must never be resident while the host loads another kernel for the first
// time — module loading synchronises the context and would wait for the spinning kernel forever.
static void preload(const void *const *fns, int n) {
  cudaFuncAttributes a;
  for (int i = 0; i < n; i++) cudaFuncGetAttributes(&a, fns[i]);
  cudaGetLastError();
}
void preload_membrane_kernels() {
  const void *const f[] = {(const void *)k_membrane_lanes, (const void *)k_membrane_finalize_fused,
                           (const void *)k_clear_membrane_buffers};
  preload(f, sizeof f / sizeof f[0]);
}
