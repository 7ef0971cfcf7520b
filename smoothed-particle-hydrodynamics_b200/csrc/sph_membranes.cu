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

// One thread per sorted particle. FUSED: also clears the displacement slots (clearMembraneBuffers) and applies
// the finalize step to LIQUID particles (legal: only liquid particles receive a displacement and no other thread
// reads a liquid particle's position in this kernel). Elastic particles — whose positions other threads read here
// as membrane vertices — are finalized by k_membrane_finalize_elastic afterwards: their displacement is +0, so
// the add can only turn a -0 coordinate into +0, but it must not race with those reads.
template <bool FUSED>
__global__ void __launch_bounds__(128) k_membrane_interaction(DevParams d, float4 *__restrict__ position,
                                                              const uint2 *__restrict__ particleIndex,
                                                              const float *__restrict__ sortedType,
                                                              const int *__restrict__ nb, const int *__restrict__ pml,
                                                              const int *__restrict__ membraneData,
                                                              const unsigned *__restrict__ elasticMask,
                                                              float4 *__restrict__ velocity) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  const int src = (int)particleIndex[id].y;
  const int type = (int)sortedType[id];
  float4 disp = f4(0.f, 0.f, 0.f, 0.f);
  bool moved = false;
  if (type == SPH_LIQUID_PARTICLE && pml != nullptr) {
    const float4 ps = position[src];
    float4 n_c_i = f4(0.f, 0.f, 0.f, 0.f);
    float w_sum = 0.f, w_second = 0.f;
    int jd_counter = 0;
    bool aborted = false;
    // fused path: the integration kernel left a mask of the slots that hold elastic particles, so the (many) liquid
    // particles without an elastic neighbour skip the scan; the visiting order of the others is unchanged
    unsigned slots = elasticMask ? elasticMask[id] : 0xffffffffu;
    while (slots && !aborted) {
      const int nc = __ffs(slots) - 1;
      slots &= slots - 1;
      int jd = nb[nb_index(id, nc)];
      if (jd == SPH_NO_ID) break;  // :1557
      if ((int)__ldg(sortedType + jd) != SPH_ELASTIC_PARTICLE) continue;
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
        if (pp.w == -1.f) { aborted = true; break; }  // :1468-1472
        float4 nrm = ps - pp;
        float len = sqrtf(nrm.x * nrm.x + nrm.y * nrm.y + nrm.z * nrm.z);
        if (len > 0.f) {
          nrm = nrm / len;
          normal = normal + nrm;
          ijk++;
        } else { aborted = true; break; }  // :1501-1505
      }
      if (aborted) break;
      if (ijk > 0) {  // :1544-1550 then :1580-1586 for this entry
        normal = normal / (float)ijk;
        float w = max0((d.r0 - dist) / d.r0);
        n_c_i = n_c_i + normal * w;
        w_sum += w;
        w_second += w * (d.r0 - dist);
        jd_counter++;
      }
    }
    if (!aborted && jd_counter > 0) {
      n_c_i.w = 0.f;
      float len = dot4(n_c_i, n_c_i);
      if (len != 0.f) {
        len = sqrtf(len);
        float4 dp = (1.0f * ((n_c_i / len) * w_second)) / w_sum;  // :1600
        disp.x = 0.f + dp.x;  // position[N+src] was cleared to 0, then += (:1601-1603)
        disp.y = 0.f + dp.y;
        disp.z = 0.f + dp.z;
        moved = true;
      }
    }
  }
  if (FUSED) {
    position[d.N + src] = disp;
    velocity[d.N + src] = f4(0.f, 0.f, 0.f, 0.f);  // clearMembraneBuffers, :1225
    if (type == SPH_LIQUID_PARTICLE) {  // finalize, :1670-1673
      float4 p = position[src];
      float4 q = p + disp;
      if (__float_as_uint(q.x) != __float_as_uint(p.x) || __float_as_uint(q.y) != __float_as_uint(p.y) ||
          __float_as_uint(q.z) != __float_as_uint(p.z) || __float_as_uint(q.w) != __float_as_uint(p.w))
        position[src] = q;
    }
  } else if (moved) {  // stage API: clearMembraneBuffers zeroed the slot just before
    position[d.N + src] = f4(disp.x, disp.y, disp.z, 0.f);
  }
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

// finalize (:1670-1673) of the fused path for everything that is neither liquid (done in the interaction kernel)
// nor boundary (never finalized): the elastic matter
__global__ void __launch_bounds__(256) k_membrane_finalize_elastic(DevParams d, float4 *__restrict__ position,
                                                                   const uint2 *__restrict__ particleIndex,
                                                                   const float *__restrict__ sortedType) {
  int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= d.N) return;
  const int type = (int)sortedType[id];
  if (type == SPH_BOUNDARY_PARTICLE || type == SPH_LIQUID_PARTICLE) return;
  const int src = (int)particleIndex[id].y;
  const float4 p = position[src];
  const float4 q = p + position[d.N + src];
  if (__float_as_uint(q.x) != __float_as_uint(p.x) || __float_as_uint(q.y) != __float_as_uint(p.y) ||
      __float_as_uint(q.z) != __float_as_uint(p.z) || __float_as_uint(q.w) != __float_as_uint(p.w))
    position[src] = q;
}

void launch_clear_membrane_buffers(sph_ctx *c) {
  k_clear_membrane_buffers<<<(c->d.N + 255) / 256, 256, 0, c->stream>>>(c->d, c->position, c->velocity);
}

void launch_membrane_interaction(sph_ctx *c, bool fuse_finalize) {
  const int *pml = c->has_pml ? c->pml : nullptr;
  if (fuse_finalize) {
    k_membrane_interaction<true><<<(c->d.N + 127) / 128, 128, 0, c->stream>>>(c->d, c->position, c->particleIndex,
                                                                            c->sortedType, c->nbIds, pml, c->membraneData, c->elasticMask, c->velocity);
    k_membrane_finalize_elastic<<<(c->d.N + 255) / 256, 256, 0, c->stream>>>(c->d, c->position, c->particleIndex, c->sortedType);
  } else
    k_membrane_interaction<false><<<(c->d.N + 127) / 128, 128, 0, c->stream>>>(c->d, c->position, c->particleIndex,
                                                                             c->sortedType, c->nbIds, pml, c->membraneData, nullptr, c->velocity);
}

void launch_membrane_finalize(sph_ctx *c) {
  k_membrane_finalize<<<(c->d.N + 255) / 256, 256, 0, c->stream>>>(c->d, c->position, c->particleIndex, c->sortedType);
}
