// ref_driver.cpp — TEST INFRASTRUCTURE (oracle/), never linked into the product.
//
// Host side of the "reference compiled here" oracle (oracle/_ref/libsphref.so):
// the kernels are the reference's own /root/reference/src/sphFluid.cl, included
// below through cl_shim.hpp; this file restates ONLY the host orchestration of
// /root/reference/src/owOpenCLSolver.cpp (buffer sizes :32-44,:70-86, kernel
// argument lists and launch order :213-687, host qsort + myCompare :255-261,
// :690-696, host index post-pass :305-319) and calcDelta
// (/root/reference/src/owPhysicsFluidSimulator.cpp:164-203), because the
// originals include <CL/cl.hpp>, which does not exist in this image.
// The physics constants come from the reference header itself
// (/root/reference/src/owPhysicsConstant.h), included unchanged.
//
// Exports the ABI of include/sph_capi.h so tests drive it like the CUDA library.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "owPhysicsConstant.h"  // the reference header, via -I/root/reference/src
#include "cl_shim.hpp"

thread_local int cl_shim_global_id = 0;
static unsigned int cl_shim_hash_mask = 0xffffu;  // sphFluid.cl:229,377 hard-code 0xffff

namespace refcl {
#include SPH_CL_INC  // sed-rewritten text of sphFluid.cl (see build_ref.py)
}

#include "sph_capi.h"

// ---- calcDelta, owPhysicsFluidSimulator.cpp:164-203 (restated line for line) -------------
static float ref_calcDelta() {
  float x[] = {1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 2, -2, 0, 0, 0, 0, 0, 0};
  float y[] = {0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 0, 2, -2, 0, 0, 0, 0};
  float z[] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 0, 0, 0, 2, -2, 1, -1};
  float sum1_x = 0.f, sum1_y = 0.f, sum1_z = 0.f;
  double sum1 = 0.0, sum2 = 0.0;
  float v_x = 0.f, v_y = 0.f, v_z = 0.f;
  float dist;
  float particleRadius = pow(mass / rho0, 1.f / 3.f);
  float h_r_2;
  for (int i = 0; i < 32; i++) {
    v_x = x[i] * 0.8f * particleRadius;
    v_y = y[i] * 0.8f * particleRadius;
    v_z = z[i] * 0.8f * particleRadius;
    dist = sqrt(v_x * v_x + v_y * v_y + v_z * v_z);
    if (dist <= h * simulationScale) {
      h_r_2 = pow((h * simulationScale - dist), 2);
      sum1_x += h_r_2 * v_x / dist;
      sum1_y += h_r_2 * v_y / dist;
      sum1_z += h_r_2 * v_z / dist;
      sum2 += h_r_2 * h_r_2;
    }
  }
  sum1 = sum1_x * sum1_x + sum1_y * sum1_y + sum1_z * sum1_z;
  double result = 1.0 / (beta * gradWspikyCoefficient * gradWspikyCoefficient * (sum1 + sum2));
  return (float)result;
}
extern const float delta = ref_calcDelta();  // owPhysicsFluidSimulator.cpp:8

static thread_local std::string g_err;
static int fail(int code, const char *msg) { g_err = msg; return code; }

struct sph_ctx {
  sph_params p;
  int N, G;
  // owOpenCLSolver.h:72-97
  std::vector<float4> acceleration, position, sortedPosition, sortedVelocity, velocity, elasticConnectionsData;
  std::vector<uint> gridCellIndex, gridCellIndexFixedUp, particleIndexBack;
  std::vector<float2> neighborMap;
  std::vector<uint2> particleIndexStore;  // one pad element in front: indexx reads [idx-1] at idx==0 (sphFluid.cl:422)
  uint2 *particleIndex;
  std::vector<float> pressure, rho, muscle_activation_signal;
  std::vector<int> membraneData, particleMembranesList;
  bool has_elastic, has_membranes, has_pml;
  int iterationCount;
};

extern "C" {

const char *sph_backend(void) { return "reference-cl-as-c++"; }
const char *sph_last_error(void) { return g_err.c_str(); }

int sph_params_set_box(sph_params *p, double mx, double my, double mz) {
  if (!p) return fail(SPH_E_INVALID, "null params");
  // `#define XMAX 30.0*h` (owPhysicsConstant.h:33): double product of a double literal and float h.
  double xmaxd = mx * p->h, ymaxd = my * p->h, zmaxd = mz * p->h;
  p->xmin = 0; p->ymin = 0; p->zmin = 0;                       // XMIN/YMIN/ZMIN 0
  p->xmax = (float)xmaxd; p->ymax = (float)ymaxd; p->zmax = (float)zmaxd;  // owOpenCLSolver.cpp:7-12
  p->grid_cells_x = (int)((xmaxd - 0) / p->h) + 1;             // owOpenCLSolver.cpp:14-16
  p->grid_cells_y = (int)((ymaxd - 0) / p->h) + 1;
  p->grid_cells_z = (int)((zmaxd - 0) / p->h) + 1;
  p->grid_cell_count = p->grid_cells_x * p->grid_cells_y * p->grid_cells_z;  // :17
  return 0;
}

int sph_params_default(sph_params *p) {
  if (!p) return fail(SPH_E_INVALID, "null params");
  memset(p, 0, sizeof(*p));
  p->muscle_count = 100;  // owWorldSimulation.cpp:31
  p->max_iteration = maxIteration;
  p->hash_mask = 0xffffu;
  p->h = h; p->hash_grid_cell_size = hashGridCellSize; p->hash_grid_cell_size_inv = hashGridCellSizeInv;
  p->simulation_scale = simulationScale; p->simulation_scale_inv = simulationScaleInv; p->r0 = r0;
  p->rho0 = rho0; p->mass = mass; p->time_step = timeStep; p->viscosity = viscosity;
  p->stiffness = stiffness; p->damping = damping;
  p->gravity_x = gravity_x; p->gravity_y = gravity_y; p->gravity_z = gravity_z;
  p->delta = delta;
  p->wpoly6_coefficient = Wpoly6Coefficient;
  p->grad_wspiky_coefficient = gradWspikyCoefficient;
  p->del2_wviscosity_coefficient = del2WviscosityCoefficient;
  // header box, computed with the header's own macros
  p->xmin = XMIN; p->xmax = XMAX; p->ymin = YMIN; p->ymax = YMAX; p->zmin = ZMIN; p->zmax = ZMAX;
  p->grid_cells_x = (int)((XMAX - XMIN) / h) + 1;
  p->grid_cells_y = (int)((YMAX - YMIN) / h) + 1;
  p->grid_cells_z = (int)((ZMAX - ZMIN) / h) + 1;
  p->grid_cell_count = p->grid_cells_x * p->grid_cells_y * p->grid_cells_z;
  return 0;
}

sph_ctx *sph_create(const sph_params *p, const float *pos4, const float *vel4, const float *elastic4,
                    const int32_t *membranes3, const int32_t *pml7, int /*device*/) {
  if (!p || !pos4 || !vel4 || p->particle_count <= 0 || p->grid_cell_count <= 0) {
    fail(SPH_E_INVALID, "sph_create: bad arguments");
    return nullptr;
  }
  sph_ctx *c = new sph_ctx();
  c->p = *p;
  const int N = c->N = p->particle_count;
  const int G = c->G = p->grid_cell_count;
  // owOpenCLSolver.cpp:32-44
  c->acceleration.resize(2 * (size_t)N);
  c->gridCellIndex.resize(G + 1);
  c->gridCellIndexFixedUp.resize(G + 1);
  c->neighborMap.resize((size_t)SPH_MAX_NEIGHBOR_COUNT * N);
  c->particleIndexStore.resize((size_t)N + 1);
  c->particleIndex = c->particleIndexStore.data() + 1;
  c->particleIndexBack.resize(N);
  c->position.resize(2 * (size_t)N);
  c->pressure.resize(N);
  c->rho.resize(2 * (size_t)N);
  c->sortedPosition.resize(2 * (size_t)N);
  c->sortedVelocity.resize(N);
  c->velocity.resize(2 * (size_t)N);
  c->muscle_activation_signal.assign(p->muscle_count > 0 ? p->muscle_count : 1, 0.f);
  memcpy(c->position.data(), pos4, sizeof(float) * 4 * (size_t)N);   // :65
  memcpy(c->velocity.data(), vel4, sizeof(float) * 4 * (size_t)N);   // :66
  c->has_membranes = membranes3 && p->num_membranes > 0;
  c->has_elastic = elastic4 && p->num_elastic > 0;
  c->has_pml = c->has_membranes && c->has_elastic && pml7;
  if (c->has_membranes) c->membraneData.assign(membranes3, membranes3 + 3 * (size_t)p->num_membranes);         // :70-71
  if (c->has_pml) c->particleMembranesList.assign(pml7, pml7 + 7 * (size_t)p->num_elastic);                    // :75-76
  if (c->has_elastic) {                                                                                        // :83-86
    c->elasticConnectionsData.resize((size_t)SPH_MAX_NEIGHBOR_COUNT * p->num_elastic);
    memcpy(c->elasticConnectionsData.data(), elastic4, sizeof(float) * 4 * SPH_MAX_NEIGHBOR_COUNT * (size_t)p->num_elastic);
  }
  c->iterationCount = 0;
  return c;
}

void sph_destroy(sph_ctx *c) { delete c; }

int sph_get_params(const sph_ctx *c, sph_params *out) {
  if (!c || !out) return fail(SPH_E_INVALID, "null");
  *out = c->p;
  return 0;
}

// owOpenCLSolver.cpp:690-696
static int myCompare(const void *v1, const void *v2) {
  const int *f1 = (const int *)v1;
  const int *f2 = (const int *)v2;
  if (f1[0] < f2[0]) return -1;
  if (f1[0] > f2[0]) return +1;
  return 0;
}

// NDRange launch: every work-item id in [0, global) runs the kernel body once.
// `#pragma omp parallel for` is legal for every kernel (each writes own slots only);
// parity runs use OMP_NUM_THREADS=1 semantics via omp `if` (see sph_ref_set_threads).
static int g_threads = 1;
#define LAUNCH(global, call)                                            \
  do {                                                                  \
    const int _g = (global);                                            \
    _Pragma("omp parallel for schedule(dynamic, 512) num_threads(g_threads)") \
    for (int _i = 0; _i < _g; ++_i) {                                   \
      cl_shim_global_id = _i;                                           \
      call;                                                             \
    }                                                                   \
  } while (0)

void sph_ref_set_threads(int n) { g_threads = n > 0 ? n : 1; }

int sph_stage(sph_ctx *c, int stage) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  const sph_params &q = c->p;
  const int N = c->N;
  const int NR = (((N - 1) / 256) + 1) * 256;  // PARTICLE_COUNT_RoundedUp, local_NDRange_size = 256
  cl_shim_hash_mask = q.hash_mask;
  float4 *acceleration = c->acceleration.data(), *position = c->position.data(), *velocity = c->velocity.data();
  float4 *sortedPosition = c->sortedPosition.data(), *sortedVelocity = c->sortedVelocity.data();
  float4 *elastic = c->elasticConnectionsData.data();
  float2 *neighborMap = c->neighborMap.data();
  uint2 *particleIndex = c->particleIndex;
  uint *particleIndexBack = c->particleIndexBack.data();
  uint *gridCellIndex = c->gridCellIndex.data(), *gridCellIndexFixedUp = c->gridCellIndexFixedUp.data();
  float *pressure = c->pressure.data(), *rho = c->rho.data();
  int *membraneData = c->membraneData.data(), *pml = c->particleMembranesList.data();
  float *muscle = c->muscle_activation_signal.data();
  const int G = c->G;
  switch (stage) {
    case SPH_STAGE_CLEAR_BUFFERS:  // :213-228
      LAUNCH(NR, refcl::clearBuffers(neighborMap, N));
      return 0;
    case SPH_STAGE_HASH_PARTICLES:  // :229-253
      LAUNCH(NR, refcl::hashParticles(position, q.grid_cells_x, q.grid_cells_y, q.grid_cells_z,
                                      q.hash_grid_cell_size_inv, q.xmin, q.ymin, q.zmin, particleIndex, N));
      return 0;
    case SPH_STAGE_SORT:  // :255-261 (the two copies are no-ops on the host)
      qsort(particleIndex, N, 2 * sizeof(int), myCompare);
      return 0;
    case SPH_STAGE_SORT_POST_PASS:  // :262-283
      LAUNCH(NR, refcl::sortPostPass(particleIndex, particleIndexBack, position, velocity, sortedPosition,
                                     sortedVelocity, N));
      return 0;
    case SPH_STAGE_INDEXX: {  // :284-304
      int GR = (((G - 1) / 256) + 1) * 256;
      LAUNCH(GR, refcl::indexx(particleIndex, G, gridCellIndex, N));
      return 0;
    }
    case SPH_STAGE_INDEX_POST_PASS: {  // :305-319
      memcpy(gridCellIndexFixedUp, gridCellIndex, sizeof(uint) * ((size_t)G + 1));
      int recentNonEmptyCell = G;
      for (int i = G; i >= 0; i--) {
        if (gridCellIndexFixedUp[i] == (uint)NO_CELL_ID) gridCellIndexFixedUp[i] = recentNonEmptyCell;
        else recentNonEmptyCell = gridCellIndexFixedUp[i];
      }
      return 0;
    }
    case SPH_STAGE_FIND_NEIGHBORS:  // :320-356
      LAUNCH(NR, refcl::findNeighbors(gridCellIndexFixedUp, sortedPosition, G, q.grid_cells_x, q.grid_cells_y,
                                      q.grid_cells_z, q.h, q.hash_grid_cell_size, q.hash_grid_cell_size_inv,
                                      q.simulation_scale, q.xmin, q.ymin, q.zmin, neighborMap, N));
      return 0;
    case SPH_STAGE_COMPUTE_DENSITY:  // :357-385
      LAUNCH(NR, refcl::pcisph_computeDensity(neighborMap, q.wpoly6_coefficient, q.h, q.mass, q.rho0,
                                              q.simulation_scale, q.stiffness, sortedPosition, pressure, rho,
                                              particleIndexBack, q.delta, N));
      return 0;
    case SPH_STAGE_COMPUTE_FORCES_INIT_PRESSURE:  // :386-418
      LAUNCH(NR, refcl::pcisph_computeForcesAndInitPressure(
                     neighborMap, rho, pressure, sortedPosition, sortedVelocity, acceleration, particleIndexBack,
                     q.wpoly6_coefficient, q.del2_wviscosity_coefficient, q.h, q.mass, q.viscosity,
                     q.simulation_scale, q.gravity_x, q.gravity_y, q.gravity_z, position, particleIndex, N));
      return 0;
    case SPH_STAGE_COMPUTE_ELASTIC_FORCES: {  // :420-452
      if (q.num_elastic == 0 || !c->has_elastic) return 0;
      int ER = (((q.num_elastic - 1) / 256) + 1) * 256;
      LAUNCH(ER, refcl::pcisph_computeElasticForces(neighborMap, sortedPosition, sortedVelocity, acceleration,
                                                    particleIndexBack, velocity, q.h, q.mass, q.simulation_scale,
                                                    q.num_elastic, elastic, q.elastic_offset, N, q.muscle_count,
                                                    muscle, position));
      return 0;
    }
    case SPH_STAGE_PREDICT_POSITIONS:  // :454-489
      LAUNCH(NR, refcl::pcisph_predictPositions(acceleration, sortedPosition, sortedVelocity, particleIndex,
                                                particleIndexBack, q.gravity_x, q.gravity_y, q.gravity_z,
                                                q.simulation_scale_inv, q.time_step, q.xmin, q.xmax, q.ymin, q.ymax,
                                                q.zmin, q.zmax, q.damping, position, velocity, q.r0, neighborMap, N));
      return 0;
    case SPH_STAGE_PREDICT_DENSITY:  // :490-518
      LAUNCH(NR, refcl::pcisph_predictDensity(neighborMap, particleIndexBack, q.wpoly6_coefficient, q.h, q.mass,
                                              q.rho0, q.simulation_scale, q.stiffness, sortedPosition, pressure, rho,
                                              q.delta, N));
      return 0;
    case SPH_STAGE_CORRECT_PRESSURE:  // :519-548
      LAUNCH(NR, refcl::pcisph_correctPressure(neighborMap, particleIndexBack, q.h, q.mass, q.rho0,
                                               q.simulation_scale, q.stiffness, sortedPosition, pressure, rho,
                                               q.delta, position, particleIndex, N));
      return 0;
    case SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION:  // :549-581
      LAUNCH(NR, refcl::pcisph_computePressureForceAcceleration(
                     neighborMap, pressure, rho, sortedPosition, sortedVelocity, particleIndexBack, q.delta,
                     q.grad_wspiky_coefficient, q.h, q.mass, q.viscosity, q.simulation_scale, acceleration, q.rho0,
                     position, particleIndex, N));
      return 0;
    case SPH_STAGE_INTEGRATE:  // :649-687
      LAUNCH(NR, refcl::pcisph_integrate(acceleration, sortedPosition, sortedVelocity, particleIndex,
                                         particleIndexBack, q.gravity_x, q.gravity_y, q.gravity_z,
                                         q.simulation_scale_inv, q.time_step, q.xmin, q.xmax, q.ymin, q.ymax, q.zmin,
                                         q.zmax, q.damping, position, velocity, rho, q.r0, neighborMap, N,
                                         c->iterationCount));
      return 0;
    case SPH_STAGE_CLEAR_MEMBRANE_BUFFERS:  // :583-600
      LAUNCH(NR, refcl::clearMembraneBuffers(position, velocity, sortedPosition, N));
      return 0;
    case SPH_STAGE_MEMBRANE_INTERACTION:  // :602-626
      LAUNCH(NR, refcl::computeInteractionWithMembranes(position, velocity, sortedPosition, particleIndex,
                                                        particleIndexBack, neighborMap, pml, membraneData, N,
                                                        q.num_elastic, q.r0));
      return 0;
    case SPH_STAGE_MEMBRANE_FINALIZE:  // :628-646
      LAUNCH(NR, refcl::computeInteractionWithMembranes_finalize(position, velocity, particleIndex,
                                                                 particleIndexBack, N));
      return 0;
  }
  return fail(SPH_E_INVALID, "unknown stage");
}

// owPhysicsFluidSimulator.cpp:88-113 + iterationCount++ (:130)
int sph_step(sph_ctx *c) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  static const int pre[] = {SPH_STAGE_CLEAR_BUFFERS, SPH_STAGE_HASH_PARTICLES, SPH_STAGE_SORT,
                            SPH_STAGE_SORT_POST_PASS, SPH_STAGE_INDEXX, SPH_STAGE_INDEX_POST_PASS,
                            SPH_STAGE_FIND_NEIGHBORS, SPH_STAGE_COMPUTE_DENSITY,
                            SPH_STAGE_COMPUTE_FORCES_INIT_PRESSURE, SPH_STAGE_COMPUTE_ELASTIC_FORCES};
  for (int s : pre) { int e = sph_stage(c, s); if (e) return e; }
  int iter = 0;
  do {
    sph_stage(c, SPH_STAGE_PREDICT_POSITIONS);
    sph_stage(c, SPH_STAGE_PREDICT_DENSITY);
    sph_stage(c, SPH_STAGE_CORRECT_PRESSURE);
    sph_stage(c, SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION);
    iter++;
  } while (iter < c->p.max_iteration);
  sph_stage(c, SPH_STAGE_INTEGRATE);
  sph_stage(c, SPH_STAGE_CLEAR_MEMBRANE_BUFFERS);
  sph_stage(c, SPH_STAGE_MEMBRANE_INTERACTION);
  sph_stage(c, SPH_STAGE_MEMBRANE_FINALIZE);
  c->iterationCount++;
  return 0;
}

int sph_sync(sph_ctx *) { return 0; }

int sph_set_muscles(sph_ctx *c, const float *signal, int n) {  // owOpenCLSolver.cpp:738-742
  if (!c || !signal || n < 0 || n > c->p.muscle_count) return fail(SPH_E_INVALID, "sph_set_muscles");
  memcpy(c->muscle_activation_signal.data(), signal, sizeof(float) * n);
  return 0;
}

static void *buf_ptr(sph_ctx *c, int id, size_t *bytes) {
  const size_t N = c->N, G = c->G;
  switch (id) {
    case SPH_BUF_POSITION: *bytes = 2 * N * 16; return c->position.data();
    case SPH_BUF_VELOCITY: *bytes = 2 * N * 16; return c->velocity.data();
    case SPH_BUF_PARTICLE_INDEX: *bytes = N * 8; return c->particleIndex;
    case SPH_BUF_PARTICLE_INDEX_BACK: *bytes = N * 4; return c->particleIndexBack.data();
    case SPH_BUF_SORTED_POSITION: *bytes = 2 * N * 16; return c->sortedPosition.data();
    case SPH_BUF_SORTED_VELOCITY: *bytes = N * 16; return c->sortedVelocity.data();
    case SPH_BUF_GRID_CELL_INDEX: *bytes = (G + 1) * 4; return c->gridCellIndex.data();
    case SPH_BUF_GRID_CELL_INDEX_FIXED_UP: *bytes = (G + 1) * 4; return c->gridCellIndexFixedUp.data();
    case SPH_BUF_NEIGHBOR_MAP: *bytes = N * 32 * 8; return c->neighborMap.data();
    case SPH_BUF_RHO: *bytes = 2 * N * 4; return c->rho.data();
    case SPH_BUF_PRESSURE: *bytes = N * 4; return c->pressure.data();
    case SPH_BUF_ACCELERATION: *bytes = 2 * N * 16; return c->acceleration.data();
    case SPH_BUF_ELASTIC_CONNECTIONS: *bytes = c->elasticConnectionsData.size() * 16; return c->elasticConnectionsData.data();
    case SPH_BUF_MEMBRANE_DATA: *bytes = c->membraneData.size() * 4; return c->membraneData.data();
    case SPH_BUF_PARTICLE_MEMBRANES_LIST: *bytes = c->particleMembranesList.size() * 4; return c->particleMembranesList.data();
    case SPH_BUF_MUSCLE_ACTIVATION_SIGNAL: *bytes = (size_t)c->p.muscle_count * 4; return c->muscle_activation_signal.data();
  }
  *bytes = 0;
  return nullptr;
}

size_t sph_buffer_bytes(const sph_ctx *c, int id) {
  size_t b = 0;
  if (c) buf_ptr(const_cast<sph_ctx *>(c), id, &b);
  return b;
}

int sph_read(sph_ctx *c, int id, void *dst, size_t bytes) {
  if (!c || !dst) return fail(SPH_E_INVALID, "sph_read: null");
  size_t have = 0;
  void *src = buf_ptr(c, id, &have);
  if (!src && bytes) return fail(SPH_E_UNSUPPORTED, "sph_read: buffer absent");
  if (bytes > have) return fail(SPH_E_SIZE, "sph_read: too many bytes");
  memcpy(dst, src, bytes);
  return 0;
}

int sph_write(sph_ctx *c, int id, const void *src, size_t bytes) {
  if (!c || !src) return fail(SPH_E_INVALID, "sph_write: null");
  if (id != SPH_BUF_POSITION && id != SPH_BUF_VELOCITY) return fail(SPH_E_UNSUPPORTED, "sph_write: buffer not writable");
  size_t have = 0;
  void *dst = buf_ptr(c, id, &have);
  if (bytes > have) return fail(SPH_E_SIZE, "sph_write: too many bytes");
  memcpy(dst, src, bytes);
  return 0;
}

}  // extern "C"
