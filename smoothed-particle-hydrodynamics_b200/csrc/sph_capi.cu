// sph_capi.cu — the C-ABI of include/sph_capi.h over the CUDA kernels (libsph_b200.so).
//
// Host orchestration that replaces owOpenCLSolver (owOpenCLSolver.cpp:27-92 buffers,
// :213-687 stage launchers) and the device half of owPhysicsFluidSimulator::simulationStep
// (owPhysicsFluidSimulator.cpp:88-113). No OpenCL, no CPU fallback: every entry point either
// runs on the CUDA device or fails with SPH_E_DEVICE.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "sph_ctx.h"

int sph_scan_tiles_for(int n);

// Ask for eager module loading when this library is loaded before CUDA is initialised (see preload() in
// sph_sort.cu for why lazily loaded kernels and flag-waiting kernels do not mix); harmless otherwise.
__attribute__((constructor)) static void sph_library_init() { setenv("CUDA_MODULE_LOADING", "EAGER", 0); }

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }

#define CU(call)                                                                                     \
  do {                                                                                               \
    cudaError_t _e = (call);                                                                         \
    if (_e != cudaSuccess)                                                                           \
      return fail(SPH_E_DEVICE, std::string(#call) + ": " + cudaGetErrorString(_e));                 \
  } while (0)

extern "C" {

const char *sph_backend(void) { return "cuda-sm_100a"; }
const char *sph_last_error(void) { return g_err.c_str(); }

// ---------------------------------------------------------------- constants
// owPhysicsConstant.h:12-76 and calcDelta (owPhysicsFluidSimulator.cpp:164-203), evaluated on the
// host with the reference's expression types: pow(float,float) is powf in C++, sqrt(float) sqrtf,
// pow(float,int) the exact double square; `beta` is an all-float (denormal) product widened last.
static float calc_delta(float mass, float rho0, float h, float simulationScale, double beta, double gradW) {
  static const float x[32] = {1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 2, -2, 0, 0, 0, 0, 0, 0};
  static const float y[32] = {0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 0, 2, -2, 0, 0, 0, 0};
  static const float z[32] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 0, 0, 0, 2, -2, 1, -1};
  float s1x = 0.f, s1y = 0.f, s1z = 0.f;
  double sum2 = 0.0;
  const float particleRadius = powf(mass / rho0, 1.f / 3.f);
  for (int i = 0; i < 32; i++) {
    float vx = x[i] * 0.8f * particleRadius, vy = y[i] * 0.8f * particleRadius, vz = z[i] * 0.8f * particleRadius;
    float dist = sqrtf(vx * vx + vy * vy + vz * vz);
    if (dist <= h * simulationScale) {
      float t = h * simulationScale - dist;
      float h_r_2 = (float)((double)t * (double)t);
      s1x += h_r_2 * vx / dist;
      s1y += h_r_2 * vy / dist;
      s1z += h_r_2 * vz / dist;
      sum2 += h_r_2 * h_r_2;
    }
  }
  double sum1 = s1x * s1x + s1y * s1y + s1z * s1z;
  return (float)(1.0 / (beta * gradW * gradW * (sum1 + sum2)));
}

int sph_params_set_box(sph_params *p, double mx, double my, double mz) {
  if (!p) return fail(SPH_E_INVALID, "null params");
  double xd = mx * p->h, yd = my * p->h, zd = mz * p->h;  // `#define XMAX 30.0*h`
  p->xmin = 0; p->ymin = 0; p->zmin = 0;
  p->xmax = (float)xd; p->ymax = (float)yd; p->zmax = (float)zd;
  p->grid_cells_x = (int)((xd - 0) / p->h) + 1;            // owOpenCLSolver.cpp:14-17
  p->grid_cells_y = (int)((yd - 0) / p->h) + 1;
  p->grid_cells_z = (int)((zd - 0) / p->h) + 1;
  p->grid_cell_count = p->grid_cells_x * p->grid_cells_y * p->grid_cells_z;
  return 0;
}

int sph_params_default(sph_params *p) {
  if (!p) return fail(SPH_E_INVALID, "null params");
  memset(p, 0, sizeof *p);
  const float rho0 = 1000.0f, mass = 3.25e-14f, timeStep = 5.0e-06f;
  const float simulationScale = 0.004f * powf(mass, 1.f / 3.f) / powf(0.00025f, 1.f / 3.f);
  const float h = 3.34f;
  p->muscle_count = 100;
  p->max_iteration = 3;
  p->hash_mask = 0xffffu;
  p->h = h;
  p->hash_grid_cell_size = 2.0f * h;
  p->hash_grid_cell_size_inv = 1.0f / p->hash_grid_cell_size;
  p->simulation_scale = simulationScale;
  p->simulation_scale_inv = 1.0f / simulationScale;
  p->r0 = 0.5f * h;
  p->rho0 = rho0; p->mass = mass; p->time_step = timeStep;
  p->stiffness = 0.75f; p->viscosity = 0.00005f; p->damping = 0.75f;
  p->gravity_x = 0.0f; p->gravity_y = -9.8f; p->gravity_z = 0.0f;
  const double beta = timeStep * timeStep * mass * mass * 2 / (rho0 * rho0);
  p->wpoly6_coefficient = 315.0 / (64.0 * M_PI * pow((double)(h * simulationScale), 9.0));
  p->grad_wspiky_coefficient = -45.0 / (M_PI * pow((double)(h * simulationScale), 6.0));
  p->del2_wviscosity_coefficient = -p->grad_wspiky_coefficient;
  p->delta = calc_delta(mass, rho0, h, simulationScale, beta, p->grad_wspiky_coefficient);
  return sph_params_set_box(p, 30.0, 20.0, 250.0);
}

}  // extern "C"

// ---------------------------------------------------------------- context
static void fill_dev_params(sph_ctx *c) {
  const sph_params &p = c->p;
  DevParams &d = c->d;
  d.N = p.particle_count;
  d.Npad = ((p.particle_count + 31) / 32) * 32;
  d.G = p.grid_cell_count;
  d.gx = p.grid_cells_x; d.gy = p.grid_cells_y; d.gz = p.grid_cells_z;
  d.mask = p.hash_mask;
  d.numElastic = p.num_elastic; d.elasticOffset = p.elastic_offset; d.muscleCount = p.muscle_count;
  d.h = p.h; d.cellSize = p.hash_grid_cell_size; d.cellSizeInv = p.hash_grid_cell_size_inv;
  d.simScale = p.simulation_scale; d.simScaleInv = p.simulation_scale_inv; d.r0 = p.r0;
  d.xmin = p.xmin; d.xmax = p.xmax; d.ymin = p.ymin; d.ymax = p.ymax; d.zmin = p.zmin; d.zmax = p.zmax;
  d.rho0 = p.rho0; d.dt = p.time_step; d.delta = p.delta;
  d.gravx = p.gravity_x; d.gravy = p.gravity_y; d.gravz = p.gravity_z;
  d.hScaled = p.h * p.simulation_scale;
  d.hScaled2 = d.hScaled * d.hScaled;
  d.hScaled6 = d.hScaled2 * d.hScaled2 * d.hScaled2;
  d.surfCoeff = (-1.5e-09f * 0.3f * (float)(p.wpoly6_coefficient * pow(d.hScaled2 / 2.0, 3.0)) * p.simulation_scale);
  d.massMu = (float)(p.mass * p.viscosity);
  d.posTimeStep = p.time_step * p.simulation_scale_inv;
  d.rho0delta = p.rho0 * p.delta;
  d.xmaxClamp = p.xmax - 0.000001f; d.ymaxClamp = p.ymax - 0.000001f; d.zmaxClamp = p.zmax - 0.000001f;
  d.massWpoly6 = ((double)p.mass) * p.wpoly6_coefficient;
  d.del2W = p.del2_wviscosity_coefficient;
  d.massGradW = ((double)p.mass) * p.grad_wspiky_coefficient;
  d.closeThr = 0.5 * (d.hScaled / 2);
  // radial-histogram bin thresholds (see DevParams::binT). The bin is a monotone step function of d2.
  auto bin_of = [&](float d2) -> int { return (int)(sqrtf(d2) * SPH_RADIUS_SEGMENTS / d.h); };
  d.binT[0] = 0.f;
  for (int j = 1; j <= SPH_RADIUS_SEGMENTS; j++) {
    uint32_t lo = 0u, hi;  // bit patterns of +0 and (2h)^2: bin_of(lo) = 0 < j <= 30 < bin_of(hi) = 60
    { float top = 4.f * d.h * d.h; memcpy(&hi, &top, 4); }
    while (hi - lo > 1u) {
      uint32_t mid = lo + (hi - lo) / 2u;
      float f; memcpy(&f, &mid, 4);
      if (bin_of(f) >= j) hi = mid; else lo = mid;
    }
    memcpy(&d.binT[j], &hi, 4);
  }
  { uint32_t inf = 0x7f800000u; memcpy(&d.binT[SPH_RADIUS_SEGMENTS + 1], &inf, 4); }
  d.binScale = (float)SPH_RADIUS_SEGMENTS / d.h;
  // smallest float d2 with sqrtf(d2) * simScale >= limit (see DevParams::d2H)
  auto d2_threshold = [&](float limit) -> float {
    uint32_t lo = 0u, hi = 0x7f800000u;  // f(+0) = 0 < limit (limits are positive), f(+inf) = +inf >= limit
    if (!(limit > 0.f)) return 0.f;
    while (hi - lo > 1u) {
      uint32_t mid = lo + (hi - lo) / 2u;
      float f; memcpy(&f, &mid, 4);
      volatile float r = sqrtf(f) * d.simScale;
      if (r >= limit) hi = mid; else lo = mid;
    }
    float out; memcpy(&out, &hi, 4);
    return out;
  };
  d.d2H = d2_threshold(d.hScaled);
  d.d2Close = d2_threshold((float)d.closeThr);
}

template <class T>
static cudaError_t dalloc(T **ptr, size_t count, bool zero = true) {
  size_t bytes = sizeof(T) * (count ? count : 1);
  cudaError_t e = cudaMalloc((void **)ptr, bytes);
  if (e == cudaSuccess && zero) e = cudaMemset(*ptr, 0, bytes);
  return e;
}

extern "C" int sph_profile_enable(sph_ctx *c, int on);

static void drop_graph(sph_ctx *c) {
  for (int h = 0; h < 2; h++)
    if (c->stepGraph[h]) { cudaGraphExecDestroy(c->stepGraph[h]); c->stepGraph[h] = nullptr; }
}


extern "C" {

void sph_destroy(sph_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  slab_destroy(c);
  void *ptrs[] = {c->position, c->velocity, c->particleIndex, c->particleIndexBack, c->sortedPosition,
                  c->sortedVelocity, c->gridCellIndex, c->cellStart, c->rho, c->pressure, c->acceleration,
                  c->elastic, c->membraneData, c->pml, c->muscle, c->nbIds, c->cellCount, c->key, c->tmpSerial,
                  c->sortedType, c->spT, c->recA, c->recB, c->cellList, c->cellListCount, c->cellCursor, c->fnDeferred, c->scanState, c->scanTicket, c->nmScratch, c->readScratch, c->elasticMask, c->soa, c->tmpRankKey, c->memList, c->memCount, c->nsStats, c->bMask};
  for (void *q : ptrs)
    if (q) cudaFree(q);
  if (c->profiler) sph_profile_enable(c, 0);
  drop_graph(c);
  if (c->copyStream) {
    cudaStreamSynchronize(c->copyStream);
    cudaStreamDestroy(c->copyStream);
    cudaEventDestroy(c->evReadReady);
    cudaEventDestroy(c->evReadDone);
  }
  if (c->hostDeferred) {
    cudaFreeHost(c->hostDeferred);
    cudaEventDestroy(c->evDeferred);
  }
  if (c->snapRho) { cudaFree(c->snapRho); cudaFree(c->snapIndex); cudaEventDestroy(c->evSnap); }
  if (c->snapVel) cudaFree(c->snapVel);
  if (c->evTimer[0]) { cudaEventDestroy(c->evTimer[0]); cudaEventDestroy(c->evTimer[1]); }
  if (c->sigRing) {
    cudaFreeHost(c->sigRing);
    for (cudaEvent_t e : c->sigDone) cudaEventDestroy(e);
  }
  if (c->ownStream) cudaStreamDestroy(c->ownStream);
  delete c;
}

sph_ctx *sph_create(const sph_params *p, const float *pos4, const float *vel4, const float *elastic4,
                    const int32_t *membranes3, const int32_t *pml7, int device) {
  if (!p || !pos4 || !vel4 || p->particle_count <= 0 || p->grid_cell_count <= 0 || p->grid_cells_x <= 0 ||
      p->grid_cells_y <= 0 || p->muscle_count < 0) {
    fail(SPH_E_INVALID, "sph_create: bad arguments");
    return nullptr;
  }
  if (p->grid_cell_count > (1 << 24)) {
    // the sorted cell id travels as a float in sortedPosition.w (the reference's layout, sphFluid.cl:457-465):
    // exact up to 2^24 only. The reference itself masks ids to 16 bits.
    fail(SPH_E_INVALID, "sph_create: grid_cell_count > 2^24 (cell ids are carried as floats in sortedPosition.w)");
    return nullptr;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0) {
    fail(SPH_E_DEVICE, std::string("sph_create: no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    return nullptr;
  }
  if (device < 0 || device >= ndev) {
    fail(SPH_E_INVALID, "sph_create: bad device ordinal");
    return nullptr;
  }
  if ((e = cudaSetDevice(device)) != cudaSuccess) {
    fail(SPH_E_DEVICE, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    return nullptr;
  }
  sph_ctx *c = new sph_ctx();
  memset(c, 0, sizeof *c);
  c->p = *p;
  c->device = device;
  fill_dev_params(c);
  const size_t N = c->d.N, G = c->d.G, Npad = c->d.Npad;
  c->has_membranes = membranes3 && p->num_membranes > 0;
  c->has_elastic = elastic4 && p->num_elastic > 0;
  c->has_pml = c->has_membranes && c->has_elastic && pml7;
  c->scanTiles = sph_scan_tiles_for((int)G + 1);
  cudaDeviceGetAttribute(&c->numSMs, cudaDevAttrMultiProcessorCount, device);
  if (c->numSMs <= 0) c->numSMs = 148;
  {  // neighbour search path: auto (tiles for large fluids), or forced for A-B runs and the parity tests
    const char *e = getenv("SPH_GENERIC_NEIGHBORS"), *q = getenv("SPH_NEIGHBOR_PATH");
    c->force_generic_neighbors = (e && e[0] == '1') || (q && !strcmp(q, "generic"));
    c->force_tiled_neighbors = q && !strcmp(q, "tiles");
    c->force_list_neighbors = q && !strcmp(q, "list");
    c->force_coop_neighbors = q && !strcmp(q, "coop");
    c->auto_tiles = q && !strcmp(q, "auto");
    const char *dg = getenv("SPH_DEEP_GATHER");
    c->deep_gather = dg ? atoi(dg) : -1;
    const char *lv = getenv("SPH_LIST_VARIANT");
    c->list_variant = lv ? atoi(lv) : 2;
    const char *gm = getenv("SPH_GRAPH");
    c->graph_mode = gm ? atoi(gm) : -1;
    const char *ln = getenv("SPH_LANES");
    c->lanes = ln ? atoi(ln) : 0;
    const char *hp = getenv("SPH_HIST_PCT");
    c->hist_pct = (hp ? (unsigned)atoi(hp) : 5u) & 255u;
    const char *hp2 = getenv("SPH_HIST2_PCT");  // histogram-ONLY first pass above this share of needy particles
    c->hist_pct |= ((hp2 ? (unsigned)atoi(hp2) : 30u) & 255u) << 8;
    const char *pd = getenv("SPH_PDL");
    c->pdl_mode = pd ? atoi(pd) : -1;
  }
  bool ok = true;
#define A(x) ok = ok && ((e = (x)) == cudaSuccess)
  A(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->ownStream = c->stream;
  c->dispDirty = true;
  A(dalloc(&c->position, 2 * N));
  A(dalloc(&c->velocity, 2 * N));
  A(dalloc(&c->particleIndex, N));
  A(dalloc(&c->particleIndexBack, N));
  A(dalloc(&c->sortedPosition, 2 * N));
  A(dalloc(&c->sortedVelocity, N));
  A(dalloc(&c->gridCellIndex, G + 1));
  A(dalloc(&c->cellStart, G + 2));
  A(dalloc(&c->rho, 2 * N));
  A(dalloc(&c->pressure, N));
  A(dalloc(&c->acceleration, 2 * N));
  A(dalloc(&c->muscle, (size_t)p->muscle_count));
  A(dalloc(&c->nbIds, Npad * SPH_NB));
  A(dalloc(&c->soa, 3 * (Npad + 32)));
  A(dalloc(&c->cellCount, G + 2));
  A(dalloc(&c->key, N));
  A(dalloc(&c->tmpSerial, N));
  A(dalloc(&c->sortedType, N));
  A(dalloc(&c->spT, N));
  A(dalloc(&c->recA, 2 * N));
  A(dalloc(&c->recB, N));
  A(dalloc(&c->bMask, N));
  A(dalloc(&c->cellList, (G < N ? G : N) + 1));
  A(dalloc(&c->cellListCount, 1));
  A(dalloc(&c->cellCursor, 2));
  A(dalloc(&c->fnDeferred, N));
  A(dalloc(&c->scanState, (size_t)c->scanTiles));
  A(dalloc(&c->scanTicket, 1));
  A(dalloc(&c->nsStats, 4));
  if (ok) A(cudaMemcpy(c->position, pos4, 16 * N, cudaMemcpyHostToDevice));   // owOpenCLSolver.cpp:65
  if (ok) A(cudaMemcpy(c->velocity, vel4, 16 * N, cudaMemcpyHostToDevice));   // :66
  if (ok && c->has_membranes) {                                                // :70-71
    A(dalloc(&c->membraneData, 3 * (size_t)p->num_membranes));
    if (ok) A(cudaMemcpy(c->membraneData, membranes3, 12 * (size_t)p->num_membranes, cudaMemcpyHostToDevice));
  }
  if (ok && c->has_pml) {                                                      // :75-76
    A(dalloc(&c->pml, 7 * (size_t)p->num_elastic));
    if (ok) A(cudaMemcpy(c->pml, pml7, 28 * (size_t)p->num_elastic, cudaMemcpyHostToDevice));
    if (ok) A(dalloc(&c->elasticMask, N));
    if (ok) A(dalloc(&c->memList, N));
    if (ok) A(dalloc(&c->memCount, 1));
  }
  if (ok && c->has_elastic) {                                                  // :83-86
    A(dalloc(&c->elastic, (size_t)SPH_NB * p->num_elastic));
    if (ok) A(cudaMemcpy(c->elastic, elastic4, 16 * (size_t)SPH_NB * p->num_elastic, cudaMemcpyHostToDevice));
  }
#undef A
  if (!ok) {
    fail(e == cudaErrorMemoryAllocation ? SPH_E_NOMEM : SPH_E_DEVICE, std::string("sph_create: ") + cudaGetErrorString(e));
    sph_destroy(c);
    return nullptr;
  }
  return c;
}

int sph_get_params(const sph_ctx *c, sph_params *out) {
  if (!c || !out) return fail(SPH_E_INVALID, "null");
  *out = c->p;
  return 0;
}

static int check_launch(sph_ctx *c, bool sync) {
  CU(cudaGetLastError());
  if (sync) CU(cudaStreamSynchronize(c->stream));
  return 0;
}

// ---------------------------------------------------------------- the 18 reference stages
int sph_stage(sph_ctx *c, int stage) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  c->pdl = false;
  switch (stage) {
    case SPH_STAGE_CLEAR_BUFFERS: launch_clear_neighbors(c); break;
    case SPH_STAGE_HASH_PARTICLES: launch_hash_count(c, true); break;
    case SPH_STAGE_SORT: launch_sort(c, false); break;
    case SPH_STAGE_SORT_POST_PASS: launch_reorder(c); break;
    case SPH_STAGE_INDEXX: launch_indexx_compat(c); break;
    case SPH_STAGE_INDEX_POST_PASS: break;  // gridCellIndexFixedUp IS the scan output of SORT
    case SPH_STAGE_FIND_NEIGHBORS: launch_find_neighbors(c, false); break;
    case SPH_STAGE_COMPUTE_DENSITY: launch_compute_density(c); break;
    case SPH_STAGE_COMPUTE_FORCES_INIT_PRESSURE: launch_forces(c, false); break;
    case SPH_STAGE_COMPUTE_ELASTIC_FORCES: launch_elastic(c); break;
    case SPH_STAGE_PREDICT_POSITIONS: launch_predict_positions(c); break;
    case SPH_STAGE_PREDICT_DENSITY: launch_predict_density(c, false); break;
    case SPH_STAGE_CORRECT_PRESSURE: launch_correct_pressure(c); break;
    case SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION: launch_pressure_force(c, 0); break;
    case SPH_STAGE_INTEGRATE: launch_integrate(c); break;
    case SPH_STAGE_CLEAR_MEMBRANE_BUFFERS: launch_clear_membrane_buffers(c); break;
    case SPH_STAGE_MEMBRANE_INTERACTION: launch_membrane_interaction(c, false); break;
    case SPH_STAGE_MEMBRANE_FINALIZE: launch_membrane_finalize(c); break;
    default: return fail(SPH_E_INVALID, "unknown stage");
  }
  return check_launch(c, true);  // QUEUE_EACH_KERNEL 1: every stage is synchronous
}

// ---------------------------------------------------------------- fused step
// Optional per-kernel timing with CUDA events on the context's own stream (bench.py roofline).
enum { PK_HASH = 0, PK_SORT, PK_NEIGHBORS, PK_FORCES, PK_ELASTIC, PK_PREDICT_DENSITY, PK_PRESSURE_FORCE,
       PK_PRESSURE_FORCE_INTEGRATE, PK_MEMBRANES, PK_SLAB_RECORDS_, PK_SLAB_ABSORB_, PK_SLAB_HALO_, PK_COUNT };
static_assert(PK_SLAB_RECORDS_ == PK_SLAB_RECORDS && PK_SLAB_ABSORB_ == PK_SLAB_ABSORB && PK_SLAB_HALO_ == PK_SLAB_HALO, "sph_ctx.h");
static const char *const kProfNames[PK_COUNT] = {
    "hash_count", "sort(scan+scatter+rank_reorder)", "find_neighbors+density", "forces+predict0+density0", "elastic",
    "predict_density+correct_pressure", "pressure_force+predict", "pressure_force+integrate", "membranes",
    "slab:classify+deliver+wait_neighbours", "slab:absorb", "slab:halo_send+wait+recv"};

struct ProfRec { int kind; cudaEvent_t a, b; };
struct Profiler {
  std::vector<ProfRec> recs;
  std::vector<cudaEvent_t> pool;
  double ms[PK_COUNT] = {0};
  long calls[PK_COUNT] = {0};
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
  void drain() {
    for (auto &r : recs) {
      float t = 0.f;
      if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { ms[r.kind] += t; calls[r.kind]++; }
      pool.push_back(r.a); pool.push_back(r.b);
    }
    recs.clear();
  }
};

struct ProfScope {
  sph_ctx *c; Profiler *p; ProfRec r;
  ProfScope(sph_ctx *c_, int kind) : c(c_), p((Profiler *)c_->profiler) {
    if (p) { r.kind = kind; r.a = p->get(); r.b = p->get(); cudaEventRecord(r.a, c->stream); }
  }
  ~ProfScope() {
    if (p) { cudaEventRecord(r.b, c->stream); p->recs.push_back(r); }
  }
};

}  // extern "C"

// the same scope for the slab kernels of sph_slab.cu
ProfMark prof_mark_begin(sph_ctx *c, int kind) {
  ProfMark m; m.on = false;
  Profiler *p = (Profiler *)c->profiler;
  if (p) { m.on = true; m.kind = kind; m.a = p->get(); m.b = p->get(); cudaEventRecord(m.a, c->stream); }
  return m;
}
void prof_mark_end(sph_ctx *c, ProfMark &m) {
  Profiler *p = (Profiler *)c->profiler;
  if (m.on && p) { cudaEventRecord(m.b, c->stream); ProfRec r; r.kind = m.kind; r.a = m.a; r.b = m.b; p->recs.push_back(r); }
}

extern "C" {

// Membrane interaction + finalize of the fused step. Without a membrane list the kernel would only zero the
// displacement half of `position` (clearMembraneBuffers) and add +0 (finalize): the integration kernel has
// folded the add in, and the displacement half is zeroed once per upload instead of once per step.
// clearMembraneBuffers zeroes both upper halves every step (sphFluid.cl:1224-1225). Nothing in the fused step writes
// them except the membrane kernels (which keep them zeroed themselves), so they are cleared once per upload.
static void clear_upper_halves_if_dirty(sph_ctx *c) {
  if (!c->dispDirty) return;
  cudaMemsetAsync(c->position + c->d.N, 0, 16 * (size_t)c->d.N, c->stream);
  cudaMemsetAsync(c->velocity + c->d.N, 0, 16 * (size_t)c->d.N, c->stream);
  c->dispDirty = false;
}

static int fused_membranes(sph_ctx *c) {
  if (c->has_pml) {
    ProfScope s(c, PK_MEMBRANES);
    launch_membrane_interaction(c, true);
    return 2;
  }
  return 0;
}

// The kernels of one fused step, in order, on c->stream, in two halves: everything up to the last density
// prediction, and the kernels that rewrite position / velocity (pressure force + integrate, membranes). A
// pipelined read-back of the previous step's positions only has to be over before the second half starts.
static int step_body(sph_ctx *c, int half) {
  const int iters = c->p.max_iteration < 1 ? 1 : c->p.max_iteration;  // do{}while: at least one pass
  const bool elastic = c->d.numElastic > 0 && c->has_elastic;
  if (half == 0) {
    { ProfScope s(c, PK_HASH); launch_hash_count(c, false); }
    { ProfScope s(c, PK_SORT); launch_sort(c, true); }
    { ProfScope s(c, PK_NEIGHBORS); launch_find_neighbors(c, true); }
    { ProfScope s(c, PK_FORCES); launch_forces(c, true); }
    if (elastic) { ProfScope s(c, PK_ELASTIC); launch_elastic(c); }
    for (int it = 0; it < iters; it++) {  // the density prediction of iteration 0 is part of the forces kernel
      if (it > 0) { ProfScope s(c, PK_PREDICT_DENSITY); launch_predict_density(c, true); }
      if (it + 1 < iters) { ProfScope s(c, PK_PRESSURE_FORCE); launch_pressure_force(c, 1); }
    }
    return 7 + (elastic ? 1 : 0) + 2 * iters - 2;
  }
  { ProfScope s(c, PK_PRESSURE_FORCE_INTEGRATE); launch_pressure_force(c, 2); }
  return 1 + fused_membranes(c);
}

#ifndef SPH_GRAPH_MAX_N
#define SPH_GRAPH_MAX_N (1 << 21)   // above this a step is milliseconds long and launch gaps do not show
#endif

static bool want_graph(const sph_ctx *c) {
  if (c->profiler || c->slab || c->auto_tiles || c->force_tiled_neighbors) return false;  // per-kernel events / host-side decisions in the step
  return c->graph_mode < 0 ? c->d.N <= SPH_GRAPH_MAX_N : c->graph_mode != 0;
}

static int capture_half(sph_ctx *c, int half) {
  cudaGraph_t g = nullptr;
  CU(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
  c->capturing = true;
  c->graphNodes[half] = step_body(c, half);
  c->capturing = false;
  cudaError_t e = cudaStreamEndCapture(c->stream, &g);
  if (e != cudaSuccess || !g) return fail(SPH_E_DEVICE, std::string("sph_step: stream capture failed: ") + cudaGetErrorString(e));
  e = cudaGraphInstantiate(&c->stepGraph[half], g, 0);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) { c->stepGraph[half] = nullptr; return fail(SPH_E_DEVICE, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e)); }
  return 0;
}

int sph_step(sph_ctx *c) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  clear_upper_halves_if_dirty(c);
  // programmatic dependent launches between the step's kernels: only on request (SPH_PDL=1). Measured with the
  // graph replay: config A 0.1269 -> 0.1282 ms, config B 0.415 -> 0.470 ms — the graph already removes the launch
  // gaps, and early-resident CTAs of the next kernel take slots from the running one. Never with per-kernel events.
  c->pdl = !c->profiler && !c->slab && c->pdl_mode > 0;
  if (!want_graph(c)) {
    c->launches += step_body(c, 0);
    c->launches += step_body(c, 1);
    return check_launch(c, false);
  }
  // Graph replay: small scenes are launch-bound (a dozen dependent kernels of a few microseconds each). What a
  // replay cannot carry sits between the two graphs: the wait for a pipelined read-back, and the once-per-upload
  // clearing of the displacement halves in front.
  if (c->stepGraph[0] && c->graphStream != c->stream) drop_graph(c);
  if (!c->cellCountClean) {   // the captured step has no memset of the cell histogram: k_rank_reorder leaves it zeroed
    CU(cudaMemsetAsync(c->cellCount, 0, sizeof(unsigned) * ((size_t)c->d.G + 1), c->stream));
    c->cellCountClean = true;
  }
  if (c->stepGraph[0] && c->graphPdl != c->pdl) drop_graph(c);
  if (!c->stepGraph[0]) {
    c->graphPdl = c->pdl;
    int rc = capture_half(c, 0);
    if (!rc) rc = capture_half(c, 1);
    if (rc) { drop_graph(c); return rc; }
    c->graphStream = c->stream;
  }
  CU(cudaGraphLaunch(c->stepGraph[0], c->stream));
  read_fence(c);
  CU(cudaGraphLaunch(c->stepGraph[1], c->stream));
  c->launches += c->graphNodes[0] + c->graphNodes[1];
  return check_launch(c, false);
}

// ---- multi-device slab extension: the fused step in the pieces between which a slab rank refreshes its
// ghosts (sph_slab.cu). part 0: hash, sort, neighbour search + rho(t)            -> refresh rho
//                       part 1: forces (+ predict #0), predictDensity + correct  -> refresh (p, rho*)
//                       part 2: pressure force + predict next                    -> refresh x*
//                       part 3: predictDensity + correctPressure                 -> refresh (p, rho*)
//                       part 4: pressure force + integrate, membranes
// A step with max_iteration = 3 is 0 | 1 | 2 | 3 | 2 | 3 | 4. Same kernels, same order as sph_step.
int sph_step_part(sph_ctx *c, int part) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  switch (part) {
    case 0:
      c->pdl = false;
      clear_upper_halves_if_dirty(c);
      { ProfScope s(c, PK_HASH); launch_hash_count(c, false); }
      { ProfScope s(c, PK_SORT); launch_sort(c, true); }
      { ProfScope s(c, PK_NEIGHBORS); launch_find_neighbors(c, true); }
      c->launches += 6;
      break;
    case 1:
      { ProfScope s(c, PK_FORCES); launch_forces(c, true); }   // includes predictDensity + correctPressure of iteration 0
      if (c->d.numElastic > 0 && c->has_elastic) { ProfScope s(c, PK_ELASTIC); launch_elastic(c); c->launches++; }
      c->launches += 1;
      break;
    case 2: { ProfScope s(c, PK_PRESSURE_FORCE); launch_pressure_force(c, 1); } c->launches++; break;
    case 3: { ProfScope s(c, PK_PREDICT_DENSITY); launch_predict_density(c, true); } c->launches++; break;
    case 4:
      { ProfScope s(c, PK_PRESSURE_FORCE_INTEGRATE); launch_pressure_force(c, 2); }
      c->launches += 1 + fused_membranes(c);
      break;
    default: return fail(SPH_E_INVALID, "sph_step_part: unknown part");
  }
  return check_launch(c, false);
}

// Run the context's work on a caller-owned CUDA stream (e.g. torch's current stream, so that collectives
// issued by the caller are ordered with the kernels); NULL restores the context's own stream.
int sph_set_stream(sph_ctx *c, void *cuda_stream) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->ownStream;
  return 0;
}

void *sph_device_alloc(size_t bytes) {
  void *p = nullptr;
  return cudaMalloc(&p, bytes ? bytes : 1) == cudaSuccess ? p : nullptr;
}
void sph_device_free(void *p) { if (p) cudaFree(p); }
// synchronous copy between any two of {host, device} buffers (unified addressing picks the direction)
int sph_device_copy(void *dst, const void *src, size_t bytes) {
  CU(cudaMemcpy(dst, src, bytes, cudaMemcpyDefault));
  return 0;
}

// ---- instrumentation (not part of the reference interface; used by bench.py / the C++ facade) ----
int sph_profile_enable(sph_ctx *c, int on) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  Profiler *p = (Profiler *)c->profiler;
  if (on && !p) c->profiler = new Profiler();
  if (!on && p) {
    cudaStreamSynchronize(c->stream);
    p->drain();
    for (auto e : p->pool) cudaEventDestroy(e);
    delete p;
    c->profiler = nullptr;
  }
  return 0;
}

// JSON object {"kernel name": {"calls": n, "ms": total}, ...}; resets the counters.
int sph_profile_report(sph_ctx *c, char *buf, size_t n) {
  if (!c || !buf || !n) return fail(SPH_E_INVALID, "sph_profile_report: null");
  Profiler *p = (Profiler *)c->profiler;
  if (!p) return fail(SPH_E_INVALID, "profiling is not enabled");
  CU(cudaStreamSynchronize(c->stream));
  p->drain();
  std::string out = "{";
  for (int k = 0; k < PK_COUNT; k++) {
    if (!p->calls[k]) continue;
    char tmp[256];
    snprintf(tmp, sizeof tmp, "%s\"%s\": {\"calls\": %ld, \"ms\": %.6f}", out.size() > 1 ? ", " : "", kProfNames[k],
             p->calls[k], p->ms[k]);
    out += tmp;
    p->calls[k] = 0; p->ms[k] = 0;
  }
  out += "}";
  if (out.size() + 1 > n) return fail(SPH_E_SIZE, "sph_profile_report: buffer too small");
  memcpy(buf, out.c_str(), out.size() + 1);
  return 0;
}

// K fused steps timed on the device with events on the context's stream (sync on both sides).
int sph_run_steps(sph_ctx *c, int steps, float *elapsed_ms) {
  if (!c || steps < 0) return fail(SPH_E_INVALID, "sph_run_steps");
  CU(cudaSetDevice(c->device));
  cudaEvent_t a, b;
  CU(cudaEventCreate(&a));
  CU(cudaEventCreate(&b));
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaEventRecord(a, c->stream));
  for (int i = 0; i < steps; i++) {
    int rc = sph_step(c);
    if (rc) return rc;
  }
  CU(cudaEventRecord(b, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, a, b));
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  if (elapsed_ms) *elapsed_ms = ms;
  return 0;
}

// Floor of a launch-bound step: a captured graph of `nodes` dependent EMPTY kernels (blocks x threads each), replayed
// `replays` times on the context's stream; *ms_per_replay is what the dependent-launch chain alone costs on this GPU.
// bench.py prints it beside the small-scene step times (why 61 k particles cannot go below ~N x this).
int sph_timer_start(sph_ctx *c);
int sph_timer_stop(sph_ctx *c, float *elapsed_ms);
__global__ void k_empty_node(int *sink) { if (sink && threadIdx.x == 0xffff) *sink = 0; }
int sph_empty_chain_ms(sph_ctx *c, int nodes, int blocks, int threads, int replays, float *ms_per_replay) {
  if (!c || nodes < 1 || blocks < 1 || threads < 1 || threads > 1024 || replays < 1 || !ms_per_replay) return fail(SPH_E_INVALID, "sph_empty_chain_ms");
  CU(cudaSetDevice(c->device));
  cudaGraph_t g = nullptr;
  cudaGraphExec_t ge = nullptr;
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
  for (int i = 0; i < nodes; i++) k_empty_node<<<blocks, threads, 0, c->stream>>>(nullptr);
  CU(cudaStreamEndCapture(c->stream, &g));
  CU(cudaGraphInstantiate(&ge, g, 0));
  cudaGraphDestroy(g);
  for (int i = 0; i < 3; i++) CU(cudaGraphLaunch(ge, c->stream));
  int rc = sph_timer_start(c);
  if (rc) return rc;
  for (int i = 0; i < replays; i++) CU(cudaGraphLaunch(ge, c->stream));
  float ms = 0.f;
  rc = sph_timer_stop(c, &ms);
  cudaGraphExecDestroy(ge);
  if (rc) return rc;
  *ms_per_replay = ms / replays;
  return 0;
}

// Device timer on the context's stream for work the caller queues itself (slab steps): start synchronises and
// records, stop records, synchronises and returns the milliseconds in between.
int sph_timer_start(sph_ctx *c) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  if (!c->evTimer[0]) { CU(cudaEventCreate(&c->evTimer[0])); CU(cudaEventCreate(&c->evTimer[1])); }
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaEventRecord(c->evTimer[0], c->stream));
  return 0;
}
int sph_timer_stop(sph_ctx *c, float *elapsed_ms) {
  if (!c || !elapsed_ms || !c->evTimer[0]) return fail(SPH_E_INVALID, "sph_timer_stop");
  CU(cudaSetDevice(c->device));
  CU(cudaEventRecord(c->evTimer[1], c->stream));
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaEventElapsedTime(elapsed_ms, c->evTimer[0], c->evTimer[1]));
  return 0;
}

long long sph_launch_count(const sph_ctx *c) { return c ? c->launches : 0; }

// {particles that needed the radial histogram in the last neighbour search, histogram-first mode of that search,
//  out-of-grid / NaN particles so far, particles that fell back to the plain two-pass scan so far}. Synchronises.
int sph_debug_counters(sph_ctx *c, unsigned out[4]) {
  if (!c || !out) return fail(SPH_E_INVALID, "sph_debug_counters");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaMemcpy(out, c->nsStats, 16, cudaMemcpyDeviceToHost));
  return 0;
}

// Particles that the hash found outside the grid or with a NaN coordinate since the context was created;
// they are binned into the last cell, where a crowd of them makes the in-cell ranking quadratic. Synchronises.
long long sph_out_of_grid_count(sph_ctx *c) {
  if (!c) return -1;
  unsigned v = 0;
  if (cudaSetDevice(c->device) != cudaSuccess || cudaStreamSynchronize(c->stream) != cudaSuccess ||
      cudaMemcpy(&v, c->nsStats + 2, 4, cudaMemcpyDeviceToHost) != cudaSuccess)
    return -1;
  return (long long)v;
}

// Pinned host memory for the read-back / upload buffers of the e2e path.
void *sph_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { fail(SPH_E_NOMEM, "cudaMallocHost failed"); return nullptr; }
  return p;
}
void sph_host_free(void *p) { if (p) cudaFreeHost(p); }

int sph_sync(sph_ctx *c) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int sph_set_muscles(sph_ctx *c, const float *signal, int n) {
  if (!c || !signal || n < 0 || n > c->p.muscle_count) return fail(SPH_E_INVALID, "sph_set_muscles: bad arguments");
  CU(cudaSetDevice(c->device));
  if (n == 0) return 0;
  // copy into the next slot of a pinned ring, upload from there in stream order: the caller's buffer is free on
  // return (the reference's blocking write, owOpenCLSolver.cpp updateMuscleActivityData) without draining the stream
  const size_t slotFloats = (size_t)(c->p.muscle_count > 0 ? c->p.muscle_count : 1);
  if (!c->sigRing) {
    CU(cudaMallocHost((void **)&c->sigRing, 4 * slotFloats * sizeof(float)));
    for (cudaEvent_t &e : c->sigDone) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  const unsigned slot = c->sigNext++ & 3u;
  CU(cudaEventSynchronize(c->sigDone[slot]));  // four uploads ago: long done
  float *stage = c->sigRing + slot * slotFloats;
  memcpy(stage, signal, 4 * (size_t)n);
  CU(cudaMemcpyAsync(c->muscle, stage, 4 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaEventRecord(c->sigDone[slot], c->stream));
  return 0;
}

// ---------------------------------------------------------------- buffer access (reference layout)
static void *dev_ptr(const sph_ctx *c, int id, size_t *bytes) {
  const size_t N = c->d.N, G = c->d.G, Ne = c->has_elastic ? c->p.num_elastic : 0;
  switch (id) {
    case SPH_BUF_POSITION: *bytes = 2 * N * 16; return c->position;
    case SPH_BUF_VELOCITY: *bytes = 2 * N * 16; return c->velocity;
    case SPH_BUF_PARTICLE_INDEX: *bytes = N * 8; return c->particleIndex;
    case SPH_BUF_PARTICLE_INDEX_BACK: *bytes = N * 4; return c->particleIndexBack;
    case SPH_BUF_SORTED_POSITION: *bytes = 2 * N * 16; return c->sortedPosition;
    case SPH_BUF_SORTED_VELOCITY: *bytes = N * 16; return c->sortedVelocity;
    case SPH_BUF_GRID_CELL_INDEX: *bytes = (G + 1) * 4; return c->gridCellIndex;
    case SPH_BUF_GRID_CELL_INDEX_FIXED_UP: *bytes = (G + 1) * 4; return c->cellStart;
    case SPH_BUF_NEIGHBOR_MAP: *bytes = N * SPH_NB * 8; return c->nbIds;  // materialised in sph_read
    case SPH_BUF_RHO: *bytes = 2 * N * 4; return c->rho;
    case SPH_BUF_PRESSURE: *bytes = N * 4; return c->pressure;
    case SPH_BUF_ACCELERATION: *bytes = 2 * N * 16; return c->acceleration;
    case SPH_BUF_ELASTIC_CONNECTIONS: *bytes = Ne * SPH_NB * 16; return c->elastic;
    case SPH_BUF_MEMBRANE_DATA: *bytes = c->has_membranes ? 12 * (size_t)c->p.num_membranes : 0; return c->membraneData;
    case SPH_BUF_PARTICLE_MEMBRANES_LIST: *bytes = c->has_pml ? 28 * (size_t)c->p.num_elastic : 0; return c->pml;
    case SPH_BUF_MUSCLE_ACTIVATION_SIGNAL: *bytes = 4 * (size_t)c->p.muscle_count; return c->muscle;
  }
  *bytes = 0;
  return nullptr;
}

size_t sph_buffer_bytes(const sph_ctx *c, int id) {
  size_t b = 0;
  if (c) dev_ptr(c, id, &b);
  return b;
}

int sph_read(sph_ctx *c, int id, void *dst, size_t bytes) {
  if (!c || !dst) return fail(SPH_E_INVALID, "sph_read: null");
  size_t have = 0;
  void *src = dev_ptr(c, id, &have);
  if (!src && bytes) return fail(SPH_E_UNSUPPORTED, "sph_read: buffer absent");
  if (bytes > have) return fail(SPH_E_SIZE, "sph_read: too many bytes");
  CU(cudaSetDevice(c->device));
  if (id == SPH_BUF_NEIGHBOR_MAP) {
    if (!c->nmScratch) {
      cudaError_t e = cudaMalloc((void **)&c->nmScratch, have);
      if (e != cudaSuccess) return fail(SPH_E_NOMEM, "sph_read: neighborMap scratch");
    }
    launch_materialize_neighbor_map(c, c->nmScratch);
    CU(cudaGetLastError());
    src = c->nmScratch;
  }
  if (id == SPH_BUF_GRID_CELL_INDEX) {  // indexx output is derived from the scan on demand
    launch_indexx_compat(c);
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

// Pipelined read-back of POSITION or VELOCITY: the copy starts when the work queued so far is done and runs on a
// second stream, so the caller can queue the next sph_step right away; the kernel of that step which rewrites the
// buffer (integration, the step's last kernel) waits for the copy. sph_read_end blocks until the bytes are in dst
// (pinned memory, or the copy is not asynchronous). One read in flight per context.
int sph_read_begin(sph_ctx *c, int id, void *dst, size_t bytes) {
  if (!c || !dst) return fail(SPH_E_INVALID, "sph_read_begin: null");
  if (id != SPH_BUF_POSITION && id != SPH_BUF_VELOCITY) return fail(SPH_E_UNSUPPORTED, "sph_read_begin: position or velocity only");
  if (c->slab) return fail(SPH_E_UNSUPPORTED, "sph_read_begin: slab ranks read through sph_slab_read_owned");
  if (c->readPending) return fail(SPH_E_INVALID, "sph_read_begin: a read is already in flight");
  size_t have = 0;
  void *src = dev_ptr(c, id, &have);
  if (bytes > have) return fail(SPH_E_SIZE, "sph_read_begin: too many bytes");
  CU(cudaSetDevice(c->device));
  if (!c->copyStream) {
    CU(cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&c->evReadReady, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->evReadDone, cudaEventDisableTiming));
  }
  CU(cudaEventRecord(c->evReadReady, c->stream));
  CU(cudaStreamWaitEvent(c->copyStream, c->evReadReady, 0));
  CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->copyStream));
  CU(cudaEventRecord(c->evReadDone, c->copyStream));
  c->readPending = 1;
  c->readSrc = src;
  return 0;
}

int sph_read_end(sph_ctx *c) {
  if (!c) return fail(SPH_E_INVALID, "sph_read_end: null");
  if (!c->readPending) return 0;
  CU(cudaSetDevice(c->device));
  CU(cudaEventSynchronize(c->evReadDone));
  c->readPending = 0;
  c->readStaged = false;
  return 0;
}

// ---- state snapshot for a caller that runs one step ahead (the drop-in owPhysicsFluidSimulator launches step
// k+1 before simulationStep k returns): rho(t) and particleIndex of the step just queued are copied device-to-
// device in stream order, so that getDensity_cpp / getParticleIndex_cpp can still deliver step k while step k+1
// overwrites the live buffers. 12 bytes per particle at HBM speed.
static int ensure_copy_stream(sph_ctx *c) {
  if (!c->copyStream) {
    CU(cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&c->evReadReady, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->evReadDone, cudaEventDisableTiming));
  }
  return 0;
}

int sph_snapshot(sph_ctx *c, int with_velocity) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  CU(cudaSetDevice(c->device));
  const size_t N = c->d.N;
  if (!c->snapRho) {
    CU(cudaMalloc((void **)&c->snapRho, 4 * N));
    CU(cudaMalloc((void **)&c->snapIndex, 8 * N));
    CU(cudaEventCreateWithFlags(&c->evSnap, cudaEventDisableTiming));
  }
  CU(cudaMemcpyAsync(c->snapRho, c->rho, 4 * N, cudaMemcpyDeviceToDevice, c->stream));
  CU(cudaMemcpyAsync(c->snapIndex, c->particleIndex, 8 * N, cudaMemcpyDeviceToDevice, c->stream));
  if (with_velocity) {
    if (!c->snapVel) CU(cudaMalloc((void **)&c->snapVel, 16 * N));
    CU(cudaMemcpyAsync(c->snapVel, c->velocity, 16 * N, cudaMemcpyDeviceToDevice, c->stream));
  }
  CU(cudaEventRecord(c->evSnap, c->stream));
  return 0;
}

// Blocking read of the last snapshot (SPH_BUF_RHO: first N floats; SPH_BUF_PARTICLE_INDEX); does not wait for
// work queued after sph_snapshot.
int sph_read_snapshot(sph_ctx *c, int id, void *dst, size_t bytes) {
  if (!c || !dst) return fail(SPH_E_INVALID, "sph_read_snapshot: null");
  if (!c->snapRho) return fail(SPH_E_INVALID, "sph_read_snapshot: no snapshot taken");
  const size_t N = c->d.N;
  const void *src;
  size_t have;
  if (id == SPH_BUF_RHO) { src = c->snapRho; have = 4 * N; }
  else if (id == SPH_BUF_PARTICLE_INDEX) { src = c->snapIndex; have = 8 * N; }
  else if (id == SPH_BUF_VELOCITY && c->snapVel) { src = c->snapVel; have = 16 * N; }
  else return fail(SPH_E_UNSUPPORTED, "sph_read_snapshot: rho, particleIndex or (if requested) velocity only");
  if (bytes > have) return fail(SPH_E_SIZE, "sph_read_snapshot: too many bytes");
  CU(cudaSetDevice(c->device));
  int rc = ensure_copy_stream(c);
  if (rc) return rc;
  if (c->readPending) return fail(SPH_E_INVALID, "sph_read_snapshot: a pipelined read is in flight");
  CU(cudaStreamWaitEvent(c->copyStream, c->evSnap, 0));
  CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->copyStream));
  CU(cudaStreamSynchronize(c->copyStream));
  return 0;
}

int sph_write(sph_ctx *c, int id, const void *src, size_t bytes) {
  if (!c || !src) return fail(SPH_E_INVALID, "sph_write: null");
  if (id != SPH_BUF_POSITION && id != SPH_BUF_VELOCITY) return fail(SPH_E_UNSUPPORTED, "sph_write: buffer not writable");
  size_t have = 0;
  void *dst = dev_ptr(c, id, &have);
  if (bytes > have) return fail(SPH_E_SIZE, "sph_write: too many bytes");
  CU(cudaSetDevice(c->device));
  read_fence(c);
  c->dispDirty = true;  // position or velocity: the upper half may have been overwritten
  CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"
