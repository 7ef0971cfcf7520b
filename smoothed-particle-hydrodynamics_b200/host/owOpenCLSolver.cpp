// owOpenCLSolver.cpp — drop-in solver class over the C-ABI (see include/ow/owOpenCLSolver.h).
// Mirrors /root/reference/src/owOpenCLSolver.cpp:27-92 (constructor: create buffers, upload the scene),
// :213-687 (one method per stage), :717-742 (blocking copies, muscle upload). Errors: the reference
// throws std::runtime_error inside and prints "ERROR: ..." + exit(-1) at the class boundary (:88-91);
// the constructor here does the same, stage methods return the non-zero status like the reference's `err`.
#include "ow/owOpenCLSolver.h"

#include <stdio.h>
#include <stdlib.h>

#include <iostream>
#include <stdexcept>
#include <string.h>
#include <vector>

int PARTICLE_COUNT = 0;
int PARTICLE_COUNT_RoundedUp = 0;
int local_NDRange_size = 256;  // owWorldSimulation.cpp:47
int MUSCLE_COUNT = 100;        // owWorldSimulation.cpp:31
int numOfLiquidP = 0, numOfElasticP = 0, numOfBoundaryP = 0, numOfMembranes = 0, numOfElasticConnections = 0;
int iterationCount = 0;
int gridCellsX = 0, gridCellsY = 0, gridCellsZ = 0, gridCellCount = 0;
float *muscle_activation_signal_cpp = 0;

static double g_box[3] = {30.0, 20.0, 250.0};  // owPhysicsConstant.h:32-37
static unsigned int g_hash_mask = 0xffffu;
static int g_device = 0;
static int g_elastic_offset = 0;  // numOfBoundaryP*(!generateInitialConfiguration), owOpenCLSolver.cpp:435

void owSetBox(double x, double y, double z) { g_box[0] = x; g_box[1] = y; g_box[2] = z; }
void owSetHashMask(unsigned int mask) { g_hash_mask = mask; }
void owSetDevice(int d) { g_device = d; }
static std::vector<int> g_slab_devices;
void owSetSlabDevices(const int *ordinals, int n) {
  g_slab_devices.clear();
  for (int i = 0; i < n && n > 1; i++) g_slab_devices.push_back(ordinals[i]);
}
void owSetElasticOffset(int off) { g_elastic_offset = off; }

void owCurrentParams(sph_params *p) {
  sph_params_default(p);
  sph_params_set_box(p, g_box[0], g_box[1], g_box[2]);
  p->particle_count = PARTICLE_COUNT;
  p->num_elastic = numOfElasticP;
  p->num_liquid = numOfLiquidP;
  p->num_boundary = numOfBoundaryP;
  p->num_membranes = numOfMembranes;
  p->muscle_count = MUSCLE_COUNT;
  p->elastic_offset = g_elastic_offset;
  p->hash_mask = g_hash_mask;
  gridCellsX = p->grid_cells_x; gridCellsY = p->grid_cells_y; gridCellsZ = p->grid_cells_z;
  gridCellCount = p->grid_cell_count;
}

owOpenCLSolver::owOpenCLSolver(const float *position_cpp, const float *velocity_cpp,
                               const float *elasticConnectionsData_cpp, const int *membraneData_cpp,
                               const int *particleMembranesList_cpp)
    : ctx(0), fromSnapshot(false), slabs(0), numSlabs(0) {
  try {
    sph_params p;
    owCurrentParams(&p);
    if (g_slab_devices.size() > 1) {
      if (elasticConnectionsData_cpp || membraneData_cpp) throw std::runtime_error("several devices: scenes with elastic matter are single-device (DESIGN.md 7)");
      createSlabs(position_cpp, velocity_cpp);
      return;
    }
    ctx = sph_create(&p, position_cpp, velocity_cpp, elasticConnectionsData_cpp, membraneData_cpp,
                     particleMembranesList_cpp, g_device);
    if (!ctx) throw std::runtime_error(std::string("Buffer creation failed: ") + sph_last_error());
    // owOpenCLSolver.cpp:80: the reference constructor frees the caller's particleMembranesList_cpp
    if (membraneData_cpp != 0 && particleMembranesList_cpp) delete[] particleMembranesList_cpp;
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

owOpenCLSolver::owOpenCLSolver(void) : ctx(0), fromSnapshot(false), slabs(0), numSlabs(0) {}

owOpenCLSolver::~owOpenCLSolver(void) {
  if (ctx) sph_destroy(ctx);
  ctx = 0;
  destroySlabs();
}

// ---- several devices: z-slabs of one fluid, linked through peer memory (include/sph_capi.h, "links") ----
static void slab_must(int rc, const char *what) {
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + sph_slab_last_error() + " / " + sph_last_error());
}

void owOpenCLSolver::createSlabs(const float *position_cpp, const float *velocity_cpp) {
  sph_params p;
  owCurrentParams(&p);
  if (p.hash_mask != 0xffffffffu) throw std::runtime_error("several devices need owSetHashMask(0xffffffff)");
  const int world = (int)g_slab_devices.size(), N = PARTICLE_COUNT, gz = p.grid_cells_z;
  // cell layer of every particle with hashParticles' arithmetic (sphFluid.cl:361-365), layer histogram, balanced cuts
  std::vector<int> cz(N), hist(gz, 0);
  for (int i = 0; i < N; i++) {
    int c = (int)(position_cpp[4 * i + 2] * p.hash_grid_cell_size_inv);
    c = c < 0 ? 0 : (c >= gz ? gz - 1 : c);
    cz[i] = c;
    hist[c]++;
  }
  std::vector<int> cuts(world + 1, 0);
  cuts[world] = gz;
  long long cum = 0;
  int layer = 0, fullest = 0;
  for (int l = 0; l < gz; l++) fullest = hist[l] > fullest ? hist[l] : fullest;
  for (int r = 1; r < world; r++) {
    const long long target = (long long)N * r / world;
    while (layer < gz && cum + hist[layer] <= target) cum += hist[layer++];
    int c = layer;
    if (c < cuts[r - 1] + 2) c = cuts[r - 1] + 2;
    if (c > gz - 2 * (world - r)) c = gz - 2 * (world - r);
    cuts[r] = c;
  }
  for (int r = 0; r < world; r++)
    if (cuts[r + 1] - cuts[r] < 2) throw std::runtime_error("grid too thin for this many slabs (>= 2 cell layers each)");
  const int haloCap = 2 * fullest + 1024, migCap = fullest + 1024;
  slabs = new Slab[world];
  numSlabs = world;
  for (int r = 0; r < world; r++) { slabs[r].ctx = 0; slabs[r].pos = 0; slabs[r].gid = 0; slabs[r].capacity = 0; }
  for (int r = 0; r < world; r++) {
    std::vector<unsigned int> gid;
    for (int i = 0; i < N; i++)
      if (cz[i] >= cuts[r] && cz[i] < cuts[r + 1]) gid.push_back((unsigned int)i);
    const int n = (int)gid.size();
    const int cap = n + 2 * haloCap + 2 * migCap + n / 8 + 1024;
    std::vector<float> pos(4 * (size_t)cap, 0.f), vel(4 * (size_t)cap, 0.f);
    for (int k = 0; k < n; k++) {
      memcpy(&pos[4 * (size_t)k], position_cpp + 4 * (size_t)gid[k], 16);
      memcpy(&vel[4 * (size_t)k], velocity_cpp + 4 * (size_t)gid[k], 16);
    }
    sph_params q = p;
    q.particle_count = cap;
    slabs[r].ctx = sph_create(&q, pos.data(), vel.data(), 0, 0, 0, g_slab_devices[r]);
    if (!slabs[r].ctx) throw std::runtime_error(std::string("Buffer creation failed: ") + sph_last_error());
    slabs[r].capacity = cap;
    slabs[r].pos = (float *)sph_host_alloc(16 * (size_t)cap);
    slabs[r].gid = (unsigned int *)sph_host_alloc(4 * (size_t)cap);
    if (!slabs[r].pos || !slabs[r].gid) throw std::runtime_error(sph_last_error());
    slab_must(sph_slab_begin(slabs[r].ctx, n, gid.data(), cuts[r], cuts[r + 1], r > 0, r + 1 < world, migCap, haloCap), "sph_slab_begin");
  }
  for (int r = 0; r < world; r++) slab_must(sph_slab_link_export(slabs[r].ctx, 0), "sph_slab_link_export");
  for (int r = 0; r < world; r++) {
    if (r > 0) slab_must(sph_slab_link_local(slabs[r].ctx, 0, slabs[r - 1].ctx), "sph_slab_link_local");
    if (r + 1 < world) slab_must(sph_slab_link_local(slabs[r].ctx, 1, slabs[r + 1].ctx), "sph_slab_link_local");
  }
}

void owOpenCLSolver::destroySlabs() {
  for (int r = 0; r < numSlabs; r++) {
    if (slabs[r].ctx) sph_destroy(slabs[r].ctx);
    if (slabs[r].pos) sph_host_free(slabs[r].pos);
    if (slabs[r].gid) sph_host_free(slabs[r].gid);
  }
  delete[] slabs;
  slabs = 0;
  numSlabs = 0;
}

// the 18 single stages exist per device only; with several devices the step is the fused one
unsigned int owOpenCLSolver::stage(int id) { return numSlabs ? (unsigned int)SPH_E_UNSUPPORTED : (unsigned int)sph_stage(ctx, id); }

unsigned int owOpenCLSolver::_runClearBuffers() { return stage(SPH_STAGE_CLEAR_BUFFERS); }
unsigned int owOpenCLSolver::_runHashParticles() { return stage(SPH_STAGE_HASH_PARTICLES); }
unsigned int owOpenCLSolver::_runSort() { return stage(SPH_STAGE_SORT); }
unsigned int owOpenCLSolver::_runSortPostPass() { return stage(SPH_STAGE_SORT_POST_PASS); }
unsigned int owOpenCLSolver::_runIndexx() { return stage(SPH_STAGE_INDEXX); }
unsigned int owOpenCLSolver::_runIndexPostPass() { return stage(SPH_STAGE_INDEX_POST_PASS); }
unsigned int owOpenCLSolver::_runFindNeighbors() { return stage(SPH_STAGE_FIND_NEIGHBORS); }
unsigned int owOpenCLSolver::_run_pcisph_computeDensity() { return stage(SPH_STAGE_COMPUTE_DENSITY); }
unsigned int owOpenCLSolver::_run_pcisph_computeForcesAndInitPressure() { return stage(SPH_STAGE_COMPUTE_FORCES_INIT_PRESSURE); }
unsigned int owOpenCLSolver::_run_pcisph_computeElasticForces() { return stage(SPH_STAGE_COMPUTE_ELASTIC_FORCES); }
unsigned int owOpenCLSolver::_run_pcisph_predictPositions() { return stage(SPH_STAGE_PREDICT_POSITIONS); }
unsigned int owOpenCLSolver::_run_pcisph_predictDensity() { return stage(SPH_STAGE_PREDICT_DENSITY); }
unsigned int owOpenCLSolver::_run_pcisph_correctPressure() { return stage(SPH_STAGE_CORRECT_PRESSURE); }
unsigned int owOpenCLSolver::_run_pcisph_computePressureForceAcceleration() { return stage(SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION); }
unsigned int owOpenCLSolver::_run_pcisph_integrate(int) { return stage(SPH_STAGE_INTEGRATE); }
unsigned int owOpenCLSolver::_run_clearMembraneBuffers() { return stage(SPH_STAGE_CLEAR_MEMBRANE_BUFFERS); }
unsigned int owOpenCLSolver::_run_computeInteractionWithMembranes() { return stage(SPH_STAGE_MEMBRANE_INTERACTION); }
unsigned int owOpenCLSolver::_run_computeInteractionWithMembranes_finalize() { return stage(SPH_STAGE_MEMBRANE_FINALIZE); }

unsigned int owOpenCLSolver::_runFusedStep() {
  if (!numSlabs) return (unsigned int)sph_step(ctx);
  for (int r = 0; r < numSlabs; r++) {  // every slab delivers before any slab waits for its neighbours
    int rc = sph_slab_step_begin(slabs[r].ctx);
    if (rc) return (unsigned int)rc;
  }
  for (int r = 0; r < numSlabs; r++) {
    int rc = sph_slab_step_end(slabs[r].ctx);
    if (rc) return (unsigned int)rc;
  }
  return 0;
}

unsigned int owOpenCLSolver::updateMuscleActivityData(float *signal) {
  if (numSlabs) return 0;  // no elastic matter, no muscles, with several devices
  return (unsigned int)sph_set_muscles(ctx, signal, MUSCLE_COUNT);
}

static void must(int rc, const char *what) {  // copy_buffer_from_device throws "Could not enqueue read" (:727-736)
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + sph_last_error());
}
void owOpenCLSolver::read_position_buffer(float *dst) {
  if (numSlabs) {  // every slab's owned particles, scattered to the caller's particle order by global id
    for (int r = 0; r < numSlabs; r++) {
      int info[6];
      must(sph_slab_read_owned(slabs[r].ctx, slabs[r].pos, 0, slabs[r].gid), "Could not enqueue read");
      must(sph_slab_info(slabs[r].ctx, info), "sph_slab_info");
      for (int k = 0; k < info[0]; k++) memcpy(dst + 4 * (size_t)slabs[r].gid[k], slabs[r].pos + 4 * (size_t)k, 16);
    }
    return;
  }
  must(sph_read(ctx, SPH_BUF_POSITION, dst, sizeof(float) * 4 * (size_t)PARTICLE_COUNT), "Could not enqueue read");
}
void owOpenCLSolver::read_velocity_buffer(float *dst) {
  if (numSlabs) {
    std::vector<float> v;
    for (int r = 0; r < numSlabs; r++) {
      int info[6];
      v.resize(4 * (size_t)slabs[r].capacity);
      must(sph_slab_read_owned(slabs[r].ctx, 0, v.data(), slabs[r].gid), "Could not enqueue read");
      must(sph_slab_info(slabs[r].ctx, info), "sph_slab_info");
      for (int k = 0; k < info[0]; k++) memcpy(dst + 4 * (size_t)slabs[r].gid[k], &v[4 * (size_t)k], 16);
    }
    return;
  }
  const size_t n = sizeof(float) * 4 * (size_t)PARTICLE_COUNT;
  must(fromSnapshot ? sph_read_snapshot(ctx, SPH_BUF_VELOCITY, dst, n) : sph_read(ctx, SPH_BUF_VELOCITY, dst, n), "Could not enqueue read");
}
void owOpenCLSolver::read_density_buffer(float *dst) {
  if (numSlabs) throw std::runtime_error("read_density_buffer: sorted-order buffers are per device; not available with several devices");
  const size_t n = sizeof(float) * (size_t)PARTICLE_COUNT;
  must(fromSnapshot ? sph_read_snapshot(ctx, SPH_BUF_RHO, dst, n) : sph_read(ctx, SPH_BUF_RHO, dst, n), "Could not enqueue read");
}
void owOpenCLSolver::read_particleIndex_buffer(unsigned int *dst) {
  if (numSlabs) throw std::runtime_error("read_particleIndex_buffer: sorted-order buffers are per device; not available with several devices");
  const size_t n = sizeof(unsigned int) * 2 * (size_t)PARTICLE_COUNT;
  must(fromSnapshot ? sph_read_snapshot(ctx, SPH_BUF_PARTICLE_INDEX, dst, n) : sph_read(ctx, SPH_BUF_PARTICLE_INDEX, dst, n),
       "Could not enqueue read");
}
void owOpenCLSolver::read_position_begin(float *dst) {
  must(sph_read_begin(ctx, SPH_BUF_POSITION, dst, sizeof(float) * 4 * (size_t)PARTICLE_COUNT), "Could not enqueue read");
}
void owOpenCLSolver::read_position_end() { must(sph_read_end(ctx), "Could not finish read"); }
void owOpenCLSolver::snapshot(bool with_velocity) { must(sph_snapshot(ctx, with_velocity ? 1 : 0), "Could not snapshot"); }
