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
    : ctx(0), fromSnapshot(false) {
  try {
    sph_params p;
    owCurrentParams(&p);
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

owOpenCLSolver::owOpenCLSolver(void) : ctx(0), fromSnapshot(false) {}

owOpenCLSolver::~owOpenCLSolver(void) {
  if (ctx) sph_destroy(ctx);
  ctx = 0;
}

unsigned int owOpenCLSolver::stage(int id) { return (unsigned int)sph_stage(ctx, id); }

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

unsigned int owOpenCLSolver::_runFusedStep() { return (unsigned int)sph_step(ctx); }

unsigned int owOpenCLSolver::updateMuscleActivityData(float *signal) {
  return (unsigned int)sph_set_muscles(ctx, signal, MUSCLE_COUNT);
}

static void must(int rc, const char *what) {  // copy_buffer_from_device throws "Could not enqueue read" (:727-736)
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + sph_last_error());
}
void owOpenCLSolver::read_position_buffer(float *dst) {
  must(sph_read(ctx, SPH_BUF_POSITION, dst, sizeof(float) * 4 * (size_t)PARTICLE_COUNT), "Could not enqueue read");
}
void owOpenCLSolver::read_velocity_buffer(float *dst) {
  const size_t n = sizeof(float) * 4 * (size_t)PARTICLE_COUNT;
  must(fromSnapshot ? sph_read_snapshot(ctx, SPH_BUF_VELOCITY, dst, n) : sph_read(ctx, SPH_BUF_VELOCITY, dst, n), "Could not enqueue read");
}
void owOpenCLSolver::read_density_buffer(float *dst) {
  const size_t n = sizeof(float) * (size_t)PARTICLE_COUNT;
  must(fromSnapshot ? sph_read_snapshot(ctx, SPH_BUF_RHO, dst, n) : sph_read(ctx, SPH_BUF_RHO, dst, n), "Could not enqueue read");
}
void owOpenCLSolver::read_particleIndex_buffer(unsigned int *dst) {
  const size_t n = sizeof(unsigned int) * 2 * (size_t)PARTICLE_COUNT;
  must(fromSnapshot ? sph_read_snapshot(ctx, SPH_BUF_PARTICLE_INDEX, dst, n) : sph_read(ctx, SPH_BUF_PARTICLE_INDEX, dst, n),
       "Could not enqueue read");
}
void owOpenCLSolver::read_position_begin(float *dst) {
  must(sph_read_begin(ctx, SPH_BUF_POSITION, dst, sizeof(float) * 4 * (size_t)PARTICLE_COUNT), "Could not enqueue read");
}
void owOpenCLSolver::read_position_end() { must(sph_read_end(ctx), "Could not finish read"); }
void owOpenCLSolver::snapshot(bool with_velocity) { must(sph_snapshot(ctx, with_velocity ? 1 : 0), "Could not snapshot"); }
