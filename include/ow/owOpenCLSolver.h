// owOpenCLSolver.h — drop-in for the reference's solver class
// (/root/reference/src/owOpenCLSolver.h:28-124): same class name, constructor and public
// methods, same `unsigned int` error returns (0 = success), every stage synchronous on return.
// The implementation has no OpenCL in it: each method is one call into the C-ABI of
// libsph_b200.so (include/sph_capi.h), i.e. hand-written sm_100a CUDA kernels.
#ifndef OW_OPENCL_SOLVER_H
#define OW_OPENCL_SOLVER_H

#include "owOpenCLConstant.h"
#include "../sph_capi.h"

// Globals the reference's callers set/read (owWorldSimulation.cpp:29-31,47;
// owPhysicsFluidSimulator.cpp:8-20; owOpenCLSolver.cpp:14-23).
extern int PARTICLE_COUNT;
extern int PARTICLE_COUNT_RoundedUp;
extern int local_NDRange_size;
extern int MUSCLE_COUNT;
extern int numOfLiquidP, numOfElasticP, numOfBoundaryP, numOfMembranes, numOfElasticConnections;
extern int iterationCount;
extern int gridCellsX, gridCellsY, gridCellsZ, gridCellCount;
extern float *muscle_activation_signal_cpp;

// The reference fixes the box with `#define XMAX 30.0*h` (owPhysicsConstant.h:32-37); the shipped
// configuration/*.txt were written for 36h x 24h x 54h, so the box is a run-time setting here.
// Call before constructing a solver/simulator. hash_mask: 0xffff in the reference (sphFluid.cl:229,377).
void owSetBox(double xmax_in_h, double ymax_in_h, double zmax_in_h);
void owSetHashMask(unsigned int mask);
void owSetDevice(int cuda_ordinal);
// Addition (the reference is single-device, owOpenCLSolver.cpp:157-182): run ONE fluid on several GPUs of this process.
// The solver cuts the hash grid into z-slabs (one per entry of `cuda_ordinals`; an ordinal may repeat), links the
// slabs through peer memory (sph_slab_link_local) and steps them with sph_slab_step_begin / _end; read_position_buffer
// assembles the global array in the caller's particle order. Needs owSetHashMask(0xffffffff) and a scene without elastic
// matter; results are bit-identical to the single-device run. n <= 1 restores the single-device solver.
void owSetSlabDevices(const int *cuda_ordinals, int n);
// Fill *p from the globals above + the current box (what every solver is created with).
void owCurrentParams(sph_params *p);

class owOpenCLSolver {
 public:
  owOpenCLSolver(const float *position_cpp, const float *velocity_cpp, const float *elasticConnectionsData_cpp = 0,
                 const int *membraneData_cpp = 0, const int *particleMembranesList_cpp = 0);
  owOpenCLSolver(void);  // declared (never defined) by the reference, owOpenCLSolver.h:32: an empty solver, no device buffers
  ~owOpenCLSolver(void);
  // neighbour search (owOpenCLSolver.h:37-43)
  unsigned int _runClearBuffers();
  unsigned int _runHashParticles();
  unsigned int _runSort();
  unsigned int _runSortPostPass();
  unsigned int _runIndexx();
  unsigned int _runIndexPostPass();
  unsigned int _runFindNeighbors();
  // PCISPH (owOpenCLSolver.h:45-52)
  unsigned int _run_pcisph_computeDensity();
  unsigned int _run_pcisph_computeForcesAndInitPressure();
  unsigned int _run_pcisph_computeElasticForces();
  unsigned int _run_pcisph_predictPositions();
  unsigned int _run_pcisph_predictDensity();
  unsigned int _run_pcisph_correctPressure();
  unsigned int _run_pcisph_computePressureForceAcceleration();
  unsigned int _run_pcisph_integrate(int iterationCount);
  // membranes (owOpenCLSolver.h:54-56)
  unsigned int _run_clearMembraneBuffers();
  unsigned int _run_computeInteractionWithMembranes();
  unsigned int _run_computeInteractionWithMembranes_finalize();
  unsigned int updateMuscleActivityData(float *_muscle_activation_signal_cpp);
  void read_position_buffer(float *position_cpp);
  void read_density_buffer(float *density_cpp);
  void read_particleIndex_buffer(unsigned int *particleIndexBuffer);
  // ---- additions (no reference counterpart) ----
  unsigned int _runFusedStep();       // the 18 stages above as one fused launch sequence, same results
  void read_velocity_buffer(float *velocity_cpp);
  // run-ahead support (owPhysicsFluidSimulator queues step k+1 before simulationStep k returns):
  void read_position_begin(float *pinned_position_cpp);  // read_position_buffer without its blocking wait ...
  void read_position_end();                              // ... which happens here
  void snapshot(bool with_velocity); // freeze rho / particleIndex (/ velocity) of the step queued so far for the read_* methods
  void setReadFromSnapshot(bool on) { fromSnapshot = on; }
  sph_ctx *context() { return ctx; }
  int slabCount() const { return numSlabs; }

 private:
  unsigned int stage(int id);
  void createSlabs(const float *position_cpp, const float *velocity_cpp);
  void destroySlabs();
  sph_ctx *ctx;
  bool fromSnapshot;
  // several devices (owSetSlabDevices): one context per slab, host staging for the read-back
  struct Slab { sph_ctx *ctx; float *pos; unsigned int *gid; int capacity; };
  Slab *slabs;
  int numSlabs;
};

#endif  // OW_OPENCL_SOLVER_H
