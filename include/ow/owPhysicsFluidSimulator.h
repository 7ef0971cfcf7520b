// owPhysicsFluidSimulator.h — drop-in for the reference's step orchestrator
// (/root/reference/src/owPhysicsFluidSimulator.h:8-36): same class name, constructor, getters and
// `double simulationStep(const bool load_to = false)`.
#ifndef OW_PHYSICS_SIMULATOR_H
#define OW_PHYSICS_SIMULATOR_H

#include "PyramidalSimulation.h"
#include "owHelper.h"
#include "owOpenCLSolver.h"

class owPhysicsFluidSimulator {
 public:
  owPhysicsFluidSimulator(void);  // declared (never defined) by the reference, owPhysicsFluidSimulator.h:11: no scene, no solver
  // Reference constructor: loads ./configuration/*.txt (the reference's `generateInitialConfiguration 0`
  // branch, owPhysicsFluidSimulator.cpp:36-37,65-66; the procedural generator is out of scope).
  owPhysicsFluidSimulator(owHelper *helper);
  // Addition: take a scene already in memory (arrays are copied). elastic: 32*Ne float4, membranes: 3*M int,
  // pml: 7*Ne int, any of them may be NULL.
  owPhysicsFluidSimulator(owHelper *helper, int particleCount, const float *position, const float *velocity,
                          int numElastic, const float *elastic, int numMembranes, const int *membranes, const int *pml);
  ~owPhysicsFluidSimulator(void);
  float *getPosition_cpp() { return position_cpp; }
  float *getvelocity_cpp() { return velocity_cpp; }
  float *getDensity_cpp() { ocl_solver->read_density_buffer(density_cpp); return density_cpp; }
  unsigned int *getParticleIndex_cpp() { ocl_solver->read_particleIndex_buffer(particleIndex_cpp); return particleIndex_cpp; }
  float *getElasticConnectionsData_cpp() { return elasticConnectionsData_cpp; }
  int *getMembraneData_cpp() { return membraneData_cpp; }
  double simulationStep(const bool load_to = false);
  // ---- additions ----
  void setFused(bool on) { fused = on; }      // default true: one fused launch sequence instead of 18 stages
  // default true (fused mode only): simulationStep k queues step k+1 on the device before it returns — every input
  // of step k+1 is known by then, the muscle signal is updated at the END of a step (owPhysicsFluidSimulator.cpp:131-139)
  // — so the position read-back of step k and whatever the caller does between two calls overlap the next step.
  // What a call returns and what the getters deliver is unchanged: the state after step k.
  void setRunAhead(bool on) { runAhead = on; }
  // solver()->read_velocity_buffer (an addition, the reference never reads velocities back) needs the velocities
  // of step k kept while step k+1 runs: 16 more bytes per particle and step, off by default
  void setKeepVelocity(bool on) { keepVelocity = on; }
  void setMuscleHook(bool on) { muscleHook = on; }
  PyramidalSimulation &muscleSimulation() { return simulation; }
  owOpenCLSolver *solver() { return ocl_solver; }

 private:
  void finishConstruction();
  void finishStepOutput(bool load_to);
  void nextMuscleSignal();
  owOpenCLSolver *ocl_solver;
  float *position_cpp;   // page-locked: read_position_buffer lands here every step
  float *velocity_cpp;
  float *elasticConnectionsData_cpp;
  int *membraneData_cpp;
  int *particleMembranesList_cpp;
  float *density_cpp;
  unsigned int *particleIndex_cpp;
  owHelper *helper;
  PyramidalSimulation simulation;
  bool fused, muscleHook, runAhead, stepInFlight, keepVelocity;
};

#endif  // OW_PHYSICS_SIMULATOR_H
