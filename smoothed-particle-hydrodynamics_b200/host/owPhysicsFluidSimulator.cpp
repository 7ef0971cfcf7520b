// owPhysicsFluidSimulator.cpp — drop-in step orchestrator (include/ow/owPhysicsFluidSimulator.h).
// Follows /root/reference/src/owPhysicsFluidSimulator.cpp:27-77 (construction), :79-149
// (simulationStep: stage order, 3 fixed PCISPH iterations, position read-back, -l_to dump on step 0
// and every 10th step, muscle signal update AFTER the read-back) and :152-162 (destructor).
#include "ow/owPhysicsFluidSimulator.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <stdexcept>

static int iter_step = 10;       // owPhysicsFluidSimulator.cpp:20
static const int maxIteration = 3;  // owPhysicsConstant.h:76
void owSetElasticOffset(int off);

void owPhysicsFluidSimulator::finishConstruction() {
  muscle_activation_signal_cpp = new float[MUSCLE_COUNT];
  for (int i = 0; i < MUSCLE_COUNT; i++) muscle_activation_signal_cpp[i] = 0.f;
  density_cpp = new float[1 * PARTICLE_COUNT];
  particleIndex_cpp = new unsigned int[PARTICLE_COUNT * 2];
  int *pml_for_solver = 0;  // the solver constructor deletes the list it is given (owOpenCLSolver.cpp:80)
  if (particleMembranesList_cpp && membraneData_cpp) {
    pml_for_solver = new int[(size_t)numOfElasticP * MAX_MEMBRANES_INCLUDING_SAME_PARTICLE];
    memcpy(pml_for_solver, particleMembranesList_cpp, sizeof(int) * (size_t)numOfElasticP * MAX_MEMBRANES_INCLUDING_SAME_PARTICLE);
  }
  if (numOfElasticP != 0)
    ocl_solver = new owOpenCLSolver(position_cpp, velocity_cpp, elasticConnectionsData_cpp, membraneData_cpp, pml_for_solver);
  else
    ocl_solver = new owOpenCLSolver(position_cpp, velocity_cpp);
  iterationCount = 0;
}

owPhysicsFluidSimulator::owPhysicsFluidSimulator(owHelper *helper_)
    : ocl_solver(0), position_cpp(0), velocity_cpp(0), elasticConnectionsData_cpp(0), membraneData_cpp(0),
      particleMembranesList_cpp(0), density_cpp(0), particleIndex_cpp(0), helper(helper_), fused(true), muscleHook(true), runAhead(true), stepInFlight(false), keepVelocity(false) {
  try {
    numOfLiquidP = numOfElasticP = numOfBoundaryP = numOfElasticConnections = numOfMembranes = 0;
    owHelper::preLoadConfiguration();  // LOAD FROM FILE branch, owPhysicsFluidSimulator.cpp:36-37
    if (PARTICLE_COUNT <= 0) throw std::runtime_error("configuration contains no particles");
    simulation.setup();
    position_cpp = (float *)sph_host_alloc(sizeof(float) * 4 * (size_t)PARTICLE_COUNT);
    velocity_cpp = new float[4 * (size_t)PARTICLE_COUNT];
    if (!position_cpp) throw std::runtime_error(sph_last_error());
    owHelper::loadConfiguration(position_cpp, velocity_cpp, elasticConnectionsData_cpp, numOfLiquidP, numOfElasticP,
                                numOfBoundaryP, numOfElasticConnections);
    owHelper::loadMembranes(membraneData_cpp, numOfMembranes, particleMembranesList_cpp, numOfElasticP);
    // shipped files list boundary particles first; elastic ones follow them (owOpenCLSolver.cpp:435)
    owSetElasticOffset(((int)position_cpp[3] == BOUNDARY_PARTICLE) ? numOfBoundaryP : 0);
    finishConstruction();
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

owPhysicsFluidSimulator::owPhysicsFluidSimulator(owHelper *helper_, int particleCount, const float *position,
                                                 const float *velocity, int numElastic, const float *elastic,
                                                 int numMembranes, const int *membranes, const int *pml)
    : ocl_solver(0), position_cpp(0), velocity_cpp(0), elasticConnectionsData_cpp(0), membraneData_cpp(0),
      particleMembranesList_cpp(0), density_cpp(0), particleIndex_cpp(0), helper(helper_), fused(true), muscleHook(true), runAhead(true), stepInFlight(false), keepVelocity(false) {
  try {
    if (particleCount <= 0 || !position || !velocity) throw std::runtime_error("empty scene");
    PARTICLE_COUNT = particleCount;
    PARTICLE_COUNT_RoundedUp = (((PARTICLE_COUNT - 1) / local_NDRange_size) + 1) * local_NDRange_size;
    numOfLiquidP = numOfElasticP = numOfBoundaryP = numOfElasticConnections = 0;
    numOfMembranes = (membranes && pml) ? numMembranes : 0;
    position_cpp = (float *)sph_host_alloc(sizeof(float) * 4 * (size_t)PARTICLE_COUNT);
    velocity_cpp = new float[4 * (size_t)PARTICLE_COUNT];
    if (!position_cpp) throw std::runtime_error(sph_last_error());
    memcpy(position_cpp, position, sizeof(float) * 4 * (size_t)PARTICLE_COUNT);
    memcpy(velocity_cpp, velocity, sizeof(float) * 4 * (size_t)PARTICLE_COUNT);
    for (int i = 0; i < PARTICLE_COUNT; i++) switch ((int)position_cpp[4 * i + 3]) {
        case LIQUID_PARTICLE: numOfLiquidP++; break;
        case ELASTIC_PARTICLE: numOfElasticP++; break;
        case BOUNDARY_PARTICLE: numOfBoundaryP++; break;
      }
    if (elastic && numElastic > 0) {
      if (numElastic != numOfElasticP) throw std::runtime_error("elastic particle count does not match the type tags");
      size_t n = 4 * (size_t)MAX_NEIGHBOR_COUNT * numElastic;
      elasticConnectionsData_cpp = new float[n];
      memcpy(elasticConnectionsData_cpp, elastic, sizeof(float) * n);
      for (size_t k = 0; k < n; k += 4) numOfElasticConnections += elastic[k] >= 0;
    } else {
      numOfElasticP = 0;  // without springs the reference takes the pure-liquid constructor
    }
    if (numOfMembranes > 0) {
      membraneData_cpp = new int[3 * (size_t)numOfMembranes];
      memcpy(membraneData_cpp, membranes, sizeof(int) * 3 * (size_t)numOfMembranes);
      particleMembranesList_cpp = new int[(size_t)MAX_MEMBRANES_INCLUDING_SAME_PARTICLE * numOfElasticP];
      memcpy(particleMembranesList_cpp, pml, sizeof(int) * (size_t)MAX_MEMBRANES_INCLUDING_SAME_PARTICLE * numOfElasticP);
    }
    owSetElasticOffset(0);  // generated scenes: elastic particles first
    simulation.setup();
    finishConstruction();
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

// :120-130: the step-time banner and the -l_to dump (step 0 with connections and membranes, then every 10th step)
void owPhysicsFluidSimulator::finishStepOutput(bool load_to) {
  if (!owHelper::quiet) {
    printf("------------------------------------\n");
    printf("_Total_step_time:\t%9.3f ms\n", helper->get_elapsedTime());
    printf("------------------------------------\n");
  }
  if (load_to) {  // formatted and written by owHelper's writer thread (the frame is copied): a 61 k-particle frame is ~20 ms of text
    if (iterationCount == 0) owHelper::loadConfigurationToFileAsync(position_cpp, elasticConnectionsData_cpp, membraneData_cpp);
    else if (iterationCount % iter_step == 0) owHelper::loadConfigurationToFileAsync(position_cpp, NULL, NULL, false);
  }
}

// :131-139: ask the Python hook for the next muscle signal and upload it
void owPhysicsFluidSimulator::nextMuscleSignal() {
  if (muscleHook && simulation.available()) {
    std::vector<float> muscle_vector = simulation.run();  // 96 values; signals 96..99 stay 0
    for (size_t index = 0; index < muscle_vector.size() && (int)index < MUSCLE_COUNT; index++)
      muscle_activation_signal_cpp[index] = muscle_vector[index];
  }
  ocl_solver->updateMuscleActivityData(muscle_activation_signal_cpp);
}

owPhysicsFluidSimulator::owPhysicsFluidSimulator(void)
    : ocl_solver(0), position_cpp(0), velocity_cpp(0), elasticConnectionsData_cpp(0), membraneData_cpp(0),
      particleMembranesList_cpp(0), density_cpp(0), particleIndex_cpp(0), helper(0), fused(true), muscleHook(true),
      runAhead(true), stepInFlight(false), keepVelocity(false) {}

double owPhysicsFluidSimulator::simulationStep(const bool load_to) {
  int iter = 0;
  helper->refreshTime();
  if (!owHelper::quiet) printf("\n[[ Step %d ]]\n", iterationCount);
  try {
    if (fused || stepInFlight || ocl_solver->slabCount() > 0) {
      // same results as the 18 stages below, as one asynchronous launch sequence
      if (!stepInFlight && ocl_solver->_runFusedStep()) throw std::runtime_error(sph_last_error());
      stepInFlight = false;
      if (fused && runAhead && ocl_solver->slabCount() == 0) {
        // Run ahead: step k is queued (or running). Freeze what the getters of step k read, start the position
        // read-back on the copy stream, produce the inputs of step k+1 — the muscle signal, which the reference
        // also updates at the end of the step, before the next one starts (:131-139) — and queue step k+1, whose
        // integration kernel waits for the read-back. Then block until the positions of step k are in position_cpp.
        ocl_solver->snapshot(keepVelocity);
        ocl_solver->setReadFromSnapshot(true);
        ocl_solver->read_position_begin(position_cpp);
        nextMuscleSignal();
        if (ocl_solver->_runFusedStep()) throw std::runtime_error(sph_last_error());
        stepInFlight = true;
        ocl_solver->read_position_end();
        helper->watch_report("_runFusedStep+readBuffer: \t%9.3f ms\n");
        finishStepOutput(load_to);
        iterationCount++;
        return helper->get_elapsedTime();
      }
      ocl_solver->setReadFromSnapshot(false);
      ocl_solver->read_position_buffer(position_cpp);
      helper->watch_report("_runFusedStep+readBuffer: \t%9.3f ms\n");
    } else {
      ocl_solver->_runClearBuffers();        helper->watch_report("_runClearBuffers: \t%9.3f ms\n");
      ocl_solver->_runHashParticles();       helper->watch_report("_runHashParticles: \t%9.3f ms\n");
      ocl_solver->_runSort();                helper->watch_report("_runSort: \t\t%9.3f ms\n");
      ocl_solver->_runSortPostPass();        helper->watch_report("_runSortPostPass: \t%9.3f ms\n");
      ocl_solver->_runIndexx();              helper->watch_report("_runIndexx: \t\t%9.3f ms\n");
      ocl_solver->_runIndexPostPass();       helper->watch_report("_runIndexPostPass: \t%9.3f ms\n");
      ocl_solver->_runFindNeighbors();       helper->watch_report("_runFindNeighbors: \t%9.3f ms\n");
      ocl_solver->_run_pcisph_computeDensity();
      ocl_solver->_run_pcisph_computeForcesAndInitPressure();
      ocl_solver->_run_pcisph_computeElasticForces();
      do {
        ocl_solver->_run_pcisph_predictPositions();
        ocl_solver->_run_pcisph_predictDensity();
        ocl_solver->_run_pcisph_correctPressure();
        ocl_solver->_run_pcisph_computePressureForceAcceleration();
        iter++;
      } while (iter < maxIteration);
      ocl_solver->_run_pcisph_integrate(iterationCount);
      helper->watch_report("_runPCISPH: \t\t%9.3f ms\t3 iteration(s)\n");
      ocl_solver->_run_clearMembraneBuffers();
      ocl_solver->_run_computeInteractionWithMembranes();
      ocl_solver->_run_computeInteractionWithMembranes_finalize();
      ocl_solver->read_position_buffer(position_cpp);
      helper->watch_report("_readBuffer: \t\t%9.3f ms\n");
    }
    finishStepOutput(load_to);
    iterationCount++;
    nextMuscleSignal();
    return helper->get_elapsedTime();
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

owPhysicsFluidSimulator::~owPhysicsFluidSimulator(void) {
  owHelper::flushRecording();  // queued frames reference elasticConnectionsData_cpp / membraneData_cpp
  if (position_cpp) sph_host_free(position_cpp);
  delete[] velocity_cpp;
  delete[] density_cpp;
  delete[] particleIndex_cpp;
  delete[] muscle_activation_signal_cpp;
  muscle_activation_signal_cpp = 0;
  delete[] membraneData_cpp;
  delete[] particleMembranesList_cpp;
  delete[] elasticConnectionsData_cpp;
  delete ocl_solver;
}
