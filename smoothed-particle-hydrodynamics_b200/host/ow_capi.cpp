// ow_capi.cpp — extern "C" handles on the drop-in C++ classes so that tests/bench (Python, ctypes)
// exercise the same code path a C++ caller of owPhysicsFluidSimulator would.
#include <string.h>

#include "ow/owPhysicsFluidSimulator.h"

extern "C" {

void ow_set_box(double x, double y, double z) { owSetBox(x, y, z); }
void ow_set_hash_mask(unsigned int m) { owSetHashMask(m); }
void ow_set_device(int d) { owSetDevice(d); }
void ow_set_slab_devices(const int *ordinals, int n) { owSetSlabDevices(ordinals, n); }
void ow_set_quiet(int q) { owHelper::quiet = q != 0; }
void ow_set_configuration_files(const char *dir, const char *pos, const char *vel) { owHelper::setConfigurationFiles(dir, pos, vel); }
void ow_set_buffers_dir(const char *dir) { owHelper::setBuffersDir(dir); }
void ow_flush_recording(void) { owHelper::flushRecording(); }

struct ow_sim { owHelper *helper; owPhysicsFluidSimulator *sim; };

ow_sim *ow_sim_create_from_files(void) {
  ow_sim *s = new ow_sim;
  s->helper = new owHelper();
  s->sim = new owPhysicsFluidSimulator(s->helper);
  s->sim->setKeepVelocity(true);  // the tests read velocities back
  return s;
}
ow_sim *ow_sim_create(int n, const float *pos, const float *vel, int ne, const float *elastic, int nm,
                      const int *membranes, const int *pml) {
  ow_sim *s = new ow_sim;
  s->helper = new owHelper();
  s->sim = new owPhysicsFluidSimulator(s->helper, n, pos, vel, ne, elastic, nm, membranes, pml);
  s->sim->setKeepVelocity(true);  // the tests read velocities back
  return s;
}
void ow_sim_destroy(ow_sim *s) { if (s) { delete s->sim; delete s->helper; delete s; } }
double ow_sim_step(ow_sim *s, int load_to) { return s->sim->simulationStep(load_to != 0); }
void ow_sim_set_fused(ow_sim *s, int on) { s->sim->setFused(on != 0); }
void ow_sim_set_run_ahead(ow_sim *s, int on) { s->sim->setRunAhead(on != 0); }
void ow_sim_set_keep_velocity(ow_sim *s, int on) { s->sim->setKeepVelocity(on != 0); }
void ow_sim_set_muscle_hook(ow_sim *s, int on) { s->sim->setMuscleHook(on != 0); }
void ow_sim_set_muscle_callback(ow_sim *s, ow_muscle_callback cb, void *user) { s->sim->muscleSimulation().setCallback(cb, user); }
const float *ow_sim_position(ow_sim *s) { return s->sim->getPosition_cpp(); }
const float *ow_sim_density(ow_sim *s) { return s->sim->getDensity_cpp(); }
const unsigned int *ow_sim_particle_index(ow_sim *s) { return s->sim->getParticleIndex_cpp(); }
void ow_sim_read_velocity(ow_sim *s, float *dst) { s->sim->solver()->read_velocity_buffer(dst); }
int ow_particle_count(void) { return PARTICLE_COUNT; }
int ow_iteration_count(void) { return iterationCount; }
const float *ow_muscle_signal(void) { return muscle_activation_signal_cpp; }
void ow_counts(int *ne, int *nl, int *nb, int *nm, int *nc) { *ne = numOfElasticP; *nl = numOfLiquidP; *nb = numOfBoundaryP; *nm = numOfMembranes; *nc = numOfElasticConnections; }

// ---- owHelper file formats without a device (CPU tests) ----
int ow_helper_preload(void) { owHelper::preLoadConfiguration(); return PARTICLE_COUNT; }
// position/velocity: 4*PARTICLE_COUNT floats (caller-allocated). Returns Ne; *connections (new[]) may stay NULL.
int ow_helper_load(float *position, float *velocity, float **connections, int *nl, int *ne, int *nb, int *nc) {
  *nl = *ne = *nb = *nc = 0;
  *connections = 0;
  owHelper::loadConfiguration(position, velocity, *connections, *nl, *ne, *nb, *nc);
  numOfLiquidP = *nl; numOfElasticP = *ne; numOfBoundaryP = *nb; numOfElasticConnections = *nc;
  return *ne;
}
void ow_helper_free_floats(float *p) { delete[] p; }
void ow_helper_set_counts(int n, int ne, int nl, int nb, int nm) {
  PARTICLE_COUNT = n; numOfElasticP = ne; numOfLiquidP = nl; numOfBoundaryP = nb; numOfMembranes = nm;
}
void ow_helper_write_buffers(float *position, float *connections, int *membranes, int first) {
  owHelper::loadConfigurationToFile(position, connections, membranes, first != 0);
}
void ow_helper_write_buffers_async(float *position, float *connections, int *membranes, int first) {
  owHelper::loadConfigurationToFileAsync(position, connections, membranes, first != 0);
}
// Reads frame `iteration`; on iteration 0 returns freshly allocated arrays through the out-pointers.
int ow_helper_read_frame(float **position, float **connections, int **membranes, int iteration) {
  owHelper::loadConfigurationFromFile(*position, *connections, *membranes, iteration);
  return iterationCount;
}

}  // extern "C"
