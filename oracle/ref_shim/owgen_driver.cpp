// owgen_driver.cpp — TEST INFRASTRUCTURE (oracle/), never linked into the product.
//
// Thin C entry points around the reference's OWN scene generator, compiled
// unchanged from /root/reference/src/owHelper.cpp (it includes no GL/CL
// headers).  It follows the call sequence of the reference constructor
// owPhysicsFluidSimulator::owPhysicsFluidSimulator
// (/root/reference/src/owPhysicsFluidSimulator.cpp:27-77): generateConfiguration
// stage 0 (counts), allocate, stage 1 (fill).  Output = BASELINE config B
// ("worm", 232 887 particles).  Built into oracle/_ref/libowgen.so by build_ref.py.
#include <stdlib.h>
#include <string.h>

#include "owHelper.h"
#include "owPhysicsConstant.h"

// globals owHelper.cpp expects (owWorldSimulation.cpp:29-31,47; owPhysicsFluidSimulator.cpp:8-20)
int PARTICLE_COUNT = 0;
int PARTICLE_COUNT_RoundedUp = 0;
int local_NDRange_size = 256;
int numOfElasticConnections = 0;
int numOfMembranes = 0;
int numOfElasticP = 0;
int numOfLiquidP = 0;
int numOfBoundaryP = 0;
int iterationCount = 0;
extern const float delta = 0.f;  // declared extern in owPhysicsConstant.h; unused by the generator

static float *g_elastic = nullptr;

extern "C" {

// stage 0: owPhysicsFluidSimulator.cpp:34
int owgen_count(int *n, int *ne, int *nl, int *nb, int *nm) {
  float *pos = nullptr, *vel = nullptr;
  int *membranes = nullptr, *pml = nullptr;
  numOfLiquidP = numOfElasticP = numOfBoundaryP = numOfElasticConnections = numOfMembranes = 0;
  g_elastic = nullptr;
  owHelper::generateConfiguration(0, pos, vel, g_elastic, membranes, numOfLiquidP, numOfElasticP, numOfBoundaryP,
                                  numOfElasticConnections, numOfMembranes, pml);
  *n = PARTICLE_COUNT; *ne = numOfElasticP; *nl = numOfLiquidP; *nb = numOfBoundaryP; *nm = numOfMembranes;
  return 0;
}

// stage 1: owPhysicsFluidSimulator.cpp:42-63.  Caller-allocated outputs:
// pos/vel 4N floats, elastic 4*32*Ne floats, membranes 3*M ints, pml 7*Ne ints.
int owgen_fill(float *pos, float *vel, float *elastic, int *membranes, int *pml) {
  if (!g_elastic && numOfElasticP > 0) return -1;
  owHelper::generateConfiguration(1, pos, vel, g_elastic, membranes, numOfLiquidP, numOfElasticP, numOfBoundaryP,
                                  numOfElasticConnections, numOfMembranes, pml);
  if (numOfElasticP > 0) memcpy(elastic, g_elastic, sizeof(float) * 4 * MAX_NEIGHBOR_COUNT * (size_t)numOfElasticP);
  delete[] g_elastic;
  g_elastic = nullptr;
  return 0;
}

}  // extern "C"
