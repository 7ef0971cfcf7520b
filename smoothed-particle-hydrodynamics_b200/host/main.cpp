// main.cpp — headless driver with the reference's CLI flags (/root/reference/src/main.cpp:8-34) and
// its headless loop (/root/reference/src/owWorldSimulation.cpp:999-1004), bounded by --steps.
//   -no_g            accepted (there is no viewer in this build; always headless)
//   -l_to            record buffers/*.txt on step 0 and every 10th step
//   -l_from          replay buffers/position_buffer.txt frame by frame (no physics), print frame count
// additions: --steps N (default 100), --box X Y Z (multiples of h; default header box 30 20 250),
//   --config DIR [--position F --velocity F], --buffers DIR, --mask HEX, --staged, --quiet, --device N, --devices a,b,.. (one z-slab per listed GPU; needs --mask ffffffff)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <vector>

#include "ow/owPhysicsFluidSimulator.h"

int main(int argc, char **argv) {
  bool load_to = false, load_from = false, staged = false;
  int steps = 100;
  const char *cfg = "./configuration", *posf = "position.txt", *velf = "velocity.txt";
  for (int i = 1; i < argc; i++) {
    if (strncmp(argv[i], "-no_g", 5) == 0) continue;
    else if (strncmp(argv[i], "-l_to", 5) == 0) { std::cout << "l_to flag, Sibernetic will save simulation results to disk\n"; load_to = true; }
    else if (strncmp(argv[i], "-l_from", 7) == 0) load_from = true;
    else if (!strcmp(argv[i], "--steps") && i + 1 < argc) steps = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--box") && i + 3 < argc) { owSetBox(atof(argv[i + 1]), atof(argv[i + 2]), atof(argv[i + 3])); i += 3; }
    else if (!strcmp(argv[i], "--config") && i + 1 < argc) cfg = argv[++i];
    else if (!strcmp(argv[i], "--position") && i + 1 < argc) posf = argv[++i];
    else if (!strcmp(argv[i], "--velocity") && i + 1 < argc) velf = argv[++i];
    else if (!strcmp(argv[i], "--buffers") && i + 1 < argc) owHelper::setBuffersDir(argv[++i]);
    else if (!strcmp(argv[i], "--mask") && i + 1 < argc) owSetHashMask((unsigned)strtoul(argv[++i], 0, 16));
    else if (!strcmp(argv[i], "--device") && i + 1 < argc) owSetDevice(atoi(argv[++i]));
    else if (!strcmp(argv[i], "--devices") && i + 1 < argc) {  // "0,1,2,3": one z-slab of the fluid per listed GPU
      std::vector<int> dv;
      for (char *t = strtok(argv[++i], ","); t; t = strtok(0, ",")) dv.push_back(atoi(t));
      owSetSlabDevices(dv.data(), (int)dv.size());
    }
    else if (!strcmp(argv[i], "--staged")) staged = true;
    else if (!strcmp(argv[i], "--quiet")) owHelper::quiet = true;
    else { fprintf(stderr, "unknown argument %s\n", argv[i]); return 2; }
  }
  owHelper::setConfigurationFiles(cfg, posf, velf);
  owHelper *helper = new owHelper();
  if (load_from) {
    float *p = 0, *c = 0; int *m = 0;
    owHelper::loadConfigurationFromFile(p, c, m, 0);
    int frames = iterationCount;
    for (int f = 1; f < frames; f++) owHelper::loadConfigurationFromFile(p, c, m, f);
    printf("replayed %d frame(s) of %d particle(s)\n", frames, PARTICLE_COUNT);
    return 0;
  }
  owPhysicsFluidSimulator *sim = new owPhysicsFluidSimulator(helper);
  sim->setFused(!staged);
  double total = 0;
  for (int s = 0; s < steps; s++) {
    total += sim->simulationStep(load_to);
    helper->refreshTime();
  }
  printf("\ntotal calculation time (%d steps) = %f ms, %.3e particle-steps/s\n", steps, total,
         total > 0 ? (double)PARTICLE_COUNT * steps / (total * 1e-3) : 0.0);
  delete sim;
  delete helper;
  return 0;
}
