// owOpenCLConstant.h — same macro names and values as the reference's
// /root/reference/src/owOpenCLConstant.h:4-18 (interface constants, not code).
#ifndef OW_OPENCL_CONSTANT_H
#define OW_OPENCL_CONSTANT_H
#define MAX_NEIGHBOR_COUNT 32
#define MAX_MEMBRANES_INCLUDING_SAME_PARTICLE 7
#define LIQUID_PARTICLE 1
#define ELASTIC_PARTICLE 2
#define BOUNDARY_PARTICLE 3
#define NO_PARTICLE_ID -1
#define NO_CELL_ID -1
#define NO_DISTANCE -1.0f
#endif
