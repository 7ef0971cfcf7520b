/*
 * sph_capi.h — C-ABI of the B200 PCISPH step (libsph_b200.so).
 *
 * The reference has no FFI: its boundary is the C++ class owOpenCLSolver
 * (/root/reference/src/owOpenCLSolver.h:28-124) driven by
 * owPhysicsFluidSimulator::simulationStep
 * (/root/reference/src/owPhysicsFluidSimulator.cpp:79-149).  This header is the
 * plain-C surface that sits *underneath* the drop-in C++ classes in
 * include/ow/ — every entry point cites the reference interface it replaces.
 *
 * The same ABI is exported by three libraries so tests can drive them with one
 * loader:
 *   smoothed-particle-hydrodynamics_b200/libsph_b200.so  CUDA sm_100a (the product)
 *   oracle/liboracle.so                                  plain-C restatement (test infrastructure)
 *   oracle/_ref/libsphref.so                             the reference's own sphFluid.cl compiled as C++
 *
 * Conventions: plain pointers and sizes, caller-owned host memory, no
 * exceptions across the boundary; every int-returning call returns 0 on
 * success (mirrors the reference's `unsigned int err`, 0 == CL_SUCCESS) and a
 * negative SPH_E_* otherwise, with sph_last_error() giving the text.
 * Not thread-safe: one context per host thread (as the reference).
 */
#ifndef SPH_CAPI_H
#define SPH_CAPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPH_MAX_NEIGHBOR_COUNT 32                  /* owOpenCLConstant.h:4  */
#define SPH_MAX_MEMBRANES_INCLUDING_SAME_PARTICLE 7 /* owOpenCLConstant.h:6  */
#define SPH_LIQUID_PARTICLE 1                      /* owOpenCLConstant.h:8  */
#define SPH_ELASTIC_PARTICLE 2                     /* owOpenCLConstant.h:9  */
#define SPH_BOUNDARY_PARTICLE 3                    /* owOpenCLConstant.h:10 */

enum sph_error {
  SPH_OK = 0,
  SPH_E_INVALID = -1,      /* bad argument / bad params                     */
  SPH_E_NOMEM = -2,        /* host or device allocation failed             */
  SPH_E_DEVICE = -3,       /* CUDA runtime error (text in sph_last_error)  */
  SPH_E_UNSUPPORTED = -4,  /* buffer/stage does not exist in this config   */
  SPH_E_SIZE = -5          /* byte count does not match the buffer         */
};

/* Everything the reference fixes at compile time in owPhysicsConstant.h /
 * owOpenCLConstant.h / owOpenCLSolver.cpp:7-17 / owWorldSimulation.cpp:29-47,
 * as run-time parameters.  sph_params_default() fills the header values
 * bit-for-bit (tests assert equality against the reference headers). */
typedef struct sph_params {
  /* counts (globals PARTICLE_COUNT, numOfElasticP, ... owPhysicsFluidSimulator.cpp:8-20) */
  int32_t particle_count;
  int32_t num_elastic;          /* elastic particles occupy original indices [0, num_elastic) */
  int32_t num_liquid;
  int32_t num_boundary;
  int32_t num_membranes;
  int32_t muscle_count;         /* MUSCLE_COUNT = 100, owWorldSimulation.cpp:31 */
  int32_t elastic_offset;       /* numOfBoundaryP*(!generateInitialConfiguration), owOpenCLSolver.cpp:435 */
  int32_t max_iteration;        /* maxIteration = 3, owPhysicsConstant.h:76 */
  /* box, owPhysicsConstant.h:32-37 and owOpenCLSolver.cpp:7-12 */
  float xmin, xmax, ymin, ymax, zmin, zmax;
  /* hash grid, owOpenCLSolver.cpp:14-17 */
  int32_t grid_cells_x, grid_cells_y, grid_cells_z, grid_cell_count;
  uint32_t hash_mask;           /* 0xffff in the reference, sphFluid.cl:229,377 */
  /* physics, owPhysicsConstant.h:12-27,62-76 */
  float h, hash_grid_cell_size, hash_grid_cell_size_inv;
  float simulation_scale, simulation_scale_inv, r0;
  float rho0, mass, time_step, viscosity, stiffness, damping;
  float gravity_x, gravity_y, gravity_z;
  float delta;                  /* calcDelta(), owPhysicsFluidSimulator.cpp:164-203 */
  double wpoly6_coefficient;
  double grad_wspiky_coefficient;
  double del2_wviscosity_coefficient;
} sph_params;

/* The reference's 18 solver stages, in simulationStep order
 * (owOpenCLSolver.h:37-59, owPhysicsFluidSimulator.cpp:88-113). */
enum sph_stage_id {
  SPH_STAGE_CLEAR_BUFFERS = 0,                       /* _runClearBuffers                              */
  SPH_STAGE_HASH_PARTICLES = 1,                      /* _runHashParticles                             */
  SPH_STAGE_SORT = 2,                                /* _runSort                                      */
  SPH_STAGE_SORT_POST_PASS = 3,                      /* _runSortPostPass                              */
  SPH_STAGE_INDEXX = 4,                              /* _runIndexx                                    */
  SPH_STAGE_INDEX_POST_PASS = 5,                     /* _runIndexPostPass                             */
  SPH_STAGE_FIND_NEIGHBORS = 6,                      /* _runFindNeighbors                             */
  SPH_STAGE_COMPUTE_DENSITY = 7,                     /* _run_pcisph_computeDensity                    */
  SPH_STAGE_COMPUTE_FORCES_INIT_PRESSURE = 8,        /* _run_pcisph_computeForcesAndInitPressure      */
  SPH_STAGE_COMPUTE_ELASTIC_FORCES = 9,              /* _run_pcisph_computeElasticForces              */
  SPH_STAGE_PREDICT_POSITIONS = 10,                  /* _run_pcisph_predictPositions                  */
  SPH_STAGE_PREDICT_DENSITY = 11,                    /* _run_pcisph_predictDensity                    */
  SPH_STAGE_CORRECT_PRESSURE = 12,                   /* _run_pcisph_correctPressure                   */
  SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION = 13,/* _run_pcisph_computePressureForceAcceleration  */
  SPH_STAGE_INTEGRATE = 14,                          /* _run_pcisph_integrate                         */
  SPH_STAGE_CLEAR_MEMBRANE_BUFFERS = 15,             /* _run_clearMembraneBuffers                     */
  SPH_STAGE_MEMBRANE_INTERACTION = 16,               /* _run_computeInteractionWithMembranes          */
  SPH_STAGE_MEMBRANE_FINALIZE = 17,                  /* _run_computeInteractionWithMembranes_finalize */
  SPH_STAGE_COUNT = 18
};

/* Device buffers of owOpenCLSolver.h:72-97, always read/written in the
 * REFERENCE layout (sizes from owOpenCLSolver.cpp:32-44,70-86), whatever the
 * library keeps internally. N = particle_count, G = grid_cell_count. */
enum sph_buffer_id {
  SPH_BUF_POSITION = 0,               /* 2N float4: [0,N) x,y,z,type (original order); [N,2N) membrane displacement */
  SPH_BUF_VELOCITY = 1,               /* 2N float4                                                                  */
  SPH_BUF_PARTICLE_INDEX = 2,         /* N uint2 (cellId, serialId)                                                 */
  SPH_BUF_PARTICLE_INDEX_BACK = 3,    /* N uint                                                                     */
  SPH_BUF_SORTED_POSITION = 4,        /* 2N float4: [0,N) sorted, w=(float)cellId; [N,2N) predicted                 */
  SPH_BUF_SORTED_VELOCITY = 5,        /* N float4                                                                   */
  SPH_BUF_GRID_CELL_INDEX = 6,        /* G+1 uint, -1 for empty cells                                               */
  SPH_BUF_GRID_CELL_INDEX_FIXED_UP = 7,/* G+1 uint                                                                  */
  SPH_BUF_NEIGHBOR_MAP = 8,           /* 32N float2 (id as float, distance in metres), sorted index                 */
  SPH_BUF_RHO = 9,                    /* 2N float: rho(t), rho*                                                     */
  SPH_BUF_PRESSURE = 10,              /* N float                                                                    */
  SPH_BUF_ACCELERATION = 11,          /* 2N float4: non-pressure, pressure                                          */
  SPH_BUF_ELASTIC_CONNECTIONS = 12,   /* 32*Ne float4                                                               */
  SPH_BUF_MEMBRANE_DATA = 13,         /* 3*M int                                                                    */
  SPH_BUF_PARTICLE_MEMBRANES_LIST = 14,/* 7*Ne int                                                                  */
  SPH_BUF_MUSCLE_ACTIVATION_SIGNAL = 15,/* muscle_count float                                                       */
  SPH_BUF_COUNT = 16
};

typedef struct sph_ctx sph_ctx;

/* Fill *p with the reference's header constants and the header box
 * (30h x 20h x 250h), counts zero.  owPhysicsConstant.h, owOpenCLSolver.cpp:7-17. */
int sph_params_default(sph_params *p);

/* Set the box to (mx*h, my*h, mz*h) exactly as `#define XMAX mx*h` would
 * (double product, then float), and recompute the grid as owOpenCLSolver.cpp:14-17. */
int sph_params_set_box(sph_params *p, double mx, double my, double mz);

/* owOpenCLSolver::owOpenCLSolver(position_cpp, velocity_cpp, elasticConnectionsData_cpp,
 * membraneData_cpp, particleMembranesList_cpp), owOpenCLSolver.cpp:27-92.
 * pos4/vel4: particle_count float4 each.  elastic4: 32*num_elastic float4 or NULL.
 * membranes3: 3*num_membranes int or NULL.  pml7: 7*num_elastic int or NULL.
 * All inputs are copied; unlike the reference the library does NOT take
 * ownership of pml7 (owOpenCLSolver.cpp:80 deletes it; the C++ facade does that).
 * device: CUDA ordinal (ignored by the CPU oracles).  NULL on failure. */
sph_ctx *sph_create(const sph_params *p, const float *pos4, const float *vel4,
                    const float *elastic4, const int32_t *membranes3,
                    const int32_t *pml7, int device);

/* owOpenCLSolver::~owOpenCLSolver, owOpenCLSolver.cpp:744-748. */
void sph_destroy(sph_ctx *ctx);

/* One reference stage, synchronous on return (QUEUE_EACH_KERNEL 1,
 * owOpenCLConstant.h:16): owOpenCLSolver::_run*, owOpenCLSolver.cpp:213-687. */
int sph_stage(sph_ctx *ctx, int stage_id);

/* The whole device part of owPhysicsFluidSimulator::simulationStep
 * (owPhysicsFluidSimulator.cpp:88-113): neighbour search, PCISPH loop
 * (max_iteration fixed iterations), integrate, membranes — fused kernels,
 * asynchronous on the context's stream. Results identical to running the
 * 18 stages in order. */
int sph_step(sph_ctx *ctx);

/* Block until all work queued on the context has finished (queue.finish()). */
int sph_sync(sph_ctx *ctx);

/* owOpenCLSolver::updateMuscleActivityData, owOpenCLSolver.cpp:738-742. */
int sph_set_muscles(sph_ctx *ctx, const float *signal, int n);

/* owOpenCLSolver::read_position_buffer / read_density_buffer /
 * read_particleIndex_buffer (owOpenCLSolver.h:60-62) generalised to every
 * buffer; blocking copy of the first `bytes` bytes (bytes <= buffer size). */
int sph_read(sph_ctx *ctx, int buffer_id, void *dst, size_t bytes);

/* copy_buffer_to_device (owOpenCLSolver.cpp:717-726) for the buffers a caller
 * may legitimately overwrite: POSITION, VELOCITY (first N float4) — used to
 * restart from a given state. */
int sph_write(sph_ctx *ctx, int buffer_id, const void *src, size_t bytes);

/* Size in bytes of a buffer in the reference layout, 0 if absent. */
size_t sph_buffer_bytes(const sph_ctx *ctx, int buffer_id);

/* Copy the parameters the context was created with. */
int sph_get_params(const sph_ctx *ctx, sph_params *out);

/* Text of the last error on this thread ("" if none). */
const char *sph_last_error(void);

/* "cuda-sm_100a" | "oracle-c" | "reference-cl-as-c++". */
const char *sph_backend(void);

/* ---- instrumentation: no counterpart in the reference (its only timing is owHelper's
 * stopwatch, owHelper.cpp:1806-1841); used by bench.py and the C++ facade. CUDA library only. ---- */

/* Run `steps` fused steps and return the device time in ms, measured with CUDA events on the
 * context's own stream with a synchronize on both sides. */
int sph_run_steps(sph_ctx *ctx, int steps, float *elapsed_ms);
/* Device timer (CUDA events on the context's stream, synchronize on both sides) around work the caller queues itself. */
int sph_timer_start(sph_ctx *ctx);
int sph_timer_stop(sph_ctx *ctx, float *elapsed_ms);
/* Cost of a dependent chain of `nodes` EMPTY kernels replayed as a CUDA graph (the floor of a launch-bound step). */
int sph_empty_chain_ms(sph_ctx *ctx, int nodes, int blocks, int threads, int replays, float *ms_per_replay);
/* Per-kernel CUDA-event timing of sph_step on/off; report = JSON {"kernel": {"calls":n,"ms":t}}. */
int sph_profile_enable(sph_ctx *ctx, int on);
int sph_profile_report(sph_ctx *ctx, char *buf, size_t n);
/* Number of kernels sph_step has launched on this context. */
long long sph_launch_count(const sph_ctx *ctx);
/* Tuning counters of the neighbour search: {needy particles of the last search, its histogram-first mode, out-of-grid /
 * NaN particles so far, fallbacks to the plain scan so far}. Synchronises. */
int sph_debug_counters(sph_ctx *ctx, unsigned out[4]);
/* Particles found outside the hash grid so far (NaN / out-of-box positions; 0 in a healthy run). Synchronises. */
long long sph_out_of_grid_count(sph_ctx *ctx);
/* Page-locked host memory for read-back / upload buffers. */
void *sph_host_alloc(size_t bytes);
void sph_host_free(void *p);

/* ---- multi-device slab extension (CUDA library only). The reference is single-device
 * (owOpenCLSolver.cpp:157-182); SURVEY.md 8e / DESIGN.md 7 define this protocol. One context per rank;
 * the rank owns hash-cell z-layers [z0, z1) and keeps ghost copies of layers z0-1 and z1. The library
 * classifies, packs and unpacks on the device; the caller moves the bytes between ranks
 * (smoothed-particle-hydrodynamics_b200/slab.py: torch.distributed / NCCL). side: 0 = down (z0-1), 1 = up (z1). ---- */

/* Bytes of one particle record (position, velocity, global id) in migration / ghost buffers. */
int sph_slab_record_bytes(void);
/* Make a context created with particle_count = CAPACITY a slab rank; its first n_owned particles are owned,
 * gid[] are their global ids (= original indices of the single-device run). Needs hash_mask 0xffffffff. */
int sph_slab_begin(sph_ctx *ctx, int n_owned, const uint32_t *gid, int z0, int z1, int has_down, int has_up,
                   int mig_cap, int halo_cap);
/* Phase 1 of a step: owned particles -> keep / migrate down / migrate up, plus copies of the kept boundary
 * layers. The four DEVICE buffers hold mig_cap / halo_cap records. counts = {mig_down, ghost_down, mig_up,
 * ghost_up}. Synchronous (the counts size the transfers). */
int sph_slab_classify(sph_ctx *ctx, void *mig_down, void *ghost_down, void *mig_up, void *ghost_up, int counts[4]);
/* Phase 2: the four received DEVICE buffers and their record counts {mig_from_down, ghost_from_down,
 * mig_from_up, ghost_from_up}; rebuilds the local particle set (owned + ghosts). Asynchronous. */
int sph_slab_absorb(sph_ctx *ctx, const void *mig_from_down, const void *ghost_from_down, const void *mig_from_up,
                    const void *ghost_from_up, const int incoming[4]);
/* out = {n_owned, n_local, n_send_down, n_send_up, n_recv_down, n_recv_up} (halo element counts). */
int sph_slab_info(const sph_ctx *ctx, int out[6]);
/* Halo refresh of field 0 = rho(t) (4 B), 1 = (pressure, rho*) (8 B), 2 = predicted position (16 B): gather the
 * values a neighbour's ghosts need into a DEVICE buffer / scatter a received buffer into the local ghosts. */
int sph_slab_pack(sph_ctx *ctx, int field, int side, void *buf);
int sph_slab_unpack(sph_ctx *ctx, int field, int side, const void *buf);
/* Owned particles (any order) and their global ids, to host memory. Synchronous. */
int sph_slab_read_owned(sph_ctx *ctx, float *pos4, float *vel4, uint32_t *gid);
/* Load balancing: move the slab's faces by at most one cell layer each; the layer that falls outside travels to the
 * neighbour as ordinary migrants in the next sph_slab_classify. Both ranks sharing a face call it before the same step. */
int sph_slab_set_cut(sph_ctx *ctx, int z0, int z1);
/* Pipelined form for the owned positions (pinned pos4): overlaps the next step; complete it with sph_read_end. */
int sph_slab_read_owned_begin(sph_ctx *ctx, float *pos4);
/* The same for packed (x, y, z) — 12 bytes per particle: the type tag never changes. Local order; it changes only in
 * steps that migrate particles (sph_slab_counts), sph_slab_read_owned(ctx, NULL, NULL, gid) gives the ids then. */
int sph_slab_read_owned_xyz_begin(sph_ctx *ctx, float *xyz3);
const char *sph_slab_last_error(void);
/* ---- links: halo exchange through PEER MEMORY (NVLink / NVSwitch P2P; north_star: "ghost-particle halos exchanged
 * by NCCL (or P2P) over NVLink each step"). Every rank owns a mailbox in device memory that its two neighbours map
 * (CUDA IPC between processes, peer access inside one process); their kernels store migrants, ghost layers, halo
 * values and delivery flags straight into it, and the consuming kernels wait on the flags. With links set up,
 * sph_slab_step runs a whole step — classify, exchange, absorb, the five parts and the six halo refreshes — with ONE
 * host synchronisation (the particle counts after migration size the launches) and no call into NCCL or the host
 * on the data path. ---- */
int sph_slab_link_handle_bytes(void);
/* Create this rank's mailbox; handle (may be NULL) receives its CUDA IPC handle for neighbours in other processes. */
int sph_slab_link_export(sph_ctx *ctx, void *handle);
/* Map the mailbox of the neighbour on `side` (0 = rank-1, 1 = rank+1): exported by another process / living in this one. */
int sph_slab_link_open(sph_ctx *ctx, int side, const void *handle);
int sph_slab_link_local(sph_ctx *ctx, int side, sph_ctx *peer);
/* One step. A process that drives several ranks calls _begin on all of them, then _end on all of them (_end waits for
 * the neighbours' deliveries); a process with one rank calls sph_slab_step. */
int sph_slab_step_begin(sph_ctx *ctx);
int sph_slab_step_end(sph_ctx *ctx);
int sph_slab_step(sph_ctx *ctx);
/* {migrants down, ghosts down, migrants up, ghosts up} sent in the last step. */
int sph_slab_counts(const sph_ctx *ctx, int out[4]);
/* The fused step in the five pieces between which a slab rank refreshes its ghosts:
 * 0 | rho | 1 | (p,rho*) | 2 | x* | 3 | (p,rho*) | 2 | x* | 3 | (p,rho*) | 4   for max_iteration = 3. */
int sph_step_part(sph_ctx *ctx, int part);
/* Run the context on a caller-owned CUDA stream (cudaStream_t), NULL = its own stream again. */
int sph_set_stream(sph_ctx *ctx, void *cuda_stream);
/* Pipelined read-back of SPH_BUF_POSITION or SPH_BUF_VELOCITY (owOpenCLSolver::read_position_buffer,
 * owOpenCLSolver.h:60, called at owPhysicsFluidSimulator.cpp:115, without its blocking wait): the copy to `dst` (pinned host memory) starts when the
 * work queued so far is done and overlaps the next sph_step, whose integration kernel waits for it.
 * sph_read_end blocks until the bytes have arrived. One read in flight per context. */
int sph_read_begin(sph_ctx *ctx, int buffer_id, void *dst, size_t bytes);
int sph_read_end(sph_ctx *ctx);
/* For a caller that runs one step ahead (the drop-in owPhysicsFluidSimulator queues step k+1 before
 * simulationStep k returns): sph_snapshot copies rho(t) and particleIndex (and, if asked, the first N rows of
 * velocity) device-to-device in stream order; sph_read_snapshot(SPH_BUF_RHO | SPH_BUF_PARTICLE_INDEX |
 * SPH_BUF_VELOCITY) is read_density_buffer / read_particleIndex_buffer
 * (owOpenCLSolver.h:61-62) on that copy: blocking, but it does not wait for work queued after the snapshot. */
int sph_snapshot(sph_ctx *ctx, int with_velocity);
int sph_read_snapshot(sph_ctx *ctx, int buffer_id, void *dst, size_t bytes);
/* Raw device memory for exchange buffers; synchronous copy between any host / device pointers. */
void *sph_device_alloc(size_t bytes);
void sph_device_free(void *p);
int sph_device_copy(void *dst, const void *src, size_t bytes);

/* ---- allocation-free entry point. NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * sph_cell_start_noalloc computes gridCellIndexFixedUp — for every hash cell the number of particles in cells with a
 * smaller id, i.e. what SPH_STAGE_HASH_PARTICLES + SPH_STAGE_SORT leave in SPH_BUF_GRID_CELL_INDEX_FIXED_UP — from
 * DEVICE positions into a DEVICE output, asynchronously on `cuda_stream` (NULL = the default stream) of the CURRENT
 * device. It allocates nothing: the cell histogram lives in a static device array of SPH_NOALLOC_MAX_CELLS words
 * (4 MiB held by every GPU that loads the library, until exit); grid_cell_count + 1 beyond that is rejected with
 * SPH_E_SIZE. All calls share the array: it is cleared at the start of each call, and calls that overlap on the
 * GPU (different streams) race. pos4_device: n float4 (x, y, z, type); cell_start_device: grid_cell_count + 1 uint32. */
#define SPH_NOALLOC_MAX_CELLS (1 << 20)
int sph_cell_start_noalloc(const sph_params *p, const float *pos4_device, int n, uint32_t *cell_start_device,
                           void *cuda_stream);
int sph_noalloc_max_cells(void);
const char *sph_noalloc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* SPH_CAPI_H */
