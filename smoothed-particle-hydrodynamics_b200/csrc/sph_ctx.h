// sph_ctx.h — host-side context of libsph_b200.so and the launcher prototypes
// implemented in the .cu files. Buffer names follow owOpenCLSolver.h:72-97.
#pragma once
#include "sph_device.cuh"

struct SlabState;

struct sph_ctx {
  sph_params p;
  DevParams d;
  int device;
  cudaStream_t stream;
  // ---- buffers kept in the reference layout ----
  float4 *position;            // 2N  [0,N) x,y,z,type original order; [N,2N) membrane displacement
  float4 *velocity;            // 2N
  uint2 *particleIndex;        // N   (cellId, serialId), sorted after SORT
  unsigned *particleIndexBack; // N
  float4 *sortedPosition;      // 2N  [N,2N) predicted positions
  float4 *sortedVelocity;      // N
  unsigned *gridCellIndex;     // G+1 (only materialised by the INDEXX stage)
  unsigned *cellStart;         // G+1 == gridCellIndexFixedUp (exclusive scan of the cell histogram)
  float *rho;                  // 2N
  float *pressure;             // N
  float4 *acceleration;        // 2N
  float4 *elastic;             // 32*Ne
  int *membraneData;           // 3*M
  int *pml;                    // 7*Ne
  float *muscle;               // muscle_count
  // ---- internal ----
  int *nbIds;                  // Npad*32 int32 neighbour ids, 32x32 tiles (see nb_index)
  unsigned *cellCount;         // G+1 (+pad) histogram, reused as scatter cursor
  unsigned *key;               // N cell id per original particle
  unsigned *tmpSerial;         // N counting-sort scatter target
  float *sortedType;           // N type tag (position.w) in sorted order
  float4 *spT;                 // N sorted (x, y, z, type tag): the copy the force kernels gather
  float4 *recA, *recB;         // gather records of the fused path (sph_device.cuh): recA 2N (32 bytes per particle), recB N (16 bytes)
  unsigned *bMask;             // N words: bit k = neighbour slot k holds a boundary particle (written by the forces kernel)
  unsigned *cellList;          // first non-empty cell of every home block (cell pair) of this step (order arbitrary)
  unsigned *cellListCount;
  unsigned *cellCursor;        // [0] work cursor of the tiled neighbour search, [1] length of fnDeferred
  int *fnDeferred;             // N: particles whose neighbourhood does not fit the tile tables (generic scan)
  int numSMs;
  bool force_generic_neighbors; // SPH_NEIGHBOR_PATH=generic (or SPH_GENERIC_NEIGHBORS=1): one thread per particle
  // auto mode: the tile kernel's deferred count of an earlier step (read back without a sync) decides whether the
  // fluid suits the tiles; above 10 % deferred particles, or below 64 particles per home block, the plain scan runs
  // for the next 64 steps, then the tiles retry
  unsigned *hostDeferred;       // pinned, 2 words: deferred particles, home blocks
  cudaEvent_t evDeferred;
  bool deferredPending;
  int genericHold;
  int nbMode;                   // 0 = not decided yet, 1 = tiles, 2 = plain scan (with a probing tile launch every 64 steps)
  bool force_list_neighbors;    // SPH_NEIGHBOR_PATH=list: single-sweep hit-list kernel
  bool force_coop_neighbors;    // SPH_NEIGHBOR_PATH=coop: eight lanes per particle (default for small scenes)
  bool auto_tiles;               // SPH_NEIGHBOR_PATH=auto: the round-1 policy (tiles for large fluids, watched by the statistics)
  int deep_gather;              // SPH_DEEP_GATHER=0|1 forces / forbids the latency-optimised gather kernels; -1 = by size
  int list_variant;             // SPH_LIST_VARIANT: sweep variant of the list kernel (experiments)
  float *soa;                   // 3 (Npad + 32): sweep copy of the sorted positions, groups of 8 as x[8] | y[8] | z[8] (k_soa_positions)
  bool force_tiled_neighbors;   // SPH_NEIGHBOR_PATH=tiles: the tile kernel whatever the particle count (grid permitting)
  unsigned long long *scanState;  // decoupled look-back tile states
  unsigned *scanTicket;
  int scanTiles;
  unsigned epoch;              // step counter for the scan state (no per-step reset needed)
  float2 *nmScratch;           // 32N float2, lazily allocated for sph_read(NEIGHBOR_MAP)
  float4 *readScratch;
  unsigned *elasticMask;       // N words, scenes with membranes only: bit k = neighbour slot k holds an elastic particle
  unsigned hist_pct;           // SPH_HIST_PCT (bits 0-7) and SPH_HIST2_PCT (bits 8-15), see k_hash_count
  unsigned *nsStats;           // [0] particles that needed the radial histogram this step, [1] this step's mode (k_hash_count)
  int *memList, *memCount;     // scenes with membranes: sorted ids of the liquid particles with an elastic neighbour (this step)
  bool has_elastic, has_membranes, has_pml;
  bool neighbors_valid;
  bool dispDirty;              // the displacement half of `position` may hold non-zero values (fresh upload)
  void *profiler;              // Profiler* when per-kernel event timing is on
  SlabState *slab;             // non-null: this context is one slab rank of a multi-device fluid (sph_slab.cu)
  unsigned *rankKey;           // in-cell ranking key per original index (global ids of a slab rank); null = the index itself
  unsigned *tmpRankKey;        // slab ranks: rankKey in scatter order (written next to tmpSerial), N words
  cudaStream_t ownStream;      // the stream created with the context (c->stream may be replaced by sph_set_stream)
  long long launches;          // kernels launched by sph_step so far
  // CUDA graph of the fused step (small scenes are launch-bound: 13 dependent kernels of a few microseconds).
  // Captured once per (context, stream); SPH_GRAPH=0|1 forbids / forces it, default: scenes below SPH_GRAPH_MAX_N.
  cudaGraphExec_t stepGraph[2]; // the two halves of the step (see step_body)
  cudaStream_t graphStream;    // the stream stepGraph was captured on
  int graphNodes[2];           // kernels in each captured half
  int graph_mode;              // -1 = by size, 0 = never, 1 = always
  bool capturing;              // inside stream capture: launchers must not touch events of other streams
  int pdl_mode;                // SPH_PDL: 1 = on, otherwise off (measured: no gain over the graph replay)
  bool graphPdl;               // the captured graphs were recorded with programmatic launches
  bool pdl;                    // programmatic dependent launch between the kernels of the step (SPH_PDL=1; default off)
  bool cellCountClean;         // the cell histogram is all zero (k_rank_reorder left it so): no memset before the next hash
  int lanes;                   // SPH_LANES: lanes per particle of the small-scene gather kernels (0 = by size)
  // pipelined read-back (sph_read_begin / sph_read_end): a device-to-host copy of position or velocity that
  // runs on its own stream while the next step computes; the integration kernel waits for it (read_fence)
  cudaStream_t copyStream;
  cudaEvent_t evReadReady, evReadDone;
  int readPending;
  const void *readSrc;         // device buffer the pending read copies from
  bool readStaged;             // the pending read copies from a staging buffer: writers of position / velocity need not wait
  float *snapRho;              // sph_snapshot: rho(t) and particleIndex of the last completed step
  uint2 *snapIndex;
  float4 *snapVel;             // only when the caller asked for it
  cudaEvent_t evSnap;
  cudaEvent_t evTimer[2];      // sph_timer_start / sph_timer_stop
  // muscle signal upload: staged through a small pinned ring so that sph_set_muscles neither waits for the
  // step in flight nor depends on the caller's buffer after it returns
  float *sigRing;
  cudaEvent_t sigDone[4];
  unsigned sigNext;
};

// Kernel launch of the fused step: with c->pdl the launch carries the programmatic-stream-serialization attribute
// (see pdl_entry, sph_device.cuh), otherwise it is a plain launch.
template <typename... KArgs, typename... Args>
static inline void launch_k(sph_ctx *c, void (*kernel)(KArgs...), dim3 grid, dim3 block, Args &&...args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = c->stream;
  cudaLaunchAttribute attr[1];
  if (c->pdl) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// order the next writer of position / velocity after a pending pipelined read of them
static inline void read_fence(sph_ctx *c) {
  if (c->readPending && !c->capturing && !c->readStaged) cudaStreamWaitEvent(c->stream, c->evReadDone, 0);
}

// per-kernel event timing (sph_profile_enable) for the slab kernels; defined in sph_capi.cu
enum { PK_SLAB_RECORDS = 9, PK_SLAB_ABSORB = 10, PK_SLAB_HALO = 11 };
struct ProfMark { int kind; cudaEvent_t a, b; bool on; };
ProfMark prof_mark_begin(sph_ctx *c, int kind);
void prof_mark_end(sph_ctx *c, ProfMark &m);
// sph_sort.cu
void launch_hash_count(sph_ctx *c, bool write_pairs);
void launch_sort(sph_ctx *c, bool fuse_reorder);   // scan + scatter + rank (+ reorder)
void launch_reorder(sph_ctx *c);                   // sortPostPass on the sorted particleIndex
void launch_indexx_compat(sph_ctx *c);
// sph_neighbors.cu
void launch_clear_neighbors(sph_ctx *c);
void launch_find_neighbors(sph_ctx *c, bool with_density);
void launch_materialize_neighbor_map(sph_ctx *c, float2 *out);
// sph_pcisph.cu
void launch_compute_density(sph_ctx *c);
void launch_forces(sph_ctx *c, bool fuse_predict);
void launch_elastic(sph_ctx *c);
void launch_predict_positions(sph_ctx *c);
void launch_predict_density(sph_ctx *c, bool fuse_correct);
void launch_correct_pressure(sph_ctx *c);
void launch_pressure_force(sph_ctx *c, int mode);  // 0 = force only, 1 = + predictPositions, 2 = + integrate
void launch_integrate(sph_ctx *c);
// sph_lanes.cu: several lanes per particle (small scenes); false = no instantiation for this L
bool launch_forces_lanes(sph_ctx *c, int L);
bool launch_predict_density_lanes(sph_ctx *c, int L);
bool launch_pressure_force_lanes(sph_ctx *c, int mode, int L);
// sph_slab.cu
void slab_destroy(sph_ctx *c);
// sph_membranes.cu
void launch_clear_membrane_buffers(sph_ctx *c);
void launch_membrane_interaction(sph_ctx *c, bool fuse_finalize);
void launch_membrane_finalize(sph_ctx *c);
