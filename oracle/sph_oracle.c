/*
 * sph_oracle.c — TEST INFRASTRUCTURE. CPU restatement (plain C99) of the reference's
 * PCISPH step. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it; the product never does.
 *
 * Parity status: PINNED. The reference ships no golden vectors, so this
 * restatement is pinned against the reference's own kernels executed here
 * (oracle/_ref/libsphref.so = /root/reference/src/sphFluid.cl compiled as C++): every
 * buffer after every stage is bit-identical on configs A (pure liquid), B (worm) and
 * the synthetic blocks (tests/test_oracle.py), and against the committed digests in
 * tests/golden/ generated from that library (tests/golden/make_golden.py).
 *
 * Every function cites the reference file:line it follows. Float expressions keep the
 * reference's association order; build with -O2 -ffp-contract=off (the reference
 * builds its kernels with -cl-opt-disable, owOpenCLSolver.cpp:196-201). Where OpenCL
 * leaves latitude the definitions are those of oracle/ref_shim/cl_shim.hpp:
 * sqrt/native_sqrt correctly rounded, IEEE division, dot() = ((xx+yy)+zz)+ww.
 *
 * Exports the ABI of include/sph_capi.h.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "sph_capi.h"

#define MAXN SPH_MAX_NEIGHBOR_COUNT
#define NO_PARTICLE_ID (-1)
#define RADIUS_SEGMENTS 30 /* sphFluid.cl:116 */

typedef struct { float x, y, z, w; } f4;
typedef struct { float x, y; } f2;
typedef struct { uint32_t x, y; } u2;

static inline f4 f4mk(float x, float y, float z, float w) { f4 r = {x, y, z, w}; return r; }
static inline f4 f4add(f4 a, f4 b) { return f4mk(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
static inline f4 f4sub(f4 a, f4 b) { return f4mk(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
static inline f4 f4mul(f4 a, float s) { return f4mk(a.x * s, a.y * s, a.z * s, a.w * s); }
static inline f4 f4lmul(float s, f4 a) { return f4mk(s * a.x, s * a.y, s * a.z, s * a.w); }
static inline f4 f4div(f4 a, float s) { return f4mk(a.x / s, a.y / s, a.z / s, a.w / s); }
static inline f4 f4neg(f4 a) { return f4mk(-a.x, -a.y, -a.z, -a.w); }
static inline float f4dot(f4 a, f4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }

static __thread char g_err[256];
static int fail(int code, const char *msg) { snprintf(g_err, sizeof g_err, "%s", msg); return code; }

struct sph_ctx {
  sph_params p;
  int N, G;
  f4 *acceleration, *position, *sortedPosition, *sortedVelocity, *velocity, *elastic;
  uint32_t *gridCellIndex, *gridCellIndexFixedUp, *particleIndexBack;
  f2 *neighborMap;
  u2 *particleIndex;
  float *pressure, *rho, *muscle;
  int32_t *membraneData, *pml;
  int has_elastic, has_membranes, has_pml;
  /* hoisted constants (computed once on the host with the reference's expressions) */
  float hScaled, hScaled2, hScaled6;
  double mass_wpoly6;   /* ((double)mass)*Wpoly6Coefficient, sphFluid.cl:516,1057 */
  float surf_coeff;     /* sphFluid.cl:662 scalar factor */
  float mass_mu;        /* (float)(mass*mu), sphFluid.cl:688 */
  double mass_gradw;    /* ((double)mass)*gradWspikyCoefficient, sphFluid.cl:1194 */
};

static int g_threads = 1;
void sph_oracle_set_threads(int n) { g_threads = n > 0 ? n : 1; }

const char *sph_backend(void) { return "oracle-c"; }
const char *sph_last_error(void) { return g_err; }

/* ---------------- constants: owPhysicsConstant.h:12-76 ---------------- */

/* calcDelta, owPhysicsFluidSimulator.cpp:164-203. In the reference (C++) pow(float,float)
 * is powf, sqrt(float) is sqrtf and pow(float,int) is the double pow, whose square is exact. */
static float oracle_calc_delta(float mass, float rho0, float h, float simulationScale, double beta, double gradW) {
  static const float x[32] = {1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 2, -2, 0, 0, 0, 0, 0, 0};
  static const float y[32] = {0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 1, 1, 1, 0, -1, -1, -1, 0, 0, 2, -2, 0, 0, 0, 0};
  static const float z[32] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 0, 0, 0, 2, -2, 1, -1};
  float sum1_x = 0.f, sum1_y = 0.f, sum1_z = 0.f;
  double sum1, sum2 = 0.0;
  float particleRadius = powf(mass / rho0, 1.f / 3.f);
  for (int i = 0; i < 32; i++) {
    float v_x = x[i] * 0.8f * particleRadius;
    float v_y = y[i] * 0.8f * particleRadius;
    float v_z = z[i] * 0.8f * particleRadius;
    float dist = sqrtf(v_x * v_x + v_y * v_y + v_z * v_z);
    if (dist <= h * simulationScale) {
      float t = h * simulationScale - dist;
      float h_r_2 = (float)((double)t * (double)t);
      sum1_x += h_r_2 * v_x / dist;
      sum1_y += h_r_2 * v_y / dist;
      sum1_z += h_r_2 * v_z / dist;
      sum2 += h_r_2 * h_r_2;
    }
  }
  sum1 = sum1_x * sum1_x + sum1_y * sum1_y + sum1_z * sum1_z;
  return (float)(1.0 / (beta * gradW * gradW * (sum1 + sum2)));
}

int sph_params_set_box(sph_params *p, double mx, double my, double mz) {
  if (!p) return fail(SPH_E_INVALID, "null params");
  /* `#define XMAX 30.0*h` (owPhysicsConstant.h:33): double literal times float h */
  double xmaxd = mx * p->h, ymaxd = my * p->h, zmaxd = mz * p->h;
  p->xmin = 0; p->ymin = 0; p->zmin = 0;
  p->xmax = (float)xmaxd; p->ymax = (float)ymaxd; p->zmax = (float)zmaxd; /* owOpenCLSolver.cpp:7-12 */
  p->grid_cells_x = (int)((xmaxd - 0) / p->h) + 1;                        /* owOpenCLSolver.cpp:14-17 */
  p->grid_cells_y = (int)((ymaxd - 0) / p->h) + 1;
  p->grid_cells_z = (int)((zmaxd - 0) / p->h) + 1;
  p->grid_cell_count = p->grid_cells_x * p->grid_cells_y * p->grid_cells_z;
  return 0;
}

int sph_params_default(sph_params *p) {
  if (!p) return fail(SPH_E_INVALID, "null params");
  memset(p, 0, sizeof *p);
  const float rho0 = 1000.0f, mass = 3.25e-14f, timeStep = 5.0e-06f;                /* :12-15 */
  const float simulationScale = 0.004f * powf(mass, 1.f / 3.f) / powf(0.00025f, 1.f / 3.f); /* :19 */
  const float h = 3.34f;                                                            /* :21 */
  p->muscle_count = 100; /* owWorldSimulation.cpp:31 */
  p->max_iteration = 3;  /* :76 */
  p->hash_mask = 0xffffu;
  p->h = h;
  p->hash_grid_cell_size = 2.0f * h;                         /* :22 */
  p->hash_grid_cell_size_inv = 1.0f / p->hash_grid_cell_size; /* :23 */
  p->simulation_scale = simulationScale;
  p->simulation_scale_inv = 1.0f / simulationScale;           /* :24 */
  p->r0 = 0.5f * h;                                           /* :27 */
  p->rho0 = rho0; p->mass = mass; p->time_step = timeStep;
  p->stiffness = 0.75f; p->viscosity = 0.00005f; p->damping = 0.75f; /* :62-64 */
  p->gravity_x = 0.0f; p->gravity_y = -9.8f; p->gravity_z = 0.0f;     /* :72-74 */
  /* :68 — the whole product is evaluated in float (denormal result), then widened */
  const double beta = timeStep * timeStep * mass * mass * 2 / (rho0 * rho0);
  p->wpoly6_coefficient = 315.0 / (64.0 * M_PI * pow((double)(h * simulationScale), 9.0)); /* :69 */
  p->grad_wspiky_coefficient = -45.0 / (M_PI * pow((double)(h * simulationScale), 6.0));   /* :70 */
  p->del2_wviscosity_coefficient = -p->grad_wspiky_coefficient;                            /* :71 */
  p->delta = oracle_calc_delta(mass, rho0, h, simulationScale, beta, p->grad_wspiky_coefficient);
  return sph_params_set_box(p, 30.0, 20.0, 250.0); /* :32-37 */
}

/* ---------------- context: owOpenCLSolver.cpp:27-92 ---------------- */

static void *xcalloc(size_t n, size_t sz) { return calloc(n ? n : 1, sz); }

sph_ctx *sph_create(const sph_params *p, const float *pos4, const float *vel4, const float *elastic4,
                    const int32_t *membranes3, const int32_t *pml7, int device) {
  (void)device;
  if (!p || !pos4 || !vel4 || p->particle_count <= 0 || p->grid_cell_count <= 0) {
    fail(SPH_E_INVALID, "sph_create: bad arguments");
    return NULL;
  }
  sph_ctx *c = (sph_ctx *)calloc(1, sizeof *c);
  c->p = *p;
  size_t N = c->N = p->particle_count, G = c->G = p->grid_cell_count;
  c->acceleration = xcalloc(2 * N, sizeof(f4));
  c->gridCellIndex = xcalloc(G + 1, 4);
  c->gridCellIndexFixedUp = xcalloc(G + 1, 4);
  c->neighborMap = xcalloc(MAXN * N, sizeof(f2));
  c->particleIndex = xcalloc(N, sizeof(u2));
  c->particleIndexBack = xcalloc(N, 4);
  c->position = xcalloc(2 * N, sizeof(f4));
  c->pressure = xcalloc(N, 4);
  c->rho = xcalloc(2 * N, 4);
  c->sortedPosition = xcalloc(2 * N, sizeof(f4));
  c->sortedVelocity = xcalloc(N, sizeof(f4));
  c->velocity = xcalloc(2 * N, sizeof(f4));
  c->muscle = xcalloc(p->muscle_count > 0 ? p->muscle_count : 1, 4);
  memcpy(c->position, pos4, 16 * N);
  memcpy(c->velocity, vel4, 16 * N);
  c->has_membranes = membranes3 && p->num_membranes > 0;
  c->has_elastic = elastic4 && p->num_elastic > 0;
  c->has_pml = c->has_membranes && c->has_elastic && pml7;
  if (c->has_membranes) {
    c->membraneData = xcalloc(3 * (size_t)p->num_membranes, 4);
    memcpy(c->membraneData, membranes3, 12 * (size_t)p->num_membranes);
  }
  if (c->has_pml) {
    c->pml = xcalloc(7 * (size_t)p->num_elastic, 4);
    memcpy(c->pml, pml7, 28 * (size_t)p->num_elastic);
  }
  if (c->has_elastic) {
    c->elastic = xcalloc(MAXN * (size_t)p->num_elastic, sizeof(f4));
    memcpy(c->elastic, elastic4, 16 * MAXN * (size_t)p->num_elastic);
  }
  /* hoisted sub-expressions */
  c->hScaled = p->h * p->simulation_scale;                   /* sphFluid.cl:495 */
  c->hScaled2 = c->hScaled * c->hScaled;                     /* :496 */
  c->hScaled6 = c->hScaled2 * c->hScaled2 * c->hScaled2;     /* :497 */
  c->mass_wpoly6 = ((double)p->mass) * p->wpoly6_coefficient;
  c->surf_coeff = (-1.5e-09f * 0.3f * (float)(p->wpoly6_coefficient * pow(c->hScaled2 / 2.0, 3.0)) * p->simulation_scale);
  c->mass_mu = (float)(p->mass * p->viscosity);
  c->mass_gradw = ((double)p->mass) * p->grad_wspiky_coefficient;
  return c;
}

void sph_destroy(sph_ctx *c) {
  if (!c) return;
  free(c->acceleration); free(c->gridCellIndex); free(c->gridCellIndexFixedUp); free(c->neighborMap);
  free(c->particleIndex); free(c->particleIndexBack); free(c->position); free(c->pressure); free(c->rho);
  free(c->sortedPosition); free(c->sortedVelocity); free(c->velocity); free(c->muscle);
  free(c->membraneData); free(c->pml); free(c->elastic);
  free(c);
}

int sph_get_params(const sph_ctx *c, sph_params *out) {
  if (!c || !out) return fail(SPH_E_INVALID, "null");
  *out = c->p;
  return 0;
}

/* ---------------- neighbour search ---------------- */

/* clearBuffers, sphFluid.cl:64-92 */
static void st_clear_buffers(sph_ctx *c) {
  size_t n = (size_t)MAXN * c->N;
  for (size_t i = 0; i < n; i++) { c->neighborMap[i].x = -1.f; c->neighborMap[i].y = -1.f; }
}

/* hashParticles + cellFactors + cellId, sphFluid.cl:346-383, :187-201, :332-342.
 * xmin/ymin/zmin are accepted by the reference but never subtracted. */
static void st_hash_particles(sph_ctx *c) {
  const sph_params *p = &c->p;
  for (int id = 0; id < c->N; id++) {
    f4 q = c->position[id];
    int cx = (int)(q.x * p->hash_grid_cell_size_inv);
    int cy = (int)(q.y * p->hash_grid_cell_size_inv);
    int cz = (int)(q.z * p->hash_grid_cell_size_inv);
    int cell = (cx + cy * p->grid_cells_x + cz * p->grid_cells_x * p->grid_cells_y) & p->hash_mask;
    c->particleIndex[id].x = (uint32_t)cell;
    c->particleIndex[id].y = (uint32_t)id;
  }
}

/* _runSort + myCompare, owOpenCLSolver.cpp:255-261, :690-696: qsort on the (signed) cell id only.
 * glibc's qsort is a stable merge sort, so ties keep ascending serial id; the oracle makes that
 * canonical order explicit by comparing (cell, serial). */
static int cmp_cell_serial(const void *a, const void *b) {
  const int32_t *f1 = (const int32_t *)a, *f2 = (const int32_t *)b;
  if (f1[0] < f2[0]) return -1;
  if (f1[0] > f2[0]) return +1;
  if (f1[1] < f2[1]) return -1;
  if (f1[1] > f2[1]) return +1;
  return 0;
}
static void st_sort(sph_ctx *c) { qsort(c->particleIndex, c->N, sizeof(u2), cmp_cell_serial); }

/* sortPostPass, sphFluid.cl:441-466 */
static void st_sort_post_pass(sph_ctx *c) {
  for (int id = 0; id < c->N; id++) {
    u2 spi = c->particleIndex[id];
    int serialId = (int)spi.y, cellId = (int)spi.x;
    f4 q = c->position[serialId];
    q.w = (float)cellId;
    c->sortedVelocity[id] = c->velocity[serialId];
    c->sortedPosition[id] = q;
    c->particleIndexBack[serialId] = (uint32_t)id;
  }
}

/* indexx, sphFluid.cl:385-439: first sorted index holding each cell id (the binary search
 * converges on the first occurrence), -1 if none; [0] = 0 and [G] = N unconditionally. */
static void st_indexx(sph_ctx *c) {
  int G = c->G, N = c->N;
  for (int g = 0; g <= G; g++) c->gridCellIndex[g] = (uint32_t)NO_PARTICLE_ID;
  for (int i = N - 1; i >= 0; i--) {
    int cell = (int)c->particleIndex[i].x;
    if (cell >= 0 && cell < G) c->gridCellIndex[cell] = (uint32_t)i;
  }
  c->gridCellIndex[0] = 0;
  c->gridCellIndex[G] = (uint32_t)N;
}

/* _runIndexPostPass, owOpenCLSolver.cpp:305-319 */
static void st_index_post_pass(sph_ctx *c) {
  int G = c->G;
  memcpy(c->gridCellIndexFixedUp, c->gridCellIndex, 4 * ((size_t)G + 1));
  uint32_t recent = (uint32_t)G;
  for (int i = G; i >= 0; i--) {
    if (c->gridCellIndexFixedUp[i] == (uint32_t)-1) c->gridCellIndexFixedUp[i] = recent;
    else recent = c->gridCellIndexFixedUp[i];
  }
}

/* searchCell, sphFluid.cl:94-112 */
static int search_cell(int cellId, int dx, int dy, int dz, int gx, int gy, int G) {
  int n = cellId + dx + dy * gx + dz * gx * gy;
  if (n < 0) n = n + G;
  if (n >= G) n = n - G;
  return n;
}

/* searchForNeighbors, sphFluid.cl:118-184 */
static int search_for_neighbors(const sph_ctx *c, int cell, f4 pos, int myId, int spaceLeft, int mode,
                                int *radius_distrib, float r_thr) {
  const sph_params *p = &c->p;
  int base = (int)c->gridCellIndexFixedUp[cell];
  int next = (int)c->gridCellIndexFixedUp[cell + 1];
  int count = next - base, found = 0;
  float r_thr_sq = r_thr * r_thr;
  if (spaceLeft > 0)
    for (int i = 0; i < count; i++) {
      int nb = base + i;
      if (myId == nb) continue;
      f4 q = c->sortedPosition[nb];
      float dx = pos.x - q.x, dy = pos.y - q.y, dz = pos.z - q.z;
      float d2 = dx * dx + dy * dy + dz * dz;
      if (d2 <= r_thr_sq) {
        float d = sqrtf(d2);
        int j = (int)(d * RADIUS_SEGMENTS / p->h);
        if (j < RADIUS_SEGMENTS) radius_distrib[j]++;
        if (mode) {
          int off = MAXN - spaceLeft + found;
          if (off >= MAXN) break;
          f2 nd = {(float)nb, d * p->simulation_scale};
          c->neighborMap[(size_t)myId * MAXN + off] = nd;
          found++;
        }
      }
    }
  return found;
}

/* findNeighbors, sphFluid.cl:207-329 */
static void st_find_neighbors(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int gx = p->grid_cells_x, gy = p->grid_cells_y, G = c->G;
#pragma omp parallel for schedule(dynamic, 256) num_threads(g_threads)
  for (int id = 0; id < c->N; id++) {
    f4 pos = c->sortedPosition[id];
    int myCell = (int)pos.w & p->hash_mask;
    int found = 0, distrib_sum = 0, hist[RADIUS_SEGMENTS] = {0};
    float r_thr = p->h;
    /* which half of its 2h cell the particle is in, per axis: :253-271 */
    float px = pos.x - p->xmin, py = pos.y - p->ymin, pz = pos.z - p->zmin;
    float cfx = (int)(pos.x * p->hash_grid_cell_size_inv) * p->hash_grid_cell_size;
    float cfy = (int)(pos.y * p->hash_grid_cell_size_inv) * p->hash_grid_cell_size;
    float cfz = (int)(pos.z * p->hash_grid_cell_size_inv) * p->hash_grid_cell_size;
    int dx = 1 + 2 * ((px - cfx) < p->h ? -1 : 0);
    int dy = 1 + 2 * ((py - cfy) < p->h ? -1 : 0);
    int dz = 1 + 2 * ((pz - cfz) < p->h ? -1 : 0);
    const int off[8][3] = {{0, 0, 0}, {dx, 0, 0}, {0, dy, 0}, {0, 0, dz}, {dx, dy, 0}, {dx, 0, dz}, {0, dy, dz}, {dx, dy, dz}};
    for (int mode = 0; mode < 2; mode++) {
      for (int k = 0; k < 8; k++) {
        int cell = k == 0 ? myCell : search_cell(myCell, off[k][0], off[k][1], off[k][2], gx, gy, G);
        found += search_for_neighbors(c, cell, pos, id, MAXN - found, mode, hist, r_thr);
      }
      if (mode == 0) { /* :310-324 */
        int j = 0;
        while (j < RADIUS_SEGMENTS) {
          distrib_sum += hist[j];
          if (distrib_sum == MAXN) break;
          if (distrib_sum > MAXN) { j--; break; }
          j++;
        }
        r_thr = (j + 1) * p->h / RADIUS_SEGMENTS;
      }
    }
  }
}

/* ---------------- PCISPH ---------------- */

static inline int is_boundary(const sph_ctx *c, int sortedId) { /* (int)position[serial].w == 3 */
  return (int)c->position[c->particleIndex[sortedId].y].w == SPH_BOUNDARY_PARTICLE;
}

/* pcisph_computeDensity, sphFluid.cl:472-518 (no r<h filter; float cubes, double sum) */
static void st_compute_density(sph_ctx *c) {
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (int id = 0; id < c->N; id++) {
    const f2 *nm = c->neighborMap + (size_t)id * MAXN;
    double density = 0.0;
    for (int nc = 0; nc < MAXN; nc++)
      if (nm[nc].x != (float)NO_PARTICLE_ID) {
        float r2 = nm[nc].y;
        r2 *= r2;
        density += (c->hScaled2 - r2) * (c->hScaled2 - r2) * (c->hScaled2 - r2);
      }
    if (density < c->hScaled6) density = c->hScaled6;
    density *= c->mass_wpoly6;
    c->rho[id] = (float)density;
  }
}

/* pcisph_computeForcesAndInitPressure, sphFluid.cl:589-708 */
static void st_compute_forces(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int N = c->N;
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (int id = 0; id < N; id++) {
    if (is_boundary(c, id)) {
      c->acceleration[id] = f4mk(0, 0, 0, 0);
      c->acceleration[N + id] = f4mk(0, 0, 0, 0);
      c->pressure[id] = 0.f;
      continue;
    }
    const f2 *nm = c->neighborMap + (size_t)id * MAXN;
    f4 sum = f4mk(0, 0, 0, 0), surf = f4mk(0, 0, 0, 0);
    for (int nc = 0; nc < MAXN; nc++) {
      int jd = (int)nm[nc].x;
      if (jd != NO_PARTICLE_ID) {
        float r_ij = nm[nc].y;
        if (r_ij < c->hScaled) {
          sum = f4add(sum, f4div(f4mul(f4sub(c->sortedVelocity[jd], c->sortedVelocity[id]), c->hScaled - r_ij), c->rho[jd]));
          surf = f4add(surf, f4lmul(c->surf_coeff, f4sub(c->sortedPosition[id], c->sortedPosition[jd])));
        }
      }
    }
    surf.w = 0.f;
    sum = f4mul(sum, c->mass_mu * (float)(p->del2_wviscosity_coefficient / c->rho[id]));
    f4 a = sum;
    a = f4add(a, f4mk(p->gravity_x, p->gravity_y, p->gravity_z, 0.0f));
    a = f4add(a, surf);
    c->acceleration[id] = a;
    c->acceleration[N + id] = f4mk(0, 0, 0, 0);
    c->pressure[id] = 0.f;
  }
}

/* pcisph_computeElasticForces, sphFluid.cl:709-815. One work-item per elastic particle
 * (original indices [0,Ne) + offset); the dead centre-of-mass block (:787-800) is omitted. */
static void st_compute_elastic(sph_ctx *c) {
  const sph_params *p = &c->p;
  if (p->num_elastic == 0 || !c->has_elastic) return; /* owOpenCLSolver.cpp:422 */
  const float k = 600000000.f;
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (int index = 0; index < p->num_elastic; index++) {
    int id = (int)c->particleIndexBack[index + p->elastic_offset];
    const f4 *ec = c->elastic + (size_t)index * MAXN;
    for (int nc = 0; nc < MAXN; nc++) {
      int jd = (int)ec[nc].x;
      if (jd == NO_PARTICLE_ID) break;
      jd = (int)c->particleIndexBack[jd];
      float r0ij = ec[nc].y;
      f4 v = f4mul(f4sub(c->sortedPosition[id], c->sortedPosition[jd]), p->simulation_scale);
      v.w = 0;
      float r_ij = sqrtf(f4dot(v, v));
      float dr = r_ij - r0ij;
      if (r_ij != 0.f) {
        c->acceleration[id] = f4add(c->acceleration[id], f4mul(f4mul(f4neg(f4div(v, r_ij)), dr), k));
        for (int i = 0; i < p->muscle_count; i++)
          if ((int)(ec[nc].z) == (i + 1))
            if (c->muscle[i] > 0.f)
              c->acceleration[id] = f4add(c->acceleration[id], f4mul(f4mul(f4neg(f4div(v, r_ij)), c->muscle[i]), 800.f));
      }
    }
  }
}

/* pcisph_predictPositions, sphFluid.cl:889-979: the predictor uses the PRESSURE acceleration only (:924) */
static void st_predict_positions(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int N = c->N;
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (int id = 0; id < N; id++) {
    f4 pos = c->sortedPosition[id];
    if (is_boundary(c, id)) { c->sortedPosition[N + id] = pos; continue; }
    f4 a = c->acceleration[N + id];
    f4 v = c->sortedVelocity[id];
    f4 nv = f4add(v, f4lmul(p->time_step, a));
    float posTimeStep = p->time_step * p->simulation_scale_inv;
    c->sortedPosition[N + id] = f4add(pos, f4lmul(posTimeStep, nv));
  }
}

/* pcisph_predictDensity, sphFluid.cl:982-1059 */
static void st_predict_density(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int N = c->N;
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (int id = 0; id < N; id++) {
    const f2 *nm = c->neighborMap + (size_t)id * MAXN;
    double density = 0.0;
    for (int nc = 0; nc < MAXN; nc++) {
      int jd = (int)nm[nc].x;
      if (jd != NO_PARTICLE_ID) {
        f4 r = f4sub(c->sortedPosition[N + id], c->sortedPosition[N + jd]);
        float r2 = (r.x * r.x + r.y * r.y + r.z * r.z) * p->simulation_scale * p->simulation_scale;
        if (r2 < c->hScaled2) density += (c->hScaled2 - r2) * (c->hScaled2 - r2) * (c->hScaled2 - r2);
      }
    }
    if (density < c->hScaled6) density = c->hScaled6;
    density *= c->mass_wpoly6;
    c->rho[N + id] = (float)density;
  }
}

/* pcisph_correctPressure, sphFluid.cl:1062-1098 */
static void st_correct_pressure(sph_ctx *c) {
  const sph_params *p = &c->p;
  for (int id = 0; id < c->N; id++) {
    float rho_err = c->rho[c->N + id] - p->rho0;
    float p_corr = rho_err * p->delta;
    if (p_corr < 0) p_corr = 0;
    c->pressure[id] += p_corr;
  }
}

/* pcisph_computePressureForceAcceleration, sphFluid.cl:1101-1212 */
static void st_pressure_force(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int N = c->N;
  const float hS = c->hScaled;
#pragma omp parallel for schedule(static) num_threads(g_threads)
  for (int id = 0; id < N; id++) {
    if (is_boundary(c, id)) { c->acceleration[N + id] = f4mk(0, 0, 0, 0); continue; }
    const f2 *nm = c->neighborMap + (size_t)id * MAXN;
    f4 result = f4mk(0, 0, 0, 0);
    for (int nc = 0; nc < MAXN; nc++) {
      int jd = (int)nm[nc].x;
      if (jd != NO_PARTICLE_ID) {
        float r_ij = nm[nc].y;
        if (r_ij < hS) {
          float value = -(hS - r_ij) * (hS - r_ij) * 0.5f * (c->pressure[id] + c->pressure[jd]) / c->rho[N + jd];
          f4 vr = f4mul(f4sub(c->sortedPosition[id], c->sortedPosition[jd]), p->simulation_scale);
          vr.w = 0;
          if (r_ij < 0.5 * (hS / 2)) /* double compare, :1166 */
            value = -(hS * 0.25f - r_ij) * (hS * 0.25f - r_ij) * 0.5f * (p->rho0 * p->delta) / c->rho[N + jd];
          result = f4add(result, f4div(f4lmul(value, vr), r_ij));
        }
      }
    }
    result = f4mul(result, (float)(c->mass_gradw / ((double)c->rho[N + id])));
    c->acceleration[N + id] = result;
  }
}

/* computeInteractionWithBoundaryParticles, sphFluid.cl:824-887 */
static void boundary_interaction(const sph_ctx *c, int id, f4 *pos_, int tangVel, f4 *vel) {
  const float r0 = c->p.r0;
  const f2 *nm = c->neighborMap + (size_t)id * MAXN;
  f4 n_c_i = f4mk(0, 0, 0, 0);
  float w_sum = 0.f, w_second = 0.f;
  for (int nc = 0; nc < MAXN; nc++) {
    int id_b = (int)nm[nc].x;
    if (id_b != NO_PARTICLE_ID) {
      int src = (int)c->particleIndex[id_b].y;
      f4 pb = c->position[src];
      if ((int)pb.w == SPH_BOUNDARY_PARTICLE) {
        float d = (pos_->x - pb.x) * (pos_->x - pb.x);
        d += (pos_->y - pb.y) * (pos_->y - pb.y);
        d += (pos_->z - pb.z) * (pos_->z - pb.z);
        d = sqrtf(d);
        float w = (0.f < (r0 - d) / r0) ? (r0 - d) / r0 : 0.f; /* max(0.f,x) := (0.f<x)?x:0.f */
        n_c_i = f4add(n_c_i, f4mul(c->velocity[src], w)); /* boundary "velocity" is the wall normal, :860 */
        w_sum += w;
        w_second += w * (r0 - d);
      }
    }
  }
  float len = f4dot(n_c_i, n_c_i);
  if (len != 0) {
    len = sqrtf(len);
    f4 dp = f4div(f4mul(f4div(n_c_i, len), w_second), w_sum);
    pos_->x += dp.x; pos_->y += dp.y; pos_->z += dp.z;
    if (tangVel) {
      float eps = 0.99f;
      float vn = n_c_i.x * vel->x + n_c_i.y * vel->y + n_c_i.z * vel->z;
      if (vn < 0) {
        vel->x -= n_c_i.x * vn; vel->y -= n_c_i.y * vn; vel->z -= n_c_i.z * vn;
        *vel = f4mul(*vel, eps);
      }
    }
  }
}

/* pcisph_integrate, sphFluid.cl:1684-1808. Sequential over sorted ids here; the only
 * cross-particle reads are of boundary particles, which are never written. */
static void st_integrate(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int N = c->N;
  for (int id = 0; id < N; id++) {
    int src = (int)c->particleIndex[id].y;
    f4 pos = c->sortedPosition[id];
    if ((int)c->position[src].w == SPH_BOUNDARY_PARTICLE) continue;
    f4 a = f4add(c->acceleration[id], c->acceleration[N + id]);
    a.w = 0.f;
    f4 v = c->sortedVelocity[id];
    f4 nv = f4add(v, f4lmul(p->time_step, a));
    float posTimeStep = p->time_step * p->simulation_scale_inv;
    f4 np = f4add(pos, f4lmul(posTimeStep, nv));
    if (np.x < p->xmin) np.x = p->xmin;
    if (np.y < p->ymin) np.y = p->ymin;
    if (np.z < p->zmin) np.z = p->zmin;
    if (np.x > p->xmax - 0.000001f) np.x = p->xmax - 0.000001f;
    if (np.y > p->ymax - 0.000001f) np.y = p->ymax - 0.000001f;
    if (np.z > p->zmax - 0.000001f) np.z = p->zmax - 0.000001f;
    float particleType = c->position[src].w;
    nv = f4mul(f4add(v, nv), 0.5f);
    boundary_interaction(c, id, &np, 1, &nv);
    c->velocity[src] = nv;
    np.w = particleType;
    c->position[src] = np;
  }
}

/* clearMembraneBuffers, sphFluid.cl:1214-1227 */
static void st_clear_membrane_buffers(sph_ctx *c) {
  memset(c->position + c->N, 0, sizeof(f4) * (size_t)c->N);
  memset(c->velocity + c->N, 0, sizeof(f4) * (size_t)c->N);
}

/* calcDeterminant3x3, sphFluid.cl:1229-1247 */
static float det3(f4 c1, f4 c2, f4 c3) {
  return c1.x * c2.y * c3.z + c1.y * c2.z * c3.x + c1.z * c2.x * c3.y - c1.z * c2.y * c3.x - c1.x * c2.z * c3.y - c1.y * c2.x * c3.z;
}

/* calculateProjectionOfPointToPlane, sphFluid.cl:1250-1308 (Cramer's rule; pm.w = -1 on a zero determinant) */
static f4 project_point_to_plane(f4 ps, f4 pa, f4 pb, f4 pc) {
  f4 pm = f4mk(0, 0, 0, 0);
  float b_1 = pa.x * ((pb.y - pa.y) * (pc.z - pa.z) - (pb.z - pa.z) * (pc.y - pa.y)) +
              pa.y * ((pb.z - pa.z) * (pc.x - pa.x) - (pb.x - pa.x) * (pc.z - pa.z)) +
              pa.z * ((pb.x - pa.x) * (pc.y - pa.y) - (pb.y - pa.y) * (pc.x - pa.x));
  float b_2 = ps.x * (pb.x - pa.x) + ps.y * (pb.y - pa.y) + ps.z * (pb.z - pa.z);
  float b_3 = ps.x * (pc.x - pa.x) + ps.y * (pc.y - pa.y) + ps.z * (pc.z - pa.z);
  f4 a_1 = f4mk((pb.y - pa.y) * (pc.z - pa.z) - (pb.z - pa.z) * (pc.y - pa.y), pb.x - pa.x, pc.x - pa.x, 0);
  f4 a_2 = f4mk((pb.z - pa.z) * (pc.x - pa.x) - (pb.x - pa.x) * (pc.z - pa.z), pb.y - pa.y, pc.y - pa.y, 0);
  f4 a_3 = f4mk((pb.x - pa.x) * (pc.y - pa.y) - (pb.y - pa.y) * (pc.x - pa.x), pb.z - pa.z, pc.z - pa.z, 0);
  /* The reference builds b = (float4)(0, b_1, b_2, b_3) (:1281) while calcDeterminant3x3 reads
   * .x,.y,.z (:1244-1245), so the determinant sees (0, b_1, b_2) and b_3 is dead. Restated as is. */
  f4 b = f4mk(0, b_1, b_2, b_3);
  float den = det3(a_1, a_2, a_3);
  if (den != 0) {
    pm.x = det3(b, a_2, a_3) / den;
    pm.y = det3(a_1, b, a_3) / den;
    pm.z = det3(a_1, a_2, b) / den;
  } else {
    pm.w = -1;
  }
  return pm;
}

/* computeInteractionWithMembranes, sphFluid.cl:1369-1650 */
static void st_membrane_interaction(sph_ctx *c) {
  const sph_params *p = &c->p;
  const int N = c->N;
  const float r0 = p->r0;
#pragma omp parallel for schedule(dynamic, 256) num_threads(g_threads)
  for (int id = 0; id < N; id++) {
    int src = (int)c->particleIndex[id].y;
    int type = (int)c->position[src].w;
    if (type == SPH_BOUNDARY_PARTICLE) continue;
    if (type != SPH_LIQUID_PARTICLE) continue;
    const f2 *nm = c->neighborMap + (size_t)id * MAXN;
    f4 normal[MAXN];
    float dist[MAXN];
    int jd_counter = 0, abort_particle = 0;
    for (int i = 0; i < MAXN; i++) normal[i] = f4mk(0, 0, 0, 0);
    for (int nc = 0; nc < MAXN && !abort_particle; nc++) {
      int jd = (int)nm[nc].x;
      if (jd == NO_PARTICLE_ID) break;
      int jsrc = (int)c->particleIndex[jd].y;
      if ((int)c->position[jsrc].w != SPH_ELASTIC_PARTICLE) continue;
      int ijk = 0;
      f4 vij = f4sub(c->position[src], c->position[jsrc]);
      vij.z = 0;
      float d_id_jd = sqrtf(f4dot(vij, vij)); /* 4-lane dot: includes the type-tag difference in .w */
      for (int mli = 0; mli < SPH_MAX_MEMBRANES_INCLUDING_SAME_PARTICLE; mli++) {
        int mdi = c->pml[jsrc * SPH_MAX_MEMBRANES_INCLUDING_SAME_PARTICLE + mli];
        if (!(mdi > -1)) break;
        f4 pi = c->position[c->membraneData[mdi * 3 + 0]];
        f4 pj = c->position[c->membraneData[mdi * 3 + 1]];
        f4 pk = c->position[c->membraneData[mdi * 3 + 2]];
        f4 pp = project_point_to_plane(c->position[src], pi, pj, pk);
        if (pp.w == -1) { abort_particle = 1; break; } /* :1468-1472 return */
        f4 nrm = f4sub(c->position[src], pp);
        float len = sqrtf(nrm.x * nrm.x + nrm.y * nrm.y + nrm.z * nrm.z);
        if (len > 0) {
          nrm = f4div(nrm, len);
          normal[jd_counter] = f4add(normal[jd_counter], nrm);
          ijk++;
        } else { abort_particle = 1; break; } /* :1501-1505 return */
      }
      if (abort_particle) break;
      if (ijk > 0) {
        normal[jd_counter] = f4div(normal[jd_counter], (float)ijk);
        dist[jd_counter] = d_id_jd;
        jd_counter++;
      }
    }
    if (abort_particle || jd_counter <= 0) continue;
    f4 n_c_i = f4mk(0, 0, 0, 0);
    float w_sum = 0.f, w_second = 0.f;
    for (int nc = 0; nc < jd_counter; nc++) { /* :1578-1591 */
      float x = dist[nc];
      float w = (0.f < (r0 - x) / r0) ? (r0 - x) / r0 : 0.f;
      n_c_i = f4add(n_c_i, f4mul(normal[nc], w));
      w_sum += w;
      w_second += w * (r0 - x);
    }
    n_c_i.w = 0;
    float len = f4dot(n_c_i, n_c_i);
    if (len != 0) {
      len = sqrtf(len);
      f4 dp = f4div(f4lmul(1.0f, f4mul(f4div(n_c_i, len), w_second)), w_sum); /* 1.0f*(...)/w, :1600 */
      c->position[N + src].x += dp.x;
      c->position[N + src].y += dp.y;
      c->position[N + src].z += dp.z;
    }
  }
}

/* computeInteractionWithMembranes_finalize, sphFluid.cl:1652-1682 */
static void st_membrane_finalize(sph_ctx *c) {
  const int N = c->N;
  for (int id = 0; id < N; id++) {
    int src = (int)c->particleIndex[id].y;
    if ((int)c->position[src].w == SPH_BOUNDARY_PARTICLE) continue;
    c->position[src] = f4add(c->position[src], c->position[N + src]);
  }
}

int sph_stage(sph_ctx *c, int stage) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  switch (stage) {
    case SPH_STAGE_CLEAR_BUFFERS: st_clear_buffers(c); return 0;
    case SPH_STAGE_HASH_PARTICLES: st_hash_particles(c); return 0;
    case SPH_STAGE_SORT: st_sort(c); return 0;
    case SPH_STAGE_SORT_POST_PASS: st_sort_post_pass(c); return 0;
    case SPH_STAGE_INDEXX: st_indexx(c); return 0;
    case SPH_STAGE_INDEX_POST_PASS: st_index_post_pass(c); return 0;
    case SPH_STAGE_FIND_NEIGHBORS: st_find_neighbors(c); return 0;
    case SPH_STAGE_COMPUTE_DENSITY: st_compute_density(c); return 0;
    case SPH_STAGE_COMPUTE_FORCES_INIT_PRESSURE: st_compute_forces(c); return 0;
    case SPH_STAGE_COMPUTE_ELASTIC_FORCES: st_compute_elastic(c); return 0;
    case SPH_STAGE_PREDICT_POSITIONS: st_predict_positions(c); return 0;
    case SPH_STAGE_PREDICT_DENSITY: st_predict_density(c); return 0;
    case SPH_STAGE_CORRECT_PRESSURE: st_correct_pressure(c); return 0;
    case SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION: st_pressure_force(c); return 0;
    case SPH_STAGE_INTEGRATE: st_integrate(c); return 0;
    case SPH_STAGE_CLEAR_MEMBRANE_BUFFERS: st_clear_membrane_buffers(c); return 0;
    case SPH_STAGE_MEMBRANE_INTERACTION:
      if (c->has_pml || c->p.num_elastic == 0) st_membrane_interaction(c);
      return 0;
    case SPH_STAGE_MEMBRANE_FINALIZE: st_membrane_finalize(c); return 0;
  }
  return fail(SPH_E_INVALID, "unknown stage");
}

/* owPhysicsFluidSimulator::simulationStep, owPhysicsFluidSimulator.cpp:88-113 */
int sph_step(sph_ctx *c) {
  if (!c) return fail(SPH_E_INVALID, "null ctx");
  for (int s = SPH_STAGE_CLEAR_BUFFERS; s <= SPH_STAGE_COMPUTE_ELASTIC_FORCES; s++) sph_stage(c, s);
  int iter = 0;
  do {
    for (int s = SPH_STAGE_PREDICT_POSITIONS; s <= SPH_STAGE_COMPUTE_PRESSURE_FORCE_ACCELERATION; s++) sph_stage(c, s);
    iter++;
  } while (iter < c->p.max_iteration);
  for (int s = SPH_STAGE_INTEGRATE; s <= SPH_STAGE_MEMBRANE_FINALIZE; s++) sph_stage(c, s);
  return 0;
}

int sph_sync(sph_ctx *c) { (void)c; return 0; }

int sph_set_muscles(sph_ctx *c, const float *signal, int n) { /* owOpenCLSolver.cpp:738-742 */
  if (!c || !signal || n < 0 || n > c->p.muscle_count) return fail(SPH_E_INVALID, "sph_set_muscles");
  memcpy(c->muscle, signal, 4 * (size_t)n);
  return 0;
}

static void *buf_ptr(const sph_ctx *c, int id, size_t *bytes) {
  const size_t N = c->N, G = c->G, Ne = c->has_elastic ? c->p.num_elastic : 0;
  switch (id) {
    case SPH_BUF_POSITION: *bytes = 2 * N * 16; return c->position;
    case SPH_BUF_VELOCITY: *bytes = 2 * N * 16; return c->velocity;
    case SPH_BUF_PARTICLE_INDEX: *bytes = N * 8; return c->particleIndex;
    case SPH_BUF_PARTICLE_INDEX_BACK: *bytes = N * 4; return c->particleIndexBack;
    case SPH_BUF_SORTED_POSITION: *bytes = 2 * N * 16; return c->sortedPosition;
    case SPH_BUF_SORTED_VELOCITY: *bytes = N * 16; return c->sortedVelocity;
    case SPH_BUF_GRID_CELL_INDEX: *bytes = (G + 1) * 4; return c->gridCellIndex;
    case SPH_BUF_GRID_CELL_INDEX_FIXED_UP: *bytes = (G + 1) * 4; return c->gridCellIndexFixedUp;
    case SPH_BUF_NEIGHBOR_MAP: *bytes = N * MAXN * 8; return c->neighborMap;
    case SPH_BUF_RHO: *bytes = 2 * N * 4; return c->rho;
    case SPH_BUF_PRESSURE: *bytes = N * 4; return c->pressure;
    case SPH_BUF_ACCELERATION: *bytes = 2 * N * 16; return c->acceleration;
    case SPH_BUF_ELASTIC_CONNECTIONS: *bytes = Ne * MAXN * 16; return c->elastic;
    case SPH_BUF_MEMBRANE_DATA: *bytes = c->has_membranes ? 12 * (size_t)c->p.num_membranes : 0; return c->membraneData;
    case SPH_BUF_PARTICLE_MEMBRANES_LIST: *bytes = c->has_pml ? 28 * (size_t)c->p.num_elastic : 0; return c->pml;
    case SPH_BUF_MUSCLE_ACTIVATION_SIGNAL: *bytes = 4 * (size_t)c->p.muscle_count; return c->muscle;
  }
  *bytes = 0;
  return NULL;
}

size_t sph_buffer_bytes(const sph_ctx *c, int id) {
  size_t b = 0;
  if (c) buf_ptr(c, id, &b);
  return b;
}

int sph_read(sph_ctx *c, int id, void *dst, size_t bytes) {
  if (!c || !dst) return fail(SPH_E_INVALID, "sph_read: null");
  size_t have = 0;
  void *src = buf_ptr(c, id, &have);
  if (!src && bytes) return fail(SPH_E_UNSUPPORTED, "sph_read: buffer absent");
  if (bytes > have) return fail(SPH_E_SIZE, "sph_read: too many bytes");
  memcpy(dst, src, bytes);
  return 0;
}

int sph_write(sph_ctx *c, int id, const void *src, size_t bytes) {
  if (!c || !src) return fail(SPH_E_INVALID, "sph_write: null");
  /* POSITION / VELOCITY as the ABI promises; the oracle (test infrastructure) also lets a test overwrite the three
   * per-stage fields a slab rank refreshes in its ghosts (tests/slab_oracle.py): rho, pressure, sortedPosition. */
  if (id != SPH_BUF_POSITION && id != SPH_BUF_VELOCITY && id != SPH_BUF_RHO && id != SPH_BUF_PRESSURE &&
      id != SPH_BUF_SORTED_POSITION)
    return fail(SPH_E_UNSUPPORTED, "sph_write: buffer not writable");
  size_t have = 0;
  void *dst = buf_ptr(c, id, &have);
  if (bytes > have) return fail(SPH_E_SIZE, "sph_write: too many bytes");
  memcpy(dst, src, bytes);
  return 0;
}
