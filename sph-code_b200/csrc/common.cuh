// Shared declarations for libsphb200 (B200 / sm_100a only).
// Device data layout, context and error plumbing.  See DESIGN.md "Data layout in HBM".
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define SPHB200_MAX_ERR 512
extern thread_local char g_sphb200_err[SPHB200_MAX_ERR];

#define SPH_FAIL(...)                                                   \
  do {                                                                  \
    snprintf(g_sphb200_err, SPHB200_MAX_ERR, __VA_ARGS__);              \
    return 1;                                                           \
  } while (0)

#define CUDA_TRY(expr)                                                                        \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      snprintf(g_sphb200_err, SPHB200_MAX_ERR, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr,  \
               cudaGetErrorString(_e));                                                       \
      return 1;                                                                               \
    }                                                                                         \
  } while (0)

#define SPH_TRY(expr)            \
  do {                           \
    int _r = (expr);             \
    if (_r) return _r;           \
  } while (0)

// ---------------------------------------------------------------------------------------------
// Packed per-particle records.  Every gather kernel reads whole 32-byte records (two LDG.128 /
// one 256-bit sector pair) so a neighbour costs a fixed number of sectors.
//   R0 = {x, y, z, h^2}                    geometry                       (all gather passes)
//   R1 = {c6, m_gas, m_num, m_dust}        c6 = 315/(64 pi h^9);  m_gas = m [t==0];
//                                          m_num = m [t==0]/(amu mu);  m_dust = m [t==2]   (density)
//   R2 = {vx, vy, vz, cs}                  cs = sqrt(gamma k T [t==0]/(mu amu))            (Pi, force)
//   R3 = {AK, PK}                          AK = 6 c6 * m k T [t==0]/(mu amu)
//                                          PK = 6 c6 * m Pi [t==0]                         (force)
// ---------------------------------------------------------------------------------------------
struct __align__(16) rec2 { double a, b; };

#ifdef __CUDACC__
// One 256-bit read-only load of a 32-byte record (LDG.E.256 on sm_100a): a gathered neighbour costs one
// L1 tag lookup per record instead of the two of a pair of 128-bit loads.  p must be 32-byte aligned.
__device__ __forceinline__ double4 ldg256(const double4* p) {
  double4 v;
  asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
#endif

struct PhysConst { double kB, amu; };

struct GridSpec {
  double ox, oy, oz;   // origin (lower corner of cell (0,0,0))
  double inv_c;        // 1 / cell edge
  double c;            // cell edge
  int nx, ny, nz;
  int ncells;
  int ngx, ngy, ngz;   // groups of 2x2x2 cells (tile.cuh)
  int ngroups;
  int n;               // particles the grid was built for
  int ok;              // 0: no finite coordinate (every particle lands in cell 0)
};

// A growable device buffer owned by the context.
#define SPHB200_MAX_SPECIES 64   // species columns of f_un the drag tables can hold (the reference has 15)

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

struct sphb200_ctx {
  int device = 0;
  PhysConst phys = {1.380649e-23, 1.66053906892e-27};   // overwritten by sphb200_set_constants
  double m_h = 1.0008 * 1.66053906892e-27;
  double m_0 = 31.622776601683793 * 1.989e30;
  // --- scratch / session state (sorted particle order unless noted) ---
  DevBuf r0;                       // sorted position records {x,y,z,guess}, N+1 entries
  DevBuf rho, pi_, aux;            // per-particle doubles
  DevBuf list;                     // (N,k) int32 neighbour list in the index space of the records
  DevBuf keys, keys_alt, perm, perm_alt, cell_start, scan_tops;
  DevBuf grid_dev;                 // GridSpec chosen on the device by the last build_cells (the hot path never
                                   // copies it back: every cell-tiled kernel reads it from here)
  DevBuf list16;                   // (N,k) uint16 neighbour rows as tile offsets (tile.cuh); session path
  DevBuf recD, recP, recK;         // per-pass records in sorted order: 48 B density, 80 B Pi/force, 8 B Pi coefficient
  DevBuf worklist;                 // k-NN targets left to the general kernel
  DevBuf group_list;               // [0]: number of groups that hold particles, [1..]: their ids (cells.cu), for the tiled kernels
  DevBuf unit_tab;                 // per-unit tile geometry of the current cell list (hydro.cu), built on first use
  int unit_tab_valid = 0;          // reset by build_cells
  DevBuf scal;                     // device scalars (doubles), see enum Scal
  DevBuf flags;                    // device ints: error / statistics counters
  DevBuf tmp0, tmp1, tmp2, tmp3;   // generic scratch
  DevBuf sbuf[48];                 // session arrays (session.cu)
  double cell_occ = 16.0;          // target mean particles per cell (tunable)
  double grid_pow = -3.0;          // cell edge from h_eff = (mean h^p)^(1/p) of the previous k-th distances (cells.cu)
  double cell_edge = 0.0;          // fixed cell edge (0 = from cell_occ)
  int knn_group = 32;              // lanes per target in the general k-NN kernel
  int tile_cfg = 0;                // pair passes: 0 = measured best split of CTAs / producer / consumer warps, 1..3 others (hydro.cu)
  int knn_lane = 0;                // session fast path: lane-per-target kernel (knn_lane.cu); 0: warp-per-target (knn_group.cu)
  int knn_mma = 0;                 // 1: tensor-core filter variant of the session fast path (knn_mma.cu; same rows, not faster - DESIGN.md 4)
  int knn_atomic = 0;              // fast path: filter survivors reserve their slot with shared-memory atomics
  int knn_fast = 1;                // 1: cell-tiled fast path (knn_cell.cu), 0: general kernel only
  double knn_margin = 0.10;        // trial radius = previous k-th distance * (1 + margin)
  GridSpec grid;                   // host copy; valid only after build_cells(..., sync_host = true)
  int max_cells = 0;               // capacity of cell_start the device-side grid choice may use
  int list_is_global = 0;          // 1: ctx->list holds int32 rows for EVERY particle (function API / restored checkpoint)
  // slab decomposition: rows whose caller index (skip_id[i]) is >= skip_from are ghosts; their owners
  // compute their sums, so the k-NN and the list-driven passes skip them as targets (nullptr: none)
  const uint32_t* skip_id = nullptr;
  uint32_t skip_from = 0;
  int64_t launches = 0;            // kernels launched through this context (for bench "gpu_launches")
  void* session = nullptr;         // the device-resident run this context holds (session.cu), or nullptr
};

int devbuf_reserve(DevBuf& b, size_t bytes);
void devbuf_free(DevBuf& b);

template <typename T>
static inline T* bufp(DevBuf& b) { return reinterpret_cast<T*>(b.p); }

// n + 2 slots (bulk copies read up to one element past a run), rounded so each array stays 16 B aligned
static inline size_t soa_stride(int64_t n) { return ((size_t)n + 2 + 1) & ~(size_t)1; }

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// enum of device scalars kept in ctx->scal (doubles)
enum Scal {
  SC_BBOX = 0,        // 6 doubles: min x,y,z, max x,y,z
  SC_CT = 6,          // crossing time (min over gas particles)
  SC_DT = 7,
  SC_D = 8,           // global length scale d
  SC_AGE = 9,
  SC_MAXDIST = 10,    // 90th percentile of |x|^2 among slow particles
  SC_NSLOW = 11,
  SC_COVER_BAD = 12,  // slab decomposition: 1 if some owned particle's k-th neighbour ball leaves the covered region
  SC_COVER_HMAX = 13, // largest smoothing length among the owned particles
  SC_COUNT = 32
};

// flags (ints)
enum Flag {
  FL_KNN_OVERFLOW = 0,   // k-NN: targets that exhausted the retry budget
  FL_KNN_RESCAN = 1,     // k-NN: number of extra scans (statistics)
  FL_KNN_WIDEN = 2,      // k-NN: number of targets that needed a wider cell block
  FL_BAD_INDEX = 3,      // neighbour index outside [0, N)
  FL_KNN_NONFINITE = 4,  // k-NN: targets with NaN/inf coordinates
  FL_KNN_NWORK = 5,      // k-NN: length of the work list of the last search
  FL_GROUP_NEXT = 6,     // dynamic group counter of the persistent cell-tiled kernels (reset by each launch)
  FL_KNN_CROWDED = 7,    // k-NN: groups whose tile exceeds the shared-memory capacity (all targets to the work list)
  FL_COUNT = 16
};
