// Cell groups and their shared-memory tiles: the geometry every cell-tiled kernel of libsphb200 shares
// (k-NN search, density, viscosity Pi, force, crossing time).
//
// The cell list (cells.cu) sorts particles by cell key (cz ny + cy) nx + cx, so the cells of one (y,z) row
// are contiguous in the sorted order.  A GROUP is a block of 2x2x2 cells; its TILE is the block grown by one
// cell on every side (4x4x4 cells, clipped at the grid faces): every particle within one cell edge of a
// particle of the group.  The tile is at most 16 ROWS, one per (y,z), each a contiguous run of sorted
// particles; rows are ordered z-major, i.e. by ascending sorted index.  Rows are widened to even bounds so
// that arrays of 8-byte elements can be staged with 16-byte-aligned bulk copies (cp.async.bulk); the at most
// two padding slots of a row hold real particles of neighbouring cells and are masked where it matters.
// A neighbour is stored as its TILE OFFSET (uint16): toff[row] + (sorted index - first padded index of the
// row).  Offsets of the three z-layers a half-group needs are one contiguous range of the tile, so a pass
// over the particles of the cells (.,.,2gz+half) stages only those rows (12 of 16).
#pragma once
#include "common.cuh"

#define TILE_MAX_ROWS 16
#define TILE_CAP 1152            // padded particles a full tile may hold (k-NN); larger groups take the general path
#define TILE_HALF_CAP 896        // padded particles of the 3 z-layers a half-group pass stages
#define TILE_MAX_TARGETS 224     // particles of the 2x2x2 cells of a group
#define TILE_HALF_TARGETS 144    // particles of the 2x2 cells of one z-half of a group
#define LIST_STRIDE(k) (((k) + 7) & ~7)   // uint16 entries per row: rows are 16-byte aligned (bulk-copyable)
#define LIST_ESCAPED 0xFFFFu     // column 0 of a uint16 row: the row lives in the global list (slot in columns 1,2)
#define LIST_MISSING 0xFFFEu     // uint16 row entry: no neighbour (fewer than k inside the distance cut)

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
// global -> shared bulk copy (TMA engine, non-tensor form; UBLKCP in SASS); bytes % 16 == 0, both sides 16 B aligned
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Geometry of one group, computed by the first 16 lanes of a warp (one row each) and kept in shared memory.
struct TileRows {
  int ps[TILE_MAX_ROWS];      // first padded sorted index of the row (even)
  int rs[TILE_MAX_ROWS];      // real run [rs, re)
  int re[TILE_MAX_ROWS];
  int toff[TILE_MAX_ROWS + 1];// tile offset of the row's first padded slot; toff[nrows] = padded total
  int nrows, ysp, zsp;        // rows = ysp x zsp
  int x0, x1, y0, z0;         // first / last cell of the rows along x, first cell along y and z
  int gx, gy, gz;             // group coordinates
  int total;                  // real particles in the tile
};

// Fill `t` for group (gx,gy,gz); call with all 32 lanes of one warp, then __syncwarp / __syncthreads.
__device__ __forceinline__ void tile_rows(const GridSpec& g, const uint32_t* __restrict__ cell_start, int gx, int gy, int gz,
                                          int lane, TileRows* t) {
  const int x0 = max(2 * gx - 1, 0), x1 = min(2 * gx + 2, g.nx - 1);
  const int y0 = max(2 * gy - 1, 0), y1 = min(2 * gy + 2, g.ny - 1);
  const int z0 = max(2 * gz - 1, 0), z1 = min(2 * gz + 2, g.nz - 1);
  const int ysp = y1 - y0 + 1, zsp = z1 - z0 + 1;
  const int nrows = ysp * zsp;
  int rs = 0, re = 0;
  if (lane < nrows) {
    const int base = ((z0 + lane / ysp) * g.ny + (y0 + lane % ysp)) * g.nx;
    rs = (int)cell_start[base + x0];
    re = (int)cell_start[base + x1 + 1];
  }
  const int ps = rs & ~1;
  const int plen = (re > rs) ? (((re + 1) & ~1) - ps) : 0;
  int inc = plen, tot = re - rs;
#pragma unroll
  for (int d = 1; d < 16; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, d);
    const int w = __shfl_up_sync(0xffffffffu, tot, d);
    if (lane >= d) { inc += v; tot += w; }
  }
  if (lane < TILE_MAX_ROWS) { t->ps[lane] = ps; t->rs[lane] = rs; t->re[lane] = re; t->toff[lane] = inc - plen; }
  if (lane == 15) { t->toff[TILE_MAX_ROWS] = inc; t->total = tot; }
  if (lane == 0) {
    t->nrows = nrows; t->ysp = ysp; t->zsp = zsp; t->x0 = x0; t->x1 = x1; t->y0 = y0; t->z0 = z0;
    t->gx = gx; t->gy = gy; t->gz = gz;
  }
}
// padded total of the tile (rows beyond nrows are empty, so toff[16] is the total for any nrows)
__device__ __forceinline__ int tile_padded_total(const TileRows* t) { return t->toff[TILE_MAX_ROWS]; }

// The particles of the group's own cells are at most four contiguous runs (cells 2gx, 2gx+1 of the same (y,z)
// are adjacent in the sorted order).  Run q = (dz, dy) in 0..3.  Returns false when the run does not exist.
__device__ __forceinline__ bool group_target_run(const GridSpec& g, const uint32_t* __restrict__ cell_start, int gx, int gy,
                                                 int gz, int q, int* begin, int* end, int* row) {
  const int cy = 2 * gy + (q & 1), cz = 2 * gz + (q >> 1);
  if (cy >= g.ny || cz >= g.nz) return false;
  const int base = (cz * g.ny + cy) * g.nx;
  const int cx0 = 2 * gx, cx1 = min(2 * gx + 1, g.nx - 1);
  *begin = (int)cell_start[base + cx0];
  *end = (int)cell_start[base + cx1 + 1];
  const int y0 = max(2 * gy - 1, 0), z0 = max(2 * gz - 1, 0);
  const int ysp = min(2 * gy + 2, g.ny - 1) - y0 + 1;
  *row = (cz - z0) * ysp + (cy - y0);
  return true;
}

// Does the group fit the shared-memory tiles?  Every cell-tiled kernel takes the SAME decision from the same
// inputs: the k-NN kernel sends all targets of a crowded group to the general kernel (global rows), the pair
// passes skip the group.  len[q] = particles of target run q (group_target_run).
__device__ __forceinline__ bool tile_crowded(const GridSpec& g, const TileRows* t, const int* len) {
  if (t->toff[TILE_MAX_ROWS] > TILE_CAP) return true;
  if (len[0] + len[1] + len[2] + len[3] > TILE_MAX_TARGETS) return true;
  if (len[0] + len[1] > TILE_HALF_TARGETS || len[2] + len[3] > TILE_HALF_TARGETS) return true;
  for (int hz = 0; hz < 2; ++hz) {
    const int cz = 2 * t->gz + hz;
    if (cz >= g.nz) continue;
    const int zlo = max(cz - 1, 0), zhi = min(cz + 1, g.nz - 1);
    const int o_lo = t->toff[(zlo - t->z0) * t->ysp], o_hi = t->toff[(zhi - t->z0 + 1) * t->ysp];
    if (o_hi - o_lo > TILE_HALF_CAP) return true;
  }
  return false;
}

__device__ __forceinline__ void group_coords(const GridSpec& g, int group, int* gx, int* gy, int* gz) {
  const int q = group / g.ngx;
  *gx = group - q * g.ngx;
  *gz = q / g.ngy;
  *gy = q - *gz * g.ngy;
}
#endif
