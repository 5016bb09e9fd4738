// Cell-list build, entirely on the device (no host synchronisation, fixed launch sizes: graph-capturable):
//   bounding box -> grid choice (one block) -> cell keys + histogram with in-cell ranks -> exclusive scan
//   -> counting-sort scatter -> in-cell ordering (deterministic) -> sorted position records.
// Replaces the cKDTree build of /root/reference/sph/navier_stokes_cleaned.py:542.
#include "common.cuh"
#include "kernels.cuh"
#include "tile.cuh"
#include <float.h>
#include <math.h>

#define BB_BLOCKS (148 * 4)
#define SCAN_TILE 2048             // elements per block of the scan (256 threads x 8)

// ---- bounding box: per-block partial min/max of the finite coordinates ----
#define BB_PART 8                  // doubles per block: min xyz, max xyz, sum of guess^-3, number of guesses
__global__ void __launch_bounds__(256) bbox_partial_kernel(const double* __restrict__ pts, const double* __restrict__ guess,
                                                           double hpow, int64_t n, double* __restrict__ part) {
  double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  double dsum = 0.0, dcnt = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const double v = pts[3 * i + a];
      if (isfinite(v)) { lo[a] = fmin(lo[a], v); hi[a] = fmax(hi[a], v); }
    }
    if (guess) {
      const double h = guess[i];
      if (h > 0.0 && h < DBL_MAX) { dsum += (hpow == -3.0) ? 1.0 / (h * h * h) : pow(h, hpow); dcnt += 1.0; }
    }
  }
  __shared__ double sh[8][BB_PART];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      lo[a] = fmin(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], d));
      hi[a] = fmax(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], d));
    }
    if (lane == 0) { sh[warp][a] = lo[a]; sh[warp][3 + a] = hi[a]; }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    dsum += __shfl_xor_sync(0xffffffffu, dsum, d);
    dcnt += __shfl_xor_sync(0xffffffffu, dcnt, d);
  }
  if (lane == 0) { sh[warp][6] = dsum; sh[warp][7] = dcnt; }
  __syncthreads();
  if (threadIdx.x < BB_PART) {
    double v = sh[0][threadIdx.x];
    for (int w = 1; w < 8; ++w)
      v = (threadIdx.x < 3) ? fmin(v, sh[w][threadIdx.x]) : (threadIdx.x < 6) ? fmax(v, sh[w][threadIdx.x]) : v + sh[w][threadIdx.x];
    part[blockIdx.x * BB_PART + threadIdx.x] = v;
  }
}

// Choose the grid for a bounding box.  `edge_hint` > 0 fixes the cell edge; otherwise the edge is set so
// that the mean occupancy of the bounding box is `occ` particles per cell.  Degenerate inputs (particles on
// a plane or a line): a collapsed direction gets one layer of cells and the occupancy is set over the
// directions that do extend, otherwise the cells come out far too small.
__device__ void choose_grid(const double* bb, int64_t n, double occ, double edge_hint, int max_cells, GridSpec* g) {
  double L[3];
  for (int a = 0; a < 3; ++a) {
    L[a] = bb[3 + a] - bb[a];
    if (!(L[a] > 0.0)) L[a] = 0.0;
  }
  double Lmax = fmax(L[0], fmax(L[1], L[2]));
  if (!(Lmax > 0.0)) Lmax = 1.0;
  double measure = 1.0;
  int dims = 0;
  for (int a = 0; a < 3; ++a) {
    if (L[a] < 1e-6 * Lmax) L[a] = 0.0;
    else { measure *= L[a]; ++dims; }
  }
  double c = edge_hint;
  if (!(c > 0.0)) {
    const double per_cell = measure * occ / (double)(n > 0 ? n : 1);
    c = (dims == 3) ? cbrt(per_cell) : (dims == 2) ? sqrt(per_cell) : (dims == 1) ? per_cell : 1.0;
  }
  if (!(c > 0.0)) c = Lmax;
  for (;;) {
    const double nx = floor(L[0] / c) + 1, ny = floor(L[1] / c) + 1, nz = floor(L[2] / c) + 1;
    if (nx * ny * nz <= (double)max_cells && nx < 2.0e6 && ny < 2.0e6 && nz < 2.0e6) {
      g->nx = (int)nx; g->ny = (int)ny; g->nz = (int)nz;
      break;
    }
    c *= 1.1;
  }
  g->c = c;
  g->inv_c = 1.0 / c;
  g->ox = bb[0]; g->oy = bb[1]; g->oz = bb[2];
  g->ncells = g->nx * g->ny * g->nz;
  g->ngx = (g->nx + 1) / 2; g->ngy = (g->ny + 1) / 2; g->ngz = (g->nz + 1) / 2;
  g->ngroups = g->ngx * g->ngy * g->ngz;
  g->n = (int)n;
}

__global__ void __launch_bounds__(256)
grid_choose_kernel(const double* __restrict__ part, int nblocks, int64_t n, double occ, double edge_hint, int knn_k, double hpow, int max_cells,
                   GridSpec* __restrict__ spec, double* __restrict__ bbox_out) {
  __shared__ double sh[256][BB_PART];
  double v[BB_PART] = {DBL_MAX, DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, -DBL_MAX, 0.0, 0.0};
  for (int b = threadIdx.x; b < nblocks; b += blockDim.x)
#pragma unroll
    for (int a = 0; a < BB_PART; ++a)
      v[a] = (a < 3) ? fmin(v[a], part[b * BB_PART + a]) : (a < 6) ? fmax(v[a], part[b * BB_PART + a]) : v[a] + part[b * BB_PART + a];
#pragma unroll
  for (int a = 0; a < BB_PART; ++a) sh[threadIdx.x][a] = v[a];
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s)
#pragma unroll
      for (int a = 0; a < BB_PART; ++a)
        sh[threadIdx.x][a] = (a < 3) ? fmin(sh[threadIdx.x][a], sh[threadIdx.x + s][a])
                           : (a < 6) ? fmax(sh[threadIdx.x][a], sh[threadIdx.x + s][a]) : sh[threadIdx.x][a] + sh[threadIdx.x + s][a];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double bb[6];
    bool ok = true;
    for (int a = 0; a < 6; ++a) bb[a] = sh[0][a];
    for (int a = 0; a < 3; ++a) if (!(bb[a] <= bb[3 + a])) ok = false;
    if (!ok) for (int a = 0; a < 6; ++a) bb[a] = 0.0;          // no finite coordinate: one cell, flagged
    // With previous k-th neighbour distances at hand the cell edge follows the particles' own mean number
    // density, n_p = mean_i k / (4 pi/3 h_i^3): a bounding box inflated by a few stragglers (or a dilute envelope
    // around a dense cloud) then no longer makes the cells of the bulk hold more than `occ` particles.
    if (!(edge_hint > 0.0) && knn_k > 0 && sh[0][7] >= 0.5 * (double)n) {
      // h_eff = (mean h^p)^(1/p): p = -3 is the particles' mean number density, larger p leans towards the dilute regions
      const double heff = pow(sh[0][6] / sh[0][7], 1.0 / hpow);
      const double dens = (double)knn_k / (4.1887902047863905 * heff * heff * heff);
      if (dens > 0.0 && dens < DBL_MAX) edge_hint = cbrt(occ / dens);
    }
    GridSpec g;
    choose_grid(bb, n, occ, edge_hint, max_cells, &g);
    g.ok = ok ? 1 : 0;
    *spec = g;
    for (int a = 0; a < 6; ++a) bbox_out[a] = bb[a];
  }
}

__device__ __forceinline__ int cell_coord(double x, double o, double inv_c, int n) {
  const double f = (x - o) * inv_c;
  return (f >= 0.0) ? ((f < (double)n) ? (int)f : n - 1) : 0;   // NaN -> 0
}

__global__ void __launch_bounds__(256) zero_counts_kernel(const GridSpec* __restrict__ spec, uint32_t* __restrict__ counts) {
  const int m = spec->ncells + 2;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < m; c += gridDim.x * blockDim.x) counts[c] = 0u;
}

// key of every particle, its arrival rank inside the cell, per-cell counts
__global__ void __launch_bounds__(256)
cell_key_kernel(const double* __restrict__ pts, int64_t n, const GridSpec* __restrict__ spec, uint32_t* __restrict__ keys,
                uint32_t* __restrict__ rank, uint32_t* __restrict__ counts) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const GridSpec g = *spec;
  const int cx = cell_coord(pts[3 * i + 0], g.ox, g.inv_c, g.nx);
  const int cy = cell_coord(pts[3 * i + 1], g.oy, g.inv_c, g.ny);
  const int cz = cell_coord(pts[3 * i + 2], g.oz, g.inv_c, g.nz);
  const uint32_t key = ((uint32_t)cz * g.ny + cy) * g.nx + cx;
  keys[i] = key;
  rank[i] = atomicAdd(&counts[key], 1u);
}

// ---- exclusive scan of counts[0 .. ncells] in place (three small kernels, fixed launch sizes) ----
__global__ void __launch_bounds__(256)
scan_tiles_kernel(const GridSpec* __restrict__ spec, int m_host, uint32_t* __restrict__ a, uint32_t* __restrict__ tops) {
  const int m = spec ? spec->ncells + 1 : m_host;
  const int base = blockIdx.x * SCAN_TILE;
  if (base >= m) return;
  __shared__ uint32_t wsum[8];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  uint32_t v[8], sum = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int idx = base + t * 8 + q;
    v[q] = (idx < m) ? a[idx] : 0u;
    sum += v[q];
  }
  uint32_t inc = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  uint32_t before = inc - sum;
  for (int w = 0; w < warp; ++w) before += wsum[w];
  if (t == 255) tops[blockIdx.x] = before + sum;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int idx = base + t * 8 + q;
    if (idx < m) a[idx] = before;
    before += v[q];
  }
}
__global__ void __launch_bounds__(1024) scan_tops_kernel(const GridSpec* __restrict__ spec, int m_host, uint32_t* __restrict__ tops) {
  const int ntiles = ((spec ? spec->ncells + 1 : m_host) + SCAN_TILE - 1) / SCAN_TILE;
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry_s;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t == 0) carry_s = 0u;
  __syncthreads();
  for (int base = 0; base < ntiles; base += 1024) {
    const int idx = base + t;
    const uint32_t v = (idx < ntiles) ? tops[idx] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = wsum[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, w, d);
        if (lane >= d) w += o;
      }
      wsum[lane] = w;                                        // inclusive totals of the warps
    }
    __syncthreads();
    const uint32_t carry = carry_s;
    const uint32_t excl = carry + (warp ? wsum[warp - 1] : 0u) + inc - v;
    if (idx < ntiles) tops[idx] = excl;
    __syncthreads();
    if (t == 1023) carry_s = excl + v;
    __syncthreads();
  }
}
__global__ void __launch_bounds__(256)
scan_add_kernel(const GridSpec* __restrict__ spec, int m_host, uint32_t* __restrict__ a, const uint32_t* __restrict__ tops) {
  const int m = spec ? spec->ncells + 1 : m_host;
  const int base = blockIdx.x * SCAN_TILE;
  if (base >= m) return;
  const uint32_t add = tops[blockIdx.x];
  for (int q = threadIdx.x; q < SCAN_TILE; q += 256) {
    const int idx = base + q;
    if (idx < m) a[idx] += add;
  }
}
// the guard slot cell_start[ncells + 1] = n (rows are read one cell past their end)
__global__ void scan_guard_kernel(const GridSpec* __restrict__ spec, uint32_t* __restrict__ a) {
  const int m = spec->ncells + 1;
  a[m] = a[m - 1];
}

// counting-sort scatter: sorted slot of particle i = cell_start[key] + rank
__global__ void __launch_bounds__(256)
scatter_kernel(int64_t n, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ rank,
               const uint32_t* __restrict__ cell_start, uint32_t* __restrict__ perm, uint32_t* __restrict__ keys_sorted) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t key = keys[i];
  const uint32_t dst = cell_start[key] + rank[i];
  perm[dst] = (uint32_t)i;
  keys_sorted[dst] = key;
}

// The arrival ranks come from atomics, so the order inside a cell differs from run to run.  One warp per cell
// puts the cell's entries into ascending source index (bitonic network over the lanes for <= 32 entries, a
// rank sort through `tmp` for up to 256; larger cells keep their arrival order): same sort order, same
// summation order, same tie-breaks every run.
__global__ void __launch_bounds__(256)
cell_order_kernel(const GridSpec* __restrict__ spec, const uint32_t* __restrict__ cell_start, uint32_t* __restrict__ perm,
                  uint32_t* __restrict__ tmp) {
  const int ncells = spec->ncells;
  const int lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < ncells; c += nwarps) {
    const int b = (int)cell_start[c], e = (int)cell_start[c + 1];
    const int m = e - b;
    if (m <= 1) continue;
    if (m <= 32) {
      uint32_t v = (lane < m) ? perm[b + lane] : 0xffffffffu;
#pragma unroll
      for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
          const uint32_t o = __shfl_xor_sync(0xffffffffu, v, j);
          const bool up = (lane & k) == 0, lower = (lane & j) == 0;
          v = (lower == up) ? min(v, o) : max(v, o);
        }
      }
      if (lane < m) perm[b + lane] = v;
    } else if (m <= 256) {
      for (int s = lane; s < m; s += 32) {
        const uint32_t v = perm[b + s];
        int r = 0;
        for (int f = 0; f < m; ++f) r += perm[b + f] < v;
        tmp[b + r] = v;
      }
      __syncwarp();
      for (int s = lane; s < m; s += 32) perm[b + s] = tmp[b + s];
    }
    __syncwarp();
  }
}

// sorted position records {x, y, z, guess}: guess = the previous k-th neighbour distance of the particle (trial
// radius of the next search; 0 = none).  Entry n is left to the caller.
__global__ void __launch_bounds__(256)
gather_pos_kernel(const double* __restrict__ pts, const uint32_t* __restrict__ perm, const double* __restrict__ guess,
                  int64_t n, double4* __restrict__ r0) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t p = perm[i];
  r0[i] = make_double4(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2], guess ? guess[p] : 0.0);
}

// in-place exclusive scan of a[0 .. m) (host-known length); a[m] is not touched
int scan_exclusive_u32(sphb200_ctx* ctx, uint32_t* a, int m, cudaStream_t st) {
  if (m <= 0) return 0;
  const int blocks = ceil_div(m, SCAN_TILE);
  SPH_TRY(devbuf_reserve(ctx->scan_tops, ((size_t)blocks + 1) * sizeof(uint32_t)));
  scan_tiles_kernel<<<blocks, 256, 0, st>>>(nullptr, m, a, bufp<uint32_t>(ctx->scan_tops));
  scan_tops_kernel<<<1, 1024, 0, st>>>(nullptr, m, bufp<uint32_t>(ctx->scan_tops));
  scan_add_kernel<<<blocks, 256, 0, st>>>(nullptr, m, a, bufp<uint32_t>(ctx->scan_tops));
  ctx->launches += 3;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int devbuf_reserve(DevBuf& b, size_t bytes) {
  if (bytes <= b.bytes) return 0;
  if (b.p) { cudaFree(b.p); b.p = nullptr; b.bytes = 0; }
  size_t want = bytes + bytes / 4 + 256;       // head-room: growing a buffer costs a cudaFree + cudaMalloc (device-wide syncs)
  CUDA_TRY(cudaMalloc(&b.p, want));
  b.bytes = want;
  return 0;
}

void devbuf_free(DevBuf& b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.bytes = 0;
}

// Build the cell list for `pts` (device, (N,3) f64).  On return (stream order) ctx->grid_dev holds the grid,
// ctx->keys the sorted cell keys, ctx->perm the sorted->source permutation, ctx->cell_start the
// (ncells + 2) exclusive scan and ctx->r0 the sorted {x,y,z,guess} records.  Nothing synchronises unless
// `sync_host` asks for the host copy ctx->grid (function-level API).
// Groups of 2x2x2 cells (tile.cuh) that hold at least one particle, in roughly ascending order (each block keeps the
// order of its chunk and appends it with one atomic).  A clustered distribution leaves most cells of its bounding
// box empty; the group-tiled kernels walk this list instead of every group.
__global__ void __launch_bounds__(256) active_groups_kernel(const GridSpec* __restrict__ spec, const uint32_t* __restrict__ cell_start,
                                                            int* __restrict__ list) {
  __shared__ int wsum[8];
  __shared__ int base_s;
  const GridSpec g = *spec;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int first = blockIdx.x * 256; first < g.ngroups; first += gridDim.x * 256) {
    const int grp = first + threadIdx.x;
    bool any = false;
    if (grp < g.ngroups) {
      const int q = grp / g.ngx;
      const int gx = grp - q * g.ngx, gz = q / g.ngy, gy = q - gz * g.ngy;
      const int cx0 = 2 * gx, cx1 = min(2 * gx + 1, g.nx - 1);
      for (int r = 0; r < 4 && !any; ++r) {
        const int cy = 2 * gy + (r & 1), cz = 2 * gz + (r >> 1);
        if (cy >= g.ny || cz >= g.nz) continue;
        const int base = (cz * g.ny + cy) * g.nx;
        any = cell_start[base + cx1 + 1] > cell_start[base + cx0];
      }
    }
    const unsigned b = __ballot_sync(0xffffffffu, any);
    if (lane == 0) wsum[w] = __popc(b);
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
      for (int i = 0; i < 8; ++i) { const int c = wsum[i]; wsum[i] = tot; tot += c; }
      base_s = tot ? atomicAdd(&list[0], tot) : 0;
    }
    __syncthreads();
    if (any) list[1 + base_s + wsum[w] + __popc(b & ((1u << lane) - 1u))] = grp;
    __syncthreads();
  }
}

int build_cells(sphb200_ctx* ctx, const double* pts, int64_t n, double edge_hint, const double* guess, int knn_k,
                bool sync_host, cudaStream_t st) {
  if (n <= 0) SPH_FAIL("build_cells: empty particle set");
  if (n >= (int64_t)0x7fffffff) SPH_FAIL("build_cells: N=%lld exceeds the 2^31-1 particle limit per GPU", (long long)n);
  const int64_t max_cells64 = 2 * n + 4096;
  const int max_cells = (int)(max_cells64 < 1000000000 ? max_cells64 : 1000000000);
  ctx->max_cells = max_cells;
  ctx->unit_tab_valid = 0;                                  // the pair passes rebuild their per-unit geometry table (hydro.cu)
  const int scan_blocks = ceil_div((int64_t)max_cells + 2, SCAN_TILE);
  SPH_TRY(devbuf_reserve(ctx->scal, SC_COUNT * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->tmp0, 64 + BB_BLOCKS * BB_PART * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->grid_dev, sizeof(GridSpec)));
  SPH_TRY(devbuf_reserve(ctx->group_list, ((size_t)max_cells + 16) * sizeof(int)));   // ngroups <= ncells
  SPH_TRY(devbuf_reserve(ctx->keys, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->keys_alt, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->perm, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->perm_alt, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->cell_start, ((size_t)max_cells + 4) * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->scan_tops, ((size_t)scan_blocks + 1) * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->r0, (size_t)(n + 1) * sizeof(double4)));
  double* part = reinterpret_cast<double*>(bufp<char>(ctx->tmp0) + 64);
  double* scal = bufp<double>(ctx->scal);
  GridSpec* spec = bufp<GridSpec>(ctx->grid_dev);
  uint32_t* cs = bufp<uint32_t>(ctx->cell_start);
  uint32_t* keys_src = bufp<uint32_t>(ctx->keys_alt);
  uint32_t* rank = bufp<uint32_t>(ctx->perm_alt);
  int bb_blocks = ceil_div(n, 256); if (bb_blocks > BB_BLOCKS) bb_blocks = BB_BLOCKS;
  const int nblk = ceil_div(n, 256);
  bbox_partial_kernel<<<bb_blocks, 256, 0, st>>>(pts, guess, ctx->grid_pow, n, part);
  grid_choose_kernel<<<1, 256, 0, st>>>(part, bb_blocks, n, ctx->cell_occ, edge_hint, knn_k, ctx->grid_pow, max_cells, spec, scal + SC_BBOX);
  zero_counts_kernel<<<148 * 4, 256, 0, st>>>(spec, cs);
  cell_key_kernel<<<nblk, 256, 0, st>>>(pts, n, spec, keys_src, rank, cs);
  scan_tiles_kernel<<<scan_blocks, 256, 0, st>>>(spec, 0, cs, bufp<uint32_t>(ctx->scan_tops));
  scan_tops_kernel<<<1, 1024, 0, st>>>(spec, 0, bufp<uint32_t>(ctx->scan_tops));
  scan_add_kernel<<<scan_blocks, 256, 0, st>>>(spec, 0, cs, bufp<uint32_t>(ctx->scan_tops));
  scan_guard_kernel<<<1, 1, 0, st>>>(spec, cs);
  scatter_kernel<<<nblk, 256, 0, st>>>(n, keys_src, rank, cs, bufp<uint32_t>(ctx->perm), bufp<uint32_t>(ctx->keys));
  cell_order_kernel<<<148 * 4, 256, 0, st>>>(spec, cs, bufp<uint32_t>(ctx->perm), rank);
  gather_pos_kernel<<<nblk, 256, 0, st>>>(pts, bufp<uint32_t>(ctx->perm), guess, n, bufp<double4>(ctx->r0));
  CUDA_TRY(cudaMemsetAsync(ctx->group_list.p, 0, sizeof(int), st));
  active_groups_kernel<<<148 * 2, 256, 0, st>>>(spec, cs, bufp<int>(ctx->group_list));
  ctx->launches += 12;
  CUDA_TRY(cudaGetLastError());
  if (sync_host) {
    CUDA_TRY(cudaMemcpyAsync(&ctx->grid, spec, sizeof(GridSpec), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (!ctx->grid.ok) SPH_FAIL("build_cells: no finite coordinates");
  }
  return 0;
}
