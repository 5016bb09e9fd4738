// Cell-list build: bounding box, cell keys + histogram, radix sort of (key, particle) pairs,
// cell-start scan, gather of positions into sorted {x,y,z,h^2} records.
// Replaces the cKDTree build of /root/reference/sph/navier_stokes_cleaned.py:542.
#include "common.cuh"
#include "kernels.cuh"
#include <cub/cub.cuh>
#include <float.h>
#include <math.h>

// ---- ordered-int encoding so that double min/max can use 64-bit integer atomics ----
__device__ __forceinline__ long long dbl_to_ordered(double v) {
  long long b = __double_as_longlong(v);
  return b >= 0 ? b : (b ^ 0x7fffffffffffffffLL);
}
__device__ __forceinline__ double ordered_to_dbl(long long b) {
  return __longlong_as_double(b >= 0 ? b : (b ^ 0x7fffffffffffffffLL));
}

__global__ void bbox_init_kernel(long long* bb) {
  int t = threadIdx.x;
  if (t < 3) bb[t] = dbl_to_ordered(DBL_MAX);
  else if (t < 6) bb[t] = dbl_to_ordered(-DBL_MAX);
}

// points: (N,3) row-major.  Non-finite coordinates are ignored (they land in a border cell).
__global__ void __launch_bounds__(256) bbox_kernel(const double* __restrict__ pts, int64_t n, long long* bb) {
  double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      double v = pts[3 * i + a];
      if (isfinite(v)) { lo[a] = fmin(lo[a], v); hi[a] = fmax(hi[a], v); }
    }
  }
  typedef cub::BlockReduce<double, 256> BR;
  __shared__ typename BR::TempStorage tmp;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    double l = BR(tmp).Reduce(lo[a], cub::Min());
    __syncthreads();
    double h = BR(tmp).Reduce(hi[a], cub::Max());
    __syncthreads();
    if (threadIdx.x == 0) {
      atomicMin(&bb[a], dbl_to_ordered(l));
      atomicMax(&bb[3 + a], dbl_to_ordered(h));
    }
  }
}

__global__ void bbox_decode_kernel(const long long* bb, double* out) {
  int t = threadIdx.x;
  if (t < 6) out[t] = ordered_to_dbl(bb[t]);
}

__device__ __forceinline__ int cell_coord(double x, double o, double inv_c, int n) {
  double f = (x - o) * inv_c;
  int c = (f >= 0.0) ? ((f < (double)n) ? (int)f : n - 1) : 0;   // NaN -> 0
  return c;
}

__global__ void __launch_bounds__(256) cell_key_kernel(const double* __restrict__ pts, int64_t n, GridSpec g,
                                                       uint32_t* __restrict__ keys, uint32_t* __restrict__ perm,
                                                       uint32_t* __restrict__ counts) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cx = cell_coord(pts[3 * i + 0], g.ox, g.inv_c, g.nx);
  int cy = cell_coord(pts[3 * i + 1], g.oy, g.inv_c, g.ny);
  int cz = cell_coord(pts[3 * i + 2], g.oz, g.inv_c, g.nz);
  uint32_t key = ((uint32_t)cz * g.ny + cy) * g.nx + cx;
  keys[i] = key;
  perm[i] = (uint32_t)i;
  atomicAdd(&counts[key], 1u);
}

// sorted position records {x,y,z,0} (entry n is left to the caller) and the same coordinates as
// three arrays (the k-NN tiles are staged from these)
__global__ void __launch_bounds__(256) gather_pos_kernel(const double* __restrict__ pts, const uint32_t* __restrict__ perm,
                                                         int64_t n, double4* __restrict__ r0, double* __restrict__ xs,
                                                         double* __restrict__ ys, double* __restrict__ zs) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t p = perm[i];
  const double x = pts[3 * p], y = pts[3 * p + 1], z = pts[3 * p + 2];
  r0[i] = make_double4(x, y, z, 0.0);
  xs[i] = x; ys[i] = y; zs[i] = z;
}

int devbuf_reserve(DevBuf& b, size_t bytes) {
  if (bytes <= b.bytes) return 0;
  if (b.p) { cudaFree(b.p); b.p = nullptr; b.bytes = 0; }
  size_t want = bytes + bytes / 8 + 256;
  CUDA_TRY(cudaMalloc(&b.p, want));
  b.bytes = want;
  return 0;
}

void devbuf_free(DevBuf& b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.bytes = 0;
}

// Choose the grid for a bounding box.  `edge_hint` > 0 fixes the cell edge; otherwise the edge is
// set so that the mean occupancy of the bounding box is `occ` particles per cell.
static void choose_grid(const double* bb, int64_t n, double occ, double edge_hint, GridSpec* g) {
  double L[3];
  for (int a = 0; a < 3; ++a) {
    L[a] = bb[3 + a] - bb[a];
    if (!(L[a] > 0.0)) L[a] = 0.0;
  }
  double Lmax = fmax(L[0], fmax(L[1], L[2]));
  if (!(Lmax > 0.0)) Lmax = 1.0;
  // Degenerate inputs (particles on a plane or a line): a collapsed direction gets one layer of cells and
  // the occupancy is set over the directions that do extend, otherwise the cells come out far too small.
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
  const double max_cells = fmin(4.0 * (double)n + 4096.0, 1.0e9);
  for (;;) {
    double nx = floor(L[0] / c) + 1, ny = floor(L[1] / c) + 1, nz = floor(L[2] / c) + 1;
    if (nx * ny * nz <= max_cells && nx < 2.0e6 && ny < 2.0e6 && nz < 2.0e6) {
      g->nx = (int)nx; g->ny = (int)ny; g->nz = (int)nz;
      break;
    }
    c *= 1.1;
  }
  g->c = c;
  g->inv_c = 1.0 / c;
  g->ox = bb[0]; g->oy = bb[1]; g->oz = bb[2];
  g->ncells = g->nx * g->ny * g->nz;
}

// Build the cell list for `pts` (device, (N,3) f64).  On return ctx->keys holds sorted keys,
// ctx->perm the sorted->original permutation, ctx->cell_start the (ncells+1) exclusive scan and
// ctx->r0 the sorted {x,y,z,0} records.
int build_cells(sphb200_ctx* ctx, const double* pts, int64_t n, double edge_hint, cudaStream_t st) {
  if (n <= 0) SPH_FAIL("build_cells: empty particle set");
  if (n >= (int64_t)0x7fffffff) SPH_FAIL("build_cells: N=%lld exceeds the 2^31-1 particle limit per GPU", (long long)n);
  SPH_TRY(devbuf_reserve(ctx->scal, SC_COUNT * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->tmp0, 64));
  long long* bb_ord = bufp<long long>(ctx->tmp0);
  double* scal = bufp<double>(ctx->scal);
  bbox_init_kernel<<<1, 32, 0, st>>>(bb_ord);
  int blocks = ceil_div(n, 256); if (blocks > 148 * 8) blocks = 148 * 8;
  bbox_kernel<<<blocks, 256, 0, st>>>(pts, n, bb_ord);
  bbox_decode_kernel<<<1, 32, 0, st>>>(bb_ord, scal + SC_BBOX);
  ctx->launches += 3;
  double bb[6];
  CUDA_TRY(cudaMemcpyAsync(bb, scal + SC_BBOX, sizeof(bb), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  for (int a = 0; a < 3; ++a)
    if (!(bb[a] <= bb[3 + a])) SPH_FAIL("build_cells: no finite coordinates");
  choose_grid(bb, n, ctx->cell_occ, edge_hint, &ctx->grid);
  const GridSpec g = ctx->grid;

  SPH_TRY(devbuf_reserve(ctx->keys, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->keys_alt, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->perm, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->perm_alt, n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->cell_start, ((size_t)g.ncells + 2) * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->r0, (size_t)(n + 1) * sizeof(double4)));
  uint32_t* cs = bufp<uint32_t>(ctx->cell_start);
  CUDA_TRY(cudaMemsetAsync(cs, 0, ((size_t)g.ncells + 2) * sizeof(uint32_t), st));
  cell_key_kernel<<<ceil_div(n, 256), 256, 0, st>>>(pts, n, g, bufp<uint32_t>(ctx->keys_alt), bufp<uint32_t>(ctx->perm_alt), cs);
  ctx->launches += 1;

  int bits = 1;
  while ((1ll << bits) < (long long)g.ncells) ++bits;
  size_t tmp_sort = 0, tmp_scan = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, bufp<uint32_t>(ctx->keys_alt), bufp<uint32_t>(ctx->keys),
                                  bufp<uint32_t>(ctx->perm_alt), bufp<uint32_t>(ctx->perm), (int)n, 0, bits, st);
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, cs, cs, g.ncells + 1, st);
  size_t tmp_bytes = tmp_sort > tmp_scan ? tmp_sort : tmp_scan;
  SPH_TRY(devbuf_reserve(ctx->cub_tmp, tmp_bytes));
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp_sort, bufp<uint32_t>(ctx->keys_alt), bufp<uint32_t>(ctx->keys),
                                           bufp<uint32_t>(ctx->perm_alt), bufp<uint32_t>(ctx->perm), (int)n, 0, bits, st));
  CUDA_TRY(cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp_scan, cs, cs, g.ncells + 1, st));
  ctx->launches += 6;   // radix sort (histogram + onesweep passes) + scan, approximate
  const size_t ns = soa_stride(n);
  SPH_TRY(devbuf_reserve(ctx->soa, 3 * ns * sizeof(double)));
  double* xs = bufp<double>(ctx->soa);
  gather_pos_kernel<<<ceil_div(n, 256), 256, 0, st>>>(pts, bufp<uint32_t>(ctx->perm), n, bufp<double4>(ctx->r0), xs, xs + ns,
                                                      xs + 2 * ns);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
