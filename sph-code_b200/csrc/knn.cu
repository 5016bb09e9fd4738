// Exact fixed-k nearest-neighbour search on the cell list.
// Replaces `kdt.query(points, N_NEIGH, eps, 2, dist)` of
// /root/reference/sph/navier_stokes_cleaned.py:544 (with eps = 0) and the row-max
// smoothing length of :548.
//
// One group of G lanes per target particle.  Lanes stride over the contiguous runs of sorted
// particles that make up the (2m+1)^3 cell block around the target (cells along x are contiguous in
// the sort order, so a block is (2m+1)^2 runs), test d^2 < R_try^2 and compact the survivors into a
// per-group shared-memory buffer with ballot/popc.  R_try starts at the radius expected to hold
// ~1.35 k particles at the block's mean density and is corrected (shrunk on overflow, grown when
// fewer than k survive) until it holds between k and CAP candidates; it never exceeds the radius
// R_safe = m*c that the searched block is guaranteed to cover, so the result is exact.  The k
// smallest are then ranked by counting (rank = number of smaller keys), which also yields the
// ascending-distance row order the reference's API returns.
//
// d^2 is evaluated as ((dx*dx + dy*dy) + dz*dz) with separately rounded products (no FMA
// contraction) — the same arithmetic as numpy and cKDTree — so that the ordering of near-equal
// distances and the returned h = sqrt(d^2_k) are bit-identical to the CPU oracle.
#include "common.cuh"
#include "kernels.cuh"
#include "knn_select.cuh"
#include <cooperative_groups.h>
#include <cooperative_groups/reduce.h>
#include <float.h>
namespace cg = cooperative_groups;

#define KNN_THREADS 128
#define KNN_MAX_ITER 96

template <int G>
__global__ void __launch_bounds__(KNN_THREADS, 8)
knn_kernel(const double4* __restrict__ r0, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ cell_start,
           const GridSpec* __restrict__ spec, int n, int k, double dist2, int cap, int use_guess, double guess_factor,
           int32_t* __restrict__ list, double* __restrict__ dists, double* __restrict__ h_out,
           int32_t* __restrict__ nontriv, int* __restrict__ flags, const int* __restrict__ worklist,
           const int* __restrict__ nwork, int by_slot, const uint32_t* __restrict__ skip_id, uint32_t skip_from) {
  const GridSpec g = *spec;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cg::thread_block blk = cg::this_thread_block();
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(blk);
  const int lane = tile.thread_rank();
  const int grp = tile.meta_group_rank();
  constexpr int GROUPS = KNN_THREADS / G;
  unsigned long long* keybuf = reinterpret_cast<unsigned long long*>(smem_raw) + (size_t)grp * cap;
  int32_t* idxbuf = reinterpret_cast<int32_t*>(smem_raw + (size_t)GROUPS * cap * sizeof(unsigned long long)) + (size_t)grp * cap;

  // targets: all particles, or the entries of a device-side work list (grid-stride)
  const int ntarget = worklist ? *nwork : n;
  for (int tgt = blockIdx.x * GROUPS + grp; tgt < ntarget; tgt += gridDim.x * GROUPS) {
  const int i = worklist ? worklist[tgt] : tgt;
  const int64_t row = (int64_t)(by_slot ? tgt : i) * k;     // session: escaped rows are stored by work-list slot
  const double4 p = r0[i];
  const uint32_t key = keys[i];
  const int cx = key % g.nx, cy = (key / g.nx) % g.ny, cz = key / (g.nx * g.ny);
  const double c3 = g.c * g.c * g.c;

  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) {   // NaN/inf target: no neighbours (cKDTree raises)
    for (int s = lane; s < k; s += G) { list[row + s] = n; if (dists) dists[row + s] = 0.0; }
    if (lane == 0) { nontriv[i] = 0; h_out[i] = 0.0; atomicAdd(&flags[FL_KNN_NONFINITE], 1); }
    continue;
  }
  if (skip_id && skip_id[i] >= skip_from) {                   // ghost of a slab decomposition: an empty, well-defined row
    for (int s = lane; s < k; s += G) { list[row + s] = n; if (dists) dists[row + s] = 0.0; }
    if (lane == 0) { nontriv[i] = 0; h_out[i] = p.w; }
    continue;
  }
  int m = 1;
  double R2try = -1.0;
  if (use_guess) { double hg = p.w * guess_factor; if (hg > 0.0) { R2try = hg * hg; m = (int)fmin(hg * g.inv_c, 1.0e6) + 1; } }
  int count = 0;
  bool done = false;
  int iter = 0, rescans = -1, widened = 0;
  for (; iter < KNN_MAX_ITER && !done; ++iter) {
    const int x0 = max(cx - m, 0), x1 = min(cx + m, g.nx - 1);
    const int y0 = max(cy - m, 0), y1 = min(cy + m, g.ny - 1);
    const int z0 = max(cz - m, 0), z1 = min(cz + m, g.nz - 1);
    const bool whole = (x0 == 0 && y0 == 0 && z0 == 0 && x1 == g.nx - 1 && y1 == g.ny - 1 && z1 == g.nz - 1);
    const double rsafe = (double)m * g.c * (1.0 - 1e-8);
    const double R2lim = whole ? dist2 : fmin(dist2, rsafe * rsafe);
    const int ysp = y1 - y0 + 1;
    const int nrows = ysp * (z1 - z0 + 1);
    if (R2try < 0.0) {
      int cnt = 0;
      for (int r = lane; r < nrows; r += G) {
        const int base = ((z0 + r / ysp) * g.ny + (y0 + r % ysp)) * g.nx;
        cnt += (int)(cell_start[base + x1 + 1] - cell_start[base + x0]);
      }
      cnt = cg::reduce(tile, cnt, cg::plus<int>());
      if (cnt < k && !whole) { ++m; widened = 1; continue; }
      const double vol = (double)(x1 - x0 + 1) * (double)nrows * c3;
      const double R3 = 1.35 * (double)k * vol / (4.1887902047863905 * (double)(cnt > 0 ? cnt : 1));
      const double Rt = cbrt(R3);
      R2try = fmin(R2lim, Rt * Rt);
    } else if (R2try > R2lim) {
      R2try = R2lim;
    }
    // ---- scan the block ----
    // The targets that reach this kernel are few (one warp each, nearly empty GPU), so the scan is bound by
    // load latency: the row bounds of up to G rows are fetched by the lanes at once, and two candidate
    // batches are in flight per trip.
    // Only the cells the trial sphere can reach are scanned: a row (y, z) whose cell column is farther than R_try
    // from the target is skipped, the others are cut to the chord of the sphere along x.  For a block of (2m+1)^3
    // cells this leaves about the sphere's share of the cube (a third of the candidates at m = 3).  Distances are
    // taken in cell units with the expression that assigned the particles to their cells, minus a safety margin;
    // the clamped boundary cells only hold particles that are farther away than their cell, so the cut is safe.
    ++rescans;
    count = 0;
    for (int rb = 0; rb < nrows; rb += G) {
      int my_s = 0, my_e = 0;
      if (rb + lane < nrows) {
        const int r = rb + lane;
        const int yy = y0 + r % ysp, zz = z0 + r / ysp;
        const int base = (zz * g.ny + yy) * g.nx;
        const double fx = (p.x - g.ox) * g.inv_c, fy = (p.y - g.oy) * g.inv_c, fz = (p.z - g.oz) * g.inv_c;
        const double Rc2 = R2try * g.inv_c * g.inv_c * (1.0 + 1e-9);
        const double dyc = fmax(fmax((double)yy - fy, fy - (double)(yy + 1)) - 1e-6, 0.0);
        const double dzc = fmax(fmax((double)zz - fz, fz - (double)(zz + 1)) - 1e-6, 0.0);
        const double rem = Rc2 - dyc * dyc - dzc * dzc;
        if (rem > 0.0) {
          const double w = sqrt(rem) + 1e-6;
          int xa = x0, xb = x1;
          if (w < 1.0e9) { xa = max(x0, (int)floor(fx - w)); xb = min(x1, (int)floor(fx + w)); }
          if (xa <= xb) { my_s = (int)cell_start[base + xa]; my_e = (int)cell_start[base + xb + 1]; }
        }
      }
      // The (cut) rows of this batch are short - a few particles each where the grid is finer than the smoothing
      // length - so they are scanned as ONE flattened sequence: an inclusive scan of the row lengths over the lanes,
      // then candidate t of the batch is found by a binary search in that prefix (5 shuffles) instead of spending a
      // trip of 2 G lanes on every row.  Candidates still arrive in ascending sorted index.
      const int len = my_e - my_s;
      int inc = len;
#pragma unroll
      for (int d = 1; d < G; d <<= 1) { const int v = tile.shfl_up(inc, d); if (lane >= d) inc += v; }
      const int T = tile.shfl(inc, G - 1);
      const int exc = inc - len;
      for (int t0 = 0; t0 < T; t0 += 2 * G) {
        int jj[2]; bool vv[2];
        double4 qq[2];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int t = t0 + half * G + lane;
          vv[half] = t < T;
          const int tc = vv[half] ? t : T - 1;
          int lo = 0;
#pragma unroll
          for (int step = G / 2; step > 0; step >>= 1) { const int v = tile.shfl(inc, lo + step - 1); if (v <= tc) lo += step; }
          jj[half] = tile.shfl(my_s, lo) + (tc - tile.shfl(exc, lo));
          qq[half] = make_double4(0, 0, 0, 0);
          if (vv[half]) qq[half] = r0[jj[half]];
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          bool ok = false;
          double d2 = 0.0;
          if (vv[half]) {
            const double4 q = qq[half];
            const double dx = q.x - p.x, dy = q.y - p.y, dz = q.z - p.z;
            d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
            ok = d2 < R2try;
          }
          const unsigned b = tile.ballot(ok);
          if (b) {
            const int pos = count + __popc(b & ((1u << lane) - 1u));
            if (ok && pos < cap) { keybuf[pos] = (unsigned long long)__double_as_longlong(d2); idxbuf[pos] = jj[half]; }
            count += __popc(b);
          }
        }
      }
    }
    tile.sync();
    if (count > cap) {                     // too many survivors: shrink the trial radius
      const double f = cbrt(0.6 * (double)cap / (double)count);
      R2try *= f * f;
      continue;
    }
    if (count >= k) { done = true; break; }
    if (R2try < R2lim) {                    // too few: grow towards the covered radius
      const double f = cbrt(1.6 * (double)k / (double)(count > 0 ? count : 1));
      R2try = fmin(R2lim, R2try * fmax(f * f, 1.2));
      continue;
    }
    if (whole || R2lim >= dist2) { done = true; break; }   // everything inside the cut was seen
    ++m; widened = 1;                       // block too small: widen, re-estimate
    R2try = -1.0;
  }
  if (!done) {
    // The radius search did not settle: more candidates share one distance than the buffer holds
    // (coincident or lattice points), so no radius separates k of them.  Select by successive minima of
    // (d^2, index) over the block instead: exact for ties, O(k x candidates) and only ever used here.
    const int x0 = max(cx - m, 0), x1 = min(cx + m, g.nx - 1);
    const int y0 = max(cy - m, 0), y1 = min(cy + m, g.ny - 1);
    const int z0 = max(cz - m, 0), z1 = min(cz + m, g.nz - 1);
    const bool whole = (x0 == 0 && y0 == 0 && z0 == 0 && x1 == g.nx - 1 && y1 == g.ny - 1 && z1 == g.nz - 1);
    const double rsafe = (double)m * g.c * (1.0 - 1e-8);
    const double R2lim = whole ? dist2 : fmin(dist2, rsafe * rsafe);
    const int ysp = y1 - y0 + 1;
    const int nrows = ysp * (z1 - z0 + 1);
    unsigned long long last_k = 0ull; int last_j = -1;
    int nsel = 0;
    for (; nsel < k; ++nsel) {
      unsigned long long best_k = ~0ull; int best_j = 0x7fffffff;
      for (int r = 0; r < nrows; ++r) {
        const int base = ((z0 + r / ysp) * g.ny + (y0 + r % ysp)) * g.nx;
        const int s = (int)cell_start[base + x0], e = (int)cell_start[base + x1 + 1];
        for (int j = s + lane; j < e; j += G) {
          const double4 q = r0[j];
          const double dx = q.x - p.x, dy = q.y - p.y, dz = q.z - p.z;
          const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
          if (!(d2 < R2lim)) continue;
          const unsigned long long kk = (unsigned long long)__double_as_longlong(d2);
          const bool after = (last_j < 0) || kk > last_k || (kk == last_k && j > last_j);
          if (after && (kk < best_k || (kk == best_k && j < best_j))) { best_k = kk; best_j = j; }
        }
      }
      const unsigned long long mk = cg::reduce(tile, best_k, cg::less<unsigned long long>());
      const int mj = cg::reduce(tile, (best_k == mk) ? best_j : 0x7fffffff, cg::less<int>());
      if (mj == 0x7fffffff) break;                          // nothing left inside the covered radius
      last_k = mk; last_j = mj;
      if (lane == 0) { list[row + nsel] = mj; if (dists) dists[row + nsel] = sqrt(__longlong_as_double((long long)mk)); }
    }
    const bool exact = (nsel == k) || whole || R2lim >= dist2;   // else the block did not reach the k-th neighbour
    for (int s = nsel + lane; s < k; s += G) { list[row + s] = n; if (dists) dists[row + s] = 0.0; }
    if (lane == 0) {
      nontriv[i] = nsel;
      h_out[i] = nsel > 0 ? sqrt(__longlong_as_double((long long)last_k)) : 0.0;
      if (!exact) atomicAdd(&flags[FL_KNN_OVERFLOW], 1);
      if (rescans > 0) atomicAdd(&flags[FL_KNN_RESCAN], rescans);
      atomicAdd(&flags[FL_KNN_WIDEN], 1);
    }
    tile.sync();
    continue;
  }
  if (lane == 0) {
    if (rescans > 0) atomicAdd(&flags[FL_KNN_RESCAN], rescans);
    if (widened) atomicAdd(&flags[FL_KNN_WIDEN], 1);
  }
  if (count > cap) count = cap;
  // ---- session rows (work list of the fast path): an unordered set with the nearest candidate in column 0 ----
  // Ranking by counting costs count^2 / 32 comparisons per lane (half of this kernel's instructions when most
  // targets of a clustered distribution arrive here); the register-resident removal selection of the fast path
  // (knn_select.cuh) needs a tenth of that.  It works on POSITIONS in the candidate buffer, which ascend with the
  // sorted index (rows are scanned in ascending order), so ties in d^2 break as the ranking would break them.
  if (G == 32 && by_slot && !dists && count >= k && count <= 32 * KG_SLICES) {
    uint16_t* rowsm = reinterpret_cast<uint16_t*>(smem_raw + (size_t)GROUPS * cap * 12) + (size_t)grp * 64;
    const int nsl = (count + 31) >> 5;
    double kr[KG_SLICES]; int ir[KG_SLICES];
    unsigned long long kmin = ~0ull;
#pragma unroll
    for (int r = 0; r < KG_SLICES; ++r) {
      const int e = 32 * r + lane;
      kr[r] = -1.0; ir[r] = -1;
      if (e < count) { const unsigned long long kb = keybuf[e]; kr[r] = __longlong_as_double((long long)kb); ir[r] = e; if (kb < kmin) kmin = kb; }
    }
    const unsigned long long kall = cg::reduce(tile, kmin, cg::less<unsigned long long>());
    int pmin = 0x7fffffff;                                    // first position that holds the smallest key
#pragma unroll
    for (int r = KG_SLICES - 1; r >= 0; --r)
      if (ir[r] >= 0 && (unsigned long long)__double_as_longlong(kr[r]) == kall) pmin = ir[r];
    pmin = cg::reduce(tile, pmin, cg::less<int>());
    KnnGroupArgs fa;
    fa.k = k; fa.list16 = rowsm;
    bool ok;
    double h2;
    if (nsl <= 2 && count - k <= KG_FAST_EXCESS) h2 = knn_select2(fa, 0, pmin, count, kr[0], kr[1], ir[0], ir[1], lane, ok);
    else h2 = knn_select16(fa, 0, pmin, count, nsl, R2try, kr, ir, lane, ok);
    tile.sync();
    if (ok) {
      for (int s2 = lane; s2 < k; s2 += G) list[row + s2] = idxbuf[rowsm[s2]];
      if (lane == 0) { nontriv[i] = k; h_out[i] = sqrt(h2); }
      tile.sync();
      continue;
    }
  }
  // ---- rank by counting; rank < k are the neighbours, in ascending (d^2, index) order ----
  const int nsel = count < k ? count : k;
  for (int e = lane; e < count; e += G) {
    const unsigned long long ke = keybuf[e];
    const int je = idxbuf[e];
    int rank = 0;
    for (int f = 0; f < count; ++f) {
      const unsigned long long kf = keybuf[f];
      rank += (kf < ke) || (kf == ke && idxbuf[f] < je);
    }
    if (rank < k) {
      list[row + rank] = je;
      const double dd = sqrt(__longlong_as_double((long long)ke));
      if (dists) dists[row + rank] = dd;
      if (rank == nsel - 1) h_out[i] = dd;
    }
  }
  for (int s = count + lane; s < k; s += G) {
    list[row + s] = n;
    if (dists) dists[row + s] = 0.0;
  }
  if (lane == 0) {
    nontriv[i] = nsel;
    if (nsel == 0) h_out[i] = 0.0;
  }
  tile.sync();
  }  // target loop
}

int knn_cap_for(int k) {
  int cap = ((3 * k + 31) / 32) * 32;
  if (cap < 64) cap = 64;
  return cap;
}

static int launch_knn_impl(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double guess_factor, int32_t* list, double* dists,
                           double* h_out, int32_t* nontriv, const int* worklist, const int* nwork, int by_slot, cudaStream_t st) {
  if (k < 1) SPH_FAIL("knn: k must be >= 1");
  const int cap = knn_cap_for(k);
  const int G = ctx->knn_group;
  const int groups = KNN_THREADS / G;
  const size_t smem = (size_t)groups * cap * 12 + (size_t)groups * 128;   // candidates + one uint16 row per group (session rows)
  if (smem > 200 * 1024) SPH_FAIL("knn: k=%d needs %zu B of shared memory per block", k, smem);
  double dist2 = dist * dist;                 // cKDTree squares the upper bound for p=2
  if (!(dist2 == dist2)) dist2 = INFINITY;
  const double4* r0 = bufp<double4>(ctx->r0);
  const uint32_t* keys = bufp<uint32_t>(ctx->keys);
  const uint32_t* cs = bufp<uint32_t>(ctx->cell_start);
  int* flags = bufp<int>(ctx->flags);
  const int blocks = worklist ? 148 * 8 : ceil_div(n, groups);
#define KNN_LAUNCH(GG)                                                                                         \
  do {                                                                                                         \
    if (smem > 48 * 1024)                                                                                      \
      CUDA_TRY(cudaFuncSetAttribute(knn_kernel<GG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));  \
    knn_kernel<GG><<<blocks, KNN_THREADS, smem, st>>>(r0, keys, cs, bufp<GridSpec>(ctx->grid_dev), n, k, dist2, cap, use_guess, \
                                                      guess_factor, list, dists, h_out, nontriv, flags, worklist, nwork, by_slot, \
                                                      ctx->skip_id, ctx->skip_from);                           \
  } while (0)
  if (G == 32) KNN_LAUNCH(32);
  else if (G == 16) KNN_LAUNCH(16);
  else if (G == 8) KNN_LAUNCH(8);
  else SPH_FAIL("knn: unsupported group size %d", G);
#undef KNN_LAUNCH
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int launch_knn(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double guess_factor, int32_t* list, double* dists,
               double* h_out, int32_t* nontriv, cudaStream_t st) {
  return launch_knn_impl(ctx, n, k, dist, use_guess, guess_factor, list, dists, h_out, nontriv, nullptr, nullptr, 0, st);
}

// use_guess: the records carry the previous k-th distances (r0[i].w); the first trial radius is guess *
// guess_factor, which sizes the block and normally settles the target in one scan.  by_slot: row of work-list
// entry t goes to list[t * k] (session: the global list holds only the escaped rows) instead of list[i * k].
int launch_knn_worklist(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double guess_factor,
                        const int* worklist, const int* nwork, int32_t* list, double* dists, double* h_out, int32_t* nontriv,
                        int by_slot, cudaStream_t st) {
  return launch_knn_impl(ctx, n, k, dist, use_guess, guess_factor, list, dists, h_out, nontriv, worklist, nwork, by_slot, st);
}

// Sorted-space rows -> the reference's return layout (original particle order, int64 indices,
// missing neighbour = N / distance 0).                                   nsc:545-548,552
__global__ void __launch_bounds__(256)
knn_unsort_kernel(const int32_t* __restrict__ list, const double* __restrict__ dists, const double* __restrict__ h,
                  const int32_t* __restrict__ nontriv, const uint32_t* __restrict__ perm, int n, int k,
                  int64_t* __restrict__ idx_out, double* __restrict__ dist_out, int64_t* __restrict__ nontriv_out,
                  double* __restrict__ h_out) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * k) return;
  const int i = (int)(t / k), s = (int)(t % k);
  const int64_t o = (int64_t)perm[i] * k + s;
  const int j = list[t];
  idx_out[o] = (j < n) ? (int64_t)perm[j] : (int64_t)n;
  dist_out[o] = dists[t];
  if (s == 0) {
    nontriv_out[perm[i]] = nontriv[i];
    h_out[perm[i]] = h[i];
  }
}

int launch_knn_unsort(sphb200_ctx* ctx, const int32_t* list, const double* dists, const double* h, const int32_t* nontriv,
                      int n, int k, int64_t* idx_out, double* dist_out, int64_t* nontriv_out, double* h_out,
                      cudaStream_t st) {
  knn_unsort_kernel<<<ceil_div((int64_t)n * k, 256), 256, 0, st>>>(list, dists, h, nontriv, bufp<uint32_t>(ctx->perm), n, k,
                                                                   idx_out, dist_out, nontriv_out, h_out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
