// Cell-tiled exact fixed-k nearest-neighbour search (the fast path of the neighbour query).
// Replaces `kdt.query(points, N_NEIGH, eps, 2, dist)` of
// /root/reference/sph/navier_stokes_cleaned.py:544 (eps = 0) and the row-max
// smoothing length of :548.
//
// One CTA (4 warps) per grid cell.  The candidates of a cell are the particles of its 3x3x3 cell
// block; cells along x are contiguous in the sort order, so the block is nine contiguous runs of the
// sorted coordinate arrays xs[], ys[], zs[] (struct-of-arrays so that the shared-memory reads below
// are conflict-free).  The nine lanes holding the runs stage them into shared memory with bulk
// asynchronous copies (cp.async.bulk -> UBLKCP in SASS, completion on an mbarrier); runs are widened
// to even bounds for the 16-byte alignment the copy engine needs and the padding entries are masked.
// Each warp then takes one target particle of the cell at a time: its 32 lanes stride over the staged
// tile, test d^2 < R_try^2 and compact the survivors into a per-warp buffer with ballot/popc.  R_try
// is the previous step's k-th distance times (1 + margin) when the caller has one, else the radius
// expected to hold ~1.3 k particles at the block's density; it is corrected (rescan) until the buffer
// holds between k and CAP entries and never exceeds the target's distance to the faces of the cell
// block, so the result is exact.  The k nearest are then picked by removing the (count - k) largest
// with warp-shuffle max reductions (session layout: unordered row, self in column 0), or ranked by
// counting (API layout: ascending distance, distances returned).  Targets whose k-th neighbour may lie
// outside the block go to a work list that the general warp-per-target kernel of knn.cu resolves.
//
// d^2 = ((dx*dx + dy*dy) + dz*dz) with separately rounded products (no FMA contraction), the same
// arithmetic as cKDTree and numpy, so neighbour sets, order and h = sqrt(d^2_k) are bit-identical to
// the CPU oracle.
#include "common.cuh"
#include "kernels.cuh"
#include <float.h>

#define KC_CAP 64          // entries of a warp's survivor buffer
#define KC_CHP 640         // padded candidates staged per chunk (multiple of 32)
#define KC_WARPS 4
#define KC_MAX_ATTEMPT 24

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
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
// global -> shared bulk copy (TMA engine, non-tensor form); bytes % 16 == 0, both sides 16 B aligned
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }


struct KnnCellArgs {
  const double* xs; const double* ys; const double* zs;   // sorted coordinates (SoA), n + 2 entries each
  const uint32_t* cell_start;
  GridSpec g;
  int n, k;
  double dist2;
  const double* h_guess;     // previous k-th distance per (sorted) particle, or nullptr
  double guess_factor;       // (1 + margin)
  int32_t* list;             // (N,k) neighbour rows, indices in sorted order; missing = n
  double* dists;             // (N,k) or nullptr (only with SORTED)
  double* h_out;             // (N)
  int32_t* nontriv;          // (N)
  int* flags;
  int* worklist;             // unresolved targets
  int* nwork;
};

// Shared-memory tile of one chunk of the cell block's candidates.
struct KnnTile {
  double* X; double* Y; double* Z;   // KC_CHP each
  int32_t* G;                        // sorted index of the entry, or -1 for alignment padding
};

// The nine runs of the 3x3x3 block, one per lane (lanes >= nrows hold empty runs).
struct KnnRuns {
  int ps, plen, poff;     // padded run: first sorted index (even), length (even), offset in the padded tile
  int rs, re;             // real run [rs, re)
  int ptotal;             // padded candidates in the block (all runs)
  int total;              // real candidates
};

__device__ __forceinline__ KnnRuns knn_runs(const KnnCellArgs& a, int lane, int x0, int x1, int y0, int z0, int ysp, int nrows) {
  KnnRuns r;
  r.rs = 0; r.re = 0;
  if (lane < nrows) {
    const int base = ((z0 + lane / ysp) * a.g.ny + (y0 + lane % ysp)) * a.g.nx;
    r.rs = (int)a.cell_start[base + x0];
    r.re = (int)a.cell_start[base + x1 + 1];
  }
  r.ps = r.rs & ~1;
  r.plen = (r.re > r.rs) ? (((r.re + 1) & ~1) - r.ps) : 0;
  int inc = r.plen, tot = r.re - r.rs;
#pragma unroll
  for (int d = 1; d < 16; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, d);
    const int w = __shfl_up_sync(0xffffffffu, tot, d);
    if (lane >= d) { inc += v; tot += w; }
  }
  r.ptotal = __shfl_sync(0xffffffffu, inc, 15);
  r.total = __shfl_sync(0xffffffffu, tot, 15);
  r.poff = inc - r.plen;
  return r;
}

// Stage padded candidates [c_begin, c_begin + c_n) of the block into the tile.  Called by warp 0 only
// (all 32 lanes); the other warps wait on the CTA barrier that follows.
__device__ __forceinline__ void knn_stage(const KnnCellArgs& a, const KnnRuns& r, const KnnTile& t, uint64_t* bar,
                                          uint32_t& phase, int lane, int c_begin, int c_n) {
  fence_proxy_async();
  if (lane == 0) mbar_expect_tx(bar, (uint32_t)c_n * 24u);
  __syncwarp();
  const int lo = max(r.poff, c_begin), hi = min(r.poff + r.plen, c_begin + c_n);   // even bounds (c_begin, c_n even)
  if (hi > lo) {
    const int src = r.ps + (lo - r.poff);
    const uint32_t bytes = (uint32_t)(hi - lo) * 8u;
    bulk_g2s(t.X + (lo - c_begin), a.xs + src, bytes, bar);
    bulk_g2s(t.Y + (lo - c_begin), a.ys + src, bytes, bar);
    bulk_g2s(t.Z + (lo - c_begin), a.zs + src, bytes, bar);
  }
  for (int run = 0; run < 9; ++run) {                     // index table, filled by the whole warp
    const int q_ps = __shfl_sync(0xffffffffu, r.ps, run), q_off = __shfl_sync(0xffffffffu, r.poff, run);
    const int q_len = __shfl_sync(0xffffffffu, r.plen, run);
    const int q_rs = __shfl_sync(0xffffffffu, r.rs, run), q_re = __shfl_sync(0xffffffffu, r.re, run);
    const int qlo = max(q_off, c_begin), qhi = min(q_off + q_len, c_begin + c_n);
    for (int c = qlo + lane; c < qhi; c += 32) {
      const int gi = q_ps + (c - q_off);
      t.G[c - c_begin] = (gi >= q_rs && gi < q_re) ? gi : -1;   // alignment padding is masked
    }
  }
  mbar_wait(bar, phase);
  phase ^= 1u;
  if (hi > lo) {                                          // padding never passes `d2 < R2` (NaN)
    if (r.ps < r.rs && r.poff >= lo) t.X[r.poff - c_begin] = nan("");
    const int last = r.poff + r.plen - 1;
    if (r.ps + r.plen > r.re && last < hi) t.X[last - c_begin] = nan("");
  }
}

// One warp scans the staged chunk for one target: survivors (d^2 < R2try) are appended to the warp's
// buffer in tile order.  Returns the updated survivor count (may exceed KC_CAP: overflow is counted,
// not stored).
__device__ __forceinline__ int knn_scan_chunk(const KnnTile& t, int c_n, double px, double py, double pz, double R2try,
                                              int count, double* keybuf, int32_t* idxbuf, int lane) {
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll 2
  for (int c0 = 0; c0 < c_n; c0 += 32) {
    const int c = c0 + lane;
    bool ok = false;
    double d2 = 0.0;
    if (c < c_n) {
      const double dx = t.X[c] - px, dy = t.Y[c] - py, dz = t.Z[c] - pz;
      d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
      ok = d2 < R2try;
    }
    const unsigned b = __ballot_sync(0xffffffffu, ok);
    if (b) {
      const int pos = count + __popc(b & lt);
      if (ok && pos < KC_CAP) { keybuf[pos] = d2; idxbuf[pos] = t.G[c]; }
      count += __popc(b);
    }
  }
  return count;
}

// Warp-wide maximum of (key, idx) pairs in lexicographic order.
__device__ __forceinline__ void warp_max_pair(double& k, int& i) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const double ko = __shfl_xor_sync(0xffffffffu, k, d);
    const int io = __shfl_xor_sync(0xffffffffu, i, d);
    if (ko > k || (ko == k && io > i)) { k = ko; i = io; }
  }
}

// The warp's buffer holds `count` (<= KC_CAP) survivors of target i; emit its row.
template <bool SORTED>
__device__ __forceinline__ void knn_emit(const KnnCellArgs& a, int i, int count, const double* keybuf, const int32_t* idxbuf,
                                         int lane) {
  const int k = a.k;
  const int nsel = min(count, k);
  const int64_t row = (int64_t)i * k;
  __syncwarp();
  if (SORTED) {
    // API layout: rank by counting; rank < k are the neighbours in ascending (d^2, index) order
    for (int e = lane; e < count; e += 32) {
      const double ke = keybuf[e]; const int je = idxbuf[e];
      int rank = 0;
      for (int f = 0; f < count; ++f) {
        const double kf = keybuf[f];
        rank += (kf < ke) || (kf == ke && idxbuf[f] < je);
      }
      if (rank < k) {
        a.list[row + rank] = je;
        const double dd = sqrt(ke);
        a.dists[row + rank] = dd;
        if (rank == nsel - 1) a.h_out[i] = dd;
      }
    }
    for (int s = nsel + lane; s < k; s += 32) { a.list[row + s] = a.n; a.dists[row + s] = 0.0; }
  } else {
    // session layout: drop the (count - k) largest, keep the rest unordered with self in column 0
    double k0 = (lane < count) ? keybuf[lane] : -1.0, k1 = (lane + 32 < count) ? keybuf[lane + 32] : -1.0;
    int i0 = (lane < count) ? idxbuf[lane] : -1, i1 = (lane + 32 < count) ? idxbuf[lane + 32] : -1;
    for (int r = count; r > nsel; --r) {
      double km; int im;
      if (k1 > k0 || (k1 == k0 && i1 > i0)) { km = k1; im = i1; } else { km = k0; im = i0; }
      warp_max_pair(km, im);
      if (k0 == km && i0 == im) { k0 = -1.0; i0 = -1; }
      else if (k1 == km && i1 == im) { k1 = -1.0; i1 = -1; }
    }
    double km = fmax(k0, k1); int im = 0;
    warp_max_pair(km, im);
    const unsigned lt = (1u << lane) - 1u;
    const unsigned b0 = __ballot_sync(0xffffffffu, i0 >= 0), b1 = __ballot_sync(0xffffffffu, i1 >= 0);
    const unsigned s0 = __ballot_sync(0xffffffffu, i0 == i), s1 = __ballot_sync(0xffffffffu, i1 == i);
    int ps = -1;                                             // position of self among the kept entries
    if (s0) ps = __popc(b0 & ((1u << (__ffs(s0) - 1)) - 1u));
    else if (s1) ps = __popc(b0) + __popc(b1 & ((1u << (__ffs(s1) - 1)) - 1u));
    if (i0 >= 0) { const int pos = __popc(b0 & lt); a.list[row + ((pos == ps) ? 0 : (pos < ps ? pos + 1 : pos))] = i0; }
    if (i1 >= 0) { const int pos = __popc(b0) + __popc(b1 & lt); a.list[row + ((pos == ps) ? 0 : (pos < ps ? pos + 1 : pos))] = i1; }
    for (int s = nsel + lane; s < k; s += 32) a.list[row + s] = a.n;
    if (lane == 0) a.h_out[i] = (nsel > 0) ? sqrt(km) : 0.0;
  }
  if (lane == 0) { a.nontriv[i] = nsel; if (SORTED && nsel == 0) a.h_out[i] = 0.0; }
  __syncwarp();
}

struct KnnBlock {          // geometry of the 3x3x3 block of one cell (warp-uniform)
  int x0, x1, y0, y1, z0, z1;
  bool whole;              // the block is the whole grid
  double vol;              // its volume
};

struct KnnTarget {
  double px, py, pz, R2lim, R2try;
  bool active, unresolved, finite;
  int count, rescans;
};

__device__ __forceinline__ KnnTarget knn_target_init(const KnnCellArgs& a, const KnnBlock& b, int total, int i, bool valid,
                                                     int lane) {
  KnnTarget t;
  t.px = t.py = t.pz = 0.0; t.R2lim = a.dist2; t.R2try = 0.0;
  t.active = valid; t.unresolved = false; t.count = 0; t.rescans = -1;
  if (valid) { t.px = a.xs[i]; t.py = a.ys[i]; t.pz = a.zs[i]; }
  if (valid && !(isfinite(t.px) && isfinite(t.py) && isfinite(t.pz))) {      // cKDTree raises on these
    t.active = false;
    for (int s = lane; s < a.k; s += 32) { a.list[(int64_t)i * a.k + s] = a.n; if (a.dists) a.dists[(int64_t)i * a.k + s] = 0.0; }
    if (lane == 0) { a.nontriv[i] = 0; a.h_out[i] = 0.0; atomicAdd(&a.flags[FL_KNN_NONFINITE], 1); }
  }
  t.finite = t.active;
  if (t.active) {
    const GridSpec& g = a.g;
    if (!b.whole) {
      // distance to the nearest face of the cell block that is not a face of the whole grid
      double rs = DBL_MAX;
      if (b.x0 > 0) rs = fmin(rs, t.px - (g.ox + (double)b.x0 * g.c));
      if (b.x1 < g.nx - 1) rs = fmin(rs, (g.ox + (double)(b.x1 + 1) * g.c) - t.px);
      if (b.y0 > 0) rs = fmin(rs, t.py - (g.oy + (double)b.y0 * g.c));
      if (b.y1 < g.ny - 1) rs = fmin(rs, (g.oy + (double)(b.y1 + 1) * g.c) - t.py);
      if (b.z0 > 0) rs = fmin(rs, t.pz - (g.oz + (double)b.z0 * g.c));
      if (b.z1 < g.nz - 1) rs = fmin(rs, (g.oz + (double)(b.z1 + 1) * g.c) - t.pz);
      rs = fmax(rs, 0.0) * (1.0 - 1e-6);
      t.R2lim = fmin(a.dist2, rs * rs);
    }
    double hg = a.h_guess ? a.h_guess[i] : 0.0;
    if (hg > 0.0) { hg *= a.guess_factor; t.R2try = hg * hg; }
    else {
      const double R3 = 1.3 * (double)a.k * b.vol / (4.1887902047863905 * (double)(total > 0 ? total : 1));
      const double Rt = cbrt(R3);
      t.R2try = Rt * Rt;
    }
    t.R2try = fmin(t.R2try, t.R2lim);
  }
  return t;
}

// after one complete scan of the block: accept, or correct the trial radius
__device__ __forceinline__ void knn_target_update(const KnnCellArgs& a, const KnnBlock& b, KnnTarget& t) {
  if (!t.active) return;
  if (t.count > KC_CAP) {                       // too many survivors: shrink the trial radius
    const double f = cbrt(0.75 * (double)KC_CAP / (double)t.count);
    t.R2try *= f * f;
  } else if (t.count >= a.k) {
    t.active = false;
  } else if (t.R2try < t.R2lim) {               // too few: grow towards the covered radius
    const double f = cbrt(1.4 * (double)a.k / (double)(t.count > 0 ? t.count : 1));
    t.R2try = fmin(t.R2lim, t.R2try * fmax(f * f, 1.1));
  } else {
    if (!(b.whole || t.R2lim >= a.dist2)) t.unresolved = true;    // k-th neighbour may lie outside the block
    t.active = false;
  }
}

template <bool SORTED>
__device__ __forceinline__ void knn_target_finish(const KnnCellArgs& a, KnnTarget& t, int i, const double* keybuf,
                                                  const int32_t* idxbuf, int lane) {
  if (!t.finite) return;
  if (t.active) t.unresolved = true;            // retry budget exhausted: leave it to the general kernel
  if (lane == 0) {
    if (t.rescans > 0) atomicAdd(&a.flags[FL_KNN_RESCAN], t.rescans);
    if (t.unresolved) {
      a.worklist[atomicAdd(a.nwork, 1)] = i;
      atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
    }
  }
  if (!t.unresolved) knn_emit<SORTED>(a, i, min(t.count, KC_CAP), keybuf, idxbuf, lane);
}

template <bool SORTED>
__global__ void __launch_bounds__(KC_WARPS * 32, 8) knn_cell_kernel(KnnCellArgs a) {
  __shared__ __align__(128) double sX[KC_CHP];
  __shared__ __align__(128) double sY[KC_CHP];
  __shared__ __align__(128) double sZ[KC_CHP];
  __shared__ int32_t sG[KC_CHP];
  __shared__ double sKey[KC_WARPS][KC_CAP];
  __shared__ int32_t sIdx[KC_WARPS][KC_CAP];
  __shared__ __align__(8) uint64_t bar;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const GridSpec& g = a.g;
  const int cell = blockIdx.x;
  const int t_begin = (int)a.cell_start[cell], t_end = (int)a.cell_start[cell + 1];
  if (t_begin == t_end) return;
  const int cx = cell % g.nx, cy = (cell / g.nx) % g.ny, cz = cell / (g.nx * g.ny);
  KnnBlock b;
  b.x0 = max(cx - 1, 0); b.x1 = min(cx + 1, g.nx - 1);
  b.y0 = max(cy - 1, 0); b.y1 = min(cy + 1, g.ny - 1);
  b.z0 = max(cz - 1, 0); b.z1 = min(cz + 1, g.nz - 1);
  const int ysp = b.y1 - b.y0 + 1;
  const int nrows = ysp * (b.z1 - b.z0 + 1);                       // <= 9
  b.whole = (b.x0 == 0 && b.y0 == 0 && b.z0 == 0 && b.x1 == g.nx - 1 && b.y1 == g.ny - 1 && b.z1 == g.nz - 1);
  b.vol = (double)(b.x1 - b.x0 + 1) * (double)nrows * g.c * g.c * g.c;
  const KnnRuns runs = knn_runs(a, lane, b.x0, b.x1, b.y0, b.z0, ysp, nrows);
  const KnnTile tile = {sX, sY, sZ, sG};
  double* keybuf = sKey[warp];
  int32_t* idxbuf = sIdx[warp];
  const int nchunks = (runs.ptotal + KC_CHP - 1) / KC_CHP;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t phase = 0;

  if (nchunks == 1) {
    // ---- common case: the whole block fits the tile; stage once, warps take targets independently ----
    if (warp == 0) knn_stage(a, runs, tile, &bar, phase, lane, 0, runs.ptotal);
    __syncthreads();
    for (int i = t_begin + warp; i < t_end; i += KC_WARPS) {
      KnnTarget t = knn_target_init(a, b, runs.total, i, true, lane);
      for (int attempt = 0; attempt < KC_MAX_ATTEMPT && t.active; ++attempt) {
        t.count = 0; ++t.rescans;
        __syncwarp();
        t.count = knn_scan_chunk(tile, runs.ptotal, t.px, t.py, t.pz, t.R2try, 0, keybuf, idxbuf, lane);
        knn_target_update(a, b, t);
      }
      knn_target_finish<SORTED>(a, t, i, keybuf, idxbuf, lane);
    }
  } else {
    // ---- crowded block: rounds of KC_WARPS targets, every round walks all chunks (CTA-synchronous) ----
    for (int base = t_begin; base < t_end; base += KC_WARPS) {
      const int i = base + warp;
      KnnTarget t = knn_target_init(a, b, runs.total, i, i < t_end, lane);
      for (int attempt = 0; attempt < KC_MAX_ATTEMPT; ++attempt) {
        if (!__syncthreads_or(t.active ? 1 : 0)) break;
        if (t.active) { t.count = 0; ++t.rescans; }
        for (int ch = 0; ch < nchunks; ++ch) {
          const int c_begin = ch * KC_CHP;
          const int c_n = min(KC_CHP, runs.ptotal - c_begin);
          __syncthreads();                                   // everyone is done with the previous chunk
          if (warp == 0) knn_stage(a, runs, tile, &bar, phase, lane, c_begin, c_n);
          __syncthreads();
          if (t.active) t.count = knn_scan_chunk(tile, c_n, t.px, t.py, t.pz, t.R2try, t.count, keybuf, idxbuf, lane);
        }
        knn_target_update(a, b, t);
      }
      if (i < t_end) knn_target_finish<SORTED>(a, t, i, keybuf, idxbuf, lane);
    }
  }
}

int knn_cell_max_k() { return KC_CAP - 8; }

// Fast path + general-kernel clean-up of the targets it could not resolve inside their cell block.
// `dists` != nullptr selects the API layout (rows sorted by distance, distances returned).
int launch_knn_cell(sphb200_ctx* ctx, int n, int k, double dist, const double* h_guess, double margin, int32_t* list,
                    double* dists, double* h_out, int32_t* nontriv, cudaStream_t st) {
  if (k < 1 || k > knn_cell_max_k()) SPH_FAIL("knn_cell: k=%d outside [1, %d]", k, knn_cell_max_k());
  SPH_TRY(devbuf_reserve(ctx->worklist, (size_t)(n + 1) * sizeof(int32_t)));
  int* nwork = bufp<int>(ctx->flags) + FL_KNN_NWORK;
  CUDA_TRY(cudaMemsetAsync(nwork, 0, sizeof(int), st));
  KnnCellArgs a;
  const size_t ns = soa_stride(n);
  a.xs = bufp<double>(ctx->soa); a.ys = a.xs + ns; a.zs = a.ys + ns;
  a.cell_start = bufp<uint32_t>(ctx->cell_start);
  a.g = ctx->grid;
  a.n = n; a.k = k;
  a.dist2 = dist * dist;
  if (!(a.dist2 == a.dist2)) a.dist2 = INFINITY;
  a.h_guess = h_guess;
  a.guess_factor = 1.0 + margin;
  a.list = list; a.dists = dists; a.h_out = h_out; a.nontriv = nontriv;
  a.flags = bufp<int>(ctx->flags);
  a.worklist = bufp<int>(ctx->worklist);
  a.nwork = nwork;
  if (dists) knn_cell_kernel<true><<<ctx->grid.ncells, KC_WARPS * 32, 0, st>>>(a);
  else knn_cell_kernel<false><<<ctx->grid.ncells, KC_WARPS * 32, 0, st>>>(a);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  // unresolved targets: general warp-per-target kernel over the work list (grid-stride, count on device)
  SPH_TRY(launch_knn_worklist(ctx, n, k, dist, bufp<int>(ctx->worklist), nwork, list, dists, h_out, nontriv, st));
  return 0;
}
