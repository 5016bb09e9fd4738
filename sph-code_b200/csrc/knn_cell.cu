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
#ifndef KC_UNROLL
#define KC_UNROLL 4
#endif
#define KC_PRAGMA_(x) _Pragma(#x)
#define KC_PRAGMA(x) KC_PRAGMA_(x)
#define KC_UNROLL_SCAN KC_PRAGMA(unroll KC_UNROLL)
#ifndef KC_MIN_CTAS
#define KC_MIN_CTAS 8          // resident CTAs per SM the register budget is set for (64 registers, 26.7 KB)
#endif

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
  const uint32_t* skip_id;   // slab decomposition: targets with skip_id[i] >= skip_from are ghosts (their owner
  int skip_from;             // searches for them); nullptr = every particle is a target
};

// Shared-memory tile of one chunk of the cell block's candidates: one base address, fixed offsets
// (X at 0, Y at KC_CHP, Z at 2 KC_CHP doubles; sorted index / -1 for alignment padding after them).
struct KnnTile { double* T; };
#define TILE_X(t) ((t).T)
#define TILE_Y(t) ((t).T + KC_CHP)
#define TILE_Z(t) ((t).T + 2 * KC_CHP)
#define TILE_G(t) (reinterpret_cast<int32_t*>((t).T + 3 * KC_CHP))

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
    bulk_g2s(TILE_X(t) + (lo - c_begin), a.xs + src, bytes, bar);
    bulk_g2s(TILE_Y(t) + (lo - c_begin), a.ys + src, bytes, bar);
    bulk_g2s(TILE_Z(t) + (lo - c_begin), a.zs + src, bytes, bar);
  }
  for (int run = 0; run < 9; ++run) {                     // index table, filled by the whole warp
    const int q_ps = __shfl_sync(0xffffffffu, r.ps, run), q_off = __shfl_sync(0xffffffffu, r.poff, run);
    const int q_len = __shfl_sync(0xffffffffu, r.plen, run);
    const int q_rs = __shfl_sync(0xffffffffu, r.rs, run), q_re = __shfl_sync(0xffffffffu, r.re, run);
    const int qlo = max(q_off, c_begin), qhi = min(q_off + q_len, c_begin + c_n);
    for (int c = qlo + lane; c < qhi; c += 32) {
      const int gi = q_ps + (c - q_off);
      TILE_G(t)[c - c_begin] = (gi >= q_rs && gi < q_re) ? gi : -1;   // alignment padding is masked
    }
  }
  for (int c = c_n + lane; c < ((c_n + 31) & ~31); c += 32) TILE_X(t)[c] = nan("");   // tail of the last 32-wide group
  mbar_wait(bar, phase);
  phase ^= 1u;
  if (hi > lo) {                                          // padding never passes `d2 < R2` (NaN)
    if (r.ps < r.rs && r.poff >= lo) TILE_X(t)[r.poff - c_begin] = nan("");
    const int last = r.poff + r.plen - 1;
    if (r.ps + r.plen > r.re && last < hi) TILE_X(t)[last - c_begin] = nan("");
  }
}

// One warp scans the staged chunk for one target: survivors (d^2 < R2try) are appended to the warp's
// buffer in tile order.  The tile is NaN-padded to a multiple of 32, so no bounds checks.  Returns the
// updated survivor count (may exceed KC_CAP: overflow is counted, not stored).
__device__ __forceinline__ int knn_scan_chunk(const KnnTile& t, int c_n, double px, double py, double pz, double R2try,
                                              int count, double* keybuf, int32_t* idxbuf, int lane) {
  const unsigned lt = (1u << lane) - 1u;
  const double* T = t.T + lane;
  KC_UNROLL_SCAN
  for (int c0 = 0; c0 < c_n; c0 += 32) {
    const double dx = T[c0] - px, dy = T[c0 + KC_CHP] - py, dz = T[c0 + 2 * KC_CHP] - pz;
    const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    const bool ok = d2 < R2try;
    const unsigned b = __ballot_sync(0xffffffffu, ok);
    if (b) {
      const int pos = count + __popc(b & lt);
      if (ok && pos < KC_CAP) { keybuf[pos] = d2; idxbuf[pos] = reinterpret_cast<const int32_t*>(T + 3 * KC_CHP - lane)[c0 + lane]; }
      count += __popc(b);
    }
  }
  return count;
}

// Warp-wide maximum of (key, idx) pairs in lexicographic order (exact; used when float keys tie).
__device__ __forceinline__ void warp_max_pair(double& k, int& i) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const double ko = __shfl_xor_sync(0xffffffffu, k, d);
    const int io = __shfl_xor_sync(0xffffffffu, i, d);
    if (ko > k || (ko == k && io > i)) { k = ko; i = io; }
  }
}

// Largest (key, idx) among the two entries each lane holds (removed / absent entries have key -1).
// Fast path: a warp REDUX on the keys rounded UP to float (monotone, so a unique float maximum is the
// exact maximum); exact pair reduction only when floats tie.  Returns the owner's (key, idx) to all.
__device__ __forceinline__ void knn_max_entry(double k0, int i0, double k1, int i1, double& km, int& im) {
  const unsigned f0 = (k0 >= 0.0) ? __float_as_uint(__double2float_ru(k0)) + 1u : 0u;
  const unsigned f1 = (k1 >= 0.0) ? __float_as_uint(__double2float_ru(k1)) + 1u : 0u;
  const unsigned m = __reduce_max_sync(0xffffffffu, max(f0, f1));
  const unsigned b0 = __ballot_sync(0xffffffffu, f0 == m), b1 = __ballot_sync(0xffffffffu, f1 == m);
  if (m != 0u && __popc(b0) + __popc(b1) == 1) {
    const int src = b0 ? (__ffs(b0) - 1) : (__ffs(b1) - 1);
    const double kk = b0 ? k0 : k1; const int ii = b0 ? i0 : i1;
    km = __shfl_sync(0xffffffffu, kk, src); im = __shfl_sync(0xffffffffu, ii, src);
    return;
  }
  double ka = (f0 == m) ? k0 : -1.0, kb = (f1 == m) ? k1 : -1.0;
  int ia = (f0 == m) ? i0 : -1, ib = (f1 == m) ? i1 : -1;
  if (kb > ka || (kb == ka && ib > ia)) { ka = kb; ia = ib; }
  warp_max_pair(ka, ia);
  km = ka; im = ia;
}

// The warp's buffer holds `count` (<= KC_CAP) survivors of target i; write its row and return the
// squared k-th distance (0 when the row is empty).
template <bool SORTED>
__device__ __forceinline__ double knn_emit(const KnnCellArgs& a, int i, int count, const double* keybuf,
                                           const int32_t* idxbuf, int lane) {
  const int k = a.k;
  const int nsel = min(count, k);
  const int64_t row = (int64_t)i * k;
  double km = -1.0; int im = -1;
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
        a.dists[row + rank] = sqrt(ke);
        if (ke > km) km = ke;
      }
    }
    for (int s = nsel + lane; s < k; s += 32) { a.list[row + s] = a.n; a.dists[row + s] = 0.0; }
    warp_max_pair(km, im);
  } else {
    // session layout: drop the (count - k) largest, keep the rest unordered with self in column 0.
    // The float keys (rounded up, so monotone in the exact key) are formed once; while the float maximum
    // is unique its owner just clears its entry: one REDUX, two ballots and no shuffle per removal.
    double k0 = (lane < count) ? keybuf[lane] : -1.0, k1 = (lane + 32 < count) ? keybuf[lane + 32] : -1.0;
    int i0 = (lane < count) ? idxbuf[lane] : -1, i1 = (lane + 32 < count) ? idxbuf[lane + 32] : -1;
    unsigned f0 = (k0 >= 0.0) ? __float_as_uint(__double2float_ru(k0)) + 1u : 0u;
    unsigned f1 = (k1 >= 0.0) ? __float_as_uint(__double2float_ru(k1)) + 1u : 0u;
    for (int r = count; r > nsel; --r) {
      const unsigned m = __reduce_max_sync(0xffffffffu, max(f0, f1));
      const unsigned b0 = __ballot_sync(0xffffffffu, f0 == m), b1 = __ballot_sync(0xffffffffu, f1 == m);
      if (__popc(b0) + __popc(b1) == 1) {
        if (f0 == m) { f0 = 0u; k0 = -1.0; i0 = -1; }
        if (f1 == m) { f1 = 0u; k1 = -1.0; i1 = -1; }
      } else {                                               // equal float keys: exact (key, index) maximum
        knn_max_entry(k0, i0, k1, i1, km, im);
        if (i0 == im && k0 == km) { f0 = 0u; k0 = -1.0; i0 = -1; }
        else if (i1 == im && k1 == km) { f1 = 0u; k1 = -1.0; i1 = -1; }
      }
    }
    knn_max_entry(k0, i0, k1, i1, km, im);
    const unsigned lt = (1u << lane) - 1u;
    const unsigned b0 = __ballot_sync(0xffffffffu, i0 >= 0), b1 = __ballot_sync(0xffffffffu, i1 >= 0);
    const unsigned s0 = __ballot_sync(0xffffffffu, i0 == i), s1 = __ballot_sync(0xffffffffu, i1 == i);
    int ps = -1;                                             // position of self among the kept entries
    if (s0) ps = __popc(b0 & ((1u << (__ffs(s0) - 1)) - 1u));
    else if (s1) ps = __popc(b0) + __popc(b1 & ((1u << (__ffs(s1) - 1)) - 1u));
    int32_t* out = a.list + row;
    if (i0 >= 0) { const int pos = __popc(b0 & lt); out[(pos == ps) ? 0 : (pos < ps ? pos + 1 : pos)] = i0; }
    if (i1 >= 0) { const int pos = __popc(b0) + __popc(b1 & lt); out[(pos == ps) ? 0 : (pos < ps ? pos + 1 : pos)] = i1; }
    for (int s = nsel + lane; s < k; s += 32) out[s] = a.n;
  }
  __syncwarp();
  return (nsel > 0 && km > 0.0) ? km : 0.0;
}

struct KnnBlock {          // geometry of the 3x3x3 block of one cell (warp-uniform)
  bool whole;              // the block is the whole grid
  double vol;              // its volume
  double lox, hix, loy, hiy, loz, hiz;   // block faces that are not faces of the whole grid (else -/+inf)
};

struct KnnTarget {         // per-lane: lane j of a warp prepares the warp's j-th target
  double px, py, pz, R2lim, R2try;
  bool active;             // finite coordinates: a search is needed
};

// `own` = this lane is the only one preparing target i (it may write the row of a non-finite target).
__device__ __forceinline__ KnnTarget knn_target_init(const KnnCellArgs& a, const KnnBlock& b, int total, int i, bool valid,
                                                     bool own) {
  KnnTarget t;
  t.px = t.py = t.pz = 0.0; t.R2lim = a.dist2; t.R2try = 0.0;
  t.active = valid;
  if (valid) { t.px = a.xs[i]; t.py = a.ys[i]; t.pz = a.zs[i]; }
  if (valid && !(isfinite(t.px) && isfinite(t.py) && isfinite(t.pz))) {      // cKDTree raises on these
    t.active = false;
    if (own) {
      for (int s = 0; s < a.k; ++s) { a.list[(int64_t)i * a.k + s] = a.n; if (a.dists) a.dists[(int64_t)i * a.k + s] = 0.0; }
      a.nontriv[i] = 0; a.h_out[i] = 0.0;
      atomicAdd(&a.flags[FL_KNN_NONFINITE], 1);
    }
  }
  if (t.active && a.skip_id && (int)a.skip_id[i] >= a.skip_from) {            // ghost: an empty, well-defined row
    t.active = false;
    if (own) {
      for (int s = 0; s < a.k; ++s) { a.list[(int64_t)i * a.k + s] = a.n; if (a.dists) a.dists[(int64_t)i * a.k + s] = 0.0; }
      a.nontriv[i] = 0; a.h_out[i] = a.h_guess ? a.h_guess[i] : 0.0;
    }
  }
  if (t.active) {
    if (!b.whole) {
      // distance to the nearest face of the cell block that is not a face of the whole grid
      double rs = fmin(fmin(t.px - b.lox, b.hix - t.px), fmin(fmin(t.py - b.loy, b.hiy - t.py), fmin(t.pz - b.loz, b.hiz - t.pz)));
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

// After one complete scan of the block with `count` survivors: 0 = accept, 1 = rescan with the
// corrected trial radius, 2 = the k-th neighbour may lie outside the block (general kernel).
__device__ __forceinline__ int knn_radius_update(const KnnCellArgs& a, bool whole, int count, double R2lim,
                                                 double& R2try) {
  if (count > KC_CAP) {                         // too many survivors: shrink the trial radius
    const double f = cbrt(0.75 * (double)KC_CAP / (double)count);
    R2try *= f * f;
    return 1;
  }
  if (count >= a.k) return 0;
  if (R2try < R2lim) {                          // too few: grow towards the covered radius
    const double f = cbrt(1.4 * (double)a.k / (double)(count > 0 ? count : 1));
    R2try = fmin(R2lim, R2try * fmax(f * f, 1.1));
    return 1;
  }
  return (whole || R2lim >= a.dist2) ? 0 : 2;
}

// shared-memory layout of knn_cell_kernel (bytes)
#define KC_OFF_KEY (28 * KC_CHP)                                  // after X, Y, Z (8 B) and G (4 B)
#define KC_OFF_IDX (KC_OFF_KEY + KC_WARPS * KC_CAP * 8)
#define KC_OFF_TGT (KC_OFF_IDX + KC_WARPS * KC_CAP * 4)           // per warp: 5 x 32 doubles + 32 ints
#define KC_TGT_BYTES (5 * 32 * 8 + 32 * 4)
#define KC_OFF_BLK (KC_OFF_TGT + KC_WARPS * KC_TGT_BYTES)         // KnnBlock
#define KC_OFF_BAR (KC_OFF_BLK + 64)
#define KC_SMEM (KC_OFF_BAR + 16)

template <bool SORTED>
__global__ void __launch_bounds__(KC_WARPS * 32, KC_MIN_CTAS) knn_cell_kernel(KnnCellArgs a) {
  __shared__ __align__(128) unsigned char smem[KC_SMEM];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x;
  const int t_begin = (int)a.cell_start[cell], t_end = (int)a.cell_start[cell + 1];
  if (t_begin == t_end) return;
  const KnnTile tile = {reinterpret_cast<double*>(smem)};
  double* keybuf = reinterpret_cast<double*>(smem + KC_OFF_KEY) + warp * KC_CAP;
  int32_t* idxbuf = reinterpret_cast<int32_t*>(smem + KC_OFF_IDX) + warp * KC_CAP;
  double* tgt = reinterpret_cast<double*>(smem + KC_OFF_TGT + warp * KC_TGT_BYTES);   // [5][32]: px py pz R2lim R2try
  int32_t* tres = reinterpret_cast<int32_t*>(tgt + 5 * 32);                          // [32]: result word per target
  KnnBlock* sb = reinterpret_cast<KnnBlock*>(smem + KC_OFF_BLK);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + KC_OFF_BAR);
  const GridSpec& g = a.g;
  const int cxy = cell / g.nx;
  const int cx = cell - cxy * g.nx, cz = cxy / g.ny, cy = cxy - cz * g.ny;
  const int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.nx - 1);
  const int y0 = max(cy - 1, 0), y1 = min(cy + 1, g.ny - 1);
  const int z0 = max(cz - 1, 0), z1 = min(cz + 1, g.nz - 1);
  const int ysp = y1 - y0 + 1;
  const int nrows = ysp * (z1 - z0 + 1);                       // <= 9
  if (threadIdx.x == 0) {
    KnnBlock b;
    b.whole = (x0 == 0 && y0 == 0 && z0 == 0 && x1 == g.nx - 1 && y1 == g.ny - 1 && z1 == g.nz - 1);
    b.vol = (double)(x1 - x0 + 1) * (double)nrows * g.c * g.c * g.c;
    b.lox = (x0 > 0) ? g.ox + (double)x0 * g.c : -INFINITY;  b.hix = (x1 < g.nx - 1) ? g.ox + (double)(x1 + 1) * g.c : INFINITY;
    b.loy = (y0 > 0) ? g.oy + (double)y0 * g.c : -INFINITY;  b.hiy = (y1 < g.ny - 1) ? g.oy + (double)(y1 + 1) * g.c : INFINITY;
    b.loz = (z0 > 0) ? g.oz + (double)z0 * g.c : -INFINITY;  b.hiz = (z1 < g.nz - 1) ? g.oz + (double)(z1 + 1) * g.c : INFINITY;
    *sb = b;
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const KnnRuns runs = knn_runs(a, lane, x0, x1, y0, z0, ysp, nrows);
  const int ptotal = runs.ptotal, total = runs.total;
  const int nchunks = (ptotal + KC_CHP - 1) / KC_CHP;
  __syncthreads();
  uint32_t phase = 0;

  if (nchunks == 1) {
    // ---- common case: the whole block fits the tile.  Stage once; each warp walks its targets
    // (t_begin + warp, + KC_WARPS, ...).  Lane j prepares the warp's j-th target (position, covered
    // radius, trial radius) into the warp's table while the tile arrives, and afterwards finishes it
    // (sqrt, scalar outputs), so the per-target scalar work is done 32 targets at a time. ----
    if (warp == 0) knn_stage(a, runs, tile, bar, phase, lane, 0, ptotal);
    const int first = t_begin + warp;
    for (int batch = first; batch < t_end || batch == first; batch += KC_WARPS * 32) {
      const int my_i = batch + lane * KC_WARPS;
      {
        const KnnTarget mine = knn_target_init(a, *sb, total, my_i, my_i < t_end, true);
        tgt[lane] = mine.px; tgt[32 + lane] = mine.py; tgt[64 + lane] = mine.pz; tgt[96 + lane] = mine.R2lim;
        tgt[128 + lane] = mine.active ? mine.R2try : -1.0;
        tres[lane] = -1;                                       // -1: nothing to write (invalid / non-finite)
      }
      if (batch == first) __syncthreads(); else __syncwarp();  // tile staged; every warp passes here with batch == first
      const int nb = min(32, (t_end - batch + KC_WARPS - 1) / KC_WARPS);
      for (int j = 0; j < nb; ++j) {
        double R2try = tgt[128 + j];
        if (R2try < 0.0) continue;
        const double px = tgt[j], py = tgt[32 + j], pz = tgt[64 + j], R2lim = tgt[96 + j];
        int count = 0, verdict = 1, scans = 0;
        for (int attempt = 0; attempt < KC_MAX_ATTEMPT && verdict == 1; ++attempt) {
          __syncwarp();
          count = knn_scan_chunk(tile, ptotal, px, py, pz, R2try, 0, keybuf, idxbuf, lane);
          verdict = knn_radius_update(a, sb->whole, count, R2lim, R2try);
          ++scans;
        }
        double h2 = 0.0;
        if (verdict == 0) h2 = knn_emit<SORTED>(a, batch + j * KC_WARPS, min(count, KC_CAP), keybuf, idxbuf, lane);
        if (lane == 0) {                                       // result word: nsel | unresolved << 8 | rescans << 16
          tgt[j] = h2;
          tres[j] = min(count, a.k) | ((verdict != 0) << 8) | ((scans - 1) << 16);
        }
      }
      __syncwarp();
      const int res = tres[lane];
      if (res >= 0) {                                          // per-target scalars, one lane per target
        if (res & 0x100) {
          a.worklist[atomicAdd(a.nwork, 1)] = my_i;
          atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
        } else {
          a.h_out[my_i] = sqrt(tgt[lane]);
          a.nontriv[my_i] = res & 0xff;
        }
        if (res >> 16) atomicAdd(&a.flags[FL_KNN_RESCAN], res >> 16);
      }
      __syncwarp();
    }
  } else {
    // ---- crowded block: rounds of KC_WARPS targets, every round walks all chunks (CTA-synchronous) ----
    for (int base = t_begin; base < t_end; base += KC_WARPS) {
      const int i = base + warp;
      const KnnTarget t = knn_target_init(a, *sb, total, i, i < t_end, lane == 0);
      double R2try = t.R2try;
      int count = 0, verdict = t.active ? 1 : 0, scans = 0;
      for (int attempt = 0; attempt < KC_MAX_ATTEMPT; ++attempt) {
        if (!__syncthreads_or(verdict == 1)) break;
        if (verdict == 1) { count = 0; ++scans; }
        for (int ch = 0; ch < nchunks; ++ch) {
          const int c_begin = ch * KC_CHP;
          const int c_n = min(KC_CHP, ptotal - c_begin);
          __syncthreads();                                   // everyone is done with the previous chunk
          if (warp == 0) knn_stage(a, runs, tile, bar, phase, lane, c_begin, c_n);
          __syncthreads();
          if (verdict == 1) count = knn_scan_chunk(tile, c_n, t.px, t.py, t.pz, R2try, count, keybuf, idxbuf, lane);
        }
        if (verdict == 1) verdict = knn_radius_update(a, sb->whole, count, t.R2lim, R2try);
      }
      if (t.active) {
        if (verdict != 0) {
          if (lane == 0) { a.worklist[atomicAdd(a.nwork, 1)] = i; atomicAdd(&a.flags[FL_KNN_WIDEN], 1); }
        } else {
          const double h2 = knn_emit<SORTED>(a, i, min(count, KC_CAP), keybuf, idxbuf, lane);
          if (lane == 0) { a.h_out[i] = sqrt(h2); a.nontriv[i] = min(count, a.k); }
        }
        if (lane == 0 && scans > 1) atomicAdd(&a.flags[FL_KNN_RESCAN], scans - 1);
      }
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
  a.skip_id = ctx->skip_id; a.skip_from = ctx->skip_from;
  if (dists) knn_cell_kernel<true><<<ctx->grid.ncells, KC_WARPS * 32, 0, st>>>(a);
  else knn_cell_kernel<false><<<ctx->grid.ncells, KC_WARPS * 32, 0, st>>>(a);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  // unresolved targets: general warp-per-target kernel over the work list (grid-stride, count on device)
  SPH_TRY(launch_knn_worklist(ctx, n, k, dist, h_guess, 1.0 + margin, bufp<int>(ctx->worklist), nwork, list, dists, h_out, nontriv, st));
  return 0;
}
