// Group-tiled exact fixed-k nearest-neighbour search (the fast path of the neighbour query).
// Replaces `kdt.query(points, N_NEIGH, eps, 2, dist)` of
// /root/reference/sph/navier_stokes_cleaned.py:544 (eps = 0) and the row-max
// smoothing length of :548.
//
// One CTA (8 warps) per GROUP of 2x2x2 cells (tile.cuh), persistent over groups.  The group's 4x4x4-cell tile of
// sorted position records {x,y,z,guess} is staged into shared memory by at most 16 bulk asynchronous copies
// (cp.async.bulk -> UBLKCP, completion on an mbarrier), then converted once into tile-relative fp32 coordinates
// in cell units (struct of arrays).  Each warp takes one target particle at a time:
//   1. FILTER (fp32): the 32 lanes stride over the three z-layers of the target's 3x3x3 cell block, two candidates
//      per lane and trip, and keep the tile offsets of the candidates whose fp32 distance is below the trial
//      radius widened by a rigorous rounding bound - a superset of the candidates inside the trial radius;
//   2. EXACT (fp64): d^2 = ((dx*dx + dy*dy) + dz*dz) with separately rounded products (no FMA contraction), the
//      arithmetic of cKDTree and numpy, for the ~50 survivors only; those with d^2 < R_try^2 are the candidates;
//   3. SELECT: the k smallest by (d^2, index): a bisection on the float-rounded keys when more than 64 are left,
//      then removal of the largest with warp REDUX-max; the row is written as uint16 TILE OFFSETS (session
//      layout: unordered, self in column 0) or ranked by counting (API layout: ascending distance + distances).
// The trial radius is the previous step's k-th distance times (1 + margin) when the caller has one, else the
// radius expected to hold ~1.3 k particles at the tile's density; it never exceeds the target's distance to the
// faces of the region that was scanned, so the result is exact.  Too few candidates -> the radius grows and the
// target is rescanned; k-th neighbour possibly outside the scanned region -> the target goes to a work list that
// the general warp-per-target kernel of knn.cu resolves (rows in the global list, index space of the sort).
#include "common.cuh"
#include "kernels.cuh"
#include "tile.cuh"
#include "knn_select.cuh"
#include <float.h>

int launch_knn_mma(sphb200_ctx* ctx, const KnnGroupArgs& a, cudaStream_t st);   // knn_mma.cu: tensor-core filter variant

#define KG_CAPF 128           // filter survivors a warp can hold for one target
#define KG_CAP 64             // candidates handed to the final selection (two per lane)
#ifndef KG_MIN_CTAS
#define KG_MIN_CTAS 3
#endif

// ---- shared-memory layout (bytes) ----
#define KG_XF_LEN (TILE_CAP + 64)                               // fp32 arrays are padded so that the last trip stays inside
#define KG_OFF_TILE 0                                           // double4 tile[TILE_CAP]
#define KG_OFF_XF (KG_OFF_TILE + TILE_CAP * 32)                 // float xf, yf, zf [KG_XF_LEN]
#define KG_OFF_F (KG_OFF_XF + 3 * KG_XF_LEN * 4)                // per warp: uint16 F[KG_CAPF]
#define KG_OFF_KEY (KG_OFF_F + KG_WARPS * KG_CAPF * 2)          // per warp: double key[KG_CAPF]
#define KG_OFF_IDX (KG_OFF_KEY + KG_WARPS * KG_CAPF * 8)        // per warp: int idx[KG_CAPF]
#define KG_OFF_TGT (KG_OFF_IDX + KG_WARPS * KG_CAPF * 4)        // target table
#define KG_TGT_BYTES (TILE_MAX_TARGETS * (8 + 8 + 4 + 4))       // R2lim | R2try/h2 | i | off+code / result
#define KG_OFF_ROWS (KG_OFF_TGT + KG_TGT_BYTES)                 // TileRows
#define KG_OFF_MISC (KG_OFF_ROWS + 384)                         // run table, segments, group id, mbarrier
#define KG_OFF_L (KG_OFF_MISC + 256)                            // per warp: uint16 L[9][32], lane-private survivor columns
#define KG_SMEM (KG_OFF_L + KG_WARPS * 9 * 32 * 2)

struct KgMisc {
  int tbegin[4], tend[4], trow[4], tbase[5];    // target runs (tile.cuh group_target_run) and their prefix
  int seg[4][3][2];                             // per run: the three z-layer segments [start, end) in tile offsets
  int group;
  int pad;
  uint64_t bar;
  int cntF[KG_WARPS];                           // filter survivors of the warp's current target (atomic variant)
};

// The warp's buffers hold `count` (<= KG_CAP) candidates (exact d^2, tile offset) of target i; write its row and
// return the squared k-th distance (0 when the row is empty).  `self` = the target's own tile offset.
template <bool SORTED>
__device__ __forceinline__ double knn_emit(const KnnGroupArgs& a, int i, int self, int count, const double* keybuf,
                                           const int* idxbuf, const int* rowbase, int lane) {
  const int k = a.k;
  const int nsel = min(count, k);
  const int64_t row = (int64_t)i * k;
  double km = -1.0; int im = -1;
  __syncwarp();
  if (SORTED) {
    // API layout: rank by counting; rank < k are the neighbours in ascending (d^2, index) order.  Tile offsets
    // ascend with the sorted index, so they break ties the way the sorted index would.
    for (int e = lane; e < count; e += 32) {
      const double ke = keybuf[e]; const int je = idxbuf[e];
      int rank = 0;
      for (int f = 0; f < count; ++f) {
        const double kf = keybuf[f];
        rank += (kf < ke) || (kf == ke && idxbuf[f] < je);
      }
      if (rank < k) {
        a.list[row + rank] = rowbase[je];                  // tile offset -> sorted index
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
    const unsigned s0 = __ballot_sync(0xffffffffu, i0 == self), s1 = __ballot_sync(0xffffffffu, i1 == self);
    int ps = -1;                                             // position of self among the kept entries
    if (s0) ps = __popc(b0 & ((1u << (__ffs(s0) - 1)) - 1u));
    else if (s1) ps = __popc(b0) + __popc(b1 & ((1u << (__ffs(s1) - 1)) - 1u));
    uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(k);
    if (i0 >= 0) { const int pos = __popc(b0 & lt); out[(pos == ps) ? 0 : (pos < ps ? pos + 1 : pos)] = (uint16_t)i0; }
    if (i1 >= 0) { const int pos = __popc(b0) + __popc(b1 & lt); out[(pos == ps) ? 0 : (pos < ps ? pos + 1 : pos)] = (uint16_t)i1; }
    for (int s = nsel + lane; s < k; s += 32) out[s] = (uint16_t)LIST_MISSING;
  }
  __syncwarp();
  return (nsel > 0 && km > 0.0) ? km : 0.0;
}

// packed fp32 pairs (FADD2 / FMUL2 / FFMA2 on sm_100a): two candidates per instruction
__device__ __forceinline__ unsigned long long f2_sub(unsigned long long a, unsigned long long b) {
  unsigned long long r; asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ unsigned long long f2_mul(unsigned long long a, unsigned long long b) {
  unsigned long long r; asm("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ unsigned long long f2_fma(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ unsigned long long f2_pack(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
}

// FILTER: candidates of the three z-layer segments whose fp32 distance to (pxf,pyf,pzf) is below T2f; their tile
// offsets go to F (first KG_CAPF kept).  Rows have even bounds, so a float2 never straddles a segment end; a pair
// past the end of its segment is switched off through the threshold (no branch).  ATOM: survivors reserve their
// slot with a shared-memory atomic (2 instructions per survivor) instead of ballot + popc prefix sums; the order
// of F then depends on the hardware's arbitration (rows are unordered sets either way).
template <bool ATOM>
__device__ __forceinline__ int knn_filter(const float* __restrict__ xf, const float* __restrict__ yf, const float* __restrict__ zf,
                                          const int (*seg)[2], float pxf, float pyf, float pzf, float T2f, uint16_t* F,
                                          int* cntF, int lane) {
  const unsigned lt = (1u << lane) - 1u;
  const unsigned long long PX = f2_pack(pxf, pxf), PY = f2_pack(pyf, pyf), PZ = f2_pack(pzf, pzf);
  int count = 0;
  if (ATOM) { if (lane == 0) *cntF = 0; __syncwarp(); }
#pragma unroll 1
  for (int s = 0; s < 3; ++s) {
    const int a = seg[s][0], b = seg[s][1];
#pragma unroll 2
    for (int base = a; base < b; base += 64) {
      const int c = base + 2 * lane;
      const float Tl = (c < b) ? T2f : -1.0f;                  // beyond the segment: never accepted
      const unsigned long long X = *reinterpret_cast<const unsigned long long*>(xf + c);
      const unsigned long long Y = *reinterpret_cast<const unsigned long long*>(yf + c);
      const unsigned long long Z = *reinterpret_cast<const unsigned long long*>(zf + c);
      const unsigned long long dx = f2_sub(X, PX), dy = f2_sub(Y, PY), dz = f2_sub(Z, PZ);
      const unsigned long long d = f2_fma(dz, dz, f2_fma(dy, dy, f2_mul(dx, dx)));
      const float d0 = __uint_as_float((unsigned)d), d1 = __uint_as_float((unsigned)(d >> 32));
      const bool ok0 = d0 < Tl, ok1 = d1 < Tl;                 // NaN (padding, non-finite particles) never passes
      if (ATOM) {
        if (ok0) { const int p0 = atomicAdd(cntF, 1); if (p0 < KG_CAPF) F[p0] = (uint16_t)c; }
        if (ok1) { const int p1 = atomicAdd(cntF, 1); if (p1 < KG_CAPF) F[p1] = (uint16_t)(c + 1); }
      } else {
        const unsigned b0 = __ballot_sync(0xffffffffu, ok0), b1 = __ballot_sync(0xffffffffu, ok1);
        if (b0 | b1) {
          const int n0 = __popc(b0);
          const int p0 = count + __popc(b0 & lt), p1 = count + n0 + __popc(b1 & lt);
          if (ok0 && p0 < KG_CAPF) F[p0] = (uint16_t)c;
          if (ok1 && p1 < KG_CAPF) F[p1] = (uint16_t)(c + 1);
          count += n0 + __popc(b1);
        }
      }
    }
  }
  if (ATOM) { __syncwarp(); count = *cntF; }
  return count;
}

// FILTER, lane-private form: every lane appends its survivors to its own column of L (no ballots, no prefix sums in
// the loop: a survivor costs its lane one predicated store and one saturating add); afterwards one warp scan of the
// per-lane counts packs the columns into F.  Returns the number of survivors, or -1 when a column overflowed (the
// caller falls back to the ballot form for this scan).
#define KG_CAPL 8
__device__ __forceinline__ int knn_filter_private(const float* __restrict__ xf, const float* __restrict__ yf,
                                                  const float* __restrict__ zf, const int (*seg)[2], float pxf, float pyf,
                                                  float pzf, float T2f, uint16_t* F, uint16_t* L, int lane) {
  const unsigned long long PX = f2_pack(pxf, pxf), PY = f2_pack(pyf, pyf), PZ = f2_pack(pzf, pzf);
  uint16_t* col = L + lane;
  int nl = 0;
#pragma unroll 1
  for (int s = 0; s < 3; ++s) {
    const int a = seg[s][0], b = seg[s][1];
#pragma unroll 2
    for (int base = a; base < b; base += 64) {
      const int c = base + 2 * lane;
      const float Tl = (c < b) ? T2f : -1.0f;
      const unsigned long long X = *reinterpret_cast<const unsigned long long*>(xf + c);
      const unsigned long long Y = *reinterpret_cast<const unsigned long long*>(yf + c);
      const unsigned long long Z = *reinterpret_cast<const unsigned long long*>(zf + c);
      const unsigned long long dx = f2_sub(X, PX), dy = f2_sub(Y, PY), dz = f2_sub(Z, PZ);
      const unsigned long long d = f2_fma(dz, dz, f2_fma(dy, dy, f2_mul(dx, dx)));
      if (__uint_as_float((unsigned)d) < Tl) { col[nl * 32] = (uint16_t)c; nl = min(nl + 1, KG_CAPL); }
      if (__uint_as_float((unsigned)(d >> 32)) < Tl) { col[nl * 32] = (uint16_t)(c + 1); nl = min(nl + 1, KG_CAPL); }
    }
  }
  if (__any_sync(0xffffffffu, nl >= KG_CAPL)) return -1;
  int inc = nl;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += v; }
  const int total = __shfl_sync(0xffffffffu, inc, 31);
  if (total <= KG_CAPF) {
    const int pre = inc - nl;
#pragma unroll
    for (int j = 0; j < KG_CAPL; ++j) if (j < nl) F[pre + j] = col[j * 32];
  }
  return total;
}

template <bool SORTED, bool ATOM>
__global__ void __launch_bounds__(KG_THREADS, KG_MIN_CTAS) knn_group_kernel(KnnGroupArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double4* tile = reinterpret_cast<double4*>(smem + KG_OFF_TILE);
  float* xf = reinterpret_cast<float*>(smem + KG_OFF_XF);
  float* yf = xf + KG_XF_LEN;
  float* zf = yf + KG_XF_LEN;
  uint16_t* F = reinterpret_cast<uint16_t*>(smem + KG_OFF_F) + warp * KG_CAPF;
  uint16_t* Lcol = reinterpret_cast<uint16_t*>(smem + KG_OFF_L) + warp * 9 * 32;
  double* keybuf = reinterpret_cast<double*>(smem + KG_OFF_KEY) + warp * KG_CAPF;
  int* idxbuf = reinterpret_cast<int*>(smem + KG_OFF_IDX) + warp * KG_CAPF;
  double* t_R2lim = reinterpret_cast<double*>(smem + KG_OFF_TGT);
  double* t_R2try = t_R2lim + TILE_MAX_TARGETS;               // in: trial radius^2 (< 0: nothing to do); out: k-th d^2
  int* t_i = reinterpret_cast<int*>(t_R2try + TILE_MAX_TARGETS);
  int* t_word = t_i + TILE_MAX_TARGETS;                       // in: tile offset | run << 16; out: result word
  TileRows* rows = reinterpret_cast<TileRows*>(smem + KG_OFF_ROWS);
  KgMisc* misc = reinterpret_cast<KgMisc*>(smem + KG_OFF_MISC);
  // SORTED only: tile offset -> sorted index (stored over the fp64 keys of warps that are idle? no: own array)
  int* rowbase = reinterpret_cast<int*>(smem + KG_SMEM);      // [TILE_CAP] ints, allocated only for SORTED
  const GridSpec g = *a.spec;
  if (tid == 0) { mbar_init(&misc->bar, 1); mbar_fence_init(); }
  uint32_t phase = 0;
  const double u24 = 5.9604644775390625e-8;                   // 2^-24
  const double eta = u24 * (2.0 * 5.0 + 5.0) + 1e-9;          // rounding of a tile-relative fp32 coordinate difference (|coord| <= 5)

  for (;;) {
    __syncthreads();                                          // previous group done with every buffer
    if (tid == 0) {
      const int idx = atomicAdd(&a.flags[FL_GROUP_NEXT], 1);
      misc->group = a.group_list ? (idx < a.group_list[0] ? a.group_list[1 + idx] : 0x7fffffff) : idx;
    }
    __syncthreads();
    const int group = misc->group;
    if (group >= g.ngroups) break;
    int gx, gy, gz;
    group_coords(g, group, &gx, &gy, &gz);
    if (warp == 0) {
      tile_rows(g, a.cell_start, gx, gy, gz, lane, rows);
      int b = 0, e = 0, r = 0;
      const bool have = (lane < 4) && group_target_run(g, a.cell_start, gx, gy, gz, lane, &b, &e, &r);
      int len = have ? (e - b) : 0;
      int inc = len;
#pragma unroll
      for (int d = 1; d < 4; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += v; }
      if (lane < 4) { misc->tbegin[lane] = b; misc->tend[lane] = e; misc->trow[lane] = r; misc->tbase[lane] = inc - len; }
      if (lane == 3) misc->tbase[4] = inc;
    }
    __syncthreads();
    const int T = tile_padded_total(rows);
    const int ntgt = misc->tbase[4];
    if (ntgt == 0) continue;                                  // (uniform: every thread reads the same shared value)
    int lens[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) lens[q] = misc->tend[q] - misc->tbegin[q];
    const bool crowded = tile_crowded(g, rows, lens);
    if (crowded) {
      // the tile does not fit: every target of the group goes to the general kernel
      if (tid == 0) atomicAdd(&a.flags[FL_KNN_CROWDED], 1);
      for (int t = tid; t < ntgt; t += KG_THREADS) {
        int q = 0;
        while (q < 3 && t >= misc->tbase[q + 1]) ++q;
        const int i = misc->tbegin[q] + (t - misc->tbase[q]);
        const bool ghost = a.skip_id && a.skip_id[i] >= a.skip_from;
        if (ghost) {
          if (a.list16) a.list16[(int64_t)i * LIST_STRIDE(a.k)] = (uint16_t)LIST_MISSING;
          if (SORTED) for (int s = 0; s < a.k; ++s) { a.list[(int64_t)i * a.k + s] = a.n; a.dists[(int64_t)i * a.k + s] = 0.0; }
          a.nontriv[i] = 0; a.h_out[i] = a.r0[i].w;
        } else {
          const int slot = atomicAdd(a.nwork, 1);
          a.worklist[slot] = i;
          if (a.list16) {
            uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(a.k);
            out[0] = (uint16_t)LIST_ESCAPED;
            if (a.k > 2) { out[1] = (uint16_t)(slot & 0xffff); out[2] = (uint16_t)(slot >> 16); }
          }
          atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
        }
      }
      continue;
    }
    // ---- stage the tile: one bulk copy per row ----
    if (warp == 0) {
      fence_proxy_async();
      const int plen = (lane < TILE_MAX_ROWS) ? (rows->toff[lane + 1 > TILE_MAX_ROWS ? TILE_MAX_ROWS : lane + 1] - rows->toff[lane]) : 0;
      unsigned bytes = (lane < TILE_MAX_ROWS) ? (unsigned)plen * 32u : 0u;
      unsigned tot = bytes;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);
      if (lane == 0) mbar_expect_tx(&misc->bar, tot);
      __syncwarp();
      if (bytes) bulk_g2s(tile + rows->toff[lane], a.r0 + rows->ps[lane], bytes, &misc->bar);
    }
    // segments of the four target runs (threads 32..43), while the tile arrives
    if (tid >= 32 && tid < 44) {
      const int q = (tid - 32) / 3, s = (tid - 32) % 3;
      const int cy = 2 * gy + (q & 1), cz = 2 * gz + (q >> 1);
      int lo = 0, hi = 0;
      const int zz = cz - 1 + s;
      if (cy < g.ny && cz < g.nz && zz >= 0 && zz < g.nz) {
        const int ylo = max(cy - 1, 0), yhi = min(cy + 1, g.ny - 1);
        const int r_lo = (zz - rows->z0) * rows->ysp + (ylo - rows->y0);
        const int r_hi = (zz - rows->z0) * rows->ysp + (yhi - rows->y0);
        lo = rows->toff[r_lo]; hi = rows->toff[r_hi + 1];
      }
      misc->seg[q][s][0] = lo; misc->seg[q][s][1] = hi;
    }
    mbar_wait(&misc->bar, phase);
    phase ^= 1u;
    // ---- tile-relative fp32 coordinates in cell units (padding slots and non-finite particles: NaN) ----
    {
      const double tx0 = (double)rows->x0, ty0 = (double)rows->y0, tz0 = (double)rows->z0;
      const int nrows = rows->nrows;
      for (int r = warp; r < nrows; r += KG_WARPS) {
        const int o = rows->toff[r], plen = rows->toff[r + 1] - o, ps = rows->ps[r], rs = rows->rs[r], re = rows->re[r];
        for (int s = lane; s < plen; s += 32) {
          const int gi = ps + s;
          const bool real = gi >= rs && gi < re;
          const double4 q = tile[o + s];
          xf[o + s] = real ? (float)((q.x - g.ox) * g.inv_c - tx0) : nanf("");
          yf[o + s] = (float)((q.y - g.oy) * g.inv_c - ty0);
          zf[o + s] = (float)((q.z - g.oz) * g.inv_c - tz0);
          if (SORTED) rowbase[o + s] = gi;
        }
      }
      for (int s = T + tid; s < T + 64; s += KG_THREADS) { xf[s] = nanf(""); yf[s] = 0.f; zf[s] = 0.f; }
    }
    // ---- target table: thread t prepares target t ----
    for (int t = tid; t < ntgt; t += KG_THREADS) {
      int q = 0;
      while (q < 3 && t >= misc->tbase[q + 1]) ++q;
      const int i = misc->tbegin[q] + (t - misc->tbase[q]);
      const int row = misc->trow[q];
      const int off = rows->toff[row] + (i - rows->ps[row]);
      const double4 p = tile[off];
      double R2try = -1.0, R2lim = a.dist2;
      const bool finite = isfinite(p.x) && isfinite(p.y) && isfinite(p.z);
      const bool ghost = a.skip_id && a.skip_id[i] >= a.skip_from;
      if (!finite || ghost) {                                 // cKDTree raises on non-finite input; ghosts: an empty, well-defined row
        if (a.list16) a.list16[(int64_t)i * LIST_STRIDE(a.k)] = (uint16_t)LIST_MISSING;
        if (SORTED) for (int s = 0; s < a.k; ++s) { a.list[(int64_t)i * a.k + s] = a.n; a.dists[(int64_t)i * a.k + s] = 0.0; }
        a.nontriv[i] = 0;
        a.h_out[i] = (ghost && finite) ? p.w : 0.0;
        if (!finite) atomicAdd(&a.flags[FL_KNN_NONFINITE], 1);
      } else {
        const int cy = 2 * gy + (q & 1), cz = 2 * gz + (q >> 1);
        // distance to the faces of the scanned region that are not faces of the whole grid
        const int sx0 = rows->x0, sx1 = rows->x1;
        const int sy0 = max(cy - 1, 0), sy1 = min(cy + 1, g.ny - 1);
        const int sz0 = max(cz - 1, 0), sz1 = min(cz + 1, g.nz - 1);
        double rs = INFINITY;
        if (sx0 > 0) rs = fmin(rs, p.x - (g.ox + (double)sx0 * g.c));
        if (sx1 < g.nx - 1) rs = fmin(rs, (g.ox + (double)(sx1 + 1) * g.c) - p.x);
        if (sy0 > 0) rs = fmin(rs, p.y - (g.oy + (double)sy0 * g.c));
        if (sy1 < g.ny - 1) rs = fmin(rs, (g.oy + (double)(sy1 + 1) * g.c) - p.y);
        if (sz0 > 0) rs = fmin(rs, p.z - (g.oz + (double)sz0 * g.c));
        if (sz1 < g.nz - 1) rs = fmin(rs, (g.oz + (double)(sz1 + 1) * g.c) - p.z);
        if (rs < INFINITY) { rs = fmax(rs, 0.0) * (1.0 - 1e-6); R2lim = fmin(a.dist2, rs * rs); }
        double hg = a.use_guess ? p.w : 0.0;
        if (hg > 0.0) { hg *= a.guess_factor; R2try = hg * hg; }
        else {
          const double vol = (double)(sx1 - sx0 + 1) * (double)rows->nrows * g.c * g.c * g.c;
          const double R3 = 1.3 * (double)a.k * vol / (4.1887902047863905 * (double)(rows->total > 0 ? rows->total : 1));
          const double Rt = cbrt(R3);
          R2try = Rt * Rt;
        }
        R2try = fmin(R2try, R2lim);
        if (!(R2try > 0.0)) R2try = 0.0;
        // The previous k-th distance already exceeds what the scanned region covers: the scan would almost surely end
        // in the work list, so the target goes there at once (the general kernel is exact whatever the reason).
        if (a.use_guess && p.w > 0.0 && p.w * p.w > R2lim && R2lim < a.dist2) R2try = -2.0;
      }
      t_R2lim[t] = R2lim; t_R2try[t] = R2try; t_i[t] = i; t_word[t] = off | (q << 16);
    }
    __syncthreads();
    // ---- the search: one warp per target ----
    for (int t = warp; t < ntgt; t += KG_WARPS) {
      double R2try = t_R2try[t];
      if (R2try == -2.0) { if (lane == 0) { t_R2try[t] = 0.0; t_word[t] = 0x100; } continue; }   // straight to the work list
      if (R2try < 0.0) { if (lane == 0) t_word[t] = -1; continue; }
      const double R2lim = t_R2lim[t];
      const int word = t_word[t];
      const int off = word & 0xffff, q = word >> 16;
      const int i = t_i[t];
      const double4 p = tile[off];
      const float pxf = xf[off], pyf = yf[off], pzf = zf[off];
      int count = 0, verdict = 1, scans = 0, nsl = 0;
      double kr[KG_SLICES]; int ir[KG_SLICES];
      for (int attempt = 0; attempt < KG_MAX_ATTEMPT && verdict == 1; ++attempt) {
        ++scans;
        __syncwarp();
        const double Rc = sqrt(R2try) * g.inv_c + 1.7320508075688772 * eta;
        const float T2f = __double2float_ru(Rc * Rc * (1.0 + 4.0 * u24));
        int nf = ATOM ? -1 : knn_filter_private(xf, yf, zf, misc->seg[q], pxf, pyf, pzf, T2f, F, Lcol, lane);
        if (nf < 0) { __syncwarp(); nf = knn_filter<ATOM>(xf, yf, zf, misc->seg[q], pxf, pyf, pzf, T2f, F, &misc->cntF[warp], lane); }
        if (nf > KG_CAPF) {                                   // too many: shrink the trial radius, scan again
          const double f = cbrt(0.75 * (double)KG_CAPF / (double)nf);
          R2try *= f * f;
          continue;
        }
        __syncwarp();
        // exact stage over the nf survivors: up to four per lane, kept in registers
        nsl = (nf + 31) >> 5;
        count = 0;
#pragma unroll
        for (int r = 0; r < KG_SLICES; ++r) {
          kr[r] = -1.0; ir[r] = -1;
          if (r < nsl) {
            const int e = 32 * r + lane;
            if (e < nf) {
              const int o = F[e];
              const double4 c = tile[o];
              const double dx = c.x - p.x, dy = c.y - p.y, dz = c.z - p.z;
              const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
              if (d2 < R2try) { kr[r] = d2; ir[r] = o; }
            }
            count += __popc(__ballot_sync(0xffffffffu, ir[r] >= 0));
          }
        }
        if (count >= a.k) { verdict = 0; break; }
        if (R2try < R2lim) {                                  // too few: grow towards the covered radius
          const double f = cbrt(1.4 * (double)a.k / (double)(count > 0 ? count : 1));
          R2try = fmin(R2lim, R2try * fmax(f * f, 1.1));
          continue;
        }
        verdict = (R2lim >= a.dist2) ? 0 : 2;                // everything inside the cut was seen / k-th neighbour may lie outside
      }
      double h2 = 0.0;
      if (verdict == 0) {
        if (SORTED) {                                          // API layout: rank the candidates through shared memory
          const unsigned lt = (1u << lane) - 1u;
          int run = 0;
          __syncwarp();
#pragma unroll
          for (int r = 0; r < KG_SLICES; ++r) {
            if (r < nsl) {
              const unsigned bb = __ballot_sync(0xffffffffu, ir[r] >= 0);
              if (ir[r] >= 0) { const int pos = run + __popc(bb & lt); keybuf[pos] = kr[r]; idxbuf[pos] = ir[r]; }
              run += __popc(bb);
            }
          }
          h2 = knn_emit<true>(a, i, off, count, keybuf, idxbuf, rowbase, lane);
        } else {
          bool ok;
          if (nsl <= 2 && count >= a.k && count - a.k <= KG_FAST_EXCESS) h2 = knn_select2(a, i, off, count, kr[0], kr[1], ir[0], ir[1], lane, ok);
          else h2 = knn_select16(a, i, off, count, nsl, R2try, kr, ir, lane, ok);
          if (!ok) { verdict = 2; count = 0; }
        }
      }
      if (lane == 0) {                                         // result word: nsel | unresolved << 8 | rescans << 16
        t_R2try[t] = h2;
        t_word[t] = min(count, a.k) | ((verdict != 0) << 8) | ((scans - 1) << 16);
      }
    }
    __syncthreads();
    // ---- per-target scalars, one thread per target ----
    for (int t = tid; t < ntgt; t += KG_THREADS) {
      const int res = t_word[t];
      if (res < 0) continue;
      const int i = t_i[t];
      if (res & 0x100) {
        const int slot = atomicAdd(a.nwork, 1);
        a.worklist[slot] = i;
        if (a.list16) {
          uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(a.k);
          out[0] = (uint16_t)LIST_ESCAPED;
          if (a.k > 2) { out[1] = (uint16_t)(slot & 0xffff); out[2] = (uint16_t)(slot >> 16); }
        }
        atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
      } else {
        a.h_out[i] = sqrt(t_R2try[t]);
        a.nontriv[i] = res & 0xff;
      }
      if (res >> 16) atomicAdd(&a.flags[FL_KNN_RESCAN], res >> 16);
    }
  }
}

int knn_group_max_k() { return KG_CAP - 8; }   // k + 3 candidates must fit two slices comfortably; 56 as before

// Fast path + general-kernel clean-up of the targets it could not resolve inside their tile.
// `dists` != nullptr selects the API layout (rows in `list` sorted by distance, distances returned);
// otherwise rows go to `list16` as tile offsets and the escaped rows to `list` by work-list slot.
int launch_knn_group(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double margin, uint16_t* list16,
                     int32_t* list, double* dists, double* h_out, int32_t* nontriv, cudaStream_t st) {
  if (k < 1 || k > knn_group_max_k()) SPH_FAIL("knn_group: k=%d outside [1, %d]", k, knn_group_max_k());
  SPH_TRY(devbuf_reserve(ctx->worklist, (size_t)(n + 1) * sizeof(int32_t)));
  ctx->unit_tab_valid = 0;                                    // the pair passes' table of active units depends on the rows
  int* flags = bufp<int>(ctx->flags);
  int* nwork = flags + FL_KNN_NWORK;
  CUDA_TRY(cudaMemsetAsync(nwork, 0, 2 * sizeof(int), st));   // FL_KNN_NWORK and FL_GROUP_NEXT are adjacent
  KnnGroupArgs a;
  a.r0 = bufp<double4>(ctx->r0);
  a.cell_start = bufp<uint32_t>(ctx->cell_start);
  a.spec = bufp<GridSpec>(ctx->grid_dev);
  a.n = n; a.k = k;
  a.dist2 = dist * dist;
  if (!(a.dist2 == a.dist2)) a.dist2 = INFINITY;
  a.use_guess = use_guess;
  a.guess_factor = 1.0 + margin;
  a.list16 = list16; a.list = list; a.dists = dists; a.h_out = h_out; a.nontriv = nontriv;
  a.flags = flags;
  a.worklist = bufp<int>(ctx->worklist);
  a.nwork = nwork;
  a.skip_id = ctx->skip_id; a.skip_from = ctx->skip_from;
  a.group_list = bufp<int>(ctx->group_list);
  const int blocks = 148 * KG_MIN_CTAS;
  if (!dists && ctx->knn_mma && list16) {
    SPH_TRY(launch_knn_mma(ctx, a, st));
  } else if (dists) {
    const int smem = KG_SMEM + TILE_CAP * 4;
    CUDA_TRY(cudaFuncSetAttribute(knn_group_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    knn_group_kernel<true, false><<<blocks, KG_THREADS, smem, st>>>(a);
  } else if (ctx->knn_atomic) {
    CUDA_TRY(cudaFuncSetAttribute(knn_group_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, KG_SMEM));
    knn_group_kernel<false, true><<<blocks, KG_THREADS, KG_SMEM, st>>>(a);
  } else {
    CUDA_TRY(cudaFuncSetAttribute(knn_group_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, KG_SMEM));
    knn_group_kernel<false, false><<<blocks, KG_THREADS, KG_SMEM, st>>>(a);
  }
  if (dists || !ctx->knn_mma || !list16) ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  // unresolved targets: general warp-per-target kernel over the work list (grid-stride, count on device)
  SPH_TRY(launch_knn_worklist(ctx, n, k, dist, use_guess, 1.0 + margin, bufp<int>(ctx->worklist), nwork, list, dists, h_out,
                              nontriv, dists ? 0 : 1, st));
  return 0;
}
