// Group-tiled exact fixed-k nearest-neighbour search, LANE-PER-TARGET form: the session's fast path.
// Replaces `kdt.query(points, N_NEIGH, eps, 2, dist)` of /root/reference/sph/navier_stokes_cleaned.py:544
// (eps = 0) and the row-max smoothing length of :548 for a search that starts from the previous
// step's k-th distances.  (knn_group.cu holds the warp-per-target form: it ranks rows by distance and retries
// with a corrected radius, which the function-level API needs; here a target that does not settle at its first
// trial radius simply goes to the general kernel.)
//
// One CTA (4 warps) per GROUP of 2x2x2 cells (tile.cuh).  The group's tile of position records is staged by bulk
// asynchronous copies (cp.async.bulk -> UBLKCP, mbarrier completion) and converted once into tile-relative fp32
// coordinates.  A warp then takes 32 targets of one cell row of the group at a time, ONE TARGET PER LANE:
//   FILTER  every lane walks the same candidate sequence (the three z-layers of the targets' 3x3x3 block) two
//           candidates per step; the candidate coordinates are warp-uniform shared-memory reads (broadcast), the
//           distance arithmetic is packed fp32 (FADD2/FFMA2), a survivor costs its lane one predicated store: no
//           ballots, no prefix sums, no per-target loop overhead;
//   EXACT   d^2 = ((dx*dx + dy*dy) + dz*dz) in fp64 with separately rounded products (the arithmetic of cKDTree
//           and numpy) for the lane's ~55 survivors; each survivor's key (float-rounded d^2, truncated) is packed
//           with its tile offset and counted into a per-lane histogram of the expected range of the k-th distance;
//   SELECT  the histogram gives the bin of the k-th smallest; candidates in lower bins are kept, the few in that
//           bin are ordered by their exact (d^2, offset);
//   EMIT    rows leave through shared memory as uint16 tile offsets, self in column 0, coalesced.
#include "common.cuh"
#include "kernels.cuh"
#include "tile.cuh"
#include <float.h>

#define KL_THREADS 128
#define KL_WARPS 4
#define KL_CAPF 88            // filter survivors a lane can hold
#define KL_CSTR 33            // column stride of the per-warp slot matrix (conflict-free both ways)
#define KL_NBIN 48            // histogram bins of the expected range (+ one below, one above)
#define KL_MAXC 4             // candidates of the k-th bin a lane resolves exactly
#define KL_DROPPED 0xffffffffu
#ifndef KL_MIN_CTAS
#define KL_MIN_CTAS 3
#endif

struct KnnLaneArgs {
  const double4* r0;           // sorted {x, y, z, guess}, n + 1 entries
  const uint32_t* cell_start;
  const GridSpec* spec;        // device
  int n, k;
  double dist2;
  int use_guess;
  double guess_factor;         // (1 + margin)
  uint16_t* list16;            // (N, LIST_STRIDE(k)) tile offsets
  double* h_out;               // (N)
  int32_t* nontriv;            // (N)
  int* flags;
  int* worklist;
  int* nwork;
  const uint32_t* skip_id;
  uint32_t skip_from;
};

// ---- shared-memory layout (bytes) ----
#define KL_XF_LEN (TILE_CAP + 64)
#define KL_OFF_XF 0                                              // float xf, yf, zf [KL_XF_LEN]
#define KL_OFF_GIDX (KL_OFF_XF + 3 * KL_XF_LEN * 4)              // int gidx[TILE_CAP]: tile offset -> sorted index
#define KL_SLOT_BYTES ((KL_CAPF + 1) * KL_CSTR * 4)              // per warp: uint32 slot[KL_CAPF + 1][KL_CSTR]
#define KL_OFF_SLOT (KL_OFF_GIDX + TILE_CAP * 4)                 // (also the landing area of the fp64 tile)
#define KL_HIST_BYTES ((KL_NBIN + 2) * 32)                       // per warp: uint8 hist[KL_NBIN + 2][32]
#define KL_OFF_HIST (KL_OFF_SLOT + KL_WARPS * KL_SLOT_BYTES)
#define KL_OFF_ROWS (KL_OFF_HIST + KL_WARPS * KL_HIST_BYTES)
#define KL_OFF_MISC (KL_OFF_ROWS + 384)
#define KL_SMEM (KL_OFF_MISC + 320)
static_assert(KL_WARPS * KL_SLOT_BYTES >= TILE_CAP * 32, "the slot area must hold the staged fp64 tile");

struct KlMisc {
  int tbegin[4], tend[4], trow[4];
  int seg[4][3][2];
  int item_q[12], item_c[12], nitems;           // (run, chunk of 32 targets)
  int group;
  uint64_t bar;
};

__device__ __forceinline__ unsigned long long kl_sub(unsigned long long a, unsigned long long b) {
  unsigned long long r; asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ unsigned long long kl_mul(unsigned long long a, unsigned long long b) {
  unsigned long long r; asm("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ unsigned long long kl_fma(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ unsigned long long kl_pack(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ double kl_d2(const double4& c, const double4& p) {
  const double dx = c.x - p.x, dy = c.y - p.y, dz = c.z - p.z;
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}
// histogram bin of a truncated float key: 0 below the range, 1 .. KL_NBIN inside, KL_NBIN + 1 above (monotone)
__device__ __forceinline__ int kl_bin(unsigned key, float lo, float scale) {
  const float f = __uint_as_float(key & ~0x7ffu);
  if (f < lo) return 0;
  const float t = (f - lo) * scale;
  return (t >= (float)KL_NBIN) ? KL_NBIN + 1 : (int)t + 1;
}

__global__ void __launch_bounds__(KL_THREADS, KL_MIN_CTAS) knn_lane_kernel(KnnLaneArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* xf = reinterpret_cast<float*>(smem + KL_OFF_XF);
  float* yf = xf + KL_XF_LEN;
  float* zf = yf + KL_XF_LEN;
  int* gidx = reinterpret_cast<int*>(smem + KL_OFF_GIDX);
  double4* stage = reinterpret_cast<double4*>(smem + KL_OFF_SLOT);              // fp64 tile while it is converted
  uint32_t* slot = reinterpret_cast<uint32_t*>(smem + KL_OFF_SLOT + warp * KL_SLOT_BYTES) + lane;   // this lane's column
  uint8_t* hist = reinterpret_cast<uint8_t*>(smem + KL_OFF_HIST + warp * KL_HIST_BYTES) + lane;
  TileRows* rows = reinterpret_cast<TileRows*>(smem + KL_OFF_ROWS);
  KlMisc* misc = reinterpret_cast<KlMisc*>(smem + KL_OFF_MISC);
  const GridSpec g = *a.spec;
  const int k = a.k, ks = LIST_STRIDE(a.k);
  if (tid == 0) { mbar_init(&misc->bar, 1); mbar_fence_init(); }
  uint32_t phase = 0;
  const double u24 = 5.9604644775390625e-8;                   // 2^-24
  const double eta = u24 * (2.0 * 5.0 + 5.0) + 1e-9;          // rounding of a tile-relative fp32 coordinate difference (|coord| <= 5)

  for (;;) {
    __syncthreads();                                          // previous group done with every buffer
    if (tid == 0) misc->group = atomicAdd(&a.flags[FL_GROUP_NEXT], 1);
    __syncthreads();
    const int group = misc->group;
    if (group >= g.ngroups) break;
    int gx, gy, gz;
    group_coords(g, group, &gx, &gy, &gz);
    if (warp == 0) {
      tile_rows(g, a.cell_start, gx, gy, gz, lane, rows);
      int b = 0, e = 0, r = 0;
      const bool have = (lane < 4) && group_target_run(g, a.cell_start, gx, gy, gz, lane, &b, &e, &r);
      if (lane < 4) { misc->tbegin[lane] = b; misc->tend[lane] = have ? e : b; misc->trow[lane] = r; }
      __syncwarp();
      if (lane == 0) {                                        // work items: (run, chunk of 32 targets)
        int ni = 0;
        for (int q = 0; q < 4; ++q)
          for (int c = 0; c * 32 < misc->tend[q] - misc->tbegin[q] && ni < 12; ++c) { misc->item_q[ni] = q; misc->item_c[ni] = c; ++ni; }
        misc->nitems = ni;
      }
    }
    __syncthreads();
    const int T = tile_padded_total(rows);
    int lens[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) lens[q] = misc->tend[q] - misc->tbegin[q];
    const int ntgt = lens[0] + lens[1] + lens[2] + lens[3];
    if (ntgt == 0) continue;
    if (tile_crowded(g, rows, lens)) {
      // the tile does not fit: every target of the group goes to the general kernel
      if (tid == 0) atomicAdd(&a.flags[FL_KNN_CROWDED], 1);
      for (int q = 0; q < 4; ++q)
        for (int i = misc->tbegin[q] + tid; i < misc->tend[q]; i += KL_THREADS) {
          uint16_t* out = a.list16 + (int64_t)i * ks;
          if (a.skip_id && a.skip_id[i] >= a.skip_from) {
            out[0] = (uint16_t)LIST_MISSING; a.nontriv[i] = 0; a.h_out[i] = a.r0[i].w;
          } else {
            const int s = atomicAdd(a.nwork, 1);
            a.worklist[s] = i;
            out[0] = (uint16_t)LIST_ESCAPED; out[1] = (uint16_t)(s & 0xffff); out[2] = (uint16_t)(s >> 16);
            atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
          }
        }
      continue;
    }
    // ---- stage the fp64 tile (one bulk copy per row) into the slot area ----
    if (warp == 0) {
      fence_proxy_async();
      const int plen = (lane < TILE_MAX_ROWS) ? (rows->toff[lane + 1] - rows->toff[lane]) : 0;
      unsigned tot = (unsigned)plen * 32u;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);
      if (lane == 0) mbar_expect_tx(&misc->bar, tot);
      __syncwarp();
      if (plen) bulk_g2s(stage + rows->toff[lane], a.r0 + rows->ps[lane], (unsigned)plen * 32u, &misc->bar);
    }
    if (tid >= 32 && tid < 44) {                               // z-layer segments of the four target runs
      const int q = (tid - 32) / 3, s = (tid - 32) % 3;
      const int cy = 2 * gy + (q & 1), cz = 2 * gz + (q >> 1);
      int lo = 0, hi = 0;
      const int zz = cz - 1 + s;
      if (cy < g.ny && cz < g.nz && zz >= 0 && zz < g.nz) {
        const int ylo = max(cy - 1, 0), yhi = min(cy + 1, g.ny - 1);
        lo = rows->toff[(zz - rows->z0) * rows->ysp + (ylo - rows->y0)];
        hi = rows->toff[(zz - rows->z0) * rows->ysp + (yhi - rows->y0) + 1];
      }
      misc->seg[q][s][0] = lo; misc->seg[q][s][1] = hi;
    }
    mbar_wait(&misc->bar, phase);
    phase ^= 1u;
    // ---- tile-relative fp32 coordinates in cell units (padding slots and non-finite particles: NaN) ----
    {
      const double tx0 = (double)rows->x0, ty0 = (double)rows->y0, tz0 = (double)rows->z0;
      const int nrows = rows->nrows;
      for (int r = warp; r < nrows; r += KL_WARPS) {
        const int o = rows->toff[r], plen = rows->toff[r + 1] - o, ps = rows->ps[r], rs = rows->rs[r], re = rows->re[r];
        for (int s = lane; s < plen; s += 32) {
          const int gi = ps + s;
          const bool real = gi >= rs && gi < re;
          const double4 q = stage[o + s];
          xf[o + s] = real ? (float)((q.x - g.ox) * g.inv_c - tx0) : nanf("");
          yf[o + s] = (float)((q.y - g.oy) * g.inv_c - ty0);
          zf[o + s] = (float)((q.z - g.oz) * g.inv_c - tz0);
          gidx[o + s] = gi;
        }
      }
      for (int s = T + tid; s < T + 64; s += KL_THREADS) { xf[s] = nanf(""); yf[s] = 0.f; zf[s] = 0.f; }
    }
    __syncthreads();                                          // floats ready; the slot area is free again
    // ---- the search: 32 targets of one run per warp pass, one target per lane ----
    uint32_t* wslot = reinterpret_cast<uint32_t*>(smem + KL_OFF_SLOT + warp * KL_SLOT_BYTES);
    for (int item = warp; item < misc->nitems; item += KL_WARPS) {
      const int q = misc->item_q[item];
      const int tb = misc->tbegin[q], len = misc->tend[q] - tb;
      const int tl = misc->item_c[item] * 32 + lane;
      const bool valid = tl < len;
      const int i = tb + (valid ? tl : 0);
      const int trow = misc->trow[q];
      const int off = rows->toff[trow] + (i - rows->ps[trow]);
      const double4 p = a.r0[i];
      const bool finite = isfinite(p.x) && isfinite(p.y) && isfinite(p.z);
      const bool ghost = a.skip_id && a.skip_id[i] >= a.skip_from;
      const bool active = valid && finite && !ghost;
      if (valid && !active) {                                 // cKDTree raises on non-finite input; ghosts: an empty, well-defined row
        a.list16[(int64_t)i * ks] = (uint16_t)LIST_MISSING;
        a.nontriv[i] = 0;
        a.h_out[i] = (ghost && finite) ? p.w : 0.0;
        if (!finite) atomicAdd(&a.flags[FL_KNN_NONFINITE], 1);
      }
      double R2lim = a.dist2, R2try = 0.0;
      float T2f = -1.0f;
      if (active) {
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
          const double R3 = 1.3 * (double)k * vol / (4.1887902047863905 * (double)(rows->total > 0 ? rows->total : 1));
          const double Rt = cbrt(R3);
          R2try = Rt * Rt;
        }
        R2try = fmin(R2try, R2lim);
        if (!(R2try > 0.0)) R2try = 0.0;
        const double Rc = sqrt(R2try) * g.inv_c + 1.7320508075688772 * eta;
        T2f = __double2float_ru(Rc * Rc * (1.0 + 4.0 * u24));
      }
      // ---- FILTER: warp-uniform candidate stream, per-lane threshold, per-lane survivor column ----
      int nf = 0;
      {
        const unsigned long long PX = kl_pack(xf[off], xf[off]), PY = kl_pack(yf[off], yf[off]), PZ = kl_pack(zf[off], zf[off]);
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
          const int sa = misc->seg[q][s][0], sb = misc->seg[q][s][1];
#pragma unroll 4
          for (int c = sa; c < sb; c += 2) {
            const unsigned long long X = *reinterpret_cast<const unsigned long long*>(xf + c);
            const unsigned long long Y = *reinterpret_cast<const unsigned long long*>(yf + c);
            const unsigned long long Z = *reinterpret_cast<const unsigned long long*>(zf + c);
            const unsigned long long dx = kl_sub(X, PX), dy = kl_sub(Y, PY), dz = kl_sub(Z, PZ);
            const unsigned long long d = kl_fma(dz, dz, kl_fma(dy, dy, kl_mul(dx, dx)));
            if (__uint_as_float((unsigned)d) < T2f) { slot[nf * KL_CSTR] = (uint32_t)c; nf = min(nf + 1, KL_CAPF); }
            if (__uint_as_float((unsigned)(d >> 32)) < T2f) { slot[nf * KL_CSTR] = (uint32_t)(c + 1); nf = min(nf + 1, KL_CAPF); }
          }
        }
      }
      const bool over = nf >= KL_CAPF;                        // the column is full: the trial radius was far too large
      if (!active || over) nf = 0;
      const int m = __reduce_max_sync(0xffffffffu, nf);
      // ---- EXACT: fp64 distances of the survivors, packed keys, histogram of the expected range of the k-th ----
#pragma unroll 1
      for (int b = 0; b < KL_NBIN + 2; ++b) hist[b * 32] = 0;
      float blo = 0.f, bscale = 0.f;
      if (nf > 0) {
        const double ratio = fmin(1.0, (double)k / (double)nf);
        const double x = R2try * cbrt(ratio * ratio);         // expected k-th d^2 for uniformly spread candidates
        const double lo = 0.70 * x, hi = fmin(1.30 * x, R2try);
        blo = (float)lo;
        bscale = (float)((double)KL_NBIN / fmax(hi - lo, 1e-300));
      }
      int cnt = 0;
      double d2max = 0.0;
#pragma unroll 1
      for (int e0 = 0; e0 < m; e0 += 4) {
        // four survivors per trip: their position records are fetched together (independent global loads in flight)
        unsigned o[4];
        double4 c[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) o[u] = (e0 + u < nf) ? slot[(e0 + u) * KL_CSTR] : (unsigned)off;
#pragma unroll
        for (int u = 0; u < 4; ++u) c[u] = a.r0[gidx[o[u]]];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (e0 + u < nf) {
            const double d2 = kl_d2(c[u], p);
            unsigned v = KL_DROPPED;
            if (d2 < R2try) {
              const unsigned key = __float_as_uint(__double2float_ru(d2)) & ~0x7ffu;
              v = key | o[u];
              hist[kl_bin(key, blo, bscale) * 32] += 1;
              ++cnt;
              d2max = fmax(d2max, d2);
            }
            slot[(e0 + u) * KL_CSTR] = v;
          }
        }
      }
      // ---- verdict per lane ----
      bool accept = false, escape = false;
      if (active) {
        if (over) escape = true;
        else if (cnt >= k) accept = true;
        else if (R2try >= R2lim && R2lim >= a.dist2) accept = true;   // everything inside the cut was seen
        else escape = true;                                            // too few at this radius: the general kernel widens
      }
      const bool selecting = accept && cnt > k;
      // ---- SELECT: bin of the k-th smallest ----
      int B = KL_NBIN + 2;
      if (selecting) {
        int cum = 0;
        B = -1;
#pragma unroll 1
        for (int b = 0; b < KL_NBIN + 2; ++b) {
          const int h = hist[b * 32];
          if (B < 0 && cum + h >= k) B = b;
          if (B < 0) cum += h;
        }
      }
      int w = 0, ps = 0, ng = 0;
      unsigned c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll 2
      for (int e = 0; e < m; ++e) {
        if (accept && e < nf) {
          const unsigned v = slot[e * KL_CSTR];
          if (v != KL_DROPPED) {
            const int b = selecting ? kl_bin(v, blo, bscale) : -1;
            if (b < B) {
              if ((int)(v & 0x7ffu) == off) ps = w;
              slot[w * KL_CSTR] = v & 0x7ffu;
              ++w;
            } else if (b == B) {
              if (ng == 0) c0 = v; else if (ng == 1) c1 = v; else if (ng == 2) c2 = v; else if (ng == 3) c3 = v;
              ++ng;
            }
          }
        }
      }
      double h2 = d2max;
      if (selecting) {
        if (ng > KL_MAXC) { accept = false; escape = true; }  // too many equal keys around the k-th: exact path elsewhere
        else {
          const int need = k - w;
          const unsigned cv[KL_MAXC] = {c0, c1, c2, c3};
          double dj[KL_MAXC];
#pragma unroll
          for (int j = 0; j < KL_MAXC; ++j) dj[j] = (j < ng) ? kl_d2(a.r0[gidx[cv[j] & 0x7ffu]], p) : INFINITY;
          h2 = 0.0;
#pragma unroll
          for (int j = 0; j < KL_MAXC; ++j) {
            if (j < ng) {
              int rank = 0;
#pragma unroll
              for (int j2 = 0; j2 < KL_MAXC; ++j2)
                if (j2 != j && j2 < ng) rank += (dj[j2] < dj[j]) || (dj[j2] == dj[j] && (cv[j2] & 0x7ffu) < (cv[j] & 0x7ffu));
              if (rank < need) {
                if ((int)(cv[j] & 0x7ffu) == off) ps = w;
                slot[w * KL_CSTR] = cv[j] & 0x7ffu;
                ++w;
                h2 = fmax(h2, dj[j]);
              }
            }
          }
        }
      }
      // ---- per-target scalars ----
      if (accept) { a.h_out[i] = sqrt(h2); a.nontriv[i] = w; }
      if (escape) {
        const int s = atomicAdd(a.nwork, 1);
        a.worklist[s] = i;
        uint16_t* out = a.list16 + (int64_t)i * ks;
        out[0] = (uint16_t)LIST_ESCAPED; out[1] = (uint16_t)(s & 0xffff); out[2] = (uint16_t)(s >> 16);
        atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
      }
      __syncwarp();
      // ---- EMIT: rows leave transposed through the slot matrix (coalesced uint16 stores), self in column 0 ----
      const int wsend = accept ? w : -1;
      for (int t = 0; t < 32; ++t) {
        const int wt = __shfl_sync(0xffffffffu, wsend, t);
        if (wt < 0) continue;
        const int it2 = __shfl_sync(0xffffffffu, i, t), pst = __shfl_sync(0xffffffffu, ps, t);
        uint16_t* out = a.list16 + (int64_t)it2 * ks;
        for (int s = lane; s < k; s += 32) {
          const unsigned val = (s < wt) ? wslot[s * KL_CSTR + t] : LIST_MISSING;
          const int pos = (s >= wt) ? s : ((s == pst) ? 0 : (s < pst ? s + 1 : s));
          out[pos] = (uint16_t)val;
        }
      }
      __syncwarp();
    }
  }
}

int launch_knn_lane(sphb200_ctx* ctx, int n, int k, double dist, int use_guess, double margin, uint16_t* list16,
                    int32_t* list, double* h_out, int32_t* nontriv, cudaStream_t st) {
  if (k < 3 || k > knn_group_max_k()) SPH_FAIL("knn_lane: k=%d outside [3, %d]", k, knn_group_max_k());
  SPH_TRY(devbuf_reserve(ctx->worklist, (size_t)(n + 1) * sizeof(int32_t)));
  ctx->unit_tab_valid = 0;                                    // the pair passes' table of active units depends on the rows
  int* flags = bufp<int>(ctx->flags);
  int* nwork = flags + FL_KNN_NWORK;
  CUDA_TRY(cudaMemsetAsync(nwork, 0, 2 * sizeof(int), st));   // FL_KNN_NWORK and FL_GROUP_NEXT are adjacent
  KnnLaneArgs a;
  a.r0 = bufp<double4>(ctx->r0);
  a.cell_start = bufp<uint32_t>(ctx->cell_start);
  a.spec = bufp<GridSpec>(ctx->grid_dev);
  a.n = n; a.k = k;
  a.dist2 = dist * dist;
  if (!(a.dist2 == a.dist2)) a.dist2 = INFINITY;
  a.use_guess = use_guess;
  a.guess_factor = 1.0 + margin;
  a.list16 = list16; a.h_out = h_out; a.nontriv = nontriv;
  a.flags = flags;
  a.worklist = bufp<int>(ctx->worklist);
  a.nwork = nwork;
  a.skip_id = ctx->skip_id; a.skip_from = ctx->skip_from;
  CUDA_TRY(cudaFuncSetAttribute(knn_lane_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, KL_SMEM));
  knn_lane_kernel<<<148 * KL_MIN_CTAS, KL_THREADS, KL_SMEM, st>>>(a);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  // targets that did not settle: general warp-per-target kernel over the work list (rows by slot in `list`)
  SPH_TRY(launch_knn_worklist(ctx, n, k, dist, use_guess, 1.0 + margin, bufp<int>(ctx->worklist), nwork, list, nullptr, h_out,
                              nontriv, 1, st));
  return 0;
}
