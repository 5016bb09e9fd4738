// Group-tiled exact k-NN search, tensor-core filter (option knn_mma=1; the default fast path is knn_group.cu).
// Same contract and the same stages as knn_group.cu (exact fp64 distances in cKDTree's arithmetic, the k smallest
// by (d^2, index), rows as uint16 tile offsets; replaces /root/reference/sph/navier_stokes_cleaned.py:544,548);
// only the FILTER that finds the ~50 candidates worth an exact distance differs:
//
//   s(b, a) = |b|^2 - 2 a.b   is a 4-term dot product of  (x_b, y_b, z_b, |b|^2)  with  (-2 x_a, -2 y_a, -2 z_a, 1),
//
// so one mma.sync.m16n8k4 (tf32 inputs, fp32 accumulation) gives s for 16 candidates x 8 targets, and a candidate is
// a survivor of target a when s < T_a - |a|^2.  A warp takes a BLOCK of 8 targets of one target run (they share the
// scanned region), streams the region's candidates through the tensor core 16 at a time and records the hits in
// bit masks (one bit per 16-candidate chunk and output position: the sign of s - threshold) - 2 shared-memory
// loads, 1 MMA, 4 subtractions and 4 funnel shifts per 128 target-candidate pairs, against ~46 instructions of the
// fp32 scalar filter.  The masks are then
// expanded into one candidate list per target and each target goes through the exact stage and the selection
// exactly as in knn_group.cu.  The filter only has to be a SUPERSET of the candidates inside the trial radius:
// coordinates are taken relative to the tile centre in cell units (|x| <= 2), rounded to tf32 (relative 2^-11),
// and the threshold is widened by a rigorous bound of the rounding (below), which lets ~2 % more candidates through.
// A target whose list overflows or holds too few candidates falls back to a scalar scan with a changed radius.
#include "common.cuh"
#include "kernels.cuh"
#include "tile.cuh"
#include "knn_select.cuh"
#include <float.h>

#define KM_CAPT 96                                              // candidates a target's list can hold
#define KM_BV_LEN (TILE_CAP + 64)
#define KM_OFF_TILE 0                                           // double4 tile[TILE_CAP]
#define KM_OFF_BV (KM_OFF_TILE + TILE_CAP * 32)                 // float4 bv[KM_BV_LEN]: tf32 (x, y, z, |b|^2), centre-relative
#define KM_OFF_F (KM_OFF_BV + KM_BV_LEN * 16)                   // per warp: uint16 F[8][KM_CAPT]
#define KM_OFF_CNT (KM_OFF_F + KG_WARPS * 8 * KM_CAPT * 2)      // per warp: int cnt[8]
#define KM_OFF_TGT (KM_OFF_CNT + KG_WARPS * 8 * 4)              // target table
#define KM_TGT_BYTES (TILE_MAX_TARGETS * (8 + 8 + 4 + 4))       // R2lim | R2try/h2 | i | off+code / result
#define KM_OFF_ROWS (KM_OFF_TGT + KM_TGT_BYTES)                 // TileRows
#define KM_OFF_MISC (KM_OFF_ROWS + 384)
#define KM_SMEM (KM_OFF_MISC + 256)

struct KmMisc {
  int tbegin[4], tend[4], trow[4], tbase[5];    // target runs (tile.cuh group_target_run) and their prefix
  int seg[4][3][2];                             // per run: the three z-layer segments [start, end) in tile offsets
  int group;
  int bbase[5];                                 // prefix of the number of 8-target blocks per run
  uint64_t bar;
};

__device__ __forceinline__ float to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

// D (16 candidates x 8 targets) = A (16 x 4, row) * B (4 x 8, col), fp32 accumulation from zero
__device__ __forceinline__ void mma_16x8x4(float a0, float a1, float b0, float& c0, float& c1, float& c2, float& c3) {
  asm("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%7,%7,%7};"
      : "=f"(c0), "=f"(c1), "=f"(c2), "=f"(c3)
      : "r"(__float_as_uint(a0)), "r"(__float_as_uint(a1)), "r"(__float_as_uint(b0)), "f"(0.0f));
}

// Rounding bound of the filter, in cell units.  A coordinate |x| <= 2.01 is rounded twice: to fp32 when the tile is
// converted (eta0) and to tf32 (2^-11 relative); a rounded point therefore lies within DELTA of the true one.  The
// tensor core forms  |b^|^2~ - 2 a^.b^  from the rounded points: the products of two tf32 numbers are exact in fp32,
// the stored norm |b^|^2~ carries one fp32 and one tf32 rounding (<= 12.2 * (2^-11 + 2^-23)) and the accumulation
// of four terms of magnitude <= 25 at most 1e-4 even if the adder truncates.
#define KM_COORD_MAX 2.01
#define KM_DELTA (1.7320508075688772 * (KM_COORD_MAX * 4.8828125e-4 + 2.0e-6))
#define KM_SLACK (12.2 * (4.8828125e-4 + 1.2e-7) + 1.0e-4 + 1.0e-4)

// Scalar fall-back filter over the three segments (one candidate per lane and trip, ballot compaction).
__device__ __forceinline__ int km_filter_scalar(const float4* __restrict__ bv, const int (*seg)[2], float4 p, float T2f,
                                                uint16_t* F, int lane) {
  const unsigned lt = (1u << lane) - 1u;
  int count = 0;
#pragma unroll 1
  for (int s = 0; s < 3; ++s) {
    const int a = seg[s][0], b = seg[s][1];
    for (int base = a; base < b; base += 32) {
      const int c = base + lane;
      const float4 q = bv[c];
      const float dx = q.x - p.x, dy = q.y - p.y, dz = q.z - p.z;
      const bool ok = (c < b) && (q.w < 1.0e29f) && (dx * dx + dy * dy + dz * dz < T2f);   // w = 1e30: padding / non-finite
      const unsigned bal = __ballot_sync(0xffffffffu, ok);
      const int pos = count + __popc(bal & lt);
      if (ok && pos < KM_CAPT) F[pos] = (uint16_t)c;
      count += __popc(bal);
    }
  }
  return count;
}

__global__ void __launch_bounds__(KG_THREADS, 3) knn_mma_kernel(KnnGroupArgs a) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;                  // mma fragment coordinates
  double4* tile = reinterpret_cast<double4*>(smem + KM_OFF_TILE);
  float4* bv = reinterpret_cast<float4*>(smem + KM_OFF_BV);
  uint16_t* Fw = reinterpret_cast<uint16_t*>(smem + KM_OFF_F) + warp * 8 * KM_CAPT;
  int* cntw = reinterpret_cast<int*>(smem + KM_OFF_CNT) + warp * 8;
  double* t_R2lim = reinterpret_cast<double*>(smem + KM_OFF_TGT);
  double* t_R2try = t_R2lim + TILE_MAX_TARGETS;               // in: trial radius^2 (< 0: nothing to do); out: k-th d^2
  int* t_i = reinterpret_cast<int*>(t_R2try + TILE_MAX_TARGETS);
  int* t_word = t_i + TILE_MAX_TARGETS;                       // in: tile offset | run << 16; out: result word
  TileRows* rows = reinterpret_cast<TileRows*>(smem + KM_OFF_ROWS);
  KmMisc* misc = reinterpret_cast<KmMisc*>(smem + KM_OFF_MISC);
  const GridSpec g = *a.spec;
  if (tid == 0) { mbar_init(&misc->bar, 1); mbar_fence_init(); }
  uint32_t phase = 0;
  const double u24 = 5.9604644775390625e-8;                   // 2^-24

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
      const int len = have ? (e - b) : 0;
      int inc = len, binc = (len + 7) >> 3;
      const int nb = binc;
#pragma unroll
      for (int d = 1; d < 4; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, d), w = __shfl_up_sync(0xffffffffu, binc, d);
        if (lane >= d) { inc += v; binc += w; }
      }
      if (lane < 4) {
        misc->tbegin[lane] = b; misc->tend[lane] = e; misc->trow[lane] = r; misc->tbase[lane] = inc - len;
        misc->bbase[lane] = binc - nb;
      }
      if (lane == 3) { misc->tbase[4] = inc; misc->bbase[4] = binc; }
    }
    __syncthreads();
    const int T = tile_padded_total(rows);
    const int ntgt = misc->tbase[4];
    if (ntgt == 0) continue;                                  // (uniform: every thread reads the same shared value)
    int lens[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) lens[q] = misc->tend[q] - misc->tbegin[q];
    if (tile_crowded(g, rows, lens)) {
      // the tile does not fit: every target of the group goes to the general kernel
      if (tid == 0) atomicAdd(&a.flags[FL_KNN_CROWDED], 1);
      for (int t = tid; t < ntgt; t += KG_THREADS) {
        int q = 0;
        while (q < 3 && t >= misc->tbase[q + 1]) ++q;
        const int i = misc->tbegin[q] + (t - misc->tbase[q]);
        const bool ghost = a.skip_id && a.skip_id[i] >= a.skip_from;
        if (ghost) {
          a.list16[(int64_t)i * LIST_STRIDE(a.k)] = (uint16_t)LIST_MISSING;
          a.nontriv[i] = 0; a.h_out[i] = a.r0[i].w;
        } else {
          const int slot = atomicAdd(a.nwork, 1);
          a.worklist[slot] = i;
          uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(a.k);
          out[0] = (uint16_t)LIST_ESCAPED;
          if (a.k > 2) { out[1] = (uint16_t)(slot & 0xffff); out[2] = (uint16_t)(slot >> 16); }
          atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
        }
      }
      continue;
    }
    // ---- stage the tile: one bulk copy per row ----
    if (warp == 0) {
      fence_proxy_async();
      const int plen = (lane < TILE_MAX_ROWS) ? (rows->toff[lane + 1] - rows->toff[lane]) : 0;
      const unsigned bytes = (unsigned)plen * 32u;
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
    // ---- tf32 vectors (x, y, z, |b|^2) relative to the tile centre, in cell units; padding slots and non-finite
    //      particles become (0, 0, 0, 1e30): s = 1e30 is above every threshold ----
    {
      const double tx0 = (double)rows->x0 + 2.0, ty0 = (double)rows->y0 + 2.0, tz0 = (double)rows->z0 + 2.0;
      const int nrows = rows->nrows;
      for (int r = warp; r < nrows; r += KG_WARPS) {
        const int o = rows->toff[r], plen = rows->toff[r + 1] - o, ps = rows->ps[r], rs = rows->rs[r], re = rows->re[r];
        for (int s = lane; s < plen; s += 32) {
          const int gi = ps + s;
          const bool real = gi >= rs && gi < re;
          const double4 q = tile[o + s];
          const float x = to_tf32((float)((q.x - g.ox) * g.inv_c - tx0));
          const float y = to_tf32((float)((q.y - g.oy) * g.inv_c - ty0));
          const float z = to_tf32((float)((q.z - g.oz) * g.inv_c - tz0));
          const float n2 = to_tf32(x * x + y * y + z * z);
          const bool fin = real && isfinite(q.x) && isfinite(q.y) && isfinite(q.z);
          bv[o + s] = fin ? make_float4(x, y, z, n2) : make_float4(0.f, 0.f, 0.f, 1.0e30f);   // never a survivor
        }
      }
      for (int s = T + tid; s < T + 64; s += KG_THREADS) bv[s] = make_float4(0.f, 0.f, 0.f, 1.0e30f);
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
        a.list16[(int64_t)i * LIST_STRIDE(a.k)] = (uint16_t)LIST_MISSING;
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
      }
      t_R2lim[t] = R2lim; t_R2try[t] = R2try; t_i[t] = i; t_word[t] = off | (q << 16);
    }
    __syncthreads();
    // ---- the search: one warp per block of 8 targets of one run ----
    const int nblocks = misc->bbase[4];
    for (int blk = warp; blk < nblocks; blk += KG_WARPS) {
      int q = 0;
      while (q < 3 && blk >= misc->bbase[q + 1]) ++q;
      const int t_first = misc->tbase[q] + 8 * (blk - misc->bbase[q]);
      const int nt = min(8, misc->tbase[q + 1] - t_first);
      // thresholds: lane j < nt prepares target t_first + j
      float th = -INFINITY;
      float4 pv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (lane < nt) {
        const double R2 = t_R2try[t_first + lane];
        if (R2 >= 0.0) {
          pv = bv[t_word[t_first + lane] & 0xffff];
          const double Rc = sqrt(R2) * g.inv_c + 2.0 * KM_DELTA;
          const double n2 = (double)pv.x * (double)pv.x + (double)pv.y * (double)pv.y + (double)pv.z * (double)pv.z;
          th = __double2float_ru((Rc * Rc + KM_SLACK - n2) + 1.0e-5);
        }
      }
      // B fragment: column gid = target t_first + gid, row tig = component tig of (-2x, -2y, -2z, 1)
      float b0;
      {
        const float bx = __shfl_sync(0xffffffffu, pv.x, gid), by = __shfl_sync(0xffffffffu, pv.y, gid);
        const float bz = __shfl_sync(0xffffffffu, pv.z, gid);
        b0 = (tig == 0) ? -2.0f * bx : (tig == 1) ? -2.0f * by : (tig == 2) ? -2.0f * bz : 1.0f;
      }
      const float th0 = __shfl_sync(0xffffffffu, th, 2 * tig), th1 = __shfl_sync(0xffffffffu, th, 2 * tig + 1);
      // scan: chunk j covers 16 candidates.  A hit is the sign bit of (s - threshold); every output position shifts
      // it into its own mask (newest chunk at bit 0), the first 32 chunks move to the "h" masks when they are full.
      unsigned m00 = 0u, m01 = 0u, m10 = 0u, m11 = 0u, h00 = 0u, h01 = 0u, h10 = 0u, h11 = 0u;
      const float* bvf = reinterpret_cast<const float*>(bv);
      int cs1 = 0, cs2 = 0, nchunk = 0;                       // first chunk of segments 1 and 2, chunks in total
      const float inval = (tig == 3) ? 1.0e30f : 0.0f;        // a candidate that is never a survivor: (0, 0, 0, 1e30)
#pragma unroll 1
      for (int s = 0; s < 3; ++s) {
        const int sa = misc->seg[q][s][0], sb = misc->seg[q][s][1];
        if (s == 1) cs1 = nchunk;
        if (s == 2) cs2 = nchunk;
#pragma unroll 2
        for (int base = sa; base < sb; base += 16) {
          if (nchunk == 32) { h00 = m00; h01 = m01; h10 = m10; h11 = m11; m00 = m01 = m10 = m11 = 0u; }
          const int c0i = base + gid, c1i = c0i + 8;
          float a0 = bvf[4 * c0i + tig], a1 = bvf[4 * c1i + tig];
          a0 = (c0i < sb) ? a0 : inval;                       // beyond the segment
          a1 = (c1i < sb) ? a1 : inval;
          float c0, c1, c2, c3;
          mma_16x8x4(a0, a1, b0, c0, c1, c2, c3);
          m00 = __funnelshift_l(__float_as_uint(c0 - th0), m00, 1);
          m01 = __funnelshift_l(__float_as_uint(c1 - th1), m01, 1);
          m10 = __funnelshift_l(__float_as_uint(c2 - th0), m10, 1);
          m11 = __funnelshift_l(__float_as_uint(c3 - th1), m11, 1);
          ++nchunk;
        }
      }
      // expand the masks into the per-target lists: the 8 lanes with the same `tig` own targets 2 tig and 2 tig + 1;
      // an exclusive scan of their hit counts (lanes 4 apart) gives every lane its place in the two lists
      {
        const int nlo = min(nchunk, 32);                      // chunks held by the h masks when nchunk > 32, else by m
        const int cA = __popc(m00) + __popc(m10) + __popc(h00) + __popc(h10);
        const int cB = __popc(m01) + __popc(m11) + __popc(h01) + __popc(h11);
        int iA = cA, iB = cB;
#pragma unroll
        for (int d = 4; d < 32; d <<= 1) {
          const int va = __shfl_up_sync(0xffffffffu, iA, d), vb = __shfl_up_sync(0xffffffffu, iB, d);
          if (lane >= d) { iA += va; iB += vb; }
        }
        if (gid == 7) { cntw[2 * tig] = iA; cntw[2 * tig + 1] = iB; }
        int posA = iA - cA, posB = iB - cB;
        const int a0s = misc->seg[q][0][0], a1s = misc->seg[q][1][0], a2s = misc->seg[q][2][0];
        uint16_t* FA = Fw + (2 * tig) * KM_CAPT;
        uint16_t* FB = FA + KM_CAPT;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
          unsigned m = (w == 0) ? m00 : (w == 1) ? m01 : (w == 2) ? m10 : (w == 3) ? m11 : (w == 4) ? h00 : (w == 5) ? h01 : (w == 6) ? h10 : h11;
          const bool hi = w >= 4;                             // h masks: chunks 0..31 (only when nchunk > 32)
          // bit position b of a mask holding chunks [j0, j0 + cnt): chunk j0 + cnt - 1 - b
          const int top = hi ? 31 : ((nchunk > 32) ? nchunk - 1 : nlo - 1);
          const int coff = gid + ((w & 2) ? 8 : 0);
          uint16_t* Fl = (w & 1) ? FB : FA;
          int pos = (w & 1) ? posB : posA;
          while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1u;
            const int j = top - b;
            const int cbase = (j >= cs2) ? a2s + 16 * (j - cs2) : (j >= cs1) ? a1s + 16 * (j - cs1) : a0s + 16 * j;
            if (pos < KM_CAPT) Fl[pos] = (uint16_t)(cbase + coff);
            ++pos;
          }
          if (w & 1) posB = pos; else posA = pos;
        }
      }
      __syncwarp();
      // ---- per target: exact stage, selection, row ----
      for (int tt = 0; tt < nt; ++tt) {
        const int t = t_first + tt;
        double R2try = t_R2try[t];
        if (R2try < 0.0) { if (lane == 0) t_word[t] = -1; continue; }
        const double R2lim = t_R2lim[t];
        const int off = t_word[t] & 0xffff;
        const int i = t_i[t];
        const double4 p = tile[off];
        const float4 pf = bv[off];
        uint16_t* F = Fw + tt * KM_CAPT;
        int count = 0, verdict = 1, scans = 0, nsl = 0;
        double kr[KG_SLICES]; int ir[KG_SLICES];
        for (int attempt = 0; attempt < KG_MAX_ATTEMPT && verdict == 1; ++attempt) {
          ++scans;
          __syncwarp();
          int nf;
          if (attempt == 0) nf = cntw[tt];
          else {
            const double Rc = sqrt(R2try) * g.inv_c + 2.0 * KM_DELTA;
            const float T2f = __double2float_ru(Rc * Rc * (1.0 + 8.0 * u24));
            nf = km_filter_scalar(bv, misc->seg[q], pf, T2f, F, lane);
          }
          if (nf > KM_CAPT) {                                 // too many: shrink the trial radius, scan again
            const double f = cbrt(0.75 * (double)KM_CAPT / (double)nf);
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
          if (R2try < R2lim) {                                // too few: grow towards the covered radius
            const double f = cbrt(1.4 * (double)a.k / (double)(count > 0 ? count : 1));
            R2try = fmin(R2lim, R2try * fmax(f * f, 1.1));
            continue;
          }
          verdict = (R2lim >= a.dist2) ? 0 : 2;              // everything inside the cut was seen / k-th neighbour may lie outside
        }
        double h2 = 0.0;
        if (verdict == 0) {
          bool ok;
          if (nsl <= 2 && count >= a.k && count - a.k <= KG_FAST_EXCESS) h2 = knn_select2(a, i, off, count, kr[0], kr[1], ir[0], ir[1], lane, ok);
          else h2 = knn_select16(a, i, off, count, nsl, R2try, kr, ir, lane, ok);
          if (!ok) { verdict = 2; count = 0; }
        }
        if (lane == 0) {                                       // result word: nsel | unresolved << 8 | rescans << 16
          t_R2try[t] = h2;
          t_word[t] = min(count, a.k) | ((verdict != 0) << 8) | ((scans - 1) << 16);
        }
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
        uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(a.k);
        out[0] = (uint16_t)LIST_ESCAPED;
        if (a.k > 2) { out[1] = (uint16_t)(slot & 0xffff); out[2] = (uint16_t)(slot >> 16); }
        atomicAdd(&a.flags[FL_KNN_WIDEN], 1);
      } else {
        a.h_out[i] = sqrt(t_R2try[t]);
        a.nontriv[i] = res & 0xff;
      }
      if (res >> 16) atomicAdd(&a.flags[FL_KNN_RESCAN], res >> 16);
    }
  }
}

// Session layout only (rows to list16); the caller (launch_knn_group) runs the general kernel over the work list.
int launch_knn_mma(sphb200_ctx* ctx, const KnnGroupArgs& a, cudaStream_t st) {
  CUDA_TRY(cudaFuncSetAttribute(knn_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, KM_SMEM));
  knn_mma_kernel<<<148 * 3, KG_THREADS, KM_SMEM, st>>>(a);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
