// Pieces the group-tiled k-NN kernels share (knn_group.cu: fp32 scalar filter; knn_mma.cu: tensor-core filter):
// the kernel arguments and the register-resident selection of the k smallest candidates.
#pragma once
#include "common.cuh"
#include "tile.cuh"

#define KG_THREADS 256
#define KG_WARPS 8
#define KG_MAX_ATTEMPT 16

struct KnnGroupArgs {
  const double4* r0;           // sorted {x, y, z, guess}, n + 1 entries
  const uint32_t* cell_start;
  const GridSpec* spec;        // device
  int n, k;
  double dist2;
  int use_guess;               // trial radius from r0[i].w (previous k-th distance) when > 0
  double guess_factor;         // (1 + margin)
  uint16_t* list16;            // session layout: (N,k) tile offsets, or nullptr
  int32_t* list;               // API layout: (N,k) sorted-space indices ascending by distance (with dists)
  double* dists;               // (N,k) or nullptr
  double* h_out;               // (N)
  int32_t* nontriv;            // (N)
  int* flags;
  int* worklist;               // unresolved targets (general kernel)
  int* nwork;
  const int* group_list;       // [0] = number of groups that hold particles, [1..] their ids (cells.cu); nullptr: every group
  const uint32_t* skip_id;     // slab decomposition: targets with skip_id[i] >= skip_from are ghosts (their owner
  uint32_t skip_from;          // searches for them); nullptr = every particle is a target
};


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


// SESSION layout, selection in registers.  Every lane holds up to four candidates (exact d^2, tile offset; absent:
// idx < 0) of which `count` are valid, count >= k.  Keeps the k smallest by (d^2, offset), writes them as the
// target's uint16 row (unordered, self in column 0) and returns the squared k-th distance; ok = false when more
// candidates tie in the float-rounded key than the removal step can separate (the caller escapes the target).
// The float keys (rounded up, so monotone in the exact key) drive everything: a bisection on the key finds a
// threshold that keeps between k and k + 3 candidates (four ballots per step), the at most three extra ones are
// removed largest-first with REDUX-max, the exact (d^2, offset) order deciding only where float keys tie.
#define KG_SLICES 4
__device__ __forceinline__ double knn_select16(const KnnGroupArgs& a, int i, int self, int count, int nsl, double R2try,
                                               double (&kr)[KG_SLICES], int (&ir)[KG_SLICES], int lane, bool& ok) {
  const int k = a.k;
  unsigned fr[KG_SLICES];
#pragma unroll
  for (int r = 0; r < KG_SLICES; ++r) fr[r] = (ir[r] >= 0) ? __float_as_uint(__double2float_ru(kr[r])) + 1u : 0u;   // 0 = absent
  ok = true;
  int kept = count;
  if (count > k + 3) {
    // keys of valid entries are in [1, hi); find tau with k <= #{0 < f < tau} <= k + 3
    unsigned lo = __float_as_uint(__double2float_rd(0.35 * R2try)) + 1u, hi = __float_as_uint(__double2float_ru(R2try)) + 2u;
    {
      int c = 0;
#pragma unroll
      for (int r = 0; r < KG_SLICES; ++r) if (r < nsl) c += __popc(__ballot_sync(0xffffffffu, fr[r] != 0u && fr[r] < lo));
      if (c >= k) lo = 1u;
    }
    unsigned tau = 0u;
    for (int it = 0; it < 40; ++it) {
      if (hi - lo <= 1u) break;
      const unsigned mid = lo + ((hi - lo) >> 1);
      int c = 0;
#pragma unroll
      for (int r = 0; r < KG_SLICES; ++r) if (r < nsl) c += __popc(__ballot_sync(0xffffffffu, fr[r] != 0u && fr[r] < mid));
      if (c < k) lo = mid;
      else if (c > k + 3) hi = mid;
      else { tau = mid; kept = c; break; }
    }
    if (tau == 0u) { ok = false; return 0.0; }
#pragma unroll
    for (int r = 0; r < KG_SLICES; ++r) if (fr[r] >= tau) { fr[r] = 0u; ir[r] = -1; }
  }
  // remove the (kept - k) largest
  for (int rm = kept; rm > k; --rm) {
    unsigned fm = 0u;
#pragma unroll
    for (int r = 0; r < KG_SLICES; ++r) if (r < nsl) fm = max(fm, fr[r]);
    const unsigned m = __reduce_max_sync(0xffffffffu, fm);
    int owners = 0;
#pragma unroll
    for (int r = 0; r < KG_SLICES; ++r) if (r < nsl) owners += __popc(__ballot_sync(0xffffffffu, fr[r] == m));
    if (owners == 1) {
#pragma unroll
      for (int r = 0; r < KG_SLICES; ++r) if (fr[r] == m) { fr[r] = 0u; ir[r] = -1; }
    } else {                                                   // equal float keys: exact (key, offset) maximum among them
      double kb = -1.0; int ib = -1;
#pragma unroll
      for (int r = 0; r < KG_SLICES; ++r)
        if (r < nsl && fr[r] == m && (kr[r] > kb || (kr[r] == kb && ir[r] > ib))) { kb = kr[r]; ib = ir[r]; }
      warp_max_pair(kb, ib);
#pragma unroll
      for (int r = 0; r < KG_SLICES; ++r) if (fr[r] == m && ir[r] == ib) { fr[r] = 0u; ir[r] = -1; }
    }
  }
  // squared k-th distance: the exact maximum of what is left
  double km = -1.0; int im = -1;
  {
    unsigned fm = 0u;
#pragma unroll
    for (int r = 0; r < KG_SLICES; ++r) if (r < nsl) fm = max(fm, fr[r]);
    const unsigned m = __reduce_max_sync(0xffffffffu, fm);
#pragma unroll
    for (int r = 0; r < KG_SLICES; ++r)
      if (r < nsl && fr[r] == m && m != 0u && (kr[r] > km || (kr[r] == km && ir[r] > im))) { km = kr[r]; im = ir[r]; }
    warp_max_pair(km, im);
  }
  // the row: kept entries in (slice, lane) order, self moved to column 0
  const unsigned lt = (1u << lane) - 1u;
  uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(k);
  unsigned bm[KG_SLICES];
  int ps = -1, run = 0;
#pragma unroll
  for (int r = 0; r < KG_SLICES; ++r) {
    bm[r] = 0u;
    if (r < nsl) {
      bm[r] = __ballot_sync(0xffffffffu, ir[r] >= 0);
      const unsigned sm = __ballot_sync(0xffffffffu, ir[r] == self);
      if (sm) ps = run + __popc(bm[r] & ((1u << (__ffs(sm) - 1)) - 1u));
      run += __popc(bm[r]);
    }
  }
  int basepos = 0;
#pragma unroll
  for (int r = 0; r < KG_SLICES; ++r) {
    if (r < nsl) {
      if (ir[r] >= 0) {
        const int pos = basepos + __popc(bm[r] & lt);
        out[(pos == ps) ? 0 : (pos < ps ? pos + 1 : pos)] = (uint16_t)ir[r];
      }
      basepos += __popc(bm[r]);
    }
  }
  for (int s2 = basepos + lane; s2 < k; s2 += 32) out[s2] = (uint16_t)LIST_MISSING;
  return (basepos > 0 && km > 0.0) ? km : 0.0;
}

// SESSION layout, the common case: at most two candidates per lane (count <= 64) and a few more than k of them.
// The (count - k) largest are removed one at a time: a warp REDUX-max over the float-rounded keys, then every lane
// clears its entries that equal the maximum - no ballots in the loop.  Two candidates whose float keys tie would be
// removed together; the count of what is left is checked afterwards and ok = false sends the target to the
// general kernel (exact by construction), so a tie can cost time but never a wrong row.
#define KG_FAST_EXCESS 28
__device__ __forceinline__ double knn_select2(const KnnGroupArgs& a, int i, int self, int count, double k0, double k1, int i0,
                                              int i1, int lane, bool& ok) {
  const int k = a.k;
  unsigned f0 = (i0 >= 0) ? __float_as_uint(__double2float_ru(k0)) + 1u : 0u;   // 0 = absent / removed
  unsigned f1 = (i1 >= 0) ? __float_as_uint(__double2float_ru(k1)) + 1u : 0u;
#pragma unroll 1
  for (int rm = count; rm > k; --rm) {
    const unsigned m = __reduce_max_sync(0xffffffffu, max(f0, f1));
    f0 = (f0 == m) ? 0u : f0;
    f1 = (f1 == m) ? 0u : f1;
  }
  const unsigned b0 = __ballot_sync(0xffffffffu, f0 != 0u), b1 = __ballot_sync(0xffffffffu, f1 != 0u);
  const int n0 = __popc(b0);
  ok = (n0 + __popc(b1) == k);
  if (!ok) return 0.0;
  // squared k-th distance: the exact maximum of what is left
  const unsigned m = __reduce_max_sync(0xffffffffu, max(f0, f1));
  const unsigned o0 = __ballot_sync(0xffffffffu, f0 == m), o1 = __ballot_sync(0xffffffffu, f1 == m);
  double km;
  if (__popc(o0) + __popc(o1) == 1) {
    km = __shfl_sync(0xffffffffu, o0 ? k0 : k1, __ffs(o0 | o1) - 1);
  } else {
    double kb = -1.0; int ib = -1;
    if (f0 == m) { kb = k0; ib = i0; }
    if (f1 == m && (k1 > kb || (k1 == kb && i1 > ib))) { kb = k1; ib = i1; }
    warp_max_pair(kb, ib);
    km = kb;
  }
  // the row: kept entries in (slice, lane) order, self swapped into column 0
  const unsigned lt = (1u << lane) - 1u;
  const unsigned s0 = __ballot_sync(0xffffffffu, f0 != 0u && i0 == self), s1 = __ballot_sync(0xffffffffu, f1 != 0u && i1 == self);
  int ps = -1;
  if (s0) ps = __popc(b0 & ((1u << (__ffs(s0) - 1)) - 1u));
  else if (s1) ps = n0 + __popc(b1 & ((1u << (__ffs(s1) - 1)) - 1u));
  uint16_t* out = a.list16 + (int64_t)i * LIST_STRIDE(k);
  if (f0 != 0u) { const int pos = __popc(b0 & lt); out[(pos == ps) ? 0 : (pos == 0 && ps > 0 ? ps : pos)] = (uint16_t)i0; }
  if (f1 != 0u) { const int pos = n0 + __popc(b1 & lt); out[(pos == ps) ? 0 : (pos == 0 && ps > 0 ? ps : pos)] = (uint16_t)i1; }
  return (km > 0.0) ? km : 0.0;
}

