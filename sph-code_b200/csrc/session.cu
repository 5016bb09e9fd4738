// Device-resident particle state and the full time step in the reference driver's order
// (/root/reference/sph/code_running.py:217-540, gas/dust only; conventions in DESIGN.md §step).
// The state lives in HBM in the CURRENT SORT ORDER (cell order of the last neighbour build) and is
// re-gathered through the new permutation each step; `id` maps back to the caller's order.
#include "common.cuh"
#include "kernels.cuh"
#include "../../include/sphb200.h"
#include "tile.cuh"
#include <float.h>
#include <math.h>
#include <new>

enum Field { F_POS = 0, F_VEL, F_ACC, F_MASS, F_H, F_T, F_E, F_MU, F_GAMMA, F_TYPE, F_RHO0, F_NFIELDS };
static const int kFieldWidth[F_NFIELDS] = {3, 3, 3, 1, 1, 1, 1, 1, 1, 1, 1};
static const char* kFieldName[F_NFIELDS] = {"points", "velocities", "total_accel", "mass", "sizes", "T", "E_internal",
                                            "mu_array", "gamma_array", "particle_type", "densities_0"};
enum Out { O_RHO = 0, O_NUM, O_DUST, O_HACC, O_VACC, O_HEAT, O_PACC, O_NOUT };
static const int kOutWidth[O_NOUT] = {1, 1, 1, 3, 3, 1, 3};
static const char* kOutName[O_NOUT] = {"densities", "num_densities", "dust_densities", "hydro_accel", "visc_accel",
                                       "visc_heat", "pressure_accel"};   // + "grav_accel" (download only, when gravity is on)
// sbuf layout: [2*f + which] state double buffers, then ids, outputs, scratch
#define SB_ID 24
#define SB_ID_ALT 25
#define SB_OUT 26            // .. 26 + O_NOUT
#define SB_NONTRIV 34
#define SB_DISTSQ 35
#define SB_KNNH 37           // k-th neighbour distance of the last search (guess for the next one)
#define SB_KNNH_ALT 38
#define SB_POS 39            // caller index -> sorted position (inverse of id), built on demand
#define SB_GM 40             // drag: mean grain mass / cross-section per particle, CALLER order (nsc:727-728)
#define SB_GC 41
#define SB_ONTO 42           // drag: accel_onto, accel_reaction (sorted order, scratch of the force phase)
#define SB_REACT 43
#define SB_GID 46            // slab decomposition: global particle id (double), travels with the particle
#define SB_GID_ALT 47

struct SessionScalars {      // host mirror of the run parameters
  double dt_0, max_age, n_int, raise_factor, au;
};

struct Session {
  int64_t n = 0;
  int k = 0;
  double dist = 0;
  int cur[F_NFIELDS];        // which of the two buffers is current
  int id_cur = 0;
  int knnh_cur = 0;
  int has_list = 0;
  int use_guess = 1;
  int pos_valid = 0;         // SB_POS matches the current id
  int drag = 0;              // dust-gas drag (net_impulse) folded into the force phase
  int grown = 0;             // particles were appended since the last kick: the old arrays have the old
                             // length, so this step takes dv = a dt (:485) and new_smoothing = 1 (:535)
  int64_t n_owned = -1;      // slab decomposition: caller indices >= n_owned are ghosts (-1: none)
  int gravity = 0;           // F2: total_accel = grav_accel + pressure_accel + visc_accel (code_running.py:477)
  double grav_G = 0, grav_eps = 0;
  int has_gid = 0;           // SB_GID holds global ids (rows carry them in column 18)
  int gid_cur = 0;
  SessionScalars sc;
};
// The session belongs to its context (ctx->session): two contexts on one GPU are two independent runs.
static Session g_no_session;   // k == 0: what the entry points see before sphb200_session_create
static Session& sess(sphb200_ctx* ctx) { return ctx->session ? *static_cast<Session*>(ctx->session) : g_no_session; }
void session_free(sphb200_ctx* ctx) {
  delete static_cast<Session*>(ctx->session);
  ctx->session = nullptr;
}
static double* fld(sphb200_ctx* ctx, int f, int alt = 0) {
  Session& s = sess(ctx);
  return bufp<double>(ctx->sbuf[2 * f + (s.cur[f] ^ alt)]);
}
static double* outp(sphb200_ctx* ctx, int o) { return bufp<double>(ctx->sbuf[SB_OUT + o]); }

__device__ __forceinline__ double nan_to_num_d(double v) {
  if (v != v) return 0.0;
  if (v > DBL_MAX) return DBL_MAX;
  if (v < -DBL_MAX) return -DBL_MAX;
  return v;
}

// ---- A10: time-step choice (device scalar)                          code_running.py:222-231 ----
__global__ void choose_dt_kernel(double* scal, SessionScalars sc) {
  const double ct = scal[SC_CT];
  const double age = scal[SC_AGE];
  double dt = (age == 0.0) ? sc.dt_0 / 10 : fmax(sc.dt_0 / 5., fmin(sc.dt_0 * 2., ct));
  if (ct > sc.max_age) dt = sc.max_age / 100.;
  scal[SC_DT] = dt;
  scal[SC_AGE] = age + dt;                                                       // :447
}

// ---- A10: position clamp + nan_to_num                                code_running.py:233-238 ----
__global__ void __launch_bounds__(256) clamp_kernel(int64_t n3, double lim, double* __restrict__ pos, double* __restrict__ vel) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= n3) return;
  double x = pos[t];
  if (x > lim) x = lim;
  if (x < -lim) x = -lim;
  // velocities[points > lim] = 0.1 is evaluated after the clamp, so it never fires (as written)
  pos[t] = nan_to_num_d(x);
  vel[t] = nan_to_num_d(vel[t]);
}

struct ReorderArgs {
  const double* src[F_NFIELDS];
  double* dst[F_NFIELDS];
  const uint32_t* id_src; uint32_t* id_dst;
  const double* knnh_src; double* knnh_dst;
  const double* gid_src; double* gid_dst;     // nullptr: the session carries no global ids
};

__global__ void __launch_bounds__(256) reorder_kernel(int n, const uint32_t* __restrict__ perm, ReorderArgs a, unsigned mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t p = perm[i];
#pragma unroll
  for (int f = 0; f < 3; ++f) {
    if (!(mask >> f & 1u)) continue;
    a.dst[f][3 * (int64_t)i + 0] = a.src[f][3 * p + 0];
    a.dst[f][3 * (int64_t)i + 1] = a.src[f][3 * p + 1];
    a.dst[f][3 * (int64_t)i + 2] = a.src[f][3 * p + 2];
  }
#pragma unroll
  for (int f = 3; f < F_NFIELDS; ++f) if (mask >> f & 1u) a.dst[f][i] = a.src[f][p];
  a.id_dst[i] = a.id_src[p];
  a.knnh_dst[i] = a.knnh_src[p];
  if (a.gid_src) a.gid_dst[i] = a.gid_src[p];
}

__global__ void __launch_bounds__(256) iota_kernel(int n, uint32_t* id, uint32_t first = 0) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) id[i] = first + (uint32_t)i;
}

// ---- A11: force assembly, kick-drift, heating                        code_running.py:460-491 ----
// also emits |x|^2 (or +inf for fast particles) for the percentile of A12 and counts slow particles.
__global__ void __launch_bounds__(256)
integrate_kernel(int n, const double* __restrict__ scal, double kB, double m_h, double* __restrict__ pos,
                 double* __restrict__ vel, double* __restrict__ acc, double* __restrict__ E, double* __restrict__ T,
                 const double* __restrict__ mass, const double* __restrict__ mu, const double* __restrict__ gamma,
                 const double* __restrict__ ptype, const double* __restrict__ hacc, double* __restrict__ vacc,
                 const double* __restrict__ heat, double* __restrict__ pacc, double* __restrict__ distsq,
                 unsigned long long* __restrict__ nslow, const uint32_t* __restrict__ id, int n_owned, int grown,
                 const double* __restrict__ grav) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  int slow = 0;
  if (i < n) {
    const double dt = scal[SC_DT];
    const double gas = (ptype[i] == 0.0) ? 1.0 : 0.0;
    double v[3], a[3], x[3], av[3], ap[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      v[c] = vel[3 * (int64_t)i + c];
      x[c] = pos[3 * (int64_t)i + c];
      av[c] = vacc[3 * (int64_t)i + c];
      ap[c] = nan_to_num_d(hacc[3 * (int64_t)i + c] * gas);                     // :460
    }
    const double vmag = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    const double amag = sqrt(av[0] * av[0] + av[1] * av[1] + av[2] * av[2]);
    if (vmag - amag * dt < 0) {                                                  // :475
#pragma unroll
      for (int c = 0; c < 3; ++c) av[c] = -v[c] / dt;
    }
    double v2 = 0.0, x2 = 0.0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      a[c] = grav ? (grav[3 * (int64_t)i + c] + ap[c]) + av[c] : ap[c] + av[c];  // :477
      const double xn = x[c] + ((a[c] * (dt * dt)) / 2. + v[c] * dt);            // :481
      const double vn = grown ? v[c] + a[c] * dt                                 // :484-485 (shapes differ)
                              : v[c] + (a[c] + acc[3 * (int64_t)i + c]) / 2. * dt;       // :482-483
      pos[3 * (int64_t)i + c] = xn;
      vel[3 * (int64_t)i + c] = vn;
      acc[3 * (int64_t)i + c] = a[c];
      vacc[3 * (int64_t)i + c] = av[c];
      pacc[3 * (int64_t)i + c] = ap[c];
      v2 += vn * vn;
      x2 += xn * xn;
    }
    const double e = nan_to_num_d(E[i]) + nan_to_num_d(heat[i] * dt);            // :490
    E[i] = e;
    T[i] = nan_to_num_d(e * (mu[i] * m_h) / (gamma[i] * mass[i] * kB));          // :491
    slow = (v2 < 80000.0 * 80000.0) && (!id || id[i] < (uint32_t)n_owned);            // :493 (ghosts belong to their owner)
    distsq[i] = slow ? x2 : INFINITY;
  }
  const int cnt = __syncthreads_count(slow);
  if (threadIdx.x == 0 && cnt) atomicAdd(nslow, (unsigned long long)cnt);
}

// ---- A12: exact percentile by radix select (no sort)                  code_running.py:515-518 ----
// np.percentile(dist_sq[slow], 90) needs two order statistics of |x|^2: ranks lo = floor(0.9 (ns-1)) and
// lo+1.  Non-negative doubles order like their bit patterns, so both are found by a most-significant-digit
// radix select: five passes of 13,13,13,13,12 bits, each one histogram over the keys that still match the
// selected prefix (shared-memory bins, flushed with one global add per non-empty bin) and one single-block
// pick.  The two ranks share a prefix until they fall into different bins; from then on the second prefix
// is tracked with direct global adds (its population is ~n/8192 by then).  Excluded particles carry +inf
// and sort after every finite key.  Between `hist` and `pick` a slab decomposition all_reduces the
// histogram (2*8192 counts + the slow count), which makes the select exact across ranks in five small
// collectives; on one GPU the ten kernels run back to back.
#define SEL_BINS 8192
#define SEL_HIST (2 * SEL_BINS + 1)                      // [A bins | B bins | slow count]
#define SEL_PASSES 5
struct SelectState {
  unsigned long long prefix[2];                          // selected high bits of the two order statistics
  unsigned long long rank[2];                            // remaining rank inside the prefix
  unsigned long long ns;                                 // slow particles (all ranks)
  double frac;                                           // 0.9 (ns-1) - lo
  int pass;                                              // next pass, 0 .. SEL_PASSES
  int split;                                             // the two prefixes differ
};
__host__ __device__ __forceinline__ int sel_width(int pass) { return pass < 4 ? 13 : 12; }
__host__ __device__ __forceinline__ int sel_consumed(int pass) { return 13 * pass; }

__global__ void select_reset_kernel(SelectState* st, long long* hist) {
  for (int b = threadIdx.x; b < SEL_HIST; b += blockDim.x) hist[b] = 0;
  if (threadIdx.x == 0) { st->prefix[0] = st->prefix[1] = 0; st->rank[0] = st->rank[1] = 0; st->ns = 0; st->frac = 0; st->pass = 0; st->split = 0; }
}

__global__ void __launch_bounds__(1024)
select_hist_kernel(int n, const double* __restrict__ keys, const SelectState* __restrict__ st,
                   const unsigned long long* __restrict__ nslow_local, long long* __restrict__ hist) {
  __shared__ unsigned int bins[SEL_BINS];
  const int pass = st->pass;
  if (pass >= SEL_PASSES) return;
  for (int b = threadIdx.x; b < SEL_BINS; b += blockDim.x) bins[b] = 0;
  __syncthreads();
  const int cons = sel_consumed(pass), w = sel_width(pass);
  const unsigned long long pa = st->prefix[0], pb = st->prefix[1];
  const int split = st->split;
  const unsigned long long dmask = (1ull << w) - 1ull;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned long long key = (unsigned long long)__double_as_longlong(keys[i]);
    const unsigned long long pre = cons ? (key >> (64 - cons)) : 0ull;
    const unsigned int dig = (unsigned int)((key >> (64 - cons - w)) & dmask);
    if (pre == pa) atomicAdd(&bins[dig], 1u);
    else if (split && pre == pb) atomicAdd((unsigned long long*)&hist[SEL_BINS + dig], 1ull);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < SEL_BINS; b += blockDim.x)
    if (bins[b]) atomicAdd((unsigned long long*)&hist[b], (unsigned long long)bins[b]);
  if (pass == 0 && blockIdx.x == 0 && threadIdx.x == 0) hist[2 * SEL_BINS] = (long long)*nslow_local;
}

// One block of 1024 threads, every thread owns 8 consecutive bins of `h`: finds the bins holding the ranks
// r0 and r1 (both looked up in the same histogram) and the ranks inside those bins.  Block scan = warp
// shuffles + one pass over the 32 warp totals.
__device__ void sel_find2(const long long* __restrict__ h, unsigned long long r0, unsigned long long r1, int* dig,
                          unsigned long long* rank_in, unsigned long long* warp_tot) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  unsigned long long loc[8], sum = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q) { loc[q] = (unsigned long long)h[t * 8 + q]; sum += loc[q]; }
  unsigned long long inc = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += v;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = warp_tot[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += v;
    }
    warp_tot[lane] = w;                                     // inclusive totals of the warps
  }
  __syncthreads();
  const unsigned long long before = (warp ? warp_tot[warp - 1] : 0ull) + inc - sum;
#pragma unroll
  for (int which = 0; which < 2; ++which) {
    const unsigned long long rank = which ? r1 : r0;
    if (rank >= before && rank < before + sum) {
      unsigned long long c = before;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        if (rank >= c && rank < c + loc[q]) { dig[which] = t * 8 + q; rank_in[which] = rank - c; }
        c += loc[q];
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(1024)
select_pick_kernel(SelectState* st, long long* __restrict__ hist, double* __restrict__ scal, double n_int) {
  __shared__ unsigned long long sh[32];
  const int pass = st->pass;
  if (pass >= SEL_PASSES) return;
  unsigned long long ns = st->ns, rA = st->rank[0], rB = st->rank[1];
  double frac = st->frac;
  if (pass == 0) {
    ns = (unsigned long long)hist[2 * SEL_BINS];
    const double pos = 0.9 * (double)(ns ? ns - 1 : 0);
    rA = (unsigned long long)floor(pos);
    frac = pos - (double)rA;
    rB = (rA + 1 < ns) ? rA + 1 : rA;
  }
  if (ns == 0) {                                           // no slow particle: np.percentile raises; report NaN
    __syncthreads();
    if (threadIdx.x == 0) { scal[SC_NSLOW] = 0.0; scal[SC_MAXDIST] = nan(""); scal[SC_D] = nan(""); st->pass = SEL_PASSES; }
    return;
  }
  const int split = st->split, w = sel_width(pass);
  __shared__ int dig[4];
  __shared__ unsigned long long rin[4];
  if (threadIdx.x < 4) { dig[threadIdx.x] = SEL_BINS - 1; rin[threadIdx.x] = 0; }
  __syncthreads();
  sel_find2(hist, rA, split ? rA : rB, dig, rin, sh);                 // one histogram serves both ranks until they split
  if (split) sel_find2(hist + SEL_BINS, rB, rB, dig + 2, rin + 2, sh);
  const int dA = dig[0], dB = split ? dig[2] : dig[1];
  const unsigned long long nrA = rin[0], nrB = split ? rin[2] : rin[1];
  const unsigned long long pA = (st->prefix[0] << w) | (unsigned long long)dA;
  const unsigned long long pB = (st->prefix[1] << w) | (unsigned long long)dB;
  __syncthreads();
  for (int b = threadIdx.x; b < SEL_HIST; b += blockDim.x) hist[b] = 0;      // ready for the next pass
  if (threadIdx.x == 0) {
    st->prefix[0] = pA; st->prefix[1] = pB; st->rank[0] = nrA; st->rank[1] = nrB;
    st->ns = ns; st->frac = frac; st->split = (pA != pB); st->pass = pass + 1;
    if (pass + 1 == SEL_PASSES) {                          // both keys complete: numpy's linear interpolation
      const double a = __longlong_as_double((long long)pA), b = __longlong_as_double((long long)pB);
      const double diff = b - a;
      double v = a + diff * frac;                          // numpy _lerp
      if (frac >= 0.5) v = b - diff * (1.0 - frac);
      scal[SC_NSLOW] = (double)ns;
      scal[SC_MAXDIST] = v;
      scal[SC_D] = cbrt(v * sqrt(v) / (double)ns / 0.9 * n_int);              // :520-521
    }
  }
}

// ---- A12: smoothing-length update                                    code_running.py:532-540 ----
// lonely = (in-degree + nontriv - 1 < 2).  Every particle lists itself, and a particle with a
// single entry inside the distance cut cannot be listed by anyone else (the cut is symmetric), so
// lonely <=> nontriv <= 1 and the O(N k) in-degree histogram of :527-530 is not needed.
__global__ void __launch_bounds__(256)
smoothing_kernel(int n, const double* __restrict__ scal, double raise_cbrt, const double* __restrict__ rho,
                 double* __restrict__ rho0, const int32_t* __restrict__ nontriv, const double* __restrict__ ptype,
                 double* __restrict__ sizes, int grown) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double d = scal[SC_D];
  const double r = rho[i];
  const double t = ptype[i];
  double h = sizes[i];
  if (t == 0.0) {
    const double ns = grown ? 1.0                       // :534-535 (len(densities_0) != len(densities))
        : (exp(nan_to_num_d(-(r - rho0[i]) / (3 * r))) * 2) / 2. + ((nontriv[i] <= 1) ? 0.1 : 0.0);             // :533
    h = ns * h;                                         // :536
    if (h > d) h = d;                                   // :537
  } else if (t == 2.0) {
    h = d / raise_cbrt;                                 // :538
  }
  sizes[i] = h;
  rho0[i] = r;                                          // :540
}

__global__ void __launch_bounds__(256) copy_kernel(int64_t n, const double* __restrict__ a, double* __restrict__ b) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t < n) b[t] = a[t];
}

// un-permute a (N,w) double array from sort order to the caller's order
__global__ void __launch_bounds__(256)
unsort_kernel(int n, int w, const uint32_t* __restrict__ id, const double* __restrict__ src, double* __restrict__ dst) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * w) return;
  const int i = (int)(t / w), c = (int)(t % w);
  dst[(int64_t)id[i] * w + c] = src[t];
}
// un-permute the (N,k) neighbour rows, every row put into the reference's order (ascending (d^2, index), nsc:544): the fast
// k-NN path keeps rows unordered because the sums do not need the order.  `r0` holds the positions
// the search ran on.  Thread per row, keys recomputed on the fly (download path, not hot).
__device__ __forceinline__ double row_key(const double4* __restrict__ r0, const double4 p, int j, int n) {
  if (j < 0 || j >= n) return INFINITY;
  const double4 q = r0[j];
  const double dx = q.x - p.x, dy = q.y - p.y, dz = q.z - p.z;
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}
__global__ void __launch_bounds__(128)
unsort_list_sorted_kernel(int n, int k, const uint32_t* __restrict__ id, const int32_t* __restrict__ list,
                          const double4* __restrict__ r0, int64_t* __restrict__ dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t* row = list + (int64_t)i * k;
  const double4 p = r0[i];
  int64_t* out = dst + (int64_t)id[i] * k;
  for (int s = 0; s < k; ++s) {
    const int js = row[s];
    const double ks = row_key(r0, p, js, n);
    int rank = 0;
    for (int f = 0; f < k; ++f) {
      const int jf = row[f];
      const double kf = row_key(r0, p, jf, n);
      rank += (kf < ks) || (kf == ks && (jf < js || (jf == js && f < s)));
    }
    out[rank] = (js >= 0 && js < n) ? (int64_t)id[js] : (int64_t)n;
  }
}
__global__ void __launch_bounds__(256)
unsort_i32_kernel(int n, const uint32_t* __restrict__ id, const int32_t* __restrict__ src, int64_t* __restrict__ dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[id[i]] = src[i];
}
__global__ void __launch_bounds__(256)
sort_in_kernel(int n, int w, const uint32_t* __restrict__ id, const double* __restrict__ src, double* __restrict__ dst) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * w) return;
  const int i = (int)(t / w), c = (int)(t % w);
  dst[t] = src[(int64_t)id[i] * w + c];
}

// Single-field access at given caller indices (ghost refresh of a slab decomposition): a handful of
// particles per call, so no full-array permutation.
__global__ void __launch_bounds__(256) invert_id_kernel(int n, const uint32_t* __restrict__ id, uint32_t* __restrict__ pos) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pos[id[i]] = (uint32_t)i;
}
__global__ void __launch_bounds__(256)
gather_at_kernel(int64_t count, int w, int n, const int64_t* __restrict__ cidx, const uint32_t* __restrict__ pos,
                 const double* __restrict__ field, double* __restrict__ out, int* __restrict__ flags) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= count * w) return;
  const int64_t e = t / w; const int q = (int)(t % w);
  const int64_t c = cidx[e];
  if (c < 0 || c >= n) { atomicAdd(&flags[FL_BAD_INDEX], 1); return; }
  out[t] = field[(int64_t)pos[c] * w + q];
}
__global__ void __launch_bounds__(256)
scatter_at_kernel(int64_t count, int w, int n, const int64_t* __restrict__ cidx, const uint32_t* __restrict__ pos,
                  const double* __restrict__ in, double* __restrict__ field, int* __restrict__ flags) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= count * w) return;
  const int64_t e = t / w; const int q = (int)(t % w);
  const int64_t c = cidx[e];
  if (c < 0 || c >= n) { atomicAdd(&flags[FL_BAD_INDEX], 1); return; }
  field[(int64_t)pos[c] * w + q] = in[t];
}

// Whole particle rows (AoS, `ncol` doubles per particle: the eleven state fields in kFieldName order,
// then the k-NN trial radius) <-> the session's sorted struct of arrays, in one pass.  The slab
// decomposition moves particles between ranks as such rows (domain.py).
struct RowArgs { double* f[F_NFIELDS]; double* knnh; const double* tag; double* gid; };   // tag: optional caller-order column 18;
                                                                                         // gid: session-order global ids <-> column 18
#define ROW_STATE_COLS 18

// One warp per 32 sorted particles: the 18 values of each particle are read field by field (coalesced)
// into shared memory, then every row goes out as one contiguous 144-byte piece (lanes 0..17).
__global__ void __launch_bounds__(256)
rows_out_kernel(int n, int64_t n_rows, int ncol, const uint32_t* __restrict__ id, RowArgs a, double* __restrict__ dst) {
  __shared__ double tile[8][32][ROW_STATE_COLS + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int base = (blockIdx.x * 8 + warp) * 32;
  if (base >= n) return;
  const int i = base + lane;
  int64_t c = -1;
  if (i < n) {
    c = id[i];
    int col = 0;
#pragma unroll
    for (int f = 0; f < 3; ++f)
      for (int q = 0; q < 3; ++q) tile[warp][lane][col++] = a.f[f][3 * (int64_t)i + q];
#pragma unroll
    for (int f = 3; f < F_NFIELDS; ++f) tile[warp][lane][col++] = a.f[f][i];
    tile[warp][lane][col] = a.knnh[i];
    if (a.gid) tile[warp][lane][ROW_STATE_COLS] = a.gid[i];
  }
  __syncwarp();
  for (int r = 0; r < 32; ++r) {
    const int64_t cr = __shfl_sync(0xffffffffu, c, r);
    if (cr >= 0 && cr < n_rows) {
      if (lane < ROW_STATE_COLS) dst[cr * ncol + lane] = tile[warp][r][lane];
      else if (lane == ROW_STATE_COLS && a.gid && ncol > ROW_STATE_COLS) dst[cr * ncol + ROW_STATE_COLS] = tile[warp][r][ROW_STATE_COLS];
      else if (lane == ROW_STATE_COLS && a.tag) dst[cr * ncol + ROW_STATE_COLS] = a.tag[cr];
    }
  }
}

__global__ void __launch_bounds__(256)
rows_in_kernel(int n, int ncol, const uint32_t* __restrict__ id, const double* __restrict__ src, RowArgs a,
               int64_t first = 0, int64_t n_rows = -1) {
  __shared__ double tile[8][32][ROW_STATE_COLS + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int base = (blockIdx.x * 8 + warp) * 32;
  if (base >= n) return;
  const int i = base + lane;
  int64_t c = (i < n) ? (id ? (int64_t)id[i] : (int64_t)i) : -1;             // id == nullptr: row r -> slot r
  if (n_rows >= 0) c = (c >= first && c < first + n_rows) ? c - first : -1;  // only the caller indices [first, first + n_rows)
  for (int r = 0; r < 32; ++r) {
    const int64_t cr = __shfl_sync(0xffffffffu, c, r);
    if (cr >= 0 && lane < ROW_STATE_COLS + ((a.gid && ncol > ROW_STATE_COLS) ? 1 : 0)) tile[warp][r][lane] = src[cr * ncol + lane];
  }
  __syncwarp();
  if (i < n && c >= 0) {
    int col = 0;
#pragma unroll
    for (int f = 0; f < 3; ++f)
      for (int q = 0; q < 3; ++q) a.f[f][3 * (int64_t)i + q] = tile[warp][lane][col++];
#pragma unroll
    for (int f = 3; f < F_NFIELDS; ++f) a.f[f][i] = tile[warp][lane][col++];
    a.knnh[i] = tile[warp][lane][col];
    if (a.gid && ncol > ROW_STATE_COLS) a.gid[i] = tile[warp][lane][ROW_STATE_COLS];
  }
}

// ---- synthetic motion (bench.py / tests): x += amplitude * u(id, seed), u uniform in [-1,1)^3 ----
// The reference's own step cannot move particles by a sizeable fraction of h and stay finite (DESIGN.md
// "Stability of the reference step"), so the measured workloads displace every owned particle between
// steps instead.  The displacement is a pure function of (particle id, seed): identical for any domain
// decomposition and reproducible on the host (ics.perturbation).  Separately rounded product and sum, so
// the host mirror is bit-identical.
__host__ __device__ __forceinline__ double perturb_unit(unsigned long long g, int c, unsigned long long seed) {
  unsigned long long z = (g * 3ull + (unsigned long long)c + 1ull) * 0x9E3779B97F4A7C15ull + seed * 0xBF58476D1CE4E5B9ull;
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
  z ^= z >> 27; z *= 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (double)(z >> 11) * (1.0 / 4503599627370496.0) - 1.0;      // (z >> 11) 2^-52 - 1, exact
}
__global__ void __launch_bounds__(256)
perturb_kernel(int n, const uint32_t* __restrict__ id, const double* __restrict__ gid, int n_owned, double amp,
               unsigned long long seed, double* __restrict__ pos) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (n_owned >= 0 && id[i] >= (uint32_t)n_owned) return;              // ghosts follow their owners
  const unsigned long long g = gid ? (unsigned long long)gid[i] : (unsigned long long)id[i];   // slabs: the GLOBAL id
#pragma unroll
  for (int c = 0; c < 3; ++c)
    pos[3 * (int64_t)i + c] = __dadd_rn(pos[3 * (int64_t)i + c], __dmul_rn(amp, perturb_unit(g, c, seed)));
}

#define SB_GRAV 36           // F2: gravitational acceleration (N,3), sorted order
#define SB_XLIST 45          // int32 rows of every particle expanded from the uint16 tile rows (download, drag)

// neighbour rows of the last search, in the form the passes take them (kernels.cuh RowsRef)
static RowsRef rows_of(sphb200_ctx* ctx) {
  RowsRef r;
  if (ctx->list_is_global) { r.list16 = nullptr; r.list = bufp<int32_t>(ctx->list); r.targets = nullptr; r.ntargets = nullptr; }
  else {
    r.list16 = bufp<uint16_t>(ctx->list16); r.list = bufp<int32_t>(ctx->list);
    r.targets = bufp<int>(ctx->worklist); r.ntargets = bufp<int>(ctx->flags) + FL_KNN_NWORK;
  }
  return r;
}
// (N,k) int32 rows of every particle in the sorted index space (unordered, missing = n): off the hot path
static int full_rows(sphb200_ctx* ctx, cudaStream_t st, const int32_t** out);


// ------------------------------------------------------------------------------------------------
// Slab decomposition on the device (domain.py drives the collectives; nothing N-sized leaves the GPU).
// Rows travel as NCOL = 19 doubles: the 17 state columns, the k-NN trial radius, the global id.
// ------------------------------------------------------------------------------------------------
#define MIG_MAX_WORLD 64
struct MigScratch { int cnt[MIG_MAX_WORLD]; int base[MIG_MAX_WORLD + 1]; int fill[MIG_MAX_WORLD]; int n_keep; int overflow; };

__device__ __forceinline__ int slab_of(double x, const double* __restrict__ bounds, int world, int self) {
  if (!(x == x)) return self;                              // NaN stays where it is
  int r = 0;
  while (r < world - 1 && x >= bounds[r + 1]) ++r;         // x == face belongs to the upper slab
  return r;
}

__global__ void mig_reset_kernel(MigScratch* m) {
  const int t = threadIdx.x;
  if (t < MIG_MAX_WORLD) { m->cnt[t] = 0; m->fill[t] = 0; }
  if (t == 0) { m->n_keep = 0; m->overflow = 0; }
}

// pass 1: keep flag of every particle (owned and still inside this rank's slab), movers counted per destination
__global__ void __launch_bounds__(256)
mig_classify_kernel(int n, const uint32_t* __restrict__ id, uint32_t n_owned, const double* __restrict__ pos,
                    const double* __restrict__ bounds, int world, int rank, uint32_t* __restrict__ flag,
                    uint32_t* __restrict__ slot, MigScratch* m) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool owned = id[i] < n_owned;
  int keep = 0;
  if (owned) {
    const int dest = slab_of(pos[3 * (int64_t)i], bounds, world, rank);
    if (dest == rank) keep = 1; else atomicAdd(&m->cnt[dest], 1);
  }
  flag[i] = (uint32_t)keep;
  slot[i] = (uint32_t)keep;                                // scanned in place afterwards
}

__global__ void mig_offsets_kernel(MigScratch* m, int world, const uint32_t* __restrict__ slot, const uint32_t* __restrict__ flag_last,
                                   int n, int* __restrict__ counts_out) {
  if (threadIdx.x == 0) {
    int run = 0;
    for (int r = 0; r < world; ++r) { m->base[r] = run; run += m->cnt[r]; if (counts_out) counts_out[r] = m->cnt[r]; }
    m->base[world] = run;
    if (slot) { m->n_keep = (int)slot[n - 1] + (int)flag_last[0]; if (counts_out) counts_out[world] = m->n_keep; }
  }
}

struct MoveArgs {
  const double* src[F_NFIELDS]; double* dst[F_NFIELDS];
  const double* knnh_src; double* knnh_dst;
  const double* gid_src; double* gid_dst;
  uint32_t* id_dst;
};

// pass 2: kept particles move to their compacted slot (new caller index = slot); movers leave as rows
__global__ void __launch_bounds__(256)
mig_move_kernel(int n, const uint32_t* __restrict__ id, uint32_t n_owned, const double* __restrict__ bounds, int world, int rank,
                const uint32_t* __restrict__ keepflag, const uint32_t* __restrict__ slot, MoveArgs a, double* __restrict__ rows_out,
                int cap_rows, int ncol, MigScratch* m) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (keepflag[i]) {
    const int64_t d = slot[i];
#pragma unroll
    for (int f = 0; f < 3; ++f)
#pragma unroll
      for (int c = 0; c < 3; ++c) a.dst[f][3 * d + c] = a.src[f][3 * (int64_t)i + c];
#pragma unroll
    for (int f = 3; f < F_NFIELDS; ++f) a.dst[f][d] = a.src[f][i];
    a.knnh_dst[d] = a.knnh_src[i];
    a.gid_dst[d] = a.gid_src[i];
    a.id_dst[d] = (uint32_t)d;
    return;
  }
  if (!(id[i] < n_owned)) return;                          // ghost of the previous step: dropped
  const int dest = slab_of(a.src[F_POS][3 * (int64_t)i], bounds, world, rank);
  const int q = m->base[dest] + atomicAdd(&m->fill[dest], 1);
  if (q >= cap_rows) { m->overflow = 1; return; }
  double* row = rows_out + (int64_t)q * ncol;
  int col = 0;
#pragma unroll
  for (int f = 0; f < 3; ++f)
#pragma unroll
    for (int c = 0; c < 3; ++c) row[col++] = a.src[f][3 * (int64_t)i + c];
#pragma unroll
  for (int f = 3; f < F_NFIELDS; ++f) row[col++] = a.src[f][i];
  row[col++] = a.knnh_src[i];
  row[col] = a.gid_src[i];
}

// halo: owned particles within w of another rank's slab, counted / packed per destination rank
__global__ void __launch_bounds__(256)
halo_count_kernel(int n, const double* __restrict__ pos, const double* __restrict__ bounds, int world, int rank, double w, MigScratch* m) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = pos[3 * (int64_t)i];
  for (int r = 0; r < world; ++r)
    if (r != rank && x >= bounds[r] - w && x < bounds[r + 1] + w) atomicAdd(&m->cnt[r], 1);
}
struct HaloArgs { const double* f[F_NFIELDS]; const double* knnh; const double* gid; };
__global__ void __launch_bounds__(256)
halo_pack_kernel(int n, const uint32_t* __restrict__ id, const double* __restrict__ bounds, int world, int rank, double w, HaloArgs a,
                 double* __restrict__ rows_out, int64_t* __restrict__ idx_out, int cap_rows, int ncol, MigScratch* m) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = a.f[F_POS][3 * (int64_t)i];
  for (int r = 0; r < world; ++r) {
    if (r == rank || !(x >= bounds[r] - w && x < bounds[r + 1] + w)) continue;
    const int q = m->base[r] + atomicAdd(&m->fill[r], 1);
    if (q >= cap_rows) { m->overflow = 1; continue; }
    double* row = rows_out + (int64_t)q * ncol;
    int col = 0;
#pragma unroll
    for (int f = 0; f < 3; ++f)
#pragma unroll
      for (int c = 0; c < 3; ++c) row[col++] = a.f[f][3 * (int64_t)i + c];
#pragma unroll
    for (int f = 3; f < F_NFIELDS; ++f) row[col++] = a.f[f][i];
    row[col++] = a.knnh[i];
    row[col] = a.gid[i];
    idx_out[q] = (int64_t)id[i];                           // caller index: what sphb200_session_gather takes
  }
}

// ------------------------------------------------------------------------------------------------
static int phase_crossing(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  double* scal = bufp<double>(ctx->scal);
  if (!s.has_list) SPH_FAIL("session: no neighbour list yet (call startup first)");
  return launch_crossing_time(ctx, (int)s.n, s.k, rows_of(ctx), fld(ctx, F_VEL), fld(ctx, F_H),
                              fld(ctx, F_TYPE), s.sc.dt_0, scal + SC_CT,
                              s.n_owned >= 0 ? bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]) : nullptr, (uint32_t)s.n_owned, st);
}

static int phase_dt_clamp(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  choose_dt_kernel<<<1, 1, 0, st>>>(bufp<double>(ctx->scal), s.sc);
  clamp_kernel<<<ceil_div(3 * s.n, 256), 256, 0, st>>>(3 * s.n, 1e11 * s.sc.au, fld(ctx, F_POS), fld(ctx, F_VEL));
  ctx->launches += 2;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// cell build + state reorder                                             (replaces nsc:542)
// `mask` selects the state fields that are carried into the new order (bit f = Field f).  The full step
// carries all of them.  A host-resident step ("sort_keys") carries only what is already on the device and
// still needed there -- positions, velocities, particle type -- and lets the remaining fields be uploaded
// straight into the new order while the search runs (uploads gather through the current `id`).
static int phase_sort(sphb200_ctx* ctx, cudaStream_t st, unsigned mask = ~0u) {
  Session& s = sess(ctx);
  // the previous k-th distances (still in the old order) ride into the sorted records as the next trial radii
  SPH_TRY(build_cells(ctx, fld(ctx, F_POS), s.n, ctx->cell_edge, s.use_guess ? bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]) : nullptr, s.k,
                      false, st));
  ReorderArgs a;
  for (int f = 0; f < F_NFIELDS; ++f) { a.src[f] = fld(ctx, f, 0); a.dst[f] = fld(ctx, f, 1); }
  a.id_src = bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]);
  a.id_dst = bufp<uint32_t>(ctx->sbuf[SB_ID + (s.id_cur ^ 1)]);
  a.knnh_src = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]);
  a.knnh_dst = bufp<double>(ctx->sbuf[SB_KNNH + (s.knnh_cur ^ 1)]);
  a.gid_src = s.has_gid ? bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]) : nullptr;
  a.gid_dst = s.has_gid ? bufp<double>(ctx->sbuf[SB_GID + (s.gid_cur ^ 1)]) : nullptr;
  reorder_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, bufp<uint32_t>(ctx->perm), a, mask);
  if (s.has_gid) s.gid_cur ^= 1;
  ctx->launches += 1;
  for (int f = 0; f < F_NFIELDS; ++f) if (mask >> f & 1u) s.cur[f] ^= 1;
  s.id_cur ^= 1;
  s.knnh_cur ^= 1;
  s.has_list = 0;
  s.pos_valid = 0;
  if (ctx->skip_id) ctx->skip_id = bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]);     // the ghost mask follows the new order
  CUDA_TRY(cudaGetLastError());
  return 0;
}

static int phase_knn(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  SPH_TRY(devbuf_reserve(ctx->list, (size_t)s.n * s.k * sizeof(int32_t)));
  double* knnh = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]);
  // previous k-th distance (in the sorted records, r0.w) as the first trial radius; zeros before start-up
  if (ctx->knn_fast && s.k >= 3 && s.k <= knn_group_max_k()) {
    SPH_TRY(devbuf_reserve(ctx->list16, (size_t)s.n * LIST_STRIDE(s.k) * sizeof(uint16_t) + 64));
    if (ctx->knn_lane)
      SPH_TRY(launch_knn_lane(ctx, (int)s.n, s.k, s.dist, s.use_guess, ctx->knn_margin, bufp<uint16_t>(ctx->list16),
                              bufp<int32_t>(ctx->list), fld(ctx, F_H), bufp<int32_t>(ctx->sbuf[SB_NONTRIV]), st));
    else
      SPH_TRY(launch_knn_group(ctx, (int)s.n, s.k, s.dist, s.use_guess, ctx->knn_margin, bufp<uint16_t>(ctx->list16),
                               bufp<int32_t>(ctx->list), nullptr, fld(ctx, F_H), bufp<int32_t>(ctx->sbuf[SB_NONTRIV]), st));
    ctx->list_is_global = 0;
  } else {
    SPH_TRY(launch_knn(ctx, (int)s.n, s.k, s.dist, s.use_guess, 1.0 + ctx->knn_margin, bufp<int32_t>(ctx->list), nullptr,
                       fld(ctx, F_H), bufp<int32_t>(ctx->sbuf[SB_NONTRIV]), st));
    ctx->list_is_global = 1;
  }
  copy_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>(s.n, fld(ctx, F_H), knnh);
  ctx->launches += 1;
  s.has_list = 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

static int phase_pack(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  PackIn in = {nullptr, fld(ctx, F_VEL), fld(ctx, F_MASS), fld(ctx, F_H), fld(ctx, F_TYPE), fld(ctx, F_T), fld(ctx, F_MU),
               fld(ctx, F_GAMMA)};
  return launch_pack(ctx, (int)s.n, in, st);
}
static int phase_density(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  return launch_density(ctx, (int)s.n, s.k, rows_of(ctx), outp(ctx, O_RHO), outp(ctx, O_DUST), outp(ctx, O_NUM), st);
}
static int phase_pi(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  return launch_pi(ctx, (int)s.n, s.k, rows_of(ctx), st);
}
static int phase_force(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  SPH_TRY(launch_force(ctx, (int)s.n, s.k, rows_of(ctx), fld(ctx, F_MASS), outp(ctx, O_HACC), outp(ctx, O_VACC),
                       outp(ctx, O_HEAT), st));
  if (!s.drag) return 0;
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ONTO], (size_t)s.n * 3 * sizeof(double)));      // code_running.py:455,462-463,473
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_REACT], (size_t)s.n * 3 * sizeof(double)));
  const int32_t* full = nullptr;
  SPH_TRY(full_rows(ctx, st, &full));
  return launch_drag(ctx, (int)s.n, s.k, full, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]),
                     bufp<double>(ctx->sbuf[SB_GM]), bufp<double>(ctx->sbuf[SB_GC]), outp(ctx, O_RHO), outp(ctx, O_DUST),
                     fld(ctx, F_TYPE), bufp<double>(ctx->sbuf[SB_ONTO]), bufp<double>(ctx->sbuf[SB_REACT]),
                     outp(ctx, O_VACC), st);
}

// A11: kick-drift + heating; also |x|^2 of the slow particles for A12
static int phase_kick(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  const int n = (int)s.n;
  SPH_TRY(devbuf_reserve(ctx->tmp0, 64));
  unsigned long long* nslow = bufp<unsigned long long>(ctx->tmp0) + 1;
  CUDA_TRY(cudaMemsetAsync(nslow, 0, sizeof(unsigned long long), st));
  integrate_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, bufp<double>(ctx->scal), ctx->phys.kB, ctx->m_h, fld(ctx, F_POS),
                                                     fld(ctx, F_VEL), fld(ctx, F_ACC), fld(ctx, F_E), fld(ctx, F_T),
                                                     fld(ctx, F_MASS), fld(ctx, F_MU), fld(ctx, F_GAMMA), fld(ctx, F_TYPE),
                                                     outp(ctx, O_HACC), outp(ctx, O_VACC), outp(ctx, O_HEAT),
                                                     outp(ctx, O_PACC), bufp<double>(ctx->sbuf[SB_DISTSQ]), nslow,
                                                     s.n_owned >= 0 ? bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]) : nullptr,
                                                     (int)s.n_owned, s.grown,
                                                     s.gravity ? bufp<double>(ctx->sbuf[SB_GRAV]) : nullptr);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// F2: gravitational acceleration of the current positions (after the search, as code_running.py:448)
static int phase_gravity(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  if (!s.gravity) return 0;
  if (s.n_owned >= 0 && s.n_owned < s.n) SPH_FAIL("session: gravity is not available on a slab with ghosts");
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_GRAV], (size_t)s.n * 3 * sizeof(double)));
  return launch_gravity_direct(ctx, (int)s.n, fld(ctx, F_POS), fld(ctx, F_MASS), s.grav_G, s.grav_eps,
                               bufp<double>(ctx->sbuf[SB_GRAV]), st);
}

// A12: global length scale d from the 90th percentile of |x|^2: radix select, see select_*_kernel.
#define SB_SEL 44            // [SEL_HIST int64 histogram | SelectState]
static long long* sel_hist(sphb200_ctx* ctx) { return bufp<long long>(ctx->sbuf[SB_SEL]); }
static SelectState* sel_state(sphb200_ctx* ctx) { return reinterpret_cast<SelectState*>(sel_hist(ctx) + SEL_HIST + 7); }

static int phase_length_begin(sphb200_ctx* ctx, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_SEL], (SEL_HIST + 8) * sizeof(long long) + sizeof(SelectState)));
  select_reset_kernel<<<1, 1024, 0, st>>>(sel_state(ctx), sel_hist(ctx));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
static int phase_length_hist(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  if (ctx->sbuf[SB_SEL].bytes == 0) SPH_FAIL("session: length_hist before length_begin");
  const unsigned long long* nslow = bufp<unsigned long long>(ctx->tmp0) + 1;
  select_hist_kernel<<<148, 1024, 0, st>>>((int)s.n, bufp<double>(ctx->sbuf[SB_DISTSQ]), sel_state(ctx), nslow, sel_hist(ctx));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
static int phase_length_pick(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  if (ctx->sbuf[SB_SEL].bytes == 0) SPH_FAIL("session: length_pick before length_begin");
  select_pick_kernel<<<1, 1024, 0, st>>>(sel_state(ctx), sel_hist(ctx), bufp<double>(ctx->scal), s.sc.n_int);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
static int phase_length(sphb200_ctx* ctx, cudaStream_t st) {
  SPH_TRY(phase_length_begin(ctx, st));
  for (int p = 0; p < SEL_PASSES; ++p) {
    SPH_TRY(phase_length_hist(ctx, st));
    SPH_TRY(phase_length_pick(ctx, st));
  }
  return 0;
}

// A12: smoothing-length update with the current d
static int phase_smooth(sphb200_ctx* ctx, cudaStream_t st) {
  Session& s = sess(ctx);
  const int n = (int)s.n;
  smoothing_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, bufp<double>(ctx->scal), cbrt(s.sc.raise_factor), outp(ctx, O_RHO),
                                                     fld(ctx, F_RHO0), bufp<int32_t>(ctx->sbuf[SB_NONTRIV]), fld(ctx, F_TYPE),
                                                     fld(ctx, F_H), s.grown);
  s.grown = 0;
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

static int phase_integrate(sphb200_ctx* ctx, cudaStream_t st) {
  SPH_TRY(phase_kick(ctx, st));
  SPH_TRY(phase_length(ctx, st));
  return phase_smooth(ctx, st);
}

// Arrays that are not driver state but that the slab-decomposed driver must read or overwrite for
// ghost particles between the passes: the owner-computed density and viscosity Pi of a ghost
// (nsc:646,651 read the NEIGHBOUR's rho and Pi), the k-NN trial radius, |x|^2 of the slow particles.
static double* aux_field(sphb200_ctx* ctx, const char* name, bool for_upload) {
  Session& s = sess(ctx);
  if (!strcmp(name, "densities")) return outp(ctx, O_RHO);
  if (!strcmp(name, "knn_h")) return bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]);
  if (!strcmp(name, "visc_pi")) return (ctx->pi_.bytes >= (size_t)s.n * sizeof(double)) ? bufp<double>(ctx->pi_) : nullptr;
  if (!for_upload && !strcmp(name, "dist_sq")) return bufp<double>(ctx->sbuf[SB_DISTSQ]);
  if (!strcmp(name, "gid")) return s.has_gid ? bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]) : nullptr;
  return nullptr;
}

// While a session call runs, the launchers skip the ghosts of a slab decomposition as targets; the
// function-level API (api.cu) never sees the pointer.  Evaluated per launch: the sort flips id_cur.
struct SkipGhosts {
  sphb200_ctx* ctx;
  explicit SkipGhosts(sphb200_ctx* c) : ctx(c) { refresh(); }
  void refresh() {
    Session& s = sess(ctx);
    ctx->skip_id = (s.n_owned >= 0 && s.n_owned < s.n) ? bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]) : nullptr;
    ctx->skip_from = (uint32_t)s.n_owned;
  }
  ~SkipGhosts() { ctx->skip_id = nullptr; }
};

static int full_rows(sphb200_ctx* ctx, cudaStream_t st, const int32_t** out) {
  Session& s = sess(ctx);
  if (ctx->list_is_global) { *out = bufp<int32_t>(ctx->list); return 0; }
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_XLIST], (size_t)s.n * s.k * sizeof(int32_t)));
  SPH_TRY(launch_expand_list(ctx, (int)s.n, s.k, bufp<uint16_t>(ctx->list16), bufp<int32_t>(ctx->list),
                             bufp<int32_t>(ctx->sbuf[SB_XLIST]), st));
  *out = bufp<int32_t>(ctx->sbuf[SB_XLIST]);
  return 0;
}

extern "C" {

int sphb200_session_create(sphb200_ctx* ctx, int64_t n, int k, double dist) {
  if (!ctx) SPH_FAIL("null context");
  if (n <= 0 || n >= (int64_t)0x7fffffff) SPH_FAIL("session: bad particle count %lld", (long long)n);
  if (k <= 0) SPH_FAIL("session: bad k=%d", k);
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (ctx->session) SPH_FAIL("session_create: this context already holds a live session of %lld particles; destroy it "
                             "(sphb200_session_destroy) or use another context", (long long)sess(ctx).n);
  ctx->session = new (std::nothrow) Session();
  if (!ctx->session) SPH_FAIL("session_create: out of host memory");
  Session& s = sess(ctx);
  s.n = n; s.k = k; s.dist = dist;
  s.sc.dt_0 = 60. * 60. * 24. * 365. * 250000.;
  s.sc.max_age = 3e7 * 60. * 60. * 24. * 365.;
  s.sc.n_int = 3000.;
  s.sc.raise_factor = 715.;
  s.sc.au = 149597870700.0;
  for (int f = 0; f < F_NFIELDS; ++f) {
    s.cur[f] = 0;
    SPH_TRY(devbuf_reserve(ctx->sbuf[2 * f], (size_t)n * kFieldWidth[f] * sizeof(double)));
    SPH_TRY(devbuf_reserve(ctx->sbuf[2 * f + 1], (size_t)n * kFieldWidth[f] * sizeof(double)));
    CUDA_TRY(cudaMemset(ctx->sbuf[2 * f].p, 0, (size_t)n * kFieldWidth[f] * sizeof(double)));
  }
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ID], n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ID_ALT], n * sizeof(uint32_t)));
  for (int o = 0; o < O_NOUT; ++o) SPH_TRY(devbuf_reserve(ctx->sbuf[SB_OUT + o], (size_t)n * kOutWidth[o] * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_NONTRIV], n * sizeof(int32_t)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_DISTSQ], n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_KNNH], n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_KNNH_ALT], n * sizeof(double)));
  CUDA_TRY(cudaMemset(ctx->sbuf[SB_KNNH].p, 0, n * sizeof(double)));
  iota_kernel<<<ceil_div(n, 256), 256>>>((int)n, bufp<uint32_t>(ctx->sbuf[SB_ID]));
  CUDA_TRY(cudaMemset(ctx->scal.p, 0, SC_COUNT * sizeof(double)));
  CUDA_TRY(cudaDeviceSynchronize());
  return 0;
}

int sphb200_session_destroy(sphb200_ctx* ctx) {
  if (!ctx) SPH_FAIL("null context");
  session_free(ctx);
  return 0;
}

// Same run, different particle set: the slab decomposition rebuilds its local set (owned + ghosts)
// every step.  Keeps k, dist, the run parameters and the device scalars (age, dt, ct, d); the state
// arrays are to be filled by sphb200_session_upload_rows (nothing is cleared, nothing synchronises).
int sphb200_session_resize(sphb200_ctx* ctx, int64_t n, int64_t n_owned, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  if (n <= 0 || n >= (int64_t)0x7fffffff) SPH_FAIL("session_resize: bad particle count %lld", (long long)n);
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Session& s = sess(ctx);
  if (s.k <= 0) SPH_FAIL("session_resize: no session to resize");
  s.n = n; s.n_owned = n_owned;
  s.has_list = 0; s.pos_valid = 0; s.id_cur = 0; s.knnh_cur = 0; s.drag = 0; s.grown = 0;
  for (int f = 0; f < F_NFIELDS; ++f) {
    s.cur[f] = 0;
    SPH_TRY(devbuf_reserve(ctx->sbuf[2 * f], (size_t)n * kFieldWidth[f] * sizeof(double)));
    SPH_TRY(devbuf_reserve(ctx->sbuf[2 * f + 1], (size_t)n * kFieldWidth[f] * sizeof(double)));
  }
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ID], n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ID_ALT], n * sizeof(uint32_t)));
  for (int o = 0; o < O_NOUT; ++o) SPH_TRY(devbuf_reserve(ctx->sbuf[SB_OUT + o], (size_t)n * kOutWidth[o] * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_NONTRIV], n * sizeof(int32_t)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_DISTSQ], n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_KNNH], n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_KNNH_ALT], n * sizeof(double)));
  iota_kernel<<<ceil_div(n, 256), 256, 0, st>>>((int)n, bufp<uint32_t>(ctx->sbuf[SB_ID]));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// device scalars addressable as "scalar:<name>" by upload / download (one double, device to device, no
// host synchronisation: the slab driver all_reduces the crossing time in place)
static int find_scalar(const char* name) {
  if (strncmp(name, "scalar:", 7)) return -1;
  name += 7;
  if (!strcmp(name, "ct")) return SC_CT;
  if (!strcmp(name, "dt")) return SC_DT;
  if (!strcmp(name, "d")) return SC_D;
  if (!strcmp(name, "age")) return SC_AGE;
  if (!strcmp(name, "n_slow")) return SC_NSLOW;
  return -1;
}

static int find_field(const char* name) {
  for (int f = 0; f < F_NFIELDS; ++f) if (!strcmp(name, kFieldName[f])) return f;
  return -1;
}

// (N,k) int64 neighbour rows in the caller's order and index space -> int32 list in the session's order
__global__ void __launch_bounds__(256)
list_in_kernel(int n, int k, const uint32_t* __restrict__ id, const uint32_t* __restrict__ pos,
               const int64_t* __restrict__ nb, int32_t* __restrict__ list, int* __restrict__ flags) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n * k) return;
  const int i = (int)(t / k), c = (int)(t % k);
  const int64_t j = nb[(int64_t)id[i] * k + c];
  if (j < 0 || j > n) { atomicAdd(&flags[FL_BAD_INDEX], 1); list[t] = n; return; }
  list[t] = (j == n) ? n : (int32_t)pos[j];            // n = the reference's "missing neighbour" index
}

// src: device pointer, caller's particle order, `count` doubles
int sphb200_session_upload(sphb200_ctx* ctx, const char* name, const void* src, int64_t count, void* stream) {
  if (!ctx || !name || !src) SPH_FAIL("session_upload: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Session& s = sess(ctx);
  const int sc = find_scalar(name);
  if (sc >= 0) {
    if (count != 1) SPH_FAIL("session_upload: '%s' is one value", name);
    CUDA_TRY(cudaMemcpyAsync(bufp<double>(ctx->scal) + sc, src, sizeof(double), cudaMemcpyDeviceToDevice, st));
    return 0;
  }
  if (!strcmp(name, "neighbor")) {      // a saved neighbour array (checkpoint restart): (N,k) int64, caller's indices
    if (count != s.n * s.k) SPH_FAIL("session_upload: neighbor needs %lld values", (long long)(s.n * s.k));
    SPH_TRY(devbuf_reserve(ctx->list, (size_t)s.n * s.k * sizeof(int32_t)));
    SPH_TRY(devbuf_reserve(ctx->sbuf[SB_POS], s.n * sizeof(uint32_t)));
    const uint32_t* idp = bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]);
    uint32_t* pos = bufp<uint32_t>(ctx->sbuf[SB_POS]);
    if (!s.pos_valid) {
      invert_id_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, idp, pos);
      s.pos_valid = 1;
      ctx->launches += 1;
    }
    list_in_kernel<<<ceil_div(s.n * s.k, 256), 256, 0, st>>>((int)s.n, s.k, idp, pos, (const int64_t*)src,
                                                            bufp<int32_t>(ctx->list), bufp<int>(ctx->flags));
    ctx->launches += 1;
    CUDA_TRY(cudaGetLastError());
    s.has_list = 1;
    ctx->list_is_global = 1;                          // restored rows are plain int32 rows: the passes take the LIST form
    return 0;
  }
  const int f = find_field(name);
  double* dstp = (f >= 0) ? fld(ctx, f) : aux_field(ctx, name, true);
  const int w = (f >= 0) ? kFieldWidth[f] : 1;
  if (!dstp) SPH_FAIL("session_upload: unknown field '%s'", name);
  if (count != s.n * w) SPH_FAIL("session_upload: '%s' needs %lld values, got %lld", name, (long long)(s.n * w), (long long)count);
  const uint32_t* id = bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]);
  sort_in_kernel<<<ceil_div(count, 256), 256, 0, st>>>((int)s.n, w, id, (const double*)src, dstp);
  ctx->launches += 1;
  if (!strcmp(name, "visc_pi")) SPH_TRY(launch_repack(ctx, (int)s.n, 1, nullptr, st));   // force-pass coefficients follow Pi
  if (!strcmp(name, "densities")) SPH_TRY(launch_repack(ctx, (int)s.n, 0, outp(ctx, O_RHO), st));   // the Pi pass reads rho from the records
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// dst: device pointer, caller's particle order
int sphb200_session_download(sphb200_ctx* ctx, const char* name, void* dst, int64_t count, void* stream) {
  if (!ctx || !name || !dst) SPH_FAIL("session_download: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const uint32_t* id = bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]);
  const int n = (int)s.n;
  const int sc = find_scalar(name);
  if (sc >= 0) {
    if (count != 1) SPH_FAIL("session_download: '%s' is one value", name);
    CUDA_TRY(cudaMemcpyAsync(dst, bufp<double>(ctx->scal) + sc, sizeof(double), cudaMemcpyDeviceToDevice, st));
    return 0;
  }
  if (!strcmp(name, "neighbor")) {
    if (!s.has_list) SPH_FAIL("session_download: no neighbour list");
    if (count != s.n * s.k) SPH_FAIL("session_download: neighbor needs %lld values", (long long)(s.n * s.k));
    const int32_t* full = nullptr;
    SPH_TRY(full_rows(ctx, st, &full));
    unsort_list_sorted_kernel<<<ceil_div(n, 128), 128, 0, st>>>(n, s.k, id, full, bufp<double4>(ctx->r0), (int64_t*)dst);
  } else if (!strcmp(name, "scalars")) {               // raw device scalars: [6] ct, [7] dt, [8] d, [9] age, [10] max_dist, [11] n_slow
    if (count < 0 || count > SC_COUNT) SPH_FAIL("session_download: scalars holds %d values", SC_COUNT);
    CUDA_TRY(cudaMemcpyAsync(dst, ctx->scal.p, (size_t)count * sizeof(double), cudaMemcpyDeviceToDevice, st));
    return 0;
  } else if (!strcmp(name, "num_nontrivial")) {
    if (count != s.n) SPH_FAIL("session_download: num_nontrivial needs %lld values", (long long)s.n);
    unsort_i32_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, id, bufp<int32_t>(ctx->sbuf[SB_NONTRIV]), (int64_t*)dst);
  } else {
    const double* src = nullptr;
    int w = 0;
    const int f = find_field(name);
    if (f >= 0) { src = fld(ctx, f); w = kFieldWidth[f]; }
    for (int o = 0; o < O_NOUT && !src; ++o) if (!strcmp(name, kOutName[o])) { src = outp(ctx, o); w = kOutWidth[o]; }
    if (!src && !strcmp(name, "grav_accel") && s.gravity) { src = bufp<double>(ctx->sbuf[SB_GRAV]); w = 3; }
    if (!src) { src = aux_field(ctx, name, false); w = 1; }
    if (!src) SPH_FAIL("session_download: unknown field '%s'", name);
    if (count != s.n * w) SPH_FAIL("session_download: '%s' needs %lld values, got %lld", name, (long long)(s.n * w), (long long)count);
    unsort_kernel<<<ceil_div(count, 256), 256, 0, st>>>(n, w, id, src, (double*)dst);
  }
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

static int find_field(const char* name);

static int field_at(sphb200_ctx* ctx, const char* name, const int64_t* idx, int64_t count, double* buf, bool scatter,
                    cudaStream_t st) {
  Session& s = sess(ctx);
  const int f = find_field(name);
  double* p = (f >= 0) ? fld(ctx, f) : aux_field(ctx, name, scatter);
  const int w = (f >= 0) ? kFieldWidth[f] : 1;
  if (!p) SPH_FAIL("session gather/scatter: unknown field '%s'", name);
  if (count <= 0) return 0;
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_POS], s.n * sizeof(uint32_t)));
  uint32_t* pos = bufp<uint32_t>(ctx->sbuf[SB_POS]);
  if (!s.pos_valid) {
    invert_id_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]), pos);
    s.pos_valid = 1;
    ctx->launches += 1;
  }
  if (scatter) scatter_at_kernel<<<ceil_div(count * w, 256), 256, 0, st>>>(count, w, (int)s.n, idx, pos, buf, p, bufp<int>(ctx->flags));
  else gather_at_kernel<<<ceil_div(count * w, 256), 256, 0, st>>>(count, w, (int)s.n, idx, pos, p, buf, bufp<int>(ctx->flags));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  if (scatter && !strcmp(name, "visc_pi")) SPH_TRY(launch_repack(ctx, (int)s.n, 1, nullptr, st));
  if (scatter && !strcmp(name, "densities")) SPH_TRY(launch_repack(ctx, (int)s.n, 0, outp(ctx, O_RHO), st));
  return 0;
}

// out (count, w) <- field at caller indices idx (device int64)
int sphb200_session_gather(sphb200_ctx* ctx, const char* field, const int64_t* idx, int64_t count, double* out, void* stream) {
  if (!ctx || !field || (count > 0 && (!idx || !out))) SPH_FAIL("session_gather: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  return field_at(ctx, field, idx, count, out, false, reinterpret_cast<cudaStream_t>(stream));
}
// field at caller indices idx <- in (count, w)
int sphb200_session_scatter(sphb200_ctx* ctx, const char* field, const int64_t* idx, int64_t count, const double* in, void* stream) {
  if (!ctx || !field || (count > 0 && (!idx || !in))) SPH_FAIL("session_scatter: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  return field_at(ctx, field, idx, count, const_cast<double*>(in), true, reinterpret_cast<cudaStream_t>(stream));
}

// rows: device pointer, (n_rows, ncol) doubles, caller order; ncol >= 18 (extra columns are the caller's)
static int download_rows_impl(sphb200_ctx* ctx, double* rows, int64_t n_rows, int ncol, const double* tag, void* stream);
int sphb200_session_download_rows(sphb200_ctx* ctx, double* rows, int64_t n_rows, int ncol, void* stream) {
  return download_rows_impl(ctx, rows, n_rows, ncol, nullptr, stream);
}
// same, and column 18 of every row is filled from `tag` (device, caller order, >= n_rows doubles): the id a
// slab decomposition sends along with a row
int sphb200_session_download_rows_tagged(sphb200_ctx* ctx, double* rows, int64_t n_rows, int ncol, const double* tag,
                                         void* stream) {
  if (!tag || ncol < ROW_STATE_COLS + 1) SPH_FAIL("session_download_rows_tagged: needs a tag array and ncol >= 19");
  return download_rows_impl(ctx, rows, n_rows, ncol, tag, stream);
}
static int download_rows_impl(sphb200_ctx* ctx, double* rows, int64_t n_rows, int ncol, const double* tag, void* stream) {
  if (!ctx || !rows) SPH_FAIL("session_download_rows: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (ncol < ROW_STATE_COLS || n_rows < 0 || n_rows > s.n) SPH_FAIL("session_download_rows: bad shape (%lld, %d)", (long long)n_rows, ncol);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  RowArgs a;
  for (int f = 0; f < F_NFIELDS; ++f) a.f[f] = fld(ctx, f);
  a.knnh = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]);
  a.tag = tag;
  a.gid = s.has_gid ? bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]) : nullptr;
  rows_out_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, n_rows, ncol, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]), a, rows);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// rows for the caller indices [first, first + n_rows)
int sphb200_session_upload_rows_at(sphb200_ctx* ctx, const double* rows, int64_t first, int64_t n_rows, int ncol, void* stream) {
  if (!ctx || (!rows && n_rows > 0)) SPH_FAIL("session_upload_rows: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (ncol < ROW_STATE_COLS || first < 0 || n_rows < 0 || first + n_rows > s.n)
    SPH_FAIL("session_upload_rows: rows [%lld, %lld) x %d do not fit %lld particles x >= 18 columns", (long long)first,
             (long long)(first + n_rows), ncol, (long long)s.n);
  if (n_rows == 0) return 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  RowArgs a;
  for (int f = 0; f < F_NFIELDS; ++f) a.f[f] = fld(ctx, f);
  a.knnh = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]);
  a.tag = nullptr;
  a.gid = s.has_gid ? bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]) : nullptr;
  rows_in_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, ncol, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]), rows, a, first, n_rows);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int sphb200_session_upload_rows(sphb200_ctx* ctx, const double* rows, int64_t n_rows, int ncol, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  if (n_rows != sess(ctx).n) SPH_FAIL("session_upload_rows: needs all %lld rows, got %lld", (long long)sess(ctx).n, (long long)n_rows);
  return sphb200_session_upload_rows_at(ctx, rows, 0, n_rows, ncol, stream);
}

// grow a buffer to `bytes`, keeping its first `keep` bytes (devbuf_reserve alone discards the content)
static int devbuf_grow(DevBuf& b, size_t keep, size_t bytes, cudaStream_t st) {
  if (bytes <= b.bytes) return 0;
  DevBuf nb;
  SPH_TRY(devbuf_reserve(nb, bytes + bytes / 4));          // head-room: injections come in bursts
  if (b.p && keep) CUDA_TRY(cudaMemcpyAsync(nb.p, b.p, keep, cudaMemcpyDeviceToDevice, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  devbuf_free(b);
  b = nb;
  return 0;
}

// F4: variable-N state (the driver grows every array when a supernova injects particles,
// code_running.py:361-381).  The new particles get caller indices N .. N+n_new-1; existing particles keep
// theirs.  rows: (n_new, ncol >= 18) device doubles in the row layout of sphb200_session_upload_rows.
// The neighbour list is invalid afterwards: run the phases "sort" and "knn" (the driver's own
// nsc.neighbors call at :386) before the next step.  The first step after an append takes
// dv = a dt and new_smoothing = 1 like the reference does when its old arrays are shorter (:484-485, :534-535).
int sphb200_session_append_rows(sphb200_ctx* ctx, const double* rows, int64_t n_new, int ncol, void* stream) {
  if (!ctx || (!rows && n_new > 0)) SPH_FAIL("session_append_rows: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (s.k <= 0) SPH_FAIL("session_append_rows: no session");
  if (n_new < 0 || ncol < ROW_STATE_COLS) SPH_FAIL("session_append_rows: needs (n_new >= 0, ncol >= 18), got (%lld, %d)", (long long)n_new, ncol);
  if (s.n_owned >= 0) SPH_FAIL("session_append_rows: not available on a slab with ghosts (re-seat the slab instead)");
  if (n_new == 0) return 0;
  const int64_t n0 = s.n, n1 = s.n + n_new;
  if (n1 >= (int64_t)0x7fffffff) SPH_FAIL("session_append_rows: too many particles");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  for (int f = 0; f < F_NFIELDS; ++f) {
    const size_t w = kFieldWidth[f] * sizeof(double);
    SPH_TRY(devbuf_grow(ctx->sbuf[2 * f + s.cur[f]], (size_t)n0 * w, (size_t)n1 * w, st));
    SPH_TRY(devbuf_reserve(ctx->sbuf[2 * f + (s.cur[f] ^ 1)], (size_t)n1 * w + (size_t)n1 * w / 4));
  }
  SPH_TRY(devbuf_grow(ctx->sbuf[SB_ID + s.id_cur], n0 * sizeof(uint32_t), n1 * sizeof(uint32_t), st));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ID + (s.id_cur ^ 1)], n1 * sizeof(uint32_t) + n1));
  SPH_TRY(devbuf_grow(ctx->sbuf[SB_KNNH + s.knnh_cur], n0 * sizeof(double), n1 * sizeof(double), st));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_KNNH + (s.knnh_cur ^ 1)], n1 * sizeof(double) + 2 * n1));
  for (int o = 0; o < O_NOUT; ++o) SPH_TRY(devbuf_reserve(ctx->sbuf[SB_OUT + o], (size_t)n1 * kOutWidth[o] * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_NONTRIV], n1 * sizeof(int32_t)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_DISTSQ], n1 * sizeof(double)));
  RowArgs a;
  for (int f = 0; f < F_NFIELDS; ++f) a.f[f] = fld(ctx, f) + (size_t)n0 * kFieldWidth[f];
  a.knnh = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]) + n0;
  a.tag = nullptr;
  a.gid = nullptr;
  rows_in_kernel<<<ceil_div(n_new, 256), 256, 0, st>>>((int)n_new, ncol, nullptr, rows, a);
  iota_kernel<<<ceil_div(n_new, 256), 256, 0, st>>>((int)n_new, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]) + n0, (uint32_t)n0);
  ctx->launches += 2;
  CUDA_TRY(cudaGetLastError());
  s.n = n1;
  s.has_list = 0; s.pos_valid = 0; s.drag = 0; s.grown = 1;
  return 0;
}

int sphb200_session_set_scalar(sphb200_ctx* ctx, const char* name, double value) {
  if (!ctx || !name) SPH_FAIL("null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  double* scal = bufp<double>(ctx->scal);
  if (!strcmp(name, "dt_0")) s.sc.dt_0 = value;
  else if (!strcmp(name, "MAX_AGE")) s.sc.max_age = value;
  else if (!strcmp(name, "N_INT_PER_PARTICLE")) s.sc.n_int = value;
  else if (!strcmp(name, "raise_factor")) s.sc.raise_factor = value;
  else if (!strcmp(name, "AU")) s.sc.au = value;
  else if (!strcmp(name, "dist")) s.dist = value;
  else if (!strcmp(name, "use_guess")) s.use_guess = (int)value;
  else if (!strcmp(name, "n_owned")) s.n_owned = (int64_t)value;
  else if (!strcmp(name, "age")) CUDA_TRY(cudaMemcpy(scal + SC_AGE, &value, sizeof(double), cudaMemcpyHostToDevice));
  else if (!strcmp(name, "d")) CUDA_TRY(cudaMemcpy(scal + SC_D, &value, sizeof(double), cudaMemcpyHostToDevice));
  else if (!strcmp(name, "dt")) CUDA_TRY(cudaMemcpy(scal + SC_DT, &value, sizeof(double), cudaMemcpyHostToDevice));
  else if (!strcmp(name, "ct")) CUDA_TRY(cudaMemcpy(scal + SC_CT, &value, sizeof(double), cudaMemcpyHostToDevice));
  else SPH_FAIL("session_set_scalar: unknown scalar '%s'", name);
  return 0;
}

int sphb200_session_scalar(sphb200_ctx* ctx, const char* name, double* out, void* stream) {
  if (!ctx || !name || !out) SPH_FAIL("null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int which = -1;
  if (!strcmp(name, "ct")) which = SC_CT;
  else if (!strcmp(name, "dt")) which = SC_DT;
  else if (!strcmp(name, "d")) which = SC_D;
  else if (!strcmp(name, "age")) which = SC_AGE;
  else if (!strcmp(name, "max_dist")) which = SC_MAXDIST;
  else if (!strcmp(name, "n_slow")) which = SC_NSLOW;
  else SPH_FAIL("session_scalar: unknown scalar '%s'", name);
  CUDA_TRY(cudaMemcpyAsync(out, bufp<double>(ctx->scal) + which, sizeof(double), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return 0;
}

// code_running.py:171-179: first neighbour query + hydro_update; densities_0 <- densities (:183)
int sphb200_session_startup(sphb200_ctx* ctx, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Session& s = sess(ctx);
  SPH_TRY(phase_sort(ctx, st));
  SPH_TRY(phase_knn(ctx, st));
  SPH_TRY(phase_pack(ctx, st));
  SPH_TRY(phase_density(ctx, st));
  SPH_TRY(phase_pi(ctx, st));
  SPH_TRY(phase_force(ctx, st));
  copy_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>(s.n, outp(ctx, O_RHO), fld(ctx, F_RHO0));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// F1: switch the dust-gas drag of code_running.py:455,462-463,473 on (f_un != NULL) or off.
int sphb200_session_set_drag(sphb200_ctx* ctx, const void* f_un, int n_species, const double* grain_mass_host,
                             const double* sigma_effective_host, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (s.k <= 0) SPH_FAIL("session_set_drag: no session");
  if (!f_un) { s.drag = 0; return 0; }
  if (!grain_mass_host || !sigma_effective_host) SPH_FAIL("session_set_drag: null table");
  if (n_species <= 0 || n_species > SPHB200_MAX_SPECIES)
    SPH_FAIL("session_set_drag: n_species=%d outside 1..%d", n_species, SPHB200_MAX_SPECIES);
  if (s.n_owned >= 0) SPH_FAIL("session_set_drag: not available on a slab with ghosts (the reaction on a ghost is not returned to its owner)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_GM], (size_t)s.n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_GC], (size_t)s.n * sizeof(double)));
  SPH_TRY(launch_grain_means(ctx, (int)s.n, n_species, static_cast<const double*>(f_un), grain_mass_host,
                             sigma_effective_host, bufp<double>(ctx->sbuf[SB_GM]), bufp<double>(ctx->sbuf[SB_GC]), st));
  s.drag = 1;
  return 0;
}

// Slab decomposition, after the search: does every owned particle's ball (x, h) stay inside [lo, hi) along x
// (the region this rank's owned + ghost particles cover), and what is the largest h?  Both land in device
// scalars (buffer "cover": [bad, hmax]) so the ranks can all_reduce(MAX) them in place.
__global__ void __launch_bounds__(256)
cover_check_kernel(int n, const double* __restrict__ pos, const double* __restrict__ h, const uint32_t* __restrict__ id,
                   int n_owned, double lo, double hi, double* __restrict__ scal) {
  __shared__ unsigned long long s_bad, s_h;
  if (threadIdx.x == 0) { s_bad = 0ull; s_h = 0ull; }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && id[i] < (uint32_t)n_owned) {
    const double x = pos[3 * (int64_t)i], hh = h[i];
    const bool ok = ((x - hh) >= lo) && ((x + hh) < hi);
    if (!ok) atomicMax(&s_bad, (unsigned long long)__double_as_longlong(1.0));
    if (hh > 0.0) atomicMax(&s_h, (unsigned long long)__double_as_longlong(hh));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_bad) atomicMax(reinterpret_cast<unsigned long long*>(scal + SC_COVER_BAD), s_bad);
    if (s_h) atomicMax(reinterpret_cast<unsigned long long*>(scal + SC_COVER_HMAX), s_h);
  }
}

int sphb200_session_cover_check(sphb200_ctx* ctx, double lo, double hi, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (s.k <= 0) SPH_FAIL("session_cover_check: no session");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  double* scal = bufp<double>(ctx->scal);
  CUDA_TRY(cudaMemsetAsync(scal + SC_COVER_BAD, 0, 2 * sizeof(double), st));
  const int n_owned = s.n_owned >= 0 ? (int)s.n_owned : (int)s.n;
  cover_check_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, fld(ctx, F_POS), fld(ctx, F_H),
                                                         bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]), n_owned, lo, hi, scal);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// Device pointer of an internal array a caller-side collective works on in place:
// "length_hist" = the radix-select histogram (int64, 2*8192 + 1 values) between the phases "length_hist"
// and "length_pick" (a slab decomposition all_reduces it there).
int sphb200_session_buffer(sphb200_ctx* ctx, const char* name, void** ptr, int64_t* count) {
  if (!ctx || !name || !ptr || !count) SPH_FAIL("session_buffer: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  if (!strcmp(name, "length_hist")) {
    SPH_TRY(devbuf_reserve(ctx->sbuf[SB_SEL], (SEL_HIST + 8) * sizeof(long long) + sizeof(SelectState)));
    *ptr = sel_hist(ctx);
    *count = SEL_HIST;
    return 0;
  }
  if (!strcmp(name, "cover")) {               // float64 [bad, hmax] written by sphb200_session_cover_check
    *ptr = bufp<double>(ctx->scal) + SC_COVER_BAD;
    *count = 2;
    return 0;
  }
  SPH_FAIL("session_buffer: unknown buffer '%s'", name);
}


// ---- slab decomposition: device side ------------------------------------------------------------
static MigScratch* mig_scratch(sphb200_ctx* ctx) { return reinterpret_cast<MigScratch*>(bufp<char>(ctx->tmp1)); }

// the session carries global particle ids (column 18 of every row) from now on
int sphb200_session_enable_gid(sphb200_ctx* ctx) {
  if (!ctx) SPH_FAIL("null context");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_GID], (size_t)s.n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_GID_ALT], (size_t)s.n * sizeof(double)));
  CUDA_TRY(cudaMemset(ctx->sbuf[SB_GID].p, 0, (size_t)s.n * sizeof(double)));
  s.has_gid = 1; s.gid_cur = 0;
  return 0;
}

// Drop the ghosts, send away the owned particles that left the slab [bounds[rank], bounds[rank+1]) and compact
// the rest (new caller index = position in the compacted order).  rows_out (device, cap_rows x ncol) receives
// the movers grouped by destination rank; counts_out (device int32[world + 1]) the movers per rank and, last, the
// number of particles kept.  The caller reads counts_out, then calls sphb200_session_set_count(kept, kept).
int sphb200_session_migrate_pack(sphb200_ctx* ctx, int phase, const double* bounds, int world, int rank, double* rows_out,
                                 int64_t cap_rows, int ncol, int* counts_out, void* stream) {
  if (!ctx || !bounds || (phase == 0 && !counts_out) || (phase == 1 && !rows_out && cap_rows > 0))
    SPH_FAIL("session_migrate_pack: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (!s.has_gid) SPH_FAIL("session_migrate_pack: call sphb200_session_enable_gid first");
  if (world < 1 || world > MIG_MAX_WORLD || rank < 0 || rank >= world || ncol < ROW_STATE_COLS + 1)
    SPH_FAIL("session_migrate_pack: bad arguments (world %d, rank %d, ncol %d)", world, rank, ncol);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int n = (int)s.n;
  SPH_TRY(devbuf_reserve(ctx->tmp1, sizeof(MigScratch)));
  SPH_TRY(devbuf_reserve(ctx->tmp2, (size_t)n * sizeof(uint32_t)));
  SPH_TRY(devbuf_reserve(ctx->tmp3, (size_t)n * sizeof(uint32_t)));
  MigScratch* m = mig_scratch(ctx);
  uint32_t* flag = bufp<uint32_t>(ctx->tmp2);
  uint32_t* slot = bufp<uint32_t>(ctx->tmp3);
  const uint32_t* id = bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]);
  const uint32_t n_owned = s.n_owned >= 0 ? (uint32_t)s.n_owned : (uint32_t)n;
  if (phase == 0) {                                        // count: the caller sizes rows_out from counts_out
    mig_reset_kernel<<<1, MIG_MAX_WORLD, 0, st>>>(m);
    mig_classify_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, id, n_owned, fld(ctx, F_POS), bounds, world, rank, flag, slot, m);
    SPH_TRY(scan_exclusive_u32(ctx, slot, n, st));
    mig_offsets_kernel<<<1, 32, 0, st>>>(m, world, slot, flag + (n - 1), n, counts_out);
    ctx->launches += 3;
    CUDA_TRY(cudaGetLastError());
    return 0;
  }
  MoveArgs a;
  for (int f = 0; f < F_NFIELDS; ++f) { a.src[f] = fld(ctx, f, 0); a.dst[f] = fld(ctx, f, 1); }
  a.knnh_src = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]); a.knnh_dst = bufp<double>(ctx->sbuf[SB_KNNH + (s.knnh_cur ^ 1)]);
  a.gid_src = bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]); a.gid_dst = bufp<double>(ctx->sbuf[SB_GID + (s.gid_cur ^ 1)]);
  a.id_dst = bufp<uint32_t>(ctx->sbuf[SB_ID + (s.id_cur ^ 1)]);
  mig_move_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, id, n_owned, bounds, world, rank, flag, slot, a, rows_out, (int)cap_rows, ncol, m);
  ctx->launches += 1;
  for (int f = 0; f < F_NFIELDS; ++f) s.cur[f] ^= 1;
  s.id_cur ^= 1; s.knnh_cur ^= 1; s.gid_cur ^= 1;
  s.has_list = 0; s.pos_valid = 0;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// particle count after a migration / before ghosts arrive (host-side bookkeeping only)
int sphb200_session_set_count(sphb200_ctx* ctx, int64_t n, int64_t n_owned) {
  if (!ctx) SPH_FAIL("null context");
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (n < 0 || n_owned > n) SPH_FAIL("session_set_count: bad counts (%lld, %lld)", (long long)n, (long long)n_owned);
  s.n = n; s.n_owned = n_owned;
  s.has_list = 0; s.pos_valid = 0;
  return 0;
}

// Append rows that arrived from other ranks (same layout, column 18 = global id) behind the current particles,
// as owned particles (movers: n_owned grows too; requires that no ghost is present) or as ghosts.
int sphb200_session_receive_rows(sphb200_ctx* ctx, const double* rows, int64_t n_new, int ncol, int as_ghosts, void* stream) {
  if (!ctx || (!rows && n_new > 0)) SPH_FAIL("session_receive_rows: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (!s.has_gid) SPH_FAIL("session_receive_rows: call sphb200_session_enable_gid first");
  if (n_new < 0 || ncol < ROW_STATE_COLS + 1) SPH_FAIL("session_receive_rows: needs (n_new >= 0, ncol >= 19)");
  const int64_t n_owned = s.n_owned >= 0 ? s.n_owned : s.n;
  if (!as_ghosts && n_owned != s.n) SPH_FAIL("session_receive_rows: owned rows cannot be appended behind ghosts");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t n0 = s.n, n1 = s.n + n_new;
  if (n1 >= (int64_t)0x7fffffff) SPH_FAIL("session_receive_rows: too many particles");
  for (int f = 0; f < F_NFIELDS; ++f) {
    const size_t w = kFieldWidth[f] * sizeof(double);
    SPH_TRY(devbuf_grow(ctx->sbuf[2 * f + s.cur[f]], (size_t)n0 * w, (size_t)n1 * w, st));
    SPH_TRY(devbuf_reserve(ctx->sbuf[2 * f + (s.cur[f] ^ 1)], (size_t)n1 * w + (size_t)n1 * w / 4));
  }
  SPH_TRY(devbuf_grow(ctx->sbuf[SB_ID + s.id_cur], n0 * sizeof(uint32_t), n1 * sizeof(uint32_t), st));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_ID + (s.id_cur ^ 1)], n1 * sizeof(uint32_t) + n1));
  SPH_TRY(devbuf_grow(ctx->sbuf[SB_KNNH + s.knnh_cur], n0 * sizeof(double), n1 * sizeof(double), st));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_KNNH + (s.knnh_cur ^ 1)], n1 * sizeof(double) + 2 * n1));
  SPH_TRY(devbuf_grow(ctx->sbuf[SB_GID + s.gid_cur], n0 * sizeof(double), n1 * sizeof(double), st));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_GID + (s.gid_cur ^ 1)], n1 * sizeof(double) + 2 * n1));
  for (int o = 0; o < O_NOUT; ++o) SPH_TRY(devbuf_reserve(ctx->sbuf[SB_OUT + o], (size_t)n1 * kOutWidth[o] * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_NONTRIV], n1 * sizeof(int32_t)));
  SPH_TRY(devbuf_reserve(ctx->sbuf[SB_DISTSQ], n1 * sizeof(double)));
  if (n_new > 0) {
    RowArgs a;
    for (int f = 0; f < F_NFIELDS; ++f) a.f[f] = fld(ctx, f) + (size_t)n0 * kFieldWidth[f];
    a.knnh = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]) + n0;
    a.gid = bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]) + n0;
    a.tag = nullptr;
    rows_in_kernel<<<ceil_div(n_new, 256), 256, 0, st>>>((int)n_new, ncol, nullptr, rows, a);
    iota_kernel<<<ceil_div(n_new, 256), 256, 0, st>>>((int)n_new, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]) + n0, (uint32_t)n0);
    ctx->launches += 2;
    CUDA_TRY(cudaGetLastError());
  }
  s.n = n1;
  s.n_owned = as_ghosts ? n_owned : n1;
  s.has_list = 0; s.pos_valid = 0;
  return 0;
}

// Select the owned particles within w of another rank's slab and pack them as rows grouped by destination rank
// (rows_out), with their caller indices (idx_out: the list later ghost refreshes gather at).  counts_out (device
// int32[world]) = rows per destination.  Call after the movers were received (every particle owned).
int sphb200_session_halo_pack(sphb200_ctx* ctx, int phase, const double* bounds, int world, int rank, double w, double* rows_out,
                              int64_t* idx_out, int64_t cap_rows, int ncol, int* counts_out, void* stream) {
  if (!ctx || !bounds || (phase == 0 && !counts_out) || (phase == 1 && cap_rows > 0 && (!rows_out || !idx_out)))
    SPH_FAIL("session_halo_pack: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  if (!s.has_gid) SPH_FAIL("session_halo_pack: call sphb200_session_enable_gid first");
  if (world < 1 || world > MIG_MAX_WORLD || rank < 0 || rank >= world || ncol < ROW_STATE_COLS + 1)
    SPH_FAIL("session_halo_pack: bad arguments (world %d, rank %d, ncol %d)", world, rank, ncol);
  const int64_t n_owned = s.n_owned >= 0 ? s.n_owned : s.n;
  if (n_owned != s.n) SPH_FAIL("session_halo_pack: ghosts present (call sphb200_session_migrate_pack first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int n = (int)s.n;
  SPH_TRY(devbuf_reserve(ctx->tmp1, sizeof(MigScratch)));
  MigScratch* m = mig_scratch(ctx);
  if (phase == 0) {
    mig_reset_kernel<<<1, MIG_MAX_WORLD, 0, st>>>(m);
    halo_count_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, fld(ctx, F_POS), bounds, world, rank, w, m);
    mig_offsets_kernel<<<1, 32, 0, st>>>(m, world, nullptr, nullptr, n, counts_out);
    ctx->launches += 3;
    CUDA_TRY(cudaGetLastError());
    return 0;
  }
  HaloArgs a;
  for (int f = 0; f < F_NFIELDS; ++f) a.f[f] = fld(ctx, f);
  a.knnh = bufp<double>(ctx->sbuf[SB_KNNH + s.knnh_cur]);
  a.gid = bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]);
  halo_pack_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]), bounds, world, rank, w, a,
                                                     rows_out, idx_out, (int)cap_rows, ncol, m);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// F2: switch self-gravity on (softening > 0: the Plummer length, the reference uses median(sizes), nsc:359) or
// off (softening <= 0).  The phase "gravity" then fills the acceleration that "integrate" adds (:477).
int sphb200_session_set_gravity(sphb200_ctx* ctx, double G, double softening) {
  if (!ctx) SPH_FAIL("null context");
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  s.gravity = softening > 0.0;
  s.grav_G = G; s.grav_eps = softening;
  return 0;
}

int sphb200_session_perturb(sphb200_ctx* ctx, double amplitude, uint64_t seed, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  Session& s = sess(ctx);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  perturb_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>((int)s.n, bufp<uint32_t>(ctx->sbuf[SB_ID + s.id_cur]),
                                                     s.has_gid ? bufp<double>(ctx->sbuf[SB_GID + s.gid_cur]) : nullptr,
                                                     (int)s.n_owned, amplitude, (unsigned long long)seed, fld(ctx, F_POS));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int sphb200_session_step(sphb200_ctx* ctx, void* stream) {
  if (!ctx) SPH_FAIL("null context");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  SkipGhosts skip(ctx);
  SPH_TRY(phase_crossing(ctx, st));
  SPH_TRY(phase_dt_clamp(ctx, st));
  SPH_TRY(phase_sort(ctx, st));
  SPH_TRY(phase_knn(ctx, st));
  SPH_TRY(phase_pack(ctx, st));
  SPH_TRY(phase_density(ctx, st));
  SPH_TRY(phase_pi(ctx, st));
  SPH_TRY(phase_force(ctx, st));
  SPH_TRY(phase_gravity(ctx, st));
  SPH_TRY(phase_integrate(ctx, st));
  return 0;
}

// run one phase of the step (per-kernel timing in bench.py / profiles)
int sphb200_session_phase(sphb200_ctx* ctx, const char* phase, void* stream) {
  if (!ctx || !phase) SPH_FAIL("null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  if (!ctx->session) SPH_FAIL("no session in this context (call sphb200_session_create first)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  SkipGhosts skip(ctx);
  if (!strcmp(phase, "crossing")) return phase_crossing(ctx, st);
  if (!strcmp(phase, "dt_clamp")) return phase_dt_clamp(ctx, st);
  if (!strcmp(phase, "sort")) return phase_sort(ctx, st);
  if (!strcmp(phase, "sort_keys")) return phase_sort(ctx, st, 1u << F_POS | 1u << F_VEL | 1u << F_TYPE);
  if (!strcmp(phase, "knn")) return phase_knn(ctx, st);
  if (!strcmp(phase, "pack")) return phase_pack(ctx, st);
  if (!strcmp(phase, "density")) return phase_density(ctx, st);
  if (!strcmp(phase, "pi")) return phase_pi(ctx, st);
  if (!strcmp(phase, "force")) return phase_force(ctx, st);
  if (!strcmp(phase, "gravity")) return phase_gravity(ctx, st);
  if (!strcmp(phase, "integrate")) return phase_integrate(ctx, st);
  if (!strcmp(phase, "kick")) return phase_kick(ctx, st);
  if (!strcmp(phase, "length")) return phase_length(ctx, st);
  if (!strcmp(phase, "length_begin")) return phase_length_begin(ctx, st);
  if (!strcmp(phase, "length_hist")) return phase_length_hist(ctx, st);
  if (!strcmp(phase, "length_pick")) return phase_length_pick(ctx, st);
  if (!strcmp(phase, "smooth")) return phase_smooth(ctx, st);
  if (!strcmp(phase, "rho0")) {               // densities_0 <- densities (end of start-up, code_running.py:183)
    Session& s = sess(ctx);
    copy_kernel<<<ceil_div(s.n, 256), 256, 0, st>>>(s.n, outp(ctx, O_RHO), fld(ctx, F_RHO0));
    ctx->launches += 1;
    CUDA_TRY(cudaGetLastError());
    return 0;
  }
  SPH_FAIL("session_phase: unknown phase '%s'", phase);
}

}  // extern "C"
