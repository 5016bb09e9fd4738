// Vectorised-generation hydro passes (nsc:556-671) and the crossing time (nsc:776-786).
//
//   pack      -> per-particle records                                  nsc:559-573, 587-592 per-particle factors
//   density   -> rho, rho_dust, n                                      nsc:588-590, 605-607
//   pi        -> per-particle artificial-viscosity Pi_i                nsc:639-649 (axis repaired)
//   force     -> hydro_accel, visc_accel, visc_heat                    nsc:591-592, 615-619, 651-654
//   crossing  -> min_i h_i / max_j |v_j - v_i|                         nsc:776-786
//   species   -> f_un_neighbor                                         nsc:624-626
//   drag      -> dust-gas drag inside the step                         nsc:719-742
//
// Every pair pass exists in two forms that share one per-pair function:
//   * TILE form (the session's hot path): one CTA per half-group of cells (tile.cuh), persistent.  The records of
//     the three z-layers the half-group needs are staged into shared memory with one bulk asynchronous copy per
//     row (cp.async.bulk -> UBLKCP, mbarrier completion); a neighbour is a uint16 offset into that tile, so every
//     neighbour read is a shared-memory read and every record leaves HBM/L2 once per tile instead of once per pair.
//     8 lanes share one target, accumulate privately in fp64 and combine with shuffles: gather formulation, no
//     atomics.  Record sizes are odd multiples of 16 B (48, 80) so that random 16-byte reads spread over all
//     eight 16-byte bank groups.
//   * LIST form: int32 rows in global memory (the function-level API, whose caller supplies arbitrary rows, and
//     the rows of targets whose k-th neighbour lies outside their tile).
// "self" follows the reference: positions/velocities of column 0 of the row (nsc:580-581), while h_i, rho_i,
// m_i, Pi_i are taken at the row index.
#include "common.cuh"
#include "kernels.cuh"
#include "tile.cuh"
#include <type_traits>
#include <float.h>

#define HY_THREADS 256
#define HY_G 8                 // lanes per target
#define HY_U 7                 // neighbours per lane the unrolled loop covers (k <= 56)

// ---------------------------------------------------------------------------------------------
// Records (sorted order in the session, caller order in the function-level API)
//   RecD (48 B) = {x, y | z, h2s | wA, wB}      h2s = h^2, sign bit set for dust;  wA = c6 m [gas or dust];
//                                               wB = c6 m [gas]/(amu mu);  c6 = 315/(64 pi h^9)          (density)
//   RecP (80 B) = {x, y | z, h2 | vx, vy | vz, AK | cs, rho}
//                                               AK = 6 c6 m k T [gas]/(mu amu); cs = sqrt(gamma k T [gas]/(mu amu));
//                                               rho is filled by the density pass                        (Pi, force)
//   recK (8 B)  = PK = 6 c6 m Pi [gas], filled by the Pi pass                                            (force)
//   aux  (8 B)  = 6 c6 m [gas] (PK = aux * Pi)
// ---------------------------------------------------------------------------------------------
struct __align__(16) RecD { double x, y, z, h2s, wA, wB; };
struct __align__(16) RecP { double x, y, z, h2, vx, vy, vz, AK, cs, rho; };

__device__ __forceinline__ double2 ld2(const double* p) { return *reinterpret_cast<const double2*>(p); }

// A record as the pair functions see it: in global memory (LIST form) or in shared memory (TILE form).  The shared
// form carries the 32-bit shared-window address and loads with ld.shared, so that the compiler never falls back to
// generic loads (LD.E on a shared address is several times slower than LDS).
struct RecG {
  const double* p;
  __device__ __forceinline__ double2 v2(int i) const { return *reinterpret_cast<const double2*>(p + i); }
  __device__ __forceinline__ double v1(int i) const { return p[i]; }
};
struct RecS {
  uint32_t a;
  __device__ __forceinline__ double2 v2(int i) const {
    double2 r;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "r"(a + 8u * (uint32_t)i));
    return r;
  }
  __device__ __forceinline__ double v1(int i) const {
    double r;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(r) : "r"(a + 8u * (uint32_t)i));
    return r;
  }
};
__device__ __forceinline__ unsigned lds_u16(uint32_t addr) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
  return (unsigned)v;
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double r;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(r) : "r"(addr));
  return r;
}

__global__ void __launch_bounds__(256)
pack_kernel(int n, PackIn in, PhysConst pc, const double4* __restrict__ r0, RecD* __restrict__ rd, RecP* __restrict__ rp,
            double* __restrict__ aux, double* __restrict__ recK) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double h = in.h[i];
  const double h2 = h * h;
  double x, y, z;
  if (in.points) { x = in.points[3 * (int64_t)i]; y = in.points[3 * (int64_t)i + 1]; z = in.points[3 * (int64_t)i + 2]; }
  else { const double4 q = r0[i]; x = q.x; y = q.y; z = q.z; }
  const double h4 = h2 * h2;
  const double c6 = 315.0 / (64.0 * 3.141592653589793 * (h4 * h4 * h));
  const double t = in.ptype[i];
  const double m = in.mass[i];
  const double gas = (t == 0.0) ? 1.0 : 0.0;
  const double mg = m * gas;
  const double mn = mg / pc.amu / in.mu[i];
  const double md = (t == 2.0) ? m : 0.0;
  RecD d;
  d.x = x; d.y = y; d.z = z;
  d.h2s = (t == 2.0) ? -h2 : h2;
  d.wA = c6 * (mg + md);
  d.wB = c6 * mn;
  rd[i] = d;
  const double kT = pc.kB * in.T[i];
  RecP p;
  p.x = x; p.y = y; p.z = z; p.h2 = h2;
  p.vx = in.vel[3 * (int64_t)i]; p.vy = in.vel[3 * (int64_t)i + 1]; p.vz = in.vel[3 * (int64_t)i + 2];
  p.AK = (6.0 * c6) * (m / in.mu[i] / pc.amu * kT * gas);            // 6 c6 A_i, A_i of nsc:615
  p.cs = sqrt(in.gamma[i] * kT / (in.mu[i] * pc.amu) * gas);
  p.rho = 0.0;
  rp[i] = p;
  aux[i] = (6.0 * c6) * mg;
  recK[i] = 0.0;
}

int launch_pack(sphb200_ctx* ctx, int n, const PackIn& in, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->recD, (size_t)(n + 2) * sizeof(RecD)));
  SPH_TRY(devbuf_reserve(ctx->recP, (size_t)(n + 2) * sizeof(RecP)));
  SPH_TRY(devbuf_reserve(ctx->recK, (size_t)(n + 2) * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->aux, (size_t)(n + 2) * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->pi_, (size_t)(n + 2) * sizeof(double)));
  pack_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, in, ctx->phys, bufp<double4>(ctx->r0), bufp<RecD>(ctx->recD),
                                                bufp<RecP>(ctx->recP), bufp<double>(ctx->aux), bufp<double>(ctx->recK));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// int64 (N,k) neighbour array of the reference API -> int32 list; out-of-range entries are flagged.
__global__ void __launch_bounds__(256)
convert_list_kernel(const int64_t* __restrict__ nb, int64_t total, int n, int32_t* __restrict__ list, int* flags) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= total) return;
  int64_t j = nb[t];
  if (j == INT64_MIN) { list[t] = -1; return; }        // padding of a ragged row (binding's sentinel)
  if (j < 0 && j >= -(int64_t)n) j += n;               // numpy negative indexing
  if (j < 0 || j >= n) { atomicAdd(&flags[FL_BAD_INDEX], 1); j = -1; }
  list[t] = (int32_t)j;
}

int launch_convert_list(sphb200_ctx* ctx, const int64_t* nb, int n, int k, int32_t* list, cudaStream_t st) {
  const int64_t total = (int64_t)n * k;
  convert_list_kernel<<<ceil_div(total, 256), 256, 0, st>>>(nb, total, n, list, bufp<int>(ctx->flags));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Pass policies.  `rec` pointers may point to shared memory (tile form) or global memory (list form).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double group_sum(double v) {                    // sum over the HY_G lanes of a target
#pragma unroll
  for (int d = HY_G / 2; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

struct DensityPass {
  static constexpr int REC = 48;
  static constexpr bool USE_K = false, HAS_FINAL = false;
  struct Params { const RecD* rec; double* rho; double* dust; double* num; RecP* rp; };
  struct Self { double x, y, z; };
  struct Acc { double rho, num, dust; };
  __device__ static const char* records(const Params& p) { return reinterpret_cast<const char*>(p.rec); }
  __device__ static void zero(Acc& a) { a.rho = a.num = a.dust = 0.0; }
  template <class R> __device__ static void self(Self& s, R pos, R, int, const Params&) {
    const double2 a = pos.v2(0), b = pos.v2(2);
    s.x = a.x; s.y = a.y; s.z = b.x;
  }
  template <class R> __device__ static void pair(Acc& a, const Self& s, R r, double) {
    const double2 xy = r.v2(0), zh = r.v2(2), w = r.v2(4);
    const double dx = xy.x - s.x, dy = xy.y - s.y, dz = zh.x - s.z;
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double t = fabs(zh.y) - r2;
    const double t3 = t * t * t;
    double PA = w.x * t3, PB = w.y * t3;                 // c6(h_j) m_j (h_j^2 - r^2)^3
    PA = (PA < 0.0) ? 0.0 : PA;                          // nsc:589 (NaN stays NaN, as in numpy)
    PB = (PB < 0.0) ? 0.0 : PB;
    const bool dust = __double2hiint(zh.y) < 0;          // sign bit of h2s
    a.rho += dust ? 0.0 : PA;
    a.dust += dust ? PA : 0.0;
    a.num += PB;
  }
  struct Final {};
  __device__ static const double* kptr(const Params&) { return nullptr; }
  __device__ static void store(Acc& a, const Self&, int i, bool lead, bool active, const Params& p, Final&) {
    a.rho = group_sum(a.rho); a.num = group_sum(a.num); a.dust = group_sum(a.dust);
    if (lead && active) { p.rho[i] = a.rho; p.dust[i] = a.dust; p.num[i] = a.num; if (p.rp) p.rp[i].rho = a.rho; }
  }
  __device__ static void final_init(Final&) {}
  __device__ static void final_flush(Final&, const Params&) {}
};

// Pi_i = sum_j -alpha/2 (2 cbar_ij - 3 w_ij) w_ij / rhobar_ij,  alpha = 1            nsc:641-649
struct PiPass {
  static constexpr int REC = 80;
  static constexpr bool USE_K = false, HAS_FINAL = false;
  struct Params { const RecP* rec; const double* aux; double* Pi; double* recK; };
  struct Self { double x, y, z, vx, vy, vz, cs_i, rho_i; };
  struct Acc { double v; };
  __device__ static const char* records(const Params& p) { return reinterpret_cast<const char*>(p.rec); }
  __device__ static void zero(Acc& a) { a.v = 0.0; }
  template <class R> __device__ static void self(Self& s, R pos, R own, int, const Params&) {
    const double2 a = pos.v2(0), b = pos.v2(2), c = pos.v2(4), d = pos.v2(6), e = own.v2(8);
    s.x = a.x; s.y = a.y; s.z = b.x; s.vx = c.x; s.vy = c.y; s.vz = d.x;
    s.cs_i = e.x; s.rho_i = e.y;
  }
  template <class R> __device__ static void pair(Acc& a, const Self& s, R r, double) {
    const double2 xy = r.v2(0), zh = r.v2(2), v01 = r.v2(4), v2a = r.v2(6), cr = r.v2(8);
    const double dx = xy.x - s.x, dy = xy.y - s.y, dz = zh.x - s.z;
    const double dvx = v01.x - s.vx, dvy = v01.y - s.vy, dvz = v2a.x - s.vz;
    const double r2 = dx * dx + dy * dy + dz * dz;
    // one rsqrt and one reciprocal instead of sqrt + two divisions (each an fp64 software sequence);
    // the extra roundings are ~1e-16 relative, the tolerance is 1e-10
    double w = (dvx * dx + dvy * dy + dvz * dz) * rsqrt(r2 + 0.01 * zh.y);   // nsc:643
    w = (w > 0.0) ? 0.0 : w;                                                // nsc:644
    const double inv_rho_avg = 2.0 * __drcp_rn(cr.y + s.rho_i);             // nsc:646
    const double cbar = 0.5 * (cr.x + s.cs_i);                              // nsc:647
    a.v += (-0.5 * (cbar * 2.0 - 3.0 * w) * w) * inv_rho_avg;               // nsc:649
  }
  struct Final {};
  __device__ static const double* kptr(const Params&) { return nullptr; }
  __device__ static void store(Acc& a, const Self&, int i, bool lead, bool active, const Params& p, Final&) {
    a.v = group_sum(a.v);
    if (lead && active) { p.Pi[i] = a.v; p.recK[i] = p.aux[i] * a.v; }
  }
  __device__ static void final_init(Final&) {}
  __device__ static void final_flush(Final&, const Params&) {}
};

// hydro_accel_i = -(1/rho_i) sum_j 1/2 (A_j gb_ij + A_i ga_ij)                       nsc:615,619
// visc_accel_i  = -sum_j 1/2 (m_j Pi_j [gas_j] gb_ij + m_i Pi_i [gas_i] ga_ij)      nsc:651-652
// visc_heat_i   = (m_i/2) sum_j 1/2 (same bracket) . dv_ij                           nsc:653-654
// with gb_ij = -6 c6(h_j) (h_j^2 - r^2)^2 dx_ij, ga_ij the same with h_i — NOT clamped outside the
// support (nsc:591-592).
struct ForcePass {
  static constexpr int REC = 80;
  static constexpr bool USE_K = true, HAS_FINAL = false;
  struct Params { const RecP* rec; const double* recK; const double* mass; double* hacc; double* vacc; double* heat; };
  struct Self { double x, y, z, vx, vy, vz, h2_i, AK_i, PK_i, rho_i, m_i; };
  struct Acc { double px, py, pz, qx, qy, qz, qh; };
  __device__ static const char* records(const Params& p) { return reinterpret_cast<const char*>(p.rec); }
  __device__ static void zero(Acc& a) { a.px = a.py = a.pz = a.qx = a.qy = a.qz = a.qh = 0.0; }
  template <class R> __device__ static void self(Self& s, R pos, R own, int i, const Params& p) {
    const double2 a = pos.v2(0), b = pos.v2(2), c = pos.v2(4), d = pos.v2(6);
    s.x = a.x; s.y = a.y; s.z = b.x; s.vx = c.x; s.vy = c.y; s.vz = d.x;
    s.h2_i = own.v1(3); s.AK_i = own.v1(7); s.rho_i = own.v1(9);
    s.PK_i = p.recK[i];
    s.m_i = p.mass[i];                                    // fetched early: its latency hides behind the pair loop
  }
  template <class R> __device__ static void pair(Acc& a, const Self& s, R r, double PK_j) {
    const double2 xy = r.v2(0), zh = r.v2(2), v01 = r.v2(4), v2a = r.v2(6);
    const double dx = xy.x - s.x, dy = xy.y - s.y, dz = zh.x - s.z;
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double tb = zh.y - r2, ta = s.h2_i - r2;
    const double tb2 = tb * tb, ta2 = ta * ta;
    const double sP = v2a.y * tb2 + s.AK_i * ta2;
    const double sV = PK_j * tb2 + s.PK_i * ta2;
    a.px += sP * dx; a.py += sP * dy; a.pz += sP * dz;
    a.qx += sV * dx; a.qy += sV * dy; a.qz += sV * dz;
    a.qh += sV * ((v01.x - s.vx) * dx + (v01.y - s.vy) * dy + (v2a.x - s.vz) * dz);
  }
  struct Final {};
  __device__ static const double* kptr(const Params& p) { return p.recK; }
  __device__ static void store(Acc& a, const Self& s, int i, bool lead, bool active, const Params& p, Final&) {
    a.px = group_sum(a.px); a.py = group_sum(a.py); a.pz = group_sum(a.pz);
    a.qx = group_sum(a.qx); a.qy = group_sum(a.qy); a.qz = group_sum(a.qz); a.qh = group_sum(a.qh);
    if (!lead) return;
    if (!active) {                                       // ghost / non-finite target: keep its outputs defined
      for (int c = 0; c < 3; ++c) { p.hacc[3 * (int64_t)i + c] = 0.0; p.vacc[3 * (int64_t)i + c] = 0.0; }
      p.heat[i] = 0.0;
      return;
    }
    const double inv = 0.5 / s.rho_i;                    // bracket = -1/2 s dx ; accel = -sum(bracket)/rho
    p.hacc[3 * (int64_t)i + 0] = a.px * inv;
    p.hacc[3 * (int64_t)i + 1] = a.py * inv;
    p.hacc[3 * (int64_t)i + 2] = a.pz * inv;
    p.vacc[3 * (int64_t)i + 0] = 0.5 * a.qx;
    p.vacc[3 * (int64_t)i + 1] = 0.5 * a.qy;
    p.vacc[3 * (int64_t)i + 2] = 0.5 * a.qz;
    p.heat[i] = -0.25 * s.m_i * a.qh;
  }
  __device__ static void final_init(Final&) {}
  __device__ static void final_flush(Final&, const Params&) {}
};

// crossing_time: min over gas i of h_i / max_j |v_j - v_i|, zeros ignored, + 1e-4; dt_0/10 if none.
// Here "self" is the row index (nsc:780 uses velocities[j]).  Records = the (N,3) velocity array itself.
struct CrossingPass {
  static constexpr int REC = 24;
  static constexpr bool USE_K = false, HAS_FINAL = true;
  struct Params { const double* vel; const double* sizes; const double* ptype; const uint32_t* id; uint32_t n_owned;
                  unsigned long long* slot; };
  struct Self { double vx, vy, vz; };
  struct Acc { double mx; int nan; };
  __device__ static const char* records(const Params& p) { return reinterpret_cast<const char*>(p.vel); }
  __device__ static void zero(Acc& a) { a.mx = 0.0; a.nan = 0; }
  template <class R> __device__ static void self(Self& s, R, R own, int, const Params&) {
    s.vx = own.v1(0); s.vy = own.v1(1); s.vz = own.v1(2);
  }
  template <class R> __device__ static void pair(Acc& a, const Self& s, R r, double) {
    const double ax = r.v1(0) - s.vx, ay = r.v1(1) - s.vy, az = r.v1(2) - s.vz;
    const double v2 = ax * ax + ay * ay + az * az;
    if (v2 != v2) a.nan = 1;                             // np.max propagates NaN
    a.mx = fmax(a.mx, v2);
  }
  struct Final { unsigned long long u; };
  __device__ static const double* kptr(const Params&) { return nullptr; }
  __device__ static void final_init(Final& f) { f.u = 0xffffffffffffffffull; }
  __device__ static void store(Acc& a, const Self&, int i, bool lead, bool active, const Params& p, Final& f) {
#pragma unroll
    for (int d = HY_G / 2; d > 0; d >>= 1) {
      a.mx = fmax(a.mx, __shfl_xor_sync(0xffffffffu, a.mx, d));
      a.nan |= __shfl_xor_sync(0xffffffffu, a.nan, d);
    }
    // ghost particles of a slab decomposition (caller index >= n_owned) belong to another rank's minimum
    if (lead && active && p.ptype[i] == 0.0 && (!p.id || p.id[i] < p.n_owned)) {
      double ct = a.nan ? 0.0 : p.sizes[i] / sqrt(a.mx);  // nan_to_num: NaN -> 0
      if (ct != ct) ct = 0.0;                             // 0/0
      if (ct > DBL_MAX) ct = DBL_MAX;                     // nan_to_num: +inf -> largest finite
      if (ct < -DBL_MAX) ct = -DBL_MAX;
      if (ct != 0.0) {
        const long long b = __double_as_longlong(ct);     // ordered encoding for signed doubles
        const unsigned long long u = (b >= 0) ? ((unsigned long long)b | 0x8000000000000000ull) : ~(unsigned long long)b;
        if (u < f.u) f.u = u;
      }
    }
  }
  __device__ static void final_flush(Final& f, const Params& p) {    // one global atomic per warp
    unsigned long long u = f.u;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      const unsigned long long o = __shfl_xor_sync(0xffffffffu, u, d);
      if (o < u) u = o;
    }
    if ((threadIdx.x & 31) == 0 && u != 0xffffffffffffffffull) atomicMin(p.slot, u);
  }
};

// ---------------------------------------------------------------------------------------------
// TILE form: producer / consumer pipeline over the half-groups of cells.
// One producer warp walks the CTA's units (a unit = the targets of one z-half of a group), computes the geometry
// of each one unit ahead and issues its bulk copies (records of the three z-layers, the 8-byte coefficients, the
// uint16 rows of the unit's targets) into a RING of shared memory: a unit takes exactly the bytes it needs
// (typically 30-50 KB of the ~220 KB ring), so four to seven units are in flight or being consumed at any time
// and the DRAM / L2 latency of a fill (~2 us) is hidden behind the arithmetic of the units before it.  Each unit
// has a descriptor slot with a "full" mbarrier (transaction count of its copies) and an "empty" mbarrier (one
// arrival per consumer warp); units are released in the order they were filled, which is what lets the
// producer reclaim ring space by waiting for the oldest unit only.
// ---------------------------------------------------------------------------------------------
struct TilePassArgs {
  const GridSpec* spec;
  const uint32_t* cell_start;
  const uint16_t* list16;      // (N, LIST_STRIDE(k)) tile offsets
  int n, k;
  int ring_bytes;              // shared-memory ring (multiple of 128)
  const int* unit_tab;         // geometry (TP_TAB_INTS ints each) of the ACTIVE units of the current cell list + rows
  const int* unit_count;       // device: number of active units (> unit_cap: the table overflowed and is not used)
  int unit_cap;                // units the table has room for
  int debug;                   // tile_cfg option: 9 = consumers skip the pair loops (measures the fill pipeline alone)
};

#define TP_SLOTS 12            // descriptor slots = units in flight at most (divisible by the producer counts 1..4)
#define TP_TEAMS 1             // consumer warps work in teams; a unit belongs to ONE team (unit sequence number mod TP_TEAMS)
#define TP_PREFETCH 4          // units whose table entry the producer has requested ahead of the one it is issuing

struct TpMeta {                // what the consumers need to know about a staged unit
  int tbegin[2], tbase[3];
  int o_lo, ntgt;
  int base;                    // byte offset of the unit's data in the ring
  int nrec;                    // records staged (padded); coefficients follow the records, rows follow those
  int pad[3];
};

struct TpShared {              // fixed part of the shared memory, after the ring
  TpMeta meta[TP_SLOTS];
  TileRows rows[4];            // producers' scratch (geometry computed in the kernel when there is no table)
  int plen4[4][4];
  int end_off[TP_SLOTS];       // producer: ring offset where the unit of a slot ends
  uint64_t full[TP_SLOTS], empty[TP_SLOTS];
};

template <class P>
__host__ __device__ constexpr int tp_unit_bytes_max(int ks) {
  return ((TILE_HALF_CAP * (P::REC + (P::USE_K ? 8 : 0)) + TILE_HALF_TARGETS * ks * 2) + 127) & ~127;
}

// geometry of one unit as the producer warp holds it in registers between "computed" and "issued"
struct TpGeom {
  int skip;                  // nothing to do for this unit (warp-uniform)
  int b0, b1, l0, l1, o_lo;  // target runs of the half-group, first staged tile offset (warp-uniform)
  int nrec;                  // padded records of the three z-layers (warp-uniform)
  int plen, o, ps;           // this lane's row: padded length, tile offset relative to o_lo, first padded sorted index
};

// Geometry of unit `unit` (all 32 lanes of one warp; rows / plen4: the warp's shared-memory scratch).
__device__ __forceinline__ TpGeom tp_geometry(const GridSpec& g, const uint32_t* __restrict__ cell_start, int unit, int nunits,
                                              int lane, TileRows* rows, int* plen4) {
  TpGeom gm;
  gm.skip = 1; gm.b0 = gm.b1 = gm.l0 = gm.l1 = gm.o_lo = gm.nrec = 0; gm.plen = gm.o = gm.ps = 0;
  if (unit >= nunits) return gm;
  const int group = unit >> 1, hz = unit & 1;
  int gx, gy, gz;
  group_coords(g, group, &gx, &gy, &gz);
  const int cz = 2 * gz + hz;
  if (cz >= g.nz) return gm;
  __syncwarp();
  tile_rows(g, cell_start, gx, gy, gz, lane, rows);
  int b = 0, e = 0, r = 0;
  const bool have = (lane < 4) && group_target_run(g, cell_start, gx, gy, gz, lane, &b, &e, &r);
  if (lane < 4) plen4[lane] = have ? (e - b) : 0;
  gm.b0 = __shfl_sync(0xffffffffu, b, 2 * hz);
  gm.b1 = __shfl_sync(0xffffffffu, b, 2 * hz + 1);
  gm.l0 = __shfl_sync(0xffffffffu, have ? e - b : 0, 2 * hz);
  gm.l1 = __shfl_sync(0xffffffffu, have ? e - b : 0, 2 * hz + 1);
  __syncwarp();
  const int zlo = max(cz - 1, 0), zhi = min(cz + 1, g.nz - 1);
  const int r_lo = (zlo - rows->z0) * rows->ysp, r_hi = (zhi - rows->z0 + 1) * rows->ysp;
  gm.o_lo = rows->toff[r_lo];
  gm.nrec = rows->toff[r_hi] - gm.o_lo;
  // the same decision the k-NN kernel took for this group: the targets of a crowded group have global rows
  gm.skip = (gm.l0 + gm.l1 == 0) || tile_crowded(g, rows, plen4);
  const int rr = r_lo + lane;
  if (rr < r_hi) { gm.plen = rows->toff[rr + 1] - rows->toff[rr]; gm.o = rows->toff[rr] - gm.o_lo; gm.ps = rows->ps[rr]; }
  __syncwarp();
  return gm;
}

// The same from the per-step table (unit_geometry_kernel): two coalesced loads per lane instead of a chain of
// dependent cell_start lookups, so one producer warp keeps up with sixteen consumer warps.
#define TP_TAB_INTS 48
__device__ __forceinline__ TpGeom tp_geometry_load(const int* __restrict__ tabp, int unit, int nunits, int lane) {
  TpGeom gm;
  gm.skip = 1; gm.b0 = gm.b1 = gm.l0 = gm.l1 = gm.o_lo = gm.nrec = 0; gm.plen = gm.o = gm.ps = 0;
  if (unit >= nunits) return gm;
  const int* e = tabp + (size_t)unit * TP_TAB_INTS;
  const int4 h0 = __ldg(reinterpret_cast<const int4*>(e)), h1 = __ldg(reinterpret_cast<const int4*>(e) + 1);
  gm.skip = h0.x; gm.b0 = h0.y; gm.b1 = h0.z; gm.l0 = h0.w; gm.l1 = h1.x; gm.o_lo = h1.y; gm.nrec = h1.z;
  if (lane < 12) { gm.plen = __ldg(e + 8 + 3 * lane); gm.o = __ldg(e + 9 + 3 * lane); gm.ps = __ldg(e + 10 + 3 * lane); }
  return gm;
}

// Table of the ACTIVE units: those with at least one target whose row is not in the global list.  Empty units (a
// clustered particle distribution leaves most cells of its bounding box empty), crowded groups and units whose
// targets all escaped to the global list cost the passes nothing.  The order of the table is the order in which the
// warps of this kernel finish (every target is computed independently, so results do not depend on it).
__global__ void __launch_bounds__(256) unit_geometry_kernel(const GridSpec* __restrict__ spec, const uint32_t* __restrict__ cell_start,
                                                            const int* __restrict__ group_list, const uint16_t* __restrict__ list16,
                                                            int ks, int* __restrict__ tabp, int* __restrict__ count, int unit_cap) {
  __shared__ TileRows rows_s[8];
  __shared__ int plen4_s[8][4];
  const GridSpec g = *spec;
  const int nunits = 2 * g.ngroups;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int nwalk = 2 * group_list[0];                                      // only the groups that hold particles (cells.cu)
  for (int u = blockIdx.x * 8 + w; u < nwalk; u += gridDim.x * 8) {
    const int unit = 2 * group_list[1 + (u >> 1)] + (u & 1);
    {                                                                       // cheap test first: any target at all?
      const int group = unit >> 1, hz = unit & 1;
      int gx, gy, gz;
      group_coords(g, group, &gx, &gy, &gz);
      int b = 0, e = 0, r = 0;
      const bool have = (lane < 2) && group_target_run(g, cell_start, gx, gy, gz, 2 * hz + lane, &b, &e, &r);
      if (!__any_sync(0xffffffffu, have && e > b)) continue;
    }
    const TpGeom gm = tp_geometry(g, cell_start, unit, nunits, lane, &rows_s[w], plen4_s[w]);
    if (gm.skip) continue;
    bool need = false;
    for (int t = lane; t < gm.l0 + gm.l1; t += 32) {
      const int i = (t < gm.l0) ? gm.b0 + t : gm.b1 + (t - gm.l0);
      need |= list16[(int64_t)i * ks] != (uint16_t)LIST_ESCAPED;
    }
    if (!__any_sync(0xffffffffu, need)) continue;
    int slot = 0;
    if (lane == 0) slot = atomicAdd(count, 1);
    slot = __shfl_sync(0xffffffffu, slot, 0);
    if (slot >= unit_cap) continue;                                         // overflow: the passes fall back to all units
    int* e = tabp + (size_t)slot * TP_TAB_INTS;
    if (lane == 0) {
      *reinterpret_cast<int4*>(e) = make_int4(gm.skip, gm.b0, gm.b1, gm.l0);
      *(reinterpret_cast<int4*>(e) + 1) = make_int4(gm.l1, gm.o_lo, gm.nrec, 0);
    }
    if (lane < 12) { e[8 + 3 * lane] = gm.plen; e[9 + 3 * lane] = gm.o; e[10 + 3 * lane] = gm.ps; }
  }
}

template <class P, int NCW, int NP, int MINB>
__global__ void __launch_bounds__(32 * (NCW + NP), MINB) tile_pass_kernel(TilePassArgs a, typename P::Params p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ks = LIST_STRIDE(a.k);
  TpShared* sh = reinterpret_cast<TpShared*>(smem + a.ring_bytes);
  const GridSpec g = *a.spec;
  const char* grec = P::records(p);
  const double* gk = P::kptr(p);
  if (tid == 0) {
    for (int s = 0; s < TP_SLOTS; ++s) { mbar_init(&sh->full[s], 1); mbar_init(&sh->empty[s], NCW / TP_TEAMS); }
    mbar_fence_init();
  }
  __syncthreads();
  const int nactive = a.unit_tab ? *a.unit_count : 0x7fffffff;
  const bool tab = nactive <= a.unit_cap;                                   // else: every unit, geometry computed here
  const int nunits = tab ? nactive : 2 * g.ngroups;

  if (warp >= NCW) {
    // ================= producer warps =================
    // A producer spends ~200 dependent instructions per unit (ring bookkeeping, descriptor, expect_tx, 14-26 bulk
    // copies), about 1.3 us next to 20 busy consumer warps, so NP of them share the work: producer r fills the
    // units r, r + NP, ... (sequence numbers within the CTA) into its own NP-th of the ring; no coordination.
    static_assert(TP_SLOTS % NP == 0, "a slot must always belong to the same producer");
    const int pr = warp - NCW;
    const int RING = (a.ring_bytes / NP) & ~127;
    unsigned char* ring0 = smem + pr * RING;
    TileRows* rows = &sh->rows[pr];
    int* plen4 = sh->plen4[pr];
    auto geometry = [&](int unit) -> TpGeom {
      if (tab) return tp_geometry_load(a.unit_tab, unit, nunits, lane);
      return tp_geometry(g, a.cell_start, unit, nunits, lane, rows, plen4);
    };
    int it = pr;
    int head = 0, tail = 0, tail_it = pr;                                   // ring state (warp-uniform)
#ifdef TP_PROFILE
    long long t_wait = 0, t_issue = 0, t_geo = 0, t_all = clock64(); int n_wait = 0;
#endif
    // table entries are requested TP_PREFETCH units ahead, so the producer never waits for a global load
    TpGeom pf[TP_PREFETCH];
    const int ustride = NP * gridDim.x;
#pragma unroll
    for (int j = 0; j < TP_PREFETCH; ++j) pf[j] = geometry(blockIdx.x + pr * gridDim.x + j * ustride);
    for (int unit = blockIdx.x + pr * gridDim.x; unit < nunits; unit += ustride, it += NP) {
      const TpGeom gm = pf[0];
      const int slot = it % TP_SLOTS;
      const int ntg = gm.l0 + gm.l1;
      const int off_k = gm.nrec * P::REC;
      const int off_l = off_k + (P::USE_K ? gm.nrec * 8 : 0);
      const int need = gm.skip ? 0 : ((off_l + ntg * ks * 2 + 127) & ~127);
      int base = 0;
#ifdef TP_PROFILE
      long long c0 = clock64();
#endif
      for (;;) {
        const int in_flight = (it - tail_it) / NP;
        if (in_flight == 0) { head = 0; tail = 0; }
        bool fits = false;
        if (in_flight < TP_SLOTS / NP) {
          if (head >= tail) {
            if (need <= RING - head) { base = head; fits = true; }
            else if (need < tail) { base = 0; fits = true; }                // wrap: the bytes up to RING stay unused this lap
          } else if (need < tail - head) { base = head; fits = true; }
        }
        if (fits) break;
        const int ts = tail_it % TP_SLOTS;
        mbar_wait(&sh->empty[ts], (uint32_t)((tail_it / TP_SLOTS) & 1));    // the oldest unit was released by every consumer warp
        tail = sh->end_off[ts];
        tail_it += NP;
#ifdef TP_PROFILE
        ++n_wait;
#endif
      }
#ifdef TP_PROFILE
      long long c1 = clock64(); t_wait += c1 - c0;
#endif
      head = base + need;
      TpMeta* meta = &sh->meta[slot];
      if (lane == 0) {
        sh->end_off[slot] = head;
        meta->tbegin[0] = gm.b0; meta->tbegin[1] = gm.b1;
        meta->tbase[0] = 0; meta->tbase[1] = gm.l0; meta->tbase[2] = ntg;
        meta->o_lo = gm.o_lo;
        meta->ntgt = gm.skip ? 0 : ntg;
        meta->base = pr * RING + base; meta->nrec = gm.nrec;
      }
      __syncwarp();
      if (gm.skip) {
        if (lane == 0) mbar_expect_tx(&sh->full[slot], 0u);                 // plain arrival: nothing to copy
      } else {
        // (no proxy fence: the ring bytes were only READ through the generic proxy, and the consumers' mbarrier
        //  arrivals that released them order those reads before the asynchronous writes issued here)
        unsigned char* sb = ring0 + base;
        const unsigned tot = (unsigned)gm.nrec * (unsigned)(P::REC + (P::USE_K ? 8 : 0)) + (unsigned)ntg * (unsigned)(ks * 2);
        if (lane == 0) mbar_expect_tx(&sh->full[slot], tot);
        __syncwarp();
        if (gm.plen) {
          bulk_g2s(sb + (size_t)gm.o * P::REC, grec + (size_t)gm.ps * P::REC, (unsigned)gm.plen * P::REC, &sh->full[slot]);
          if (P::USE_K) bulk_g2s(sb + off_k + (size_t)gm.o * 8, gk + gm.ps, (unsigned)gm.plen * 8u, &sh->full[slot]);
        }
        uint16_t* rowsm = reinterpret_cast<uint16_t*>(sb + off_l);
        if (lane == 0 && gm.l0) bulk_g2s(rowsm, a.list16 + (int64_t)gm.b0 * ks, (unsigned)gm.l0 * ks * 2u, &sh->full[slot]);
        if (lane == 1 && gm.l1) bulk_g2s(rowsm + (size_t)gm.l0 * ks, a.list16 + (int64_t)gm.b1 * ks, (unsigned)gm.l1 * ks * 2u, &sh->full[slot]);
      }
#ifdef TP_PROFILE
      long long c2 = clock64(); t_issue += c2 - c1;
#endif
#pragma unroll
      for (int j = 0; j + 1 < TP_PREFETCH; ++j) pf[j] = pf[j + 1];
      pf[TP_PREFETCH - 1] = geometry(unit + TP_PREFETCH * ustride);
#ifdef TP_PROFILE
      t_geo += clock64() - c2;
#endif
    }
#ifdef TP_PROFILE
    if (lane == 0 && (blockIdx.x == 0 || blockIdx.x == 77))
      printf("TP_PROFILE cta %d rec %d ncw %d: units %d total %lld cyc, ring wait %lld (%d waits), issue %lld, prefetch %lld\n",
             (int)blockIdx.x, P::REC, NCW, it, clock64() - t_all, t_wait, n_wait, t_issue, t_geo);
#endif
    return;
  }
  // ================= consumer warps =================
  typename P::Final fin;
  P::final_init(fin);
  const int sub = lane & (HY_G - 1);
  constexpr int TPW = 32 / HY_G;                                            // targets a warp works on at a time
  // Every mbarrier hand-over costs each participating warp ~40 cycles, so a unit is consumed by ONE team of
  // WPT warps (the other teams work on the units before and after it).  A unit holds ~45 targets = ~11 quads
  // for the team's WPT * TPW slots per round: the quads are dealt round-robin over the team's warps ACROSS its
  // units (`rot` = quads dealt so far, mod WPT), so the warp that got one quad less is served first next time.
  static_assert(NCW % TP_TEAMS == 0, "consumer warps must split evenly into teams");
  constexpr int WPT = NCW / TP_TEAMS;
  const int team = warp / WPT, wt = warp - team * WPT;
  int rot = 0;
  for (int it = team; blockIdx.x + (long long)it * gridDim.x < nunits; it += TP_TEAMS) {
    const int slot = it % TP_SLOTS;
    mbar_wait(&sh->full[slot], (uint32_t)((it / TP_SLOTS) & 1));            // the unit is filled
    const TpMeta* meta = &sh->meta[slot];
    const int ntgt = meta->ntgt;
    if (ntgt > 0 && a.debug < 90) {
      const unsigned char* sb = smem + meta->base;
      const int nrec = meta->nrec;
      const uint32_t trec = smem_u32(sb);
      const uint32_t tk = trec + (uint32_t)(nrec * P::REC);
      const uint32_t rowsm = tk + (P::USE_K ? (uint32_t)(nrec * 8) : 0u);
      const int o_lo = meta->o_lo, tb1 = meta->tbase[1], b0 = meta->tbegin[0], b1 = meta->tbegin[1];
      const int nq = (ntgt + TPW - 1) / TPW;
      int q0 = wt - rot;
      if (q0 < 0) q0 += WPT;
      rot = (rot + nq) % WPT;
      for (int q = q0; q < nq; q += WPT) {
        const int t = q * TPW + (lane / HY_G);
        const bool valid = t < ntgt;
        const int i = valid ? ((t >= tb1) ? b1 + (t - tb1) : b0 + t) : 0;
        const uint32_t row = rowsm + (uint32_t)((valid ? t : 0) * ks) * 2u;
        const unsigned s0 = lds_u16(row);
        const bool active = valid && s0 < LIST_MISSING;
        typename P::Self self;
        typename P::Acc acc;
        P::zero(acc);
        if (active) {
          const RecS rs = {trec + (uint32_t)(s0 - o_lo) * (uint32_t)P::REC};
          P::self(self, rs, rs, i, p);
          if (lds_u16(row + 2u * (uint32_t)(a.k - 1)) < LIST_MISSING) {
            // full row (missing entries sit at the end): branch-free, unrolled, so that the shared-memory reads of
            // up to HY_U neighbours are in flight together
            unsigned oo[HY_U];
#pragma unroll
            for (int u = 0; u < HY_U; ++u) oo[u] = (sub + HY_G * u < a.k) ? lds_u16(row + 2u * (uint32_t)(sub + HY_G * u)) : s0;
#pragma unroll
            for (int u = 0; u < HY_U; ++u) {
              if (sub + HY_G * u < a.k) {
                const RecS r = {trec + (uint32_t)(oo[u] - o_lo) * (uint32_t)P::REC};
                P::pair(acc, self, r, P::USE_K ? lds_f64(tk + 8u * (uint32_t)(oo[u] - o_lo)) : 0.0);
              }
            }
          } else {
            for (int e = sub; e < a.k; e += HY_G) {
              const unsigned o = lds_u16(row + 2u * (uint32_t)e);
              if (o >= LIST_MISSING) continue;
              const RecS r = {trec + (uint32_t)(o - o_lo) * (uint32_t)P::REC};
              P::pair(acc, self, r, P::USE_K ? lds_f64(tk + 8u * (uint32_t)(o - o_lo)) : 0.0);
            }
          }
        }
        const bool escaped = valid && s0 == LIST_ESCAPED;      // its row is in the global list: the LIST form writes it
        P::store(acc, self, i, sub == 0 && valid && !escaped, active, p, fin);
      }
    }
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&sh->empty[slot])) : "memory");
  }
  if (P::HAS_FINAL) P::final_flush(fin, p);
}

// MINB CTAs per SM, each with its own producer warp, NCW consumer warps and a ring of 1/MINB of the SM's shared
// memory.  More CTAs = more producer warps (a producer spends ~200 dependent instructions per unit, which bounds a
// CTA at about one unit per 1.3 us) but shorter rings; NCW is what the pass's register count allows.
template <class P, int NCW, int NP, int MINB>
static int launch_tile_pass_cfg(sphb200_ctx* ctx, int n, int k, const uint16_t* list16, const typename P::Params& p, cudaStream_t st) {
  const int ks = LIST_STRIDE(k);
  const int fixed = ((int)sizeof(TpShared) + 127) & ~127;
  const int ring = (((227 * 1024) / MINB - fixed - 1024) / 128) * 128;
  if (tp_unit_bytes_max<P>(ks) > ((ring / NP) & ~127)) SPH_FAIL("tile pass: k=%d needs %d B of shared memory per unit", k, tp_unit_bytes_max<P>(ks));
  TilePassArgs a;
  a.spec = bufp<GridSpec>(ctx->grid_dev);
  a.cell_start = bufp<uint32_t>(ctx->cell_start);
  a.list16 = list16;
  a.n = n; a.k = k;
  a.ring_bytes = ring;
  // per-unit geometry table of the current cell list, built by the first pass after a cell build
  const int64_t cap64 = (int64_t)n / 8 + 4096;
  const int unit_cap = (int)(cap64 < 40000000 ? cap64 : 40000000);
  {
    const void* before = ctx->unit_tab.p;
    SPH_TRY(devbuf_reserve(ctx->unit_tab, 64 + (size_t)unit_cap * TP_TAB_INTS * sizeof(int)));   // [0]: count of active units
    if (ctx->unit_tab.p != before) ctx->unit_tab_valid = 0;
  }
  if (!ctx->unit_tab_valid) {
    const int blocks = (unit_cap + 7) / 8 < 148 * 8 ? (unit_cap + 7) / 8 : 148 * 8;
    CUDA_TRY(cudaMemsetAsync(ctx->unit_tab.p, 0, 64, st));
    unit_geometry_kernel<<<blocks, 256, 0, st>>>(a.spec, a.cell_start, bufp<int>(ctx->group_list), list16, ks, bufp<int>(ctx->unit_tab) + 16,
                                                 bufp<int>(ctx->unit_tab), unit_cap);
    ctx->launches += 1;
    CUDA_TRY(cudaGetLastError());
    ctx->unit_tab_valid = 1;
  }
  a.unit_tab = bufp<int>(ctx->unit_tab) + 16;
  a.unit_count = bufp<int>(ctx->unit_tab);
  a.unit_cap = unit_cap;
  a.debug = ctx->tile_cfg;
  const int smem = ring + fixed;
  CUDA_TRY(cudaFuncSetAttribute(tile_pass_kernel<P, NCW, NP, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  tile_pass_kernel<P, NCW, NP, MINB><<<148 * MINB, 32 * (NCW + NP), smem, st>>>(a, p);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

template <class P> struct TpSmall { static constexpr bool V = false; };       // records of <= 48 B: three rings fit an SM
template <> struct TpSmall<DensityPass> { static constexpr bool V = true; };
template <> struct TpSmall<CrossingPass> { static constexpr bool V = true; };

// Measured at N = 10^6 (B200, profiles/README.md): the passes with small records are bound by the producers (most
// CTAs / producers win), Pi and force by the consumers' shared-memory gathers (one CTA with a long ring wins).
// tile_cfg 1..3 select other splits for experiments.
template <class P>
static int launch_tile_pass(sphb200_ctx* ctx, int n, int k, const uint16_t* list16, const typename P::Params& p, cudaStream_t st) {
  const int cfg = ctx->tile_cfg % 10;       // (tile_cfg >= 90: consumers skip the pair loops, to time the fill pipeline alone)
  constexpr bool S = TpSmall<P>::V;
  if (cfg == 1) return launch_tile_pass_cfg<P, S ? 24 : 20, 1, 1>(ctx, n, k, list16, p, st);
  if (cfg == 2) return launch_tile_pass_cfg<P, S ? 24 : 20, 2, 1>(ctx, n, k, list16, p, st);
  if (cfg == 3) return launch_tile_pass_cfg<P, 10, 1, 2>(ctx, n, k, list16, p, st);
  if (std::is_same<P, CrossingPass>::value) return launch_tile_pass_cfg<P, 8, 1, 3>(ctx, n, k, list16, p, st);
  if (std::is_same<P, DensityPass>::value) return launch_tile_pass_cfg<P, 12, 2, 2>(ctx, n, k, list16, p, st);
  return launch_tile_pass_cfg<P, 20, 1, 1>(ctx, n, k, list16, p, st);
}

// ---------------------------------------------------------------------------------------------
// LIST form: int32 rows in global memory.  targets == nullptr: every particle i, row i; otherwise the
// particles targets[0 .. *ntargets) with row t for entry t (escaped rows of a session, stored by slot).
// ---------------------------------------------------------------------------------------------
struct ListPassArgs {
  const int32_t* list;
  const int* targets;
  const int* ntargets;
  int n, k;
  const uint32_t* skip_id;     // slab decomposition: ghosts are not targets
  uint32_t skip_from;
};

template <class P>
__global__ void __launch_bounds__(HY_THREADS) list_pass_kernel(ListPassArgs a, typename P::Params p) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane & (HY_G - 1);
  constexpr int TPW = 32 / HY_G;                              // targets per warp
  const int ntarget = a.targets ? *a.ntargets : a.n;
  const char* grec = P::records(p);
  const double* gk = P::kptr(p);
  typename P::Final fin;
  P::final_init(fin);
  const int wstride = gridDim.x * (HY_THREADS / 32) * TPW;
  for (int base = (blockIdx.x * (HY_THREADS / 32) + warp) * TPW; base < ntarget; base += wstride) {
    const int t = base + lane / HY_G;
    const bool valid = t < ntarget;
    const int i = valid ? (a.targets ? a.targets[t] : t) : 0;
    const int32_t* row = a.list + (int64_t)(valid ? (a.targets ? t : i) : 0) * a.k;
    const bool active = valid && !(a.skip_id && a.skip_id[i] >= a.skip_from);
    typename P::Self self;
    typename P::Acc acc;
    P::zero(acc);
    if (active) {
      const int js = row[0];
      __align__(16) double zrec[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
      const RecG own = {reinterpret_cast<const double*>(grec + (size_t)i * P::REC)};
      const RecG pos = {(js >= 0 && js < a.n) ? reinterpret_cast<const double*>(grec + (size_t)js * P::REC) : zrec};
      P::self(self, pos, own, i, p);
      for (int s = sub; s < a.k; s += HY_G) {
        const int j = row[s];
        if (j < 0 || j >= a.n) continue;
        const RecG r = {reinterpret_cast<const double*>(grec + (size_t)j * P::REC)};
        P::pair(acc, self, r, P::USE_K ? gk[j] : 0.0);
      }
    }
    P::store(acc, self, i, sub == 0 && valid, active, p, fin);
  }
  if (P::HAS_FINAL) P::final_flush(fin, p);
}

template <class P>
static int launch_list_pass(sphb200_ctx* ctx, int n, int k, const int32_t* list, const int* targets, const int* ntargets,
                            const typename P::Params& p, cudaStream_t st) {
  ListPassArgs a;
  a.list = list; a.targets = targets; a.ntargets = ntargets; a.n = n; a.k = k;
  a.skip_id = ctx->skip_id; a.skip_from = ctx->skip_from;
  const int per_block = HY_THREADS / HY_G;
  // work list: grid-stride; measured on the clustered cloud, the force pass is faster with 8 blocks per SM queued
  // (0.998 vs 1.075 ms at N = 2e6), the lighter passes with 4
  int blocks = targets ? 148 * (std::is_same<P, ForcePass>::value ? 8 : 4) : ceil_div(n, per_block);
  if (blocks > 148 * 64) blocks = 148 * 64;                  // grid-stride beyond that
  list_pass_kernel<P><<<blocks, HY_THREADS, 0, st>>>(a, p);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// Run one pass over the rows `r`: the tile form over the uint16 rows plus the list form over the escaped
// targets, or the list form over everything.
template <class P>
static int run_pass(sphb200_ctx* ctx, int n, int k, const RowsRef& r, const typename P::Params& p, cudaStream_t st) {
  if (r.list16) {
    SPH_TRY(launch_tile_pass<P>(ctx, n, k, r.list16, p, st));
    return launch_list_pass<P>(ctx, n, k, r.list, r.targets, r.ntargets, p, st);
  }
  return launch_list_pass<P>(ctx, n, k, r.list, nullptr, nullptr, p, st);
}

int launch_density(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, double* rho, double* dust, double* num, cudaStream_t st) {
  DensityPass::Params p = {bufp<RecD>(ctx->recD), rho, dust, num, bufp<RecP>(ctx->recP)};
  return run_pass<DensityPass>(ctx, n, k, rows, p, st);
}

int launch_pi(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, cudaStream_t st) {
  PiPass::Params p = {bufp<RecP>(ctx->recP), bufp<double>(ctx->aux), bufp<double>(ctx->pi_), bufp<double>(ctx->recK)};
  return run_pass<PiPass>(ctx, n, k, rows, p, st);
}

int launch_force(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, const double* mass, double* hydro_accel,
                 double* visc_accel, double* visc_heat, cudaStream_t st) {
  ForcePass::Params p = {bufp<RecP>(ctx->recP), bufp<double>(ctx->recK), mass, hydro_accel, visc_accel, visc_heat};
  return run_pass<ForcePass>(ctx, n, k, rows, p, st);
}

// Owner-computed values of ghost particles arrive after a pass (slab decomposition): put them where the next pass
// reads them.  what = 0: rho[] -> RecP.rho;  1: Pi[] -> recK = aux * Pi.
__global__ void __launch_bounds__(256)
repack_kernel(int n, int what, const double* __restrict__ rho, const double* __restrict__ Pi, const double* __restrict__ aux,
              RecP* __restrict__ rp, double* __restrict__ recK) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (what == 0) rp[i].rho = rho[i];
  else recK[i] = aux[i] * Pi[i];
}

int launch_repack(sphb200_ctx* ctx, int n, int what, const double* rho, cudaStream_t st) {
  repack_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, what, rho, bufp<double>(ctx->pi_), bufp<double>(ctx->aux),
                                                 bufp<RecP>(ctx->recP), bufp<double>(ctx->recK));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---- crossing time -----------------------------------------------------------------------------
__global__ void crossing_init_kernel(unsigned long long* slot) { *slot = 0xffffffffffffffffull; }

__global__ void crossing_final_kernel(const unsigned long long* slot, double dt_0, double* out) {
  const unsigned long long u = *slot;
  if (u == 0xffffffffffffffffull) { *out = dt_0 / 10.0; return; }
  const long long b = (u & 0x8000000000000000ull) ? (long long)(u & 0x7fffffffffffffffull) : (long long)~u;
  *out = __longlong_as_double(b) + 0.0001;
}

int launch_crossing_time(sphb200_ctx* ctx, int n, int k, const RowsRef& rows, const double* vel, const double* sizes,
                         const double* ptype, double dt_0, double* out_dev, const uint32_t* id, uint32_t n_owned, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->tmp0, 64));
  unsigned long long* slot = bufp<unsigned long long>(ctx->tmp0);
  crossing_init_kernel<<<1, 1, 0, st>>>(slot);
  CrossingPass::Params p = {vel, sizes, ptype, id, n_owned, slot};
  SPH_TRY(run_pass<CrossingPass>(ctx, n, k, rows, p, st));
  crossing_final_kernel<<<1, 1, 0, st>>>(slot, dt_0, out_dev);
  ctx->launches += 2;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// f_un_neighbor[s,i] = sum_j m_num_j f_un[j,s] W_ij                                  nsc:624-626
// 8 lanes per target (each takes every 8th neighbour); the weight of a pair is computed once and applied to
// up to 16 species held in registers; the lanes' partial sums are combined with shuffles.
// ---------------------------------------------------------------------------------------------
#define SP_CHUNK 16
__global__ void __launch_bounds__(HY_THREADS)
species_kernel(int n, int k, int S, const int32_t* __restrict__ list, const RecD* __restrict__ rec,
               const double* __restrict__ f_un, double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (HY_G - 1);
  const int t = (blockIdx.x * HY_THREADS + threadIdx.x) / HY_G;
  const bool valid = t < n;
  const int i = valid ? t : 0;
  const int32_t* row = list + (int64_t)i * k;
  const int js = valid ? row[0] : -1;
  double sx = 0, sy = 0, sz = 0;
  if (js >= 0 && js < n) { sx = rec[js].x; sy = rec[js].y; sz = rec[js].z; }
  for (int s0 = 0; s0 < S; s0 += SP_CHUNK) {
    double acc[SP_CHUNK];
#pragma unroll
    for (int c = 0; c < SP_CHUNK; ++c) acc[c] = 0.0;
    if (valid) {
      for (int s = sub; s < k; s += HY_G) {
        const int j = row[s];
        if (j < 0 || j >= n) continue;
        const double* r = reinterpret_cast<const double*>(rec + j);
        const double2 xy = ld2(r), zh = ld2(r + 2), w = ld2(r + 4);
        const double dx = xy.x - sx, dy = xy.y - sy, dz = zh.x - sz;
        const double tt = fabs(zh.y) - (dx * dx + dy * dy + dz * dz);
        double PB = w.y * (tt * tt * tt);                  // m_num_j c6(h_j) (h_j^2 - r^2)^3
        PB = (PB < 0.0) ? 0.0 : PB;
        const double* f = f_un + (int64_t)j * S + s0;
#pragma unroll
        for (int c = 0; c < SP_CHUNK; ++c) if (s0 + c < S) acc[c] += f[c] * PB;
      }
    }
#pragma unroll
    for (int c = 0; c < SP_CHUNK; ++c) {
      const double v = group_sum(acc[c]);
      if (valid && sub == 0 && s0 + c < S) out[(int64_t)(s0 + c) * n + i] = v;
    }
  }
}

int launch_species(sphb200_ctx* ctx, int n, int k, int S, const int32_t* list, const double* f_un, double* out,
                   cudaStream_t st) {
  species_kernel<<<ceil_div((int64_t)n * HY_G, HY_THREADS), HY_THREADS, 0, st>>>(n, k, S, list, bufp<RecD>(ctx->recD), f_un, out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// F1 in the device-resident step: dust-gas drag on the packed records          nsc:719-742
// (code_running.py:455,462-463,473).  Only dust neighbours carry a term, so the per-particle MRN
// means stay in the caller's order and are fetched through `id` for those neighbours alone; the
// reaction on the dust particle is a sparse fp64 RED.  LIST form over int32 rows of every particle.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(HY_THREADS)
drag_kernel(int n, int k, const int32_t* __restrict__ list, const RecD* __restrict__ rd, const RecP* __restrict__ rp,
            const uint32_t* __restrict__ id, const double* __restrict__ mg, const double* __restrict__ mc,
            double* __restrict__ onto, double* __restrict__ reaction) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (HY_G - 1);
  const int t = (blockIdx.x * HY_THREADS + threadIdx.x) / HY_G;
  const bool valid = t < n;
  const int i = valid ? t : 0;
  const int32_t* row = list + (int64_t)i * k;
  const RecP ps = rp[i];
  double ax = 0, ay = 0, az = 0;
  if (valid) {
    for (int s = sub; s < k; s += HY_G) {
      const int j = row[s];
      if (j < 0 || j >= n) continue;
      const RecD q = rd[j];
      if (__double2hiint(q.h2s) >= 0) continue;             // [type_j == 2]: the dust flag of the record
      const double dx = q.x - ps.x, dy = q.y - ps.y, dz = q.z - ps.z;
      const double r2_ = dx * dx + dy * dy + dz * dz;
      const double tt = fabs(q.h2s) - r2_;
      const double wf = q.wA * (tt * tt * tt);              // Weigh2_dust: m c6(h_j) (h_j^2 - r^2)^3, nsc:678-681
      if (!(wf > 0.0)) continue;                            // [weight_factor > 0]
      const RecP v = rp[j];
      const double du = v.vx - ps.vx, dv = v.vy - ps.vy, dw = v.vz - ps.vz;
      const double speed = sqrt(du * du + dv * dv + dw * dw);
      const uint32_t g = id ? id[j] : (uint32_t)j;
      const double c = wf / mg[g] * mc[g] * speed;          // nsc:738
      const double gx = c * du, gy = c * dv, gz = c * dw;
      ax += gx; ay += gy; az += gz;
      if (j != i) {                                         // nsc:741
        if (gx != 0.0) atomicAdd(reaction + 3 * (int64_t)j, -gx);
        if (gy != 0.0) atomicAdd(reaction + 3 * (int64_t)j + 1, -gy);
        if (gz != 0.0) atomicAdd(reaction + 3 * (int64_t)j + 2, -gz);
      }
    }
  }
  ax = group_sum(ax); ay = group_sum(ay); az = group_sum(az);
  if (valid && sub == 0) { onto[3 * (int64_t)i] = ax; onto[3 * (int64_t)i + 1] = ay; onto[3 * (int64_t)i + 2] = az; }
}

__device__ __forceinline__ double drag_nan_to_num(double v) {
  if (v != v) return 0.0;
  if (v > DBL_MAX) return DBL_MAX;
  if (v < -DBL_MAX) return -DBL_MAX;
  return v;
}

// visc_accel = drag_accel_gas + drag_accel_dust + av[0]                    code_running.py:462-463,473
__global__ void __launch_bounds__(256)
drag_apply_kernel(int n, const double* __restrict__ onto, const double* __restrict__ reaction,
                  const double* __restrict__ rho, const double* __restrict__ rho_dust, const double* __restrict__ ptype,
                  double* __restrict__ visc_accel) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double gas = (ptype[i] == 0.0) ? 1.0 : 0.0;
  const double rd = rho_dust[i], r = rho[i];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const double dg = drag_nan_to_num(onto[3 * (int64_t)i + c] * rd / r * gas);
    const double dd = drag_nan_to_num(reaction[3 * (int64_t)i + c]);
    visc_accel[3 * (int64_t)i + c] = (dg + dd) + visc_accel[3 * (int64_t)i + c];
  }
}

int launch_drag(sphb200_ctx* ctx, int n, int k, const int32_t* list, const uint32_t* id, const double* mg,
                const double* mc, const double* rho, const double* rho_dust, const double* ptype, double* onto,
                double* reaction, double* visc_accel, cudaStream_t st) {
  CUDA_TRY(cudaMemsetAsync(reaction, 0, (size_t)n * 3 * sizeof(double), st));
  drag_kernel<<<ceil_div((int64_t)n * HY_G, HY_THREADS), HY_THREADS, 0, st>>>(
      n, k, list, bufp<RecD>(ctx->recD), bufp<RecP>(ctx->recP), id, mg, mc, onto, reaction);
  drag_apply_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, onto, reaction, rho, rho_dust, ptype, visc_accel);
  ctx->launches += 2;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// uint16 tile-offset rows (+ the escaped rows of the global list, by slot) -> int32 rows in the sorted index
// space, unordered, missing = n.  For the consumers off the hot path: download of "neighbor", drag, checkpoints.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
expand_list_kernel(int n, int k, const GridSpec* __restrict__ spec, const uint32_t* __restrict__ cell_start,
                   const uint32_t* __restrict__ keys, const uint16_t* __restrict__ list16, const int32_t* __restrict__ glist,
                   int32_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint16_t* row = list16 + (int64_t)i * LIST_STRIDE(k);
  int32_t* o = out + (int64_t)i * k;
  const unsigned s0 = row[0];
  if (s0 == LIST_ESCAPED) {
    const int slot = (int)row[1] | ((int)row[2] << 16);
    for (int s = 0; s < k; ++s) o[s] = glist[(int64_t)slot * k + s];
    return;
  }
  if (s0 == LIST_MISSING) { for (int s = 0; s < k; ++s) o[s] = n; return; }
  const GridSpec g = *spec;
  const uint32_t key = keys[i];
  const int cx = key % g.nx, cy = (key / g.nx) % g.ny, cz = key / (g.nx * g.ny);
  const int gx = cx >> 1, gy = cy >> 1, gz = cz >> 1;
  const int x0 = max(2 * gx - 1, 0), x1 = min(2 * gx + 2, g.nx - 1);
  const int y0 = max(2 * gy - 1, 0), y1 = min(2 * gy + 2, g.ny - 1);
  const int z0 = max(2 * gz - 1, 0), z1 = min(2 * gz + 2, g.nz - 1);
  const int ysp = y1 - y0 + 1, nrows = ysp * (z1 - z0 + 1);
  int ps[TILE_MAX_ROWS], toff[TILE_MAX_ROWS + 1];
  int run = 0;
  for (int r = 0; r < TILE_MAX_ROWS; ++r) {
    int rs = 0, re = 0;
    if (r < nrows) {
      const int base = ((z0 + r / ysp) * g.ny + (y0 + r % ysp)) * g.nx;
      rs = (int)cell_start[base + x0]; re = (int)cell_start[base + x1 + 1];
    }
    ps[r] = rs & ~1;
    toff[r] = run;
    run += (re > rs) ? (((re + 1) & ~1) - ps[r]) : 0;
  }
  toff[TILE_MAX_ROWS] = run;
  for (int s = 0; s < k; ++s) {
    const int off = row[s];
    if (off >= (int)LIST_MISSING) { o[s] = n; continue; }
    int r = 0;
    while (r < TILE_MAX_ROWS - 1 && off >= toff[r + 1]) ++r;
    o[s] = ps[r] + (off - toff[r]);
  }
}

int launch_expand_list(sphb200_ctx* ctx, int n, int k, const uint16_t* list16, const int32_t* glist, int32_t* out, cudaStream_t st) {
  expand_list_kernel<<<ceil_div(n, 128), 128, 0, st>>>(n, k, bufp<GridSpec>(ctx->grid_dev), bufp<uint32_t>(ctx->cell_start),
                                                      bufp<uint32_t>(ctx->keys), list16, glist, out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
