// Vectorised-generation hydro passes over a stored neighbour list:
//   pack      -> per-particle records (see common.cuh)             nsc:559-573, 587-592 per-particle factors
//   density   -> rho, rho_dust, n                                  nsc:588-590, 605-607
//   pi        -> per-particle artificial-viscosity Pi_i            nsc:639-649 (axis repaired)
//   force     -> hydro_accel, visc_accel, visc_heat                nsc:591-592, 615-619, 651-654
//   species   -> f_un_neighbor                                     nsc:624-626
//   crossing  -> min_i h_i / max_j |v_j - v_i|                     nsc:776-786
// G lanes share one target particle: each lane takes neighbours s = lane, lane+G, ... of the row,
// accumulates privately in fp64 and the G partial sums are combined with warp shuffles.  No atomics;
// each particle's sums are produced by exactly one group (gather formulation).
// "self" follows the reference: positions/velocities of column 0 of the row (nsc:580-581), while
// h_i, rho_i, m_i, Pi_i are taken at the row index.
#include "common.cuh"
#include "kernels.cuh"
#include <cooperative_groups.h>
#include <cooperative_groups/reduce.h>
#include <float.h>
namespace cg = cooperative_groups;

#define HY_THREADS 128
#define HY_G 8


// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
pack_kernel(int n, PackIn in, PhysConst pc, double4* __restrict__ r0, double4* __restrict__ r1, double4* __restrict__ r2,
            double* __restrict__ coefA) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double h = in.h[i];
  const double h2 = h * h;
  if (in.points) r0[i] = make_double4(in.points[3 * i], in.points[3 * i + 1], in.points[3 * i + 2], h2);
  else { double4 q = r0[i]; q.w = h2; r0[i] = q; }
  const double h4 = h2 * h2;
  const double c6 = 315.0 / (64.0 * 3.141592653589793 * (h4 * h4 * h));
  const double t = in.ptype[i];
  const double m = in.mass[i];
  const double gas = (t == 0.0) ? 1.0 : 0.0;
  const double mg = m * gas;
  const double mn = mg / pc.amu / in.mu[i];
  r1[i] = make_double4(c6, mg, mn, (t == 2.0) ? m : 0.0);
  const double kT = pc.kB * in.T[i];
  const double cs = sqrt(in.gamma[i] * kT / (in.mu[i] * pc.amu) * gas);
  r2[i] = make_double4(in.vel[3 * i], in.vel[3 * i + 1], in.vel[3 * i + 2], cs);
  coefA[i] = m / in.mu[i] / pc.amu * kT * gas;          // A_i of nsc:615
}

int launch_pack(sphb200_ctx* ctx, int n, const PackIn& in, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->r0, (size_t)(n + 1) * sizeof(double4)));
  SPH_TRY(devbuf_reserve(ctx->r1, (size_t)(n + 1) * sizeof(double4)));
  SPH_TRY(devbuf_reserve(ctx->r2, (size_t)(n + 1) * sizeof(double4)));
  SPH_TRY(devbuf_reserve(ctx->r3, (size_t)(n + 1) * sizeof(rec2)));
  SPH_TRY(devbuf_reserve(ctx->aux, (size_t)(n + 1) * sizeof(double)));
  pack_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, in, ctx->phys, bufp<double4>(ctx->r0), bufp<double4>(ctx->r1),
                                                bufp<double4>(ctx->r2), bufp<double>(ctx->aux));
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
template <int G>
__global__ void __launch_bounds__(HY_THREADS)
density_kernel(int n, int k, const int32_t* __restrict__ list, const double4* __restrict__ r0,
               const double4* __restrict__ r1, double* __restrict__ rho, double* __restrict__ dust,
               double* __restrict__ num, const uint32_t* __restrict__ skip_id, int skip_from) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (HY_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  if (skip_id && (int)skip_id[i] >= skip_from) return;     // ghost: its owner computes it
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const int js = row[0];
  double4 ps = make_double4(0, 0, 0, 0);
  if (js >= 0) ps = r0[js];
  double a_rho = 0.0, a_dust = 0.0, a_num = 0.0;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0 || j >= n) continue;
    const double4 q = ldg256(r0 + j);
    const double4 w = ldg256(r1 + j);
    const double dx = q.x - ps.x, dy = q.y - ps.y, dz = q.z - ps.z;
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double t = q.w - r2;
    double W = w.x * (t * t * t);                       // c6(h_j) (h_j^2 - r^2)^3
    W = (W < 0.0) ? 0.0 : W;                            // nsc:589 (NaN stays NaN, as in numpy)
    a_rho += w.y * W;
    a_num += w.z * W;
    a_dust += w.w * W;
  }
  a_rho = cg::reduce(tile, a_rho, cg::plus<double>());
  a_num = cg::reduce(tile, a_num, cg::plus<double>());
  a_dust = cg::reduce(tile, a_dust, cg::plus<double>());
  if (lane == 0) { rho[i] = a_rho; dust[i] = a_dust; num[i] = a_num; }
}

int launch_density(sphb200_ctx* ctx, int n, int k, const int32_t* list, double* rho, double* dust, double* num,
                   cudaStream_t st) {
  density_kernel<HY_G><<<ceil_div(n, HY_THREADS / HY_G), HY_THREADS, 0, st>>>(
      n, k, list, bufp<double4>(ctx->r0), bufp<double4>(ctx->r1), rho, dust, num, ctx->skip_id, ctx->skip_from);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Pi_i = sum_j -alpha/2 (2 cbar_ij - 3 w_ij) w_ij / rhobar_ij,  alpha = 1            nsc:641-649
// then r3[i] = {6 c6_i A_i, 6 c6_i m_i Pi_i [gas_i]} for the force pass.
template <int G>
__global__ void __launch_bounds__(HY_THREADS)
pi_kernel(int n, int k, const int32_t* __restrict__ list, const double4* __restrict__ r0, const double4* __restrict__ r1,
          const double4* __restrict__ r2, const double* __restrict__ rho, const double* __restrict__ coefA,
          double* __restrict__ Pi, rec2* __restrict__ r3, const uint32_t* __restrict__ skip_id, int skip_from) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (HY_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  if (skip_id && (int)skip_id[i] >= skip_from) return;     // ghost: Pi and r3 arrive from its owner
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const int js = row[0];
  double4 ps = make_double4(0, 0, 0, 0), vs = make_double4(0, 0, 0, 0);
  if (js >= 0) { ps = r0[js]; vs = r2[js]; }
  const double rho_i = rho[i];
  const double cs_i = r2[i].w;
  double acc = 0.0;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0 || j >= n) continue;
    const double4 q = ldg256(r0 + j);
    const double4 v = ldg256(r2 + j);
    const double dx = q.x - ps.x, dy = q.y - ps.y, dz = q.z - ps.z;
    const double dvx = v.x - vs.x, dvy = v.y - vs.y, dvz = v.z - vs.z;
    const double r2_ = dx * dx + dy * dy + dz * dz;
    // one rsqrt and one reciprocal instead of sqrt + two divisions (each an fp64 software sequence);
    // the extra roundings are ~1e-16 relative, the tolerance is 1e-10
    double w = (dvx * dx + dvy * dy + dvz * dz) * rsqrt(r2_ + 0.01 * q.w);  // nsc:643
    w = (w > 0.0) ? 0.0 : w;                                               // nsc:644
    const double inv_rho_avg = 2.0 * __drcp_rn(rho[j] + rho_i);            // nsc:646
    const double cbar = 0.5 * (v.w + cs_i);                                // nsc:647
    acc += (-0.5 * (cbar * 2.0 - 3.0 * w) * w) * inv_rho_avg;              // nsc:649
  }
  acc = cg::reduce(tile, acc, cg::plus<double>());
  if (lane == 0) {
    Pi[i] = acc;
    const double4 w = r1[i];
    const double K = 6.0 * w.x;
    rec2 o; o.a = K * coefA[i]; o.b = K * (w.y * acc);
    r3[i] = o;
  }
}

// r3[i] = {6 c6_i A_i, 6 c6_i m_i Pi_i [gas_i]} from the current Pi array (after the ghosts' Pi were
// overwritten with their owners' values; the pi pass itself skips ghost rows)
__global__ void __launch_bounds__(256)
repack_pi_kernel(int n, const double4* __restrict__ r1, const double* __restrict__ coefA, const double* __restrict__ Pi,
                 rec2* __restrict__ r3) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double4 w = r1[i];
  const double K = 6.0 * w.x;
  rec2 o; o.a = K * coefA[i]; o.b = K * (w.y * Pi[i]);
  r3[i] = o;
}

int launch_repack_pi(sphb200_ctx* ctx, int n, cudaStream_t st) {
  repack_pi_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, bufp<double4>(ctx->r1), bufp<double>(ctx->aux), bufp<double>(ctx->pi_),
                                                     bufp<rec2>(ctx->r3));
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int launch_pi(sphb200_ctx* ctx, int n, int k, const int32_t* list, const double* rho, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->pi_, (size_t)(n + 1) * sizeof(double)));
  pi_kernel<HY_G><<<ceil_div(n, HY_THREADS / HY_G), HY_THREADS, 0, st>>>(
      n, k, list, bufp<double4>(ctx->r0), bufp<double4>(ctx->r1), bufp<double4>(ctx->r2), rho, bufp<double>(ctx->aux),
      bufp<double>(ctx->pi_), bufp<rec2>(ctx->r3), ctx->skip_id, ctx->skip_from);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// hydro_accel_i = -(1/rho_i) sum_j 1/2 (A_j gb_ij + A_i ga_ij)                       nsc:615,619
// visc_accel_i  = -sum_j 1/2 (m_j Pi_j [gas_j] gb_ij + m_i Pi_i [gas_i] ga_ij)      nsc:651-652
// visc_heat_i   = (m_i/2) sum_j 1/2 (same bracket) . dv_ij                           nsc:653-654
// with gb_ij = -6 c6(h_j) (h_j^2 - r^2)^2 dx_ij, ga_ij the same with h_i — NOT clamped outside the
// support (nsc:591-592).
template <int G>
__global__ void __launch_bounds__(HY_THREADS)
force_kernel(int n, int k, const int32_t* __restrict__ list, const double4* __restrict__ r0,
             const double4* __restrict__ r2, const rec2* __restrict__ r3, const double* __restrict__ rho,
             const double* __restrict__ mass, double* __restrict__ hydro_accel, double* __restrict__ visc_accel,
             double* __restrict__ visc_heat, const uint32_t* __restrict__ skip_id, int skip_from) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (HY_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  if (skip_id && (int)skip_id[i] >= skip_from) {            // ghost: not advanced here; keep its outputs defined
    if (tile.thread_rank() == 0) {
      for (int c = 0; c < 3; ++c) { hydro_accel[3 * (int64_t)i + c] = 0.0; visc_accel[3 * (int64_t)i + c] = 0.0; }
      visc_heat[i] = 0.0;
    }
    return;
  }
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const int js = row[0];
  double4 ps = make_double4(0, 0, 0, 0), vs = make_double4(0, 0, 0, 0);
  if (js >= 0) { ps = r0[js]; vs = r2[js]; }
  const double h2_i = r0[i].w;
  const rec2 ci = r3[i];
  double px = 0, py = 0, pz = 0, qx = 0, qy = 0, qz = 0, qh = 0;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0 || j >= n) continue;
    const double4 q = ldg256(r0 + j);
    const double4 v = ldg256(r2 + j);
    const rec2 cj = r3[j];
    const double dx = q.x - ps.x, dy = q.y - ps.y, dz = q.z - ps.z;
    const double r2_ = dx * dx + dy * dy + dz * dz;
    const double tb = q.w - r2_, ta = h2_i - r2_;
    const double tb2 = tb * tb, ta2 = ta * ta;
    const double sP = cj.a * tb2 + ci.a * ta2;
    const double sV = cj.b * tb2 + ci.b * ta2;
    px += sP * dx; py += sP * dy; pz += sP * dz;
    qx += sV * dx; qy += sV * dy; qz += sV * dz;
    qh += sV * ((v.x - vs.x) * dx + (v.y - vs.y) * dy + (v.z - vs.z) * dz);
  }
  px = cg::reduce(tile, px, cg::plus<double>()); py = cg::reduce(tile, py, cg::plus<double>());
  pz = cg::reduce(tile, pz, cg::plus<double>()); qx = cg::reduce(tile, qx, cg::plus<double>());
  qy = cg::reduce(tile, qy, cg::plus<double>()); qz = cg::reduce(tile, qz, cg::plus<double>());
  qh = cg::reduce(tile, qh, cg::plus<double>());
  if (lane == 0) {
    const double inv = 0.5 / rho[i];                    // bracket = -1/2 s dx ; accel = -sum(bracket)/rho
    hydro_accel[3 * (int64_t)i + 0] = px * inv;
    hydro_accel[3 * (int64_t)i + 1] = py * inv;
    hydro_accel[3 * (int64_t)i + 2] = pz * inv;
    visc_accel[3 * (int64_t)i + 0] = 0.5 * qx;
    visc_accel[3 * (int64_t)i + 1] = 0.5 * qy;
    visc_accel[3 * (int64_t)i + 2] = 0.5 * qz;
    visc_heat[i] = -0.25 * mass[i] * qh;
  }
}

int launch_force(sphb200_ctx* ctx, int n, int k, const int32_t* list, const double* rho, const double* mass,
                 double* hydro_accel, double* visc_accel, double* visc_heat, cudaStream_t st) {
  force_kernel<HY_G><<<ceil_div(n, HY_THREADS / HY_G), HY_THREADS, 0, st>>>(
      n, k, list, bufp<double4>(ctx->r0), bufp<double4>(ctx->r2), bufp<rec2>(ctx->r3), rho, mass, hydro_accel, visc_accel,
      visc_heat, ctx->skip_id, ctx->skip_from);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// f_un_neighbor[s,i] = sum_j m_num_j f_un[j,s] W_ij                                  nsc:624-626
// One lane per (target, species block); S <= 32 species handled by the lanes of one warp.
__global__ void __launch_bounds__(HY_THREADS)
species_kernel(int n, int k, int S, const int32_t* __restrict__ list, const double4* __restrict__ r0,
               const double4* __restrict__ r1, const double* __restrict__ f_un, double* __restrict__ out) {
  const int warp = (blockIdx.x * HY_THREADS + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= n) return;
  const int i = warp;
  const int32_t* row = list + (int64_t)i * k;
  const int js = row[0];
  double4 ps = make_double4(0, 0, 0, 0);
  if (js >= 0) ps = r0[js];
  for (int s0 = 0; s0 < S; s0 += 32) {
    const int sp = s0 + lane;
    double acc = 0.0;
    for (int s = 0; s < k; ++s) {
      const int j = row[s];
      if (j < 0 || j >= n) continue;
      const double4 q = ldg256(r0 + j);
      const double4 w = ldg256(r1 + j);
      const double dx = q.x - ps.x, dy = q.y - ps.y, dz = q.z - ps.z;
      const double t = q.w - (dx * dx + dy * dy + dz * dz);
      double W = w.x * (t * t * t);
      W = (W < 0.0) ? 0.0 : W;
      if (sp < S) acc += w.z * f_un[(int64_t)j * S + sp] * W;
    }
    if (sp < S) out[(int64_t)sp * n + i] = acc;
  }
}

int launch_species(sphb200_ctx* ctx, int n, int k, int S, const int32_t* list, const double* f_un, double* out,
                   cudaStream_t st) {
  species_kernel<<<ceil_div((int64_t)n * 32, HY_THREADS), HY_THREADS, 0, st>>>(n, k, S, list, bufp<double4>(ctx->r0),
                                                                            bufp<double4>(ctx->r1), f_un, out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// crossing_time: min over gas i of h_i / max_j |v_j - v_i|, zeros ignored, + 1e-4; dt_0/10 if none.
// Here "self" is the row index (nsc:780 uses velocities[j]).
__global__ void crossing_init_kernel(unsigned long long* slot) { *slot = 0xffffffffffffffffull; }

// velocities (N,3) -> 32-byte records so that a neighbour costs one sector pair instead of three
// strided 8-byte loads
__global__ void __launch_bounds__(256) vel_pack_kernel(int n, const double* __restrict__ vel, double4* __restrict__ rec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rec[i] = make_double4(vel[3 * (int64_t)i], vel[3 * (int64_t)i + 1], vel[3 * (int64_t)i + 2], 0.0);
}

template <int G>
__global__ void __launch_bounds__(HY_THREADS)
crossing_kernel(int n, int k, const int32_t* __restrict__ list, const double4* __restrict__ vrec,
                const double* __restrict__ sizes, const double* __restrict__ ptype, const uint32_t* __restrict__ id,
                int n_owned, unsigned long long* slot) {
  __shared__ unsigned long long smin;
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = min(blockIdx.x * (HY_THREADS / G) + (int)tile.meta_group_rank(), n - 1);   // tail groups repeat n-1
  if (threadIdx.x == 0) smin = 0xffffffffffffffffull;
  __syncthreads();
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const double4 vi = vrec[i];
  double mx = 0.0;
  bool nan = false;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0 || j >= n) continue;
    const double4 vj = ldg256(vrec + j);
    const double ax = vj.x - vi.x, ay = vj.y - vi.y, az = vj.z - vi.z;
    const double v2 = ax * ax + ay * ay + az * az;
    if (v2 != v2) nan = true;                           // np.max propagates NaN
    mx = fmax(mx, v2);
  }
  mx = cg::reduce(tile, mx, cg::greater<double>());
  nan = tile.any(nan);
  unsigned long long u = 0xffffffffffffffffull;
  // ghost particles of a slab decomposition (caller index >= n_owned) belong to another rank's minimum
  if (lane == 0 && ptype[i] == 0.0 && (!id || (int)id[i] < n_owned)) {
    double ct = nan ? 0.0 : sizes[i] / sqrt(mx);        // nan_to_num: NaN -> 0
    if (ct != ct) ct = 0.0;                             // 0/0
    if (ct > DBL_MAX) ct = DBL_MAX;                     // nan_to_num: +inf -> largest finite
    if (ct < -DBL_MAX) ct = -DBL_MAX;
    if (ct != 0.0) {
      // ordered encoding for signed doubles
      long long b = __double_as_longlong(ct);
      u = (b >= 0) ? ((unsigned long long)b | 0x8000000000000000ull) : ~(unsigned long long)b;
    }
  }
  // one global atomic per CTA, not per particle (a single hot address serialises in L2)
  if (u != 0xffffffffffffffffull) atomicMin(&smin, u);
  __syncthreads();
  if (threadIdx.x == 0 && smin != 0xffffffffffffffffull) atomicMin(slot, smin);
}

__global__ void crossing_final_kernel(const unsigned long long* slot, double dt_0, double* out) {
  const unsigned long long u = *slot;
  if (u == 0xffffffffffffffffull) { *out = dt_0 / 10.0; return; }
  const long long b = (u & 0x8000000000000000ull) ? (long long)(u & 0x7fffffffffffffffull) : (long long)~u;
  *out = __longlong_as_double(b) + 0.0001;
}

int launch_crossing_time(sphb200_ctx* ctx, int n, int k, const int32_t* list, const double* vel, const double* sizes,
                         const double* ptype, double dt_0, double* out_dev, const uint32_t* id, int n_owned, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->tmp0, 64));
  unsigned long long* slot = bufp<unsigned long long>(ctx->tmp0);
  SPH_TRY(devbuf_reserve(ctx->r2, (size_t)(n + 1) * sizeof(double4)));   // scratch here; pack() rewrites it later
  double4* vrec = bufp<double4>(ctx->r2);
  crossing_init_kernel<<<1, 1, 0, st>>>(slot);
  vel_pack_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, vel, vrec);
  crossing_kernel<HY_G><<<ceil_div(n, HY_THREADS / HY_G), HY_THREADS, 0, st>>>(n, k, list, vrec, sizes, ptype, id, n_owned, slot);
  crossing_final_kernel<<<1, 1, 0, st>>>(slot, dt_0, out_dev);
  ctx->launches += 4;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// F1 in the device-resident step: dust-gas drag on the packed records          nsc:719-742
// (code_running.py:455,462-463,473).  Only dust neighbours carry a term, so the per-particle MRN
// means stay in the caller's order and are fetched through `id` for those neighbours alone; the
// reaction on the dust particle is a sparse fp64 RED.
template <int G>
__global__ void __launch_bounds__(HY_THREADS)
drag_kernel(int n, int k, const int32_t* __restrict__ list, const double4* __restrict__ r0,
            const double4* __restrict__ r1, const double4* __restrict__ r2, const uint32_t* __restrict__ id,
            const double* __restrict__ mg, const double* __restrict__ mc, double* __restrict__ onto,
            double* __restrict__ reaction) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (HY_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const double4 ps = r0[i];
  const double4 vs = r2[i];
  double ax = 0, ay = 0, az = 0;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0 || j >= n) continue;
    const double4 w = ldg256(r1 + j);
    if (w.w == 0.0) continue;                             // [type_j == 2] (m_dust of the record)
    const double4 q = ldg256(r0 + j);
    const double4 v = ldg256(r2 + j);
    const double dx = q.x - ps.x, dy = q.y - ps.y, dz = q.z - ps.z;
    const double r2_ = dx * dx + dy * dy + dz * dz;
    const double t = q.w - r2_;
    const double wf = w.w * w.x * (t * t * t);            // Weigh2_dust: m c6(h_j) (h_j^2 - r^2)^3, nsc:678-681
    if (!(wf > 0.0)) continue;                            // [weight_factor > 0]
    const double du = v.x - vs.x, dv = v.y - vs.y, dw = v.z - vs.z;
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
  ax = cg::reduce(tile, ax, cg::plus<double>());
  ay = cg::reduce(tile, ay, cg::plus<double>());
  az = cg::reduce(tile, az, cg::plus<double>());
  if (lane == 0) { onto[3 * (int64_t)i] = ax; onto[3 * (int64_t)i + 1] = ay; onto[3 * (int64_t)i + 2] = az; }
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
  drag_kernel<HY_G><<<ceil_div(n, HY_THREADS / HY_G), HY_THREADS, 0, st>>>(
      n, k, list, bufp<double4>(ctx->r0), bufp<double4>(ctx->r1), bufp<double4>(ctx->r2), id, mg, mc, onto, reaction);
  drag_apply_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, onto, reaction, rho, rho_dust, ptype, visc_accel);
  ctx->launches += 2;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
