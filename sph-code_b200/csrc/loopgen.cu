// Loop-generation SPH sums of the reference (the functions its time loop calls every step):
//   Weigh2 / Weigh2_dust / grad_weight          navier_stokes_cleaned.py:673-690
//   density, dust_density, num_dens             :693-702, :704-717, :744-753
//   del_pressure                                :755-774
//   artificial_viscosity                        :788-816
// Smoothing length here is mass based, h_j = (m_j/m_0)^(1/3) d with the caller-set global `d`.
// Particles stay in the caller's order; G lanes share one row and combine with shuffles.
#include "common.cuh"
#include "kernels.cuh"
#include <cooperative_groups.h>
#include <cooperative_groups/reduce.h>
#include <float.h>
namespace cg = cooperative_groups;

#define LG_THREADS 128
#define LG_G 8


__device__ __forceinline__ double nan_to_num(double v) {
  if (v != v) return 0.0;
  if (v > DBL_MAX) return DBL_MAX;
  if (v < -DBL_MAX) return -DBL_MAX;
  return v;
}

// mode 0: density, 1: dust_density, 2: num_dens
template <int G>
__global__ void __launch_bounds__(LG_THREADS)
loop_density_kernel(int n, int k, const int32_t* __restrict__ list, LoopIn in, int mode, double m_0, double m_h,
                    double* __restrict__ out) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (LG_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const double x0 = in.points[3 * (int64_t)i], y0 = in.points[3 * (int64_t)i + 1], z0 = in.points[3 * (int64_t)i + 2];
  const double d2 = in.d * in.d;
  const double d4 = d2 * d2;
  const double d9 = d4 * d4 * in.d;
  double acc = 0.0;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0) continue;
    const double dx = in.points[3 * (int64_t)j] - x0, dy = in.points[3 * (int64_t)j + 1] - y0,
                 dz = in.points[3 * (int64_t)j + 2] - z0;
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double m = in.mass[j];
    double w;
    if (mode == 1) {
      const double ds = in.sizes[j];
      const double ds2 = ds * ds, ds4 = ds2 * ds2;
      const double t = ds2 - r2;
      w = m * 315.0 * (t * t * t) / (64.0 * 3.141592653589793 * (ds4 * ds4 * ds)) * ((in.ptype[j] == 2.0) ? 1.0 : 0.0);
    } else {
      const double q = m_0 / m;
      const double t = pow(m / m_0, 2.0 / 3.0) * d2 - r2;
      w = m * 315.0 * (q * q * q) * (t * t * t) / (64.0 * 3.141592653589793 * d9);
      if (mode == 0) w *= (in.ptype[j] == 0.0) ? 1.0 : 0.0;
      else w = w / (in.mu[j] * m_h);
    }
    if (w > 0.0) acc += w;                              // np.sum(rho[rho > 0])
  }
  acc = cg::reduce(tile, acc, cg::plus<double>());
  if (lane == 0) out[i] = acc;
}

int launch_loop_density(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, int mode, double* out,
                        cudaStream_t st) {
  loop_density_kernel<LG_G><<<ceil_div(n, LG_THREADS / LG_G), LG_THREADS, 0, st>>>(n, k, list, in, mode, ctx->m_0, ctx->m_h, out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// grad_weight scalar factor (without the (x - x_0) vector)                            nsc:683-690
__device__ __forceinline__ double grad_weight_scalar(double m, double m_0, double d2, double d9, double r2, bool gas) {
  const double q = m_0 / m;
  const double t = pow(m / m_0, 2.0 / 3.0) * d2 - r2;
  return -315.0 * 6.0 * (q * q * q) / (64.0 * 3.141592653589793 * d9) * (t * t) * (gas ? 1.0 : 0.0);
}

template <int G>
__global__ void __launch_bounds__(LG_THREADS)
loop_del_pressure_kernel(int n, int k, const int32_t* __restrict__ list, LoopIn in, double m_0,
                         double* __restrict__ out) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (LG_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  double ax = 0, ay = 0, az = 0;
  if (in.ptype[i] == 0.0) {                              // nsc:758
    const double x0 = in.points[3 * (int64_t)i], y0 = in.points[3 * (int64_t)i + 1], z0 = in.points[3 * (int64_t)i + 2];
    const double d2 = in.d * in.d, d4 = d2 * d2, d9 = d4 * d4 * in.d;
    const double Ei = in.E[i];
    for (int s = lane; s < k; s += G) {
      const int j = row[s];
      if (j < 0) continue;
      const double dx = in.points[3 * (int64_t)j] - x0, dy = in.points[3 * (int64_t)j + 1] - y0,
                   dz = in.points[3 * (int64_t)j + 2] - z0;
      const double r2 = dx * dx + dy * dy + dz * dz;
      const double m = in.mass[j];
      const double w = grad_weight_scalar(m, m_0, d2, d9, r2, in.ptype[j] == 0.0);
      const double inside = (pow(m / m_0, 2.0 / 3.0) * d2 - r2 > 0.0) ? 1.0 : 0.0;
      const double f = 0.5 * (in.E[j] + Ei) / in.gamma[j];
      ax += nan_to_num(w * dx * inside) * f;
      ay += nan_to_num(w * dy * inside) * f;
      az += nan_to_num(w * dz * inside) * f;
    }
  }
  ax = cg::reduce(tile, ax, cg::plus<double>());
  ay = cg::reduce(tile, ay, cg::plus<double>());
  az = cg::reduce(tile, az, cg::plus<double>());
  if (lane == 0) { out[3 * (int64_t)i] = ax; out[3 * (int64_t)i + 1] = ay; out[3 * (int64_t)i + 2] = az; }
}

int launch_loop_del_pressure(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, double* out,
                             cudaStream_t st) {
  loop_del_pressure_kernel<LG_G><<<ceil_div(n, LG_THREADS / LG_G), LG_THREADS, 0, st>>>(n, k, list, in, ctx->m_0, out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

template <int G>
__global__ void __launch_bounds__(LG_THREADS)
loop_viscosity_kernel(int n, int k, const int32_t* __restrict__ list, LoopIn in, PhysConst pc, double m_0,
                      double* __restrict__ accel, double* __restrict__ heat) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (LG_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  double ax = 0, ay = 0, az = 0, hh = 0;
  if (in.ptype[i] == 0.0) {                              // nsc:792 (rows with no gas neighbour sum to zero anyway)
    const double x0 = in.points[3 * (int64_t)i], y0 = in.points[3 * (int64_t)i + 1], z0 = in.points[3 * (int64_t)i + 2];
    const double u0 = in.vel[3 * (int64_t)i], v0 = in.vel[3 * (int64_t)i + 1], w0 = in.vel[3 * (int64_t)i + 2];
    const double d2 = in.d * in.d, d4 = d2 * d2, d9 = d4 * d4 * in.d;
    const double mi = in.mass[i];
    const double css_i = nan_to_num(sqrt(in.gamma[i] * pc.kB * in.T[i] / (in.mu[i] * pc.amu)));
    const double rho_i = in.dens[i];
    const double hi2 = pow(mi / m_0, 2.0 / 3.0) * d2;
    for (int s = lane; s < k; s += G) {
      const int j = row[s];
      if (j < 0) continue;
      if (in.ptype[j] != 0.0) continue;                 // nsc:811-812 keep gas neighbours only
      const double dx = in.points[3 * (int64_t)j] - x0, dy = in.points[3 * (int64_t)j + 1] - y0,
                   dz = in.points[3 * (int64_t)j + 2] - z0;
      const double du = in.vel[3 * (int64_t)j] - u0, dv = in.vel[3 * (int64_t)j + 1] - v0, dw = in.vel[3 * (int64_t)j + 2] - w0;
      const double r2 = dx * dx + dy * dy + dz * dz;
      double wij = (du * dx + dv * dy + dw * dz) / sqrt(r2);          // nsc:794
      if (wij > 0.0) wij = 0.0;
      wij = nan_to_num(wij);
      const double css_j = nan_to_num(sqrt(in.gamma[j] * pc.kB * in.T[j] / (in.mu[j] * pc.amu)));
      const double vsig = css_j + css_i - 3.0 * wij;                  // nsc:797-798
      const double rho = (in.dens[j] + rho_i) / 2.0;
      const double PI = -0.5 * vsig * wij / rho;                      // nsc:802
      const double g = grad_weight_scalar(mi, m_0, d2, d9, r2, true); // nsc:804: m = mass[j-row], gas mask = 1 here
      const double inside = (hi2 - r2 > 0.0) ? 1.0 : 0.0;
      const double gx = nan_to_num(g * dx * inside), gy = nan_to_num(g * dy * inside), gz = nan_to_num(g * dz * inside);
      const double mm = (in.mass[j] + mi) / 2.0;
      ax += mm * PI * gx; ay += mm * PI * gy; az += mm * PI * gz;     // nsc:808
      hh += 0.5 * mm * PI * (du * gx + dv * gy + dw * gz);            // nsc:809
    }
  }
  ax = cg::reduce(tile, ax, cg::plus<double>());
  ay = cg::reduce(tile, ay, cg::plus<double>());
  az = cg::reduce(tile, az, cg::plus<double>());
  hh = cg::reduce(tile, hh, cg::plus<double>());
  if (lane == 0) {
    accel[3 * (int64_t)i] = ax; accel[3 * (int64_t)i + 1] = ay; accel[3 * (int64_t)i + 2] = az;
    heat[i] = hh;
  }
}

int launch_loop_viscosity(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, double* accel,
                          double* heat, cudaStream_t st) {
  loop_viscosity_kernel<LG_G><<<ceil_div(n, LG_THREADS / LG_G), LG_THREADS, 0, st>>>(n, k, list, in, ctx->phys, ctx->m_0, accel, heat);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ---- F1: dust-gas drag, net_impulse                                                 nsc:719-742 ----
// Per-particle MRN means first (nsc:727-728: sum over species of meff*f_un and seff*f_un; the row sum
// follows numpy's order for a contiguous row: 8 running sums combined as a tree, then the tail), then
// a gather over the row for `accel_onto` and a scatter for `accel_reaction`.  The scatter is sparse:
// only dust neighbours with a positive kernel weight (or a non-finite product) carry a term, so it
// is done with RED.ADD.F64 on those entries alone; no symmetric neighbour list is needed.
struct GrainTable { double meff[SPHB200_MAX_SPECIES]; double seff[SPHB200_MAX_SPECIES]; };

__device__ __forceinline__ double row_sum_numpy(const double* __restrict__ f, const double* __restrict__ w, int ns) {
  if (ns < 8) {
    double r = 0.0;
    for (int s = 0; s < ns; ++s) r += __dmul_rn(w[s], f[s]);
    return r;
  }
  double r[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) r[q] = __dmul_rn(w[q], f[q]);
  const int full = ns - (ns % 8);
  for (int s = 8; s < full; s += 8) {
#pragma unroll
    for (int q = 0; q < 8; ++q) r[q] = __dadd_rn(r[q], __dmul_rn(w[s + q], f[s + q]));
  }
  double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                         __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
  for (int s = full; s < ns; ++s) res = __dadd_rn(res, __dmul_rn(w[s], f[s]));
  return res;
}

__global__ void __launch_bounds__(256)
grain_means_kernel(int n, int ns, const double* __restrict__ f_un, GrainTable tab, double* __restrict__ mg,
                   double* __restrict__ mc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* f = f_un + (int64_t)i * ns;
  mg[i] = row_sum_numpy(f, tab.meff, ns);
  mc[i] = row_sum_numpy(f, tab.seff, ns);
}

template <int G>
__global__ void __launch_bounds__(LG_THREADS)
loop_net_impulse_kernel(int n, int k, const int32_t* __restrict__ list, LoopIn in, const double* __restrict__ mg,
                        const double* __restrict__ mc, double* __restrict__ onto, double* __restrict__ reaction) {
  cg::thread_block_tile<G> tile = cg::tiled_partition<G>(cg::this_thread_block());
  const int i = blockIdx.x * (LG_THREADS / G) + tile.meta_group_rank();
  if (i >= n) return;
  const int lane = tile.thread_rank();
  const int32_t* row = list + (int64_t)i * k;
  const double x0 = in.points[3 * (int64_t)i], y0 = in.points[3 * (int64_t)i + 1], z0 = in.points[3 * (int64_t)i + 2];
  const double u0 = in.vel[3 * (int64_t)i], v0 = in.vel[3 * (int64_t)i + 1], w0 = in.vel[3 * (int64_t)i + 2];
  double ax = 0, ay = 0, az = 0;
  for (int s = lane; s < k; s += G) {
    const int j = row[s];
    if (j < 0) continue;
    const double dx = in.points[3 * (int64_t)j] - x0, dy = in.points[3 * (int64_t)j + 1] - y0,
                 dz = in.points[3 * (int64_t)j + 2] - z0;
    const double du = in.vel[3 * (int64_t)j] - u0, dv = in.vel[3 * (int64_t)j + 1] - v0, dw = in.vel[3 * (int64_t)j + 2] - w0;
    const double r2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    const double ds = in.sizes[j];
    const double ds2 = ds * ds, ds4 = ds2 * ds2;
    const double t = ds2 - r2;
    const double wf = in.mass[j] * 315.0 * (t * t * t) / (64.0 * 3.141592653589793 * (ds4 * ds4 * ds));   // nsc:678-681
    const double speed = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(du, du), __dmul_rn(dv, dv)), __dmul_rn(dw, dw)));
    const double dust = (in.ptype[j] == 2.0) ? 1.0 : 0.0, pos = (wf > 0.0) ? 1.0 : 0.0;
    const double c = wf / mg[j] * mc[j] * speed;                                                         // nsc:738
    const double gx = c * du * dust * pos, gy = c * dv * dust * pos, gz = c * dw * dust * pos;
    ax += gx; ay += gy; az += gz;                                                                        // nsc:740
    if (j != i) {                                                                                        // nsc:741
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

int launch_grain_means(sphb200_ctx* ctx, int n, int ns, const double* f_un, const double* meff, const double* seff,
                       double* mg, double* mc, cudaStream_t st) {
  GrainTable tab = {};
  for (int s = 0; s < ns; ++s) { tab.meff[s] = meff[s]; tab.seff[s] = seff[s]; }
  grain_means_kernel<<<ceil_div(n, 256), 256, 0, st>>>(n, ns, f_un, tab, mg, mc);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int launch_loop_net_impulse(sphb200_ctx* ctx, int n, int k, const int32_t* list, const LoopIn& in, const double* mg,
                            const double* mc, double* onto, double* reaction, cudaStream_t st) {
  CUDA_TRY(cudaMemsetAsync(reaction, 0, (size_t)n * 3 * sizeof(double), st));
  loop_net_impulse_kernel<LG_G><<<ceil_div(n, LG_THREADS / LG_G), LG_THREADS, 0, st>>>(n, k, list, in, mg, mc, onto, reaction);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
