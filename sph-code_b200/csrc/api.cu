// extern "C" entry points of libsphb200 (declared in include/sphb200.h).
#include "common.cuh"
#include "kernels.cuh"
#include "../../include/sphb200.h"
#include <new>

thread_local char g_sphb200_err[SPHB200_MAX_ERR] = {0};

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// The function-level calls use the context's scratch buffers (neighbour list, records, cell list) that a live
// session also owns: interleaving them would silently corrupt the run, so it is refused.
static int check_no_session(sphb200_ctx* ctx, const char* who) {
  if (ctx && ctx->session) SPH_FAIL("%s: this context holds a live session; call the function-level API on another context", who);
  return 0;
}

static int check_nk(int64_t n, int k) {
  if (n <= 0) SPH_FAIL("empty particle set (N=%lld)", (long long)n);
  if (n >= (int64_t)0x7fffffff) SPH_FAIL("N=%lld exceeds 2^31-1 particles per GPU", (long long)n);
  if (k <= 0) SPH_FAIL("neighbour count k=%d must be positive", k);
  return 0;
}

extern "C" {

const char* sphb200_last_error(void) { return g_sphb200_err; }
int sphb200_version(void) { return 100; }

int sphb200_create(sphb200_ctx** out, int device) {
  if (!out) SPH_FAIL("sphb200_create: null output pointer");
  int count = 0;
  CUDA_TRY(cudaGetDeviceCount(&count));
  if (device < 0 || device >= count) SPH_FAIL("sphb200_create: device %d not available (%d visible)", device, count);
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) SPH_FAIL("sphb200_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  sphb200_ctx* ctx = new (std::nothrow) sphb200_ctx();
  if (!ctx) SPH_FAIL("sphb200_create: out of host memory");
  ctx->device = device;
  if (devbuf_reserve(ctx->flags, FL_COUNT * sizeof(int)) || devbuf_reserve(ctx->scal, SC_COUNT * sizeof(double)) ||
      cudaMemset(ctx->flags.p, 0, FL_COUNT * sizeof(int)) != cudaSuccess ||
      cudaMemset(ctx->scal.p, 0, SC_COUNT * sizeof(double)) != cudaSuccess) {
    if (!g_sphb200_err[0]) snprintf(g_sphb200_err, SPHB200_MAX_ERR, "sphb200_create: clearing the device scalars failed");
    devbuf_free(ctx->flags); devbuf_free(ctx->scal);
    delete ctx;
    return 1;
  }
  *out = ctx;
  return 0;
}

int sphb200_destroy(sphb200_ctx* ctx) {
  if (!ctx) return 0;
  cudaSetDevice(ctx->device);
  session_free(ctx);
  DevBuf* all[] = {&ctx->r0, &ctx->recD, &ctx->recP, &ctx->recK, &ctx->rho, &ctx->pi_, &ctx->aux, &ctx->list, &ctx->list16,
                   &ctx->keys, &ctx->keys_alt, &ctx->perm, &ctx->perm_alt, &ctx->cell_start, &ctx->scan_tops, &ctx->grid_dev,
                   &ctx->scal, &ctx->flags, &ctx->worklist, &ctx->group_list, &ctx->unit_tab, &ctx->tmp0, &ctx->tmp1, &ctx->tmp2, &ctx->tmp3};
  for (DevBuf* b : all) devbuf_free(*b);
  for (DevBuf& b : ctx->sbuf) devbuf_free(b);
  delete ctx;
  return 0;
}

int sphb200_set_constants(sphb200_ctx* ctx, double k_boltzmann, double amu, double m_h, double m_0) {
  if (!ctx) SPH_FAIL("null context");
  ctx->phys.kB = k_boltzmann; ctx->phys.amu = amu; ctx->m_h = m_h; ctx->m_0 = m_0;
  return 0;
}

int sphb200_set_option(sphb200_ctx* ctx, const char* name, double value) {
  if (!ctx || !name) SPH_FAIL("null argument");
  if (!strcmp(name, "cell_occ")) { if (!(value > 0)) SPH_FAIL("cell_occ must be > 0"); ctx->cell_occ = value; }
  else if (!strcmp(name, "cell_edge")) ctx->cell_edge = value > 0 ? value : 0.0;
  else if (!strcmp(name, "knn_group")) {
    int g = (int)value;
    if (g != 8 && g != 16 && g != 32) SPH_FAIL("knn_group must be 8, 16 or 32");
    ctx->knn_group = g;
  } else if (!strcmp(name, "knn_fast")) ctx->knn_fast = value != 0.0;
  else if (!strcmp(name, "knn_atomic")) ctx->knn_atomic = value != 0.0;
  else if (!strcmp(name, "knn_lane")) ctx->knn_lane = value != 0.0;
  else if (!strcmp(name, "tile_cfg")) ctx->tile_cfg = (int)value;
  else if (!strcmp(name, "knn_mma")) ctx->knn_mma = (int)value;
  else if (!strcmp(name, "grid_pow")) { if (value == 0.0) SPH_FAIL("grid_pow must not be 0"); ctx->grid_pow = value; }
  else if (!strcmp(name, "knn_margin")) { if (!(value >= 0)) SPH_FAIL("knn_margin must be >= 0"); ctx->knn_margin = value; }
  else SPH_FAIL("unknown option '%s'", name);
  return 0;
}

static int read_flag(sphb200_ctx* ctx, int which, int64_t* out, cudaStream_t st) {
  int v = 0;
  int* f = bufp<int>(ctx->flags) + which;
  CUDA_TRY(cudaMemcpyAsync(&v, f, sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemsetAsync(f, 0, sizeof(int), st));
  CUDA_TRY(cudaStreamSynchronize(st));
  *out = v;
  return 0;
}

int sphb200_get_counter(sphb200_ctx* ctx, const char* name, int64_t* value_host, void* stream) {
  if (!ctx || !name || !value_host) SPH_FAIL("null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  if (!strcmp(name, "launches")) { *value_host = ctx->launches; return 0; }
  if (!strcmp(name, "grid_ncells")) { *value_host = ctx->grid.ncells; return 0; }
  if (!strcmp(name, "grid_nx")) { *value_host = ctx->grid.nx; return 0; }
  if (!strcmp(name, "grid_ny")) { *value_host = ctx->grid.ny; return 0; }
  if (!strcmp(name, "grid_nz")) { *value_host = ctx->grid.nz; return 0; }
  if (!strcmp(name, "knn_worklist")) {          // length of the last search's work list (not reset)
    int v = 0;
    CUDA_TRY(cudaMemcpyAsync(&v, bufp<int>(ctx->flags) + FL_KNN_NWORK, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    *value_host = v;
    return 0;
  }
  if (!strcmp(name, "knn_overflow")) return read_flag(ctx, FL_KNN_OVERFLOW, value_host, st);
  if (!strcmp(name, "knn_rescan")) return read_flag(ctx, FL_KNN_RESCAN, value_host, st);
  if (!strcmp(name, "knn_widen")) return read_flag(ctx, FL_KNN_WIDEN, value_host, st);
  if (!strcmp(name, "knn_crowded")) return read_flag(ctx, FL_KNN_CROWDED, value_host, st);
  if (!strcmp(name, "knn_nonfinite")) return read_flag(ctx, FL_KNN_NONFINITE, value_host, st);
  if (!strcmp(name, "bad_index")) return read_flag(ctx, FL_BAD_INDEX, value_host, st);
  SPH_FAIL("unknown counter '%s'", name);
}

// ---- fp64 pipe rate (roofline denominator of the pair passes) ---------------------------------------
__global__ void __launch_bounds__(512) fp64_chain_kernel(int iters, double seed, double* out) {
  double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 1.0000001, c = 1e-9;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
  }
  const double r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (r == 12345.678) out[0] = r;                  // never true: keeps the chain alive
}

int sphb200_fp64_peak(sphb200_ctx* ctx, double* dfma_per_second_host, void* stream) {
  if (!ctx || !dfma_per_second_host) SPH_FAIL("sphb200_fp64_peak: null argument");
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  cudaEvent_t e0, e1;
  CUDA_TRY(cudaEventCreate(&e0));
  CUDA_TRY(cudaEventCreate(&e1));
  const int blocks = 148 * 4, threads = 512, iters = 2048;
  double* out = bufp<double>(ctx->scal) + SC_COUNT - 1;
  fp64_chain_kernel<<<blocks, threads, 0, st>>>(64, 1.0, out);             // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 3; ++rep) {
    CUDA_TRY(cudaEventRecord(e0, st));
    fp64_chain_kernel<<<blocks, threads, 0, st>>>(iters, 1.0, out);
    CUDA_TRY(cudaEventRecord(e1, st));
    CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  ctx->launches += 4;
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  CUDA_TRY(cudaGetLastError());
  *dfma_per_second_host = (double)blocks * threads * iters * 64.0 / ((double)best * 1e-3);
  return 0;
}

// ---- A1 ----------------------------------------------------------------------------------------
int sphb200_neighbors(sphb200_ctx* ctx, const double* points, int64_t n, int k, double dist, int64_t* idx,
                      double* dists, int64_t* nontriv, double* new_sizes, void* stream) {
  if (!ctx || !points || !idx || !dists || !nontriv || !new_sizes) SPH_FAIL("sphb200_neighbors: null argument");
  SPH_TRY(check_no_session(ctx, "sphb200_neighbors"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(build_cells(ctx, points, n, ctx->cell_edge, nullptr, k, true, st));
  SPH_TRY(devbuf_reserve(ctx->list, (size_t)n * k * sizeof(int32_t)));
  SPH_TRY(devbuf_reserve(ctx->tmp1, (size_t)n * k * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->tmp2, (size_t)n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->tmp3, (size_t)n * sizeof(int32_t)));
  if (ctx->knn_fast && k <= knn_group_max_k())
    SPH_TRY(launch_knn_group(ctx, (int)n, k, dist, 0, 0.0, nullptr, bufp<int32_t>(ctx->list), bufp<double>(ctx->tmp1),
                             bufp<double>(ctx->tmp2), bufp<int32_t>(ctx->tmp3), st));
  else
    SPH_TRY(launch_knn(ctx, (int)n, k, dist, 0, 1.0, bufp<int32_t>(ctx->list), bufp<double>(ctx->tmp1),
                       bufp<double>(ctx->tmp2), bufp<int32_t>(ctx->tmp3), st));
  SPH_TRY(launch_knn_unsort(ctx, bufp<int32_t>(ctx->list), bufp<double>(ctx->tmp1), bufp<double>(ctx->tmp2),
                            bufp<int32_t>(ctx->tmp3), (int)n, k, idx, dists, nontriv, new_sizes, st));
  return 0;
}

static int prep_list(sphb200_ctx* ctx, const int64_t* neighbor, int64_t n, int k, cudaStream_t st) {
  SPH_TRY(devbuf_reserve(ctx->list, (size_t)n * k * sizeof(int32_t)));
  return launch_convert_list(ctx, neighbor, (int)n, k, bufp<int32_t>(ctx->list), st);
}

// ---- A2-A6 -------------------------------------------------------------------------------------
int sphb200_hydro_update(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                         const double* mass, const double* sizes, const double* f_un, int n_species,
                         const double* particle_type, const double* T, const double* mu_array,
                         const double* gamma_array, const double* velocities, double* hydro_accel,
                         double* visc_accel, double* visc_heat, double* density, double* num_density,
                         double* f_un_neighbor, double* dust_density, void* stream) {
  if (!ctx || !neighbor || !points || !mass || !sizes || !particle_type || !T || !mu_array || !gamma_array ||
      !velocities || !hydro_accel || !visc_accel || !visc_heat || !density || !num_density || !dust_density)
    SPH_FAIL("sphb200_hydro_update: null argument");
  SPH_TRY(check_no_session(ctx, "sphb200_hydro_update"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(prep_list(ctx, neighbor, n, k, st));
  PackIn in = {points, velocities, mass, sizes, particle_type, T, mu_array, gamma_array};
  SPH_TRY(launch_pack(ctx, (int)n, in, st));
  const int32_t* list = bufp<int32_t>(ctx->list);
  const RowsRef rows = {nullptr, list, nullptr, nullptr};
  SPH_TRY(launch_density(ctx, (int)n, k, rows, density, dust_density, num_density, st));
  if (f_un && f_un_neighbor && n_species > 0)
    SPH_TRY(launch_species(ctx, (int)n, k, n_species, list, f_un, f_un_neighbor, st));
  SPH_TRY(launch_pi(ctx, (int)n, k, rows, st));
  SPH_TRY(launch_force(ctx, (int)n, k, rows, mass, hydro_accel, visc_accel, visc_heat, st));
  return 0;
}

// ---- A9 ----------------------------------------------------------------------------------------
int sphb200_crossing_time(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* velocities,
                          const double* sizes, const double* particle_type, double dt_0, double* result_host,
                          void* stream) {
  if (!ctx || !neighbor || !velocities || !sizes || !particle_type || !result_host) SPH_FAIL("sphb200_crossing_time: null argument");
  SPH_TRY(check_no_session(ctx, "sphb200_crossing_time"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(prep_list(ctx, neighbor, n, k, st));
  double* out = bufp<double>(ctx->scal) + SC_CT;
  const RowsRef rows = {nullptr, bufp<int32_t>(ctx->list), nullptr, nullptr};
  SPH_TRY(launch_crossing_time(ctx, (int)n, k, rows, velocities, sizes, particle_type, dt_0, out, nullptr, (uint32_t)n, st));
  CUDA_TRY(cudaMemcpyAsync(result_host, out, sizeof(double), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return 0;
}

// ---- A7-A8 -------------------------------------------------------------------------------------
int sphb200_loop_density(sphb200_ctx* ctx, int mode, int64_t n, int k, const int64_t* neighbor, const double* points,
                         const double* mass, const double* particle_type, const double* sizes,
                         const double* mu_array, double d, double* out, void* stream) {
  if (!ctx || !neighbor || !points || !mass || !out) SPH_FAIL("sphb200_loop_density: null argument");
  if (mode < 0 || mode > 2) SPH_FAIL("sphb200_loop_density: mode %d", mode);
  if ((mode == 0 && !particle_type) || (mode == 1 && (!particle_type || !sizes)) || (mode == 2 && !mu_array))
    SPH_FAIL("sphb200_loop_density: missing array for mode %d", mode);
  SPH_TRY(check_no_session(ctx, "sphb200_loop_density"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(prep_list(ctx, neighbor, n, k, st));
  LoopIn in = {};
  in.points = points; in.mass = mass; in.ptype = particle_type; in.sizes = sizes; in.mu = mu_array; in.d = d;
  return launch_loop_density(ctx, (int)n, k, bufp<int32_t>(ctx->list), in, mode, out, st);
}

int sphb200_del_pressure(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                         const double* mass, const double* particle_type, const double* E_internal,
                         const double* gamma_array, double d, double* grad_pressure, void* stream) {
  if (!ctx || !neighbor || !points || !mass || !particle_type || !E_internal || !gamma_array || !grad_pressure)
    SPH_FAIL("sphb200_del_pressure: null argument");
  SPH_TRY(check_no_session(ctx, "sphb200_del_pressure"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(prep_list(ctx, neighbor, n, k, st));
  LoopIn in = {};
  in.points = points; in.mass = mass; in.ptype = particle_type; in.E = E_internal; in.gamma = gamma_array; in.d = d;
  return launch_loop_del_pressure(ctx, (int)n, k, bufp<int32_t>(ctx->list), in, grad_pressure, st);
}

int sphb200_artificial_viscosity(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                                 const double* particle_type, const double* sizes, const double* mass,
                                 const double* densities, const double* velocities, const double* T,
                                 const double* gamma_array, const double* mu_array, double d, double* visc_accel,
                                 double* visc_heat, void* stream) {
  if (!ctx || !neighbor || !points || !particle_type || !mass || !densities || !velocities || !T || !gamma_array ||
      !mu_array || !visc_accel || !visc_heat)
    SPH_FAIL("sphb200_artificial_viscosity: null argument");
  (void)sizes;   // accepted and unused, as in the reference (nsc:788)
  SPH_TRY(check_no_session(ctx, "sphb200_artificial_viscosity"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(prep_list(ctx, neighbor, n, k, st));
  LoopIn in = {};
  in.points = points; in.vel = velocities; in.mass = mass; in.ptype = particle_type; in.T = T; in.mu = mu_array;
  in.gamma = gamma_array; in.dens = densities; in.d = d;
  return launch_loop_viscosity(ctx, (int)n, k, bufp<int32_t>(ctx->list), in, visc_accel, visc_heat, st);
}

// ---- F2 ----------------------------------------------------------------------------------------
int sphb200_gravity(sphb200_ctx* ctx, int64_t n, const double* mass, const double* points, double G, double softening,
                    double* accel, void* stream) {
  if (!ctx || !mass || !points || !accel) SPH_FAIL("sphb200_gravity: null argument");
  if (n <= 0 || n >= (int64_t)0x7fffffff) SPH_FAIL("sphb200_gravity: bad particle count %lld", (long long)n);
  CUDA_TRY(cudaSetDevice(ctx->device));
  return launch_gravity_direct(ctx, (int)n, points, mass, G, softening, accel, as_stream(stream));
}

// ---- F1 ----------------------------------------------------------------------------------------
int sphb200_net_impulse(sphb200_ctx* ctx, int64_t n, int k, const int64_t* neighbor, const double* points,
                        const double* mass, const double* sizes, const double* velocities,
                        const double* particle_type, const double* f_un, int n_species, const double* grain_mass_host,
                        const double* sigma_effective_host, double* accel_onto, double* accel_reaction, void* stream) {
  if (!ctx || !neighbor || !points || !mass || !sizes || !velocities || !particle_type || !f_un || !grain_mass_host ||
      !sigma_effective_host || !accel_onto || !accel_reaction)
    SPH_FAIL("sphb200_net_impulse: null argument");
  if (n_species <= 0 || n_species > SPHB200_MAX_SPECIES)
    SPH_FAIL("sphb200_net_impulse: n_species=%d outside 1..%d", n_species, SPHB200_MAX_SPECIES);
  SPH_TRY(check_no_session(ctx, "sphb200_net_impulse"));
  SPH_TRY(check_nk(n, k));
  CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t st = as_stream(stream);
  SPH_TRY(prep_list(ctx, neighbor, n, k, st));
  SPH_TRY(devbuf_reserve(ctx->tmp1, (size_t)n * sizeof(double)));
  SPH_TRY(devbuf_reserve(ctx->tmp2, (size_t)n * sizeof(double)));
  SPH_TRY(launch_grain_means(ctx, (int)n, n_species, f_un, grain_mass_host, sigma_effective_host, bufp<double>(ctx->tmp1),
                             bufp<double>(ctx->tmp2), st));
  LoopIn in = {};
  in.points = points; in.vel = velocities; in.mass = mass; in.sizes = sizes; in.ptype = particle_type;
  return launch_loop_net_impulse(ctx, (int)n, k, bufp<int32_t>(ctx->list), in, bufp<double>(ctx->tmp1),
                                 bufp<double>(ctx->tmp2), accel_onto, accel_reaction, st);
}

}  // extern "C"
