// F2: self-gravity.  The reference's grav_force_calculation_new
// (/root/reference/sph/navier_stokes_cleaned.py:252-415, called at sph/code_running.py:188,448 and summed into
// total_accel at :477) bins the cloud with a KD bisection and lets every bin act on every particle through the
// Plummer-softened monopole law  a_i = G M (x_c - x_i) / (|x_c - x_i|^2 + eps^2)^(3/2),  eps = median(sizes)
// (:359, :392).  The routine itself cannot run (it walks attributes of scipy's removed pure-Python KDTree), and
// its bin structure is an approximation of the direct sum, not a specification; what is built here is that law
// summed DIRECTLY over all particles in fp64 (exact limit of the reference's approximation), as a tiled
// all-pairs kernel: one target per thread, sources staged through shared memory in tiles of 256.  O(N^2): meant
// for N up to a few 10^5 (10^5 particles: 10^10 pairs, ~10 ms on a B200).
#include "common.cuh"
#include "kernels.cuh"

#define GR_THREADS 256

__global__ void __launch_bounds__(GR_THREADS)
gravity_direct_kernel(int n, const double* __restrict__ pos, const double* __restrict__ mass, double G, double eps2,
                      double* __restrict__ out) {
  __shared__ double4 src[GR_THREADS];
  const int i = blockIdx.x * GR_THREADS + threadIdx.x;
  double x = 0, y = 0, z = 0;
  if (i < n) { x = pos[3 * (int64_t)i]; y = pos[3 * (int64_t)i + 1]; z = pos[3 * (int64_t)i + 2]; }
  double ax = 0, ay = 0, az = 0;
  for (int base = 0; base < n; base += GR_THREADS) {
    const int j = base + threadIdx.x;
    src[threadIdx.x] = (j < n) ? make_double4(pos[3 * (int64_t)j], pos[3 * (int64_t)j + 1], pos[3 * (int64_t)j + 2], mass[j])
                               : make_double4(0, 0, 0, 0);
    __syncthreads();
    const int m = min(GR_THREADS, n - base);
#pragma unroll 4
    for (int t = 0; t < m; ++t) {
      const double4 s = src[t];
      const double dx = s.x - x, dy = s.y - y, dz = s.z - z;
      const double r2 = dx * dx + dy * dy + dz * dz + eps2;
      const double inv = rsqrt(r2);
      const double w = s.w * (inv * inv * inv);          // m_j / (r^2 + eps^2)^(3/2); the self term has dx = 0
      ax += w * dx; ay += w * dy; az += w * dz;
    }
    __syncthreads();
  }
  if (i < n) { out[3 * (int64_t)i] = G * ax; out[3 * (int64_t)i + 1] = G * ay; out[3 * (int64_t)i + 2] = G * az; }
}

int launch_gravity_direct(sphb200_ctx* ctx, int n, const double* pos, const double* mass, double G, double eps, double* out,
                          cudaStream_t st) {
  gravity_direct_kernel<<<ceil_div(n, GR_THREADS), GR_THREADS, 0, st>>>(n, pos, mass, G, eps * eps, out);
  ctx->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}
