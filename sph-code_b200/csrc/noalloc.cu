// sphb200_crossing_time_noalloc: the crossing time of /root/reference/sph/navier_stokes_cleaned.py:776-786 for
// callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  It has no counterpart in the reference.
//
// The ordinary entry point (sphb200_crossing_time, api.cu) keeps two pieces of scratch in buffers the context
// grows with cudaMalloc: the int32 copy of the caller's int64 neighbour array and the 8-byte slot of the atomic
// minimum.  Here both live in ONE file-scope device array of fixed size (below), so the call allocates nothing:
// no cudaMalloc / cudaFree, no context, no torch.  Every GPU that loads the library holds the array (16 MiB + 256 B)
// until the process exits.
//
// Sharing: every call on a device uses the same array.  The header (minimum slot, bad-index count) is reset by the
// first kernel of each call and the list is overwritten before it is read, so consecutive calls on one stream are
// independent; calls that OVERLAP on the GPU (different streams, or graphs replayed concurrently) race on it.
#include "common.cuh"
#include <float.h>
#include <mutex>

#define NA_MAX_ENTRIES (4 << 20)        // n * k <= 4 194 304 neighbour entries (k = 40: n <= 104 857 particles)
#define NA_HEADER_BYTES 256             // [0] u64 minimum slot, [8] int bad-index count
#define NA_G 8                          // lanes per target
#define NA_THREADS 256
#define NA_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_na_scratch[NA_HEADER_BYTES + (size_t)NA_MAX_ENTRIES * sizeof(int32_t)];

// In host code `g_na_scratch` is only a stand-in for the symbol: the device address is looked up once per device.
static void* g_na_ptr[NA_MAX_DEVICES] = {};
static std::mutex g_na_mutex;

static int na_scratch_ptr(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[device]) {
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_na_scratch));          // the calling thread's current device is `device`
    if (!p) SPH_FAIL("sphb200_crossing_time_noalloc: no device address for the scratch array on device %d", device);
    g_na_ptr[device] = p;
  }
  *out = reinterpret_cast<unsigned char*>(g_na_ptr[device]);
  return 0;
}

__global__ void na_reset_kernel(unsigned char* scratch) {
  *reinterpret_cast<unsigned long long*>(scratch) = 0xffffffffffffffffull;
  *reinterpret_cast<int*>(scratch + 8) = 0;
}

// int64 (N,k) neighbour array -> int32 list, same rules as convert_list_kernel (hydro.cu): INT64_MIN = padding of a
// ragged row, [-n, 0) = numpy negative indexing, anything else outside [0, n) is counted and skipped.
__global__ void __launch_bounds__(NA_THREADS)
na_convert_kernel(const int64_t* __restrict__ nb, int64_t total, int n, unsigned char* scratch) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= total) return;
  int32_t* list = reinterpret_cast<int32_t*>(scratch + NA_HEADER_BYTES);
  int64_t j = nb[t];
  if (j == INT64_MIN) { list[t] = -1; return; }
  if (j < 0 && j >= -(int64_t)n) j += n;
  if (j < 0 || j >= n) { atomicAdd(reinterpret_cast<int*>(scratch + 8), 1); j = -1; }
  list[t] = (int32_t)j;
}

// min over gas i of h_i / max_j |v_j - v_i|, zeros ignored (the arithmetic of CrossingPass, hydro.cu): 8 lanes per
// target, each takes every 8th neighbour; one atomicMin per warp on the ordered encoding of the double.
__global__ void __launch_bounds__(NA_THREADS)
na_crossing_kernel(int n, int k, const double* __restrict__ vel, const double* __restrict__ sizes,
                   const double* __restrict__ ptype, unsigned char* scratch) {
  const int32_t* list = reinterpret_cast<const int32_t*>(scratch + NA_HEADER_BYTES);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & (NA_G - 1);
  constexpr int TPW = 32 / NA_G;
  unsigned long long best = 0xffffffffffffffffull;
  const int wstride = gridDim.x * (NA_THREADS / 32) * TPW;
  for (int base = (blockIdx.x * (NA_THREADS / 32) + warp) * TPW; base < n; base += wstride) {
    const int i = base + lane / NA_G;
    const bool valid = i < n;
    double mx = 0.0; int nan = 0;
    if (valid) {
      const double vx = vel[3 * (int64_t)i], vy = vel[3 * (int64_t)i + 1], vz = vel[3 * (int64_t)i + 2];
      const int32_t* row = list + (int64_t)i * k;
      for (int s = sub; s < k; s += NA_G) {
        const int j = row[s];
        if (j < 0 || j >= n) continue;
        const double ax = vel[3 * (int64_t)j] - vx, ay = vel[3 * (int64_t)j + 1] - vy, az = vel[3 * (int64_t)j + 2] - vz;
        const double v2 = ax * ax + ay * ay + az * az;
        if (v2 != v2) nan = 1;                               // np.max propagates NaN
        mx = fmax(mx, v2);
      }
    }
#pragma unroll
    for (int d = NA_G / 2; d > 0; d >>= 1) {
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, d));
      nan |= __shfl_xor_sync(0xffffffffu, nan, d);
    }
    if (sub == 0 && valid && ptype[i] == 0.0) {
      double ct = nan ? 0.0 : sizes[i] / sqrt(mx);           // nan_to_num: NaN -> 0
      if (ct != ct) ct = 0.0;
      if (ct > DBL_MAX) ct = DBL_MAX;
      if (ct < -DBL_MAX) ct = -DBL_MAX;
      if (ct != 0.0) {
        const long long b = __double_as_longlong(ct);
        const unsigned long long u = (b >= 0) ? ((unsigned long long)b | 0x8000000000000000ull) : ~(unsigned long long)b;
        if (u < best) best = u;
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, d);
    if (o < best) best = o;
  }
  if (lane == 0 && best != 0xffffffffffffffffull) atomicMin(reinterpret_cast<unsigned long long*>(scratch), best);
}

__global__ void na_final_kernel(const unsigned char* scratch, double dt_0, double* result, int* bad_index) {
  const unsigned long long u = *reinterpret_cast<const unsigned long long*>(scratch);
  if (bad_index) *bad_index = *reinterpret_cast<const int*>(scratch + 8);
  if (u == 0xffffffffffffffffull) { *result = dt_0 / 10.0; return; }
  const long long b = (u & 0x8000000000000000ull) ? (long long)(u & 0x7fffffffffffffffull) : (long long)~u;
  *result = __longlong_as_double(b) + 0.0001;
}

extern "C" int sphb200_crossing_time_noalloc(int device, int64_t n, int k, const int64_t* neighbor, const double* velocities,
                                             const double* sizes, const double* particle_type, double dt_0,
                                             double* result_dev, int* bad_index_dev, void* stream) {
  if (!neighbor || !velocities || !sizes || !particle_type || !result_dev) SPH_FAIL("sphb200_crossing_time_noalloc: null argument");
  if (device < 0 || device >= NA_MAX_DEVICES) SPH_FAIL("sphb200_crossing_time_noalloc: device %d outside [0, %d)", device, NA_MAX_DEVICES);
  if (n < 1 || k < 1) SPH_FAIL("sphb200_crossing_time_noalloc: n=%lld k=%d must be positive", (long long)n, k);
  if (n > (int64_t)NA_MAX_ENTRIES || n * (int64_t)k > (int64_t)NA_MAX_ENTRIES)
    SPH_FAIL("sphb200_crossing_time_noalloc: n*k = %lld exceeds the fixed scratch of %d neighbour entries "
             "(use sphb200_crossing_time for larger inputs)", (long long)(n * (int64_t)k), (int)NA_MAX_ENTRIES);
  CUDA_TRY(cudaSetDevice(device));
  unsigned char* scratch = nullptr;
  SPH_TRY(na_scratch_ptr(device, &scratch));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t total = n * (int64_t)k;
  na_reset_kernel<<<1, 1, 0, st>>>(scratch);
  na_convert_kernel<<<(unsigned)((total + NA_THREADS - 1) / NA_THREADS), NA_THREADS, 0, st>>>(neighbor, total, (int)n, scratch);
  const int per_block = NA_THREADS / NA_G;
  int64_t blocks = (n + per_block - 1) / per_block;
  if (blocks > 148 * 16) blocks = 148 * 16;                   // grid-stride beyond that
  na_crossing_kernel<<<(unsigned)blocks, NA_THREADS, 0, st>>>((int)n, k, velocities, sizes, particle_type, scratch);
  na_final_kernel<<<1, 1, 0, st>>>(scratch, dt_0, result_dev, bad_index_dev);
  CUDA_TRY(cudaGetLastError());
  return 0;
}
