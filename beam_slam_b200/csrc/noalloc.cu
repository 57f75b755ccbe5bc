// bsnls_reprojection_cost_noalloc — the robust reprojection cost of a set of observations, computed without
// allocating anything.
//
// Who needs this: code that runs where device memory cannot be requested — a real-time optimisation thread whose
// latency budget forbids the implicit device synchronisation cudaMalloc/cudaFree perform, a process that handed its
// whole GPU memory budget to a fixed arena at start-up, or a region being recorded into a stream capture, where
// allocation calls are illegal. Such a caller already owns its input and output buffers; the only thing the cost
// kernel needs beyond them is room for one partial sum per thread block. Here that room is a fixed array inside
// the library image.
//
// Consequences a user must know:
//  * Limit: at most BSNLS_NOALLOC_MAX_OBSERVATIONS observations per call (the array holds one double per 256
//    observations: 16384 doubles = 128 KiB). Larger requests are refused with BSNLS_ERR_INVALID.
//  * Every GPU that loads this library carries those 128 KiB until the process exits.
//  * The first call on each device looks up the array's device address (cudaGetSymbolAddress). Make that first
//    call before starting any stream capture on that device.
//  * All calls on a device share the one array and it keeps whatever the previous call left in it. Nothing has to
//    be cleared: a call overwrites exactly the entries it later reads. But two calls whose kernels can overlap on
//    the GPU (different streams, no ordering between them) race on it — order them yourself.
//  * The call only enqueues work on `stream`; the result is in *d_cost once the stream has caught up.
// The arithmetic is the candidate-cost evaluation of the solver (reprojection_cost_only in device_math.cuh):
// 0.5 * rho(|r|^2) per observation, block sums by a fixed tree, then one block adds the block sums in a fixed order.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "../../include/bsnls.h"
#include "device_math.cuh"

namespace {

constexpr int kBlock = 256;
constexpr int kMaxPartials = BSNLS_NOALLOC_MAX_OBSERVATIONS / kBlock;

// The scratch. In host code this name is only a handle for the runtime, not an address a kernel could use.
static __device__ __align__(256) double g_noalloc_partials[kMaxPartials];

constexpr int kMaxDevices = 64;
double* g_partials_addr[kMaxDevices] = {};
std::mutex g_setup_mutex;

__global__ void __launch_bounds__(kBlock) noalloc_cost_kernel(const double* __restrict__ x, const int* __restrict__ q_off,
                                                              const int* __restrict__ t_off, const int* __restrict__ p_off,
                                                              const int* __restrict__ cam_index, const double* __restrict__ pixel,
                                                              const bsnls::CameraModel* __restrict__ cams, int n,
                                                              double* __restrict__ partials) {
  __shared__ double sm[kBlock];
  const int o = blockIdx.x * kBlock + threadIdx.x;
  double cost = 0.0;
  if (o < n) {
    const bsnls::CameraModel cam = cams[cam_index[o]];
    const double px[2] = {pixel[2 * (size_t)o], pixel[2 * (size_t)o + 1]};
    const int qo = q_off[o], to = t_off[o], po = p_off[o];
    const double q[4] = {x[qo], x[qo + 1], x[qo + 2], x[qo + 3]};
    const double t[3] = {x[to], x[to + 1], x[to + 2]};
    const double P[3] = {x[po], x[po + 1], x[po + 2]};
    cost = bsnls::reprojection_cost_only(cam, px, q, t, P);
  }
  sm[threadIdx.x] = cost; __syncthreads();
  for (int s = kBlock / 2; s > 0; s >>= 1) { if (threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) partials[blockIdx.x] = sm[0];
}

__global__ void __launch_bounds__(kBlock) noalloc_sum_kernel(const double* __restrict__ partials, int n, double* __restrict__ out) {
  __shared__ double sm[kBlock];
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += kBlock) v += partials[i];
  sm[threadIdx.x] = v; __syncthreads();
  for (int s = kBlock / 2; s > 0; s >>= 1) { if (threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) *out = sm[0];
}

int Fail(char* msg, int cap, int code, const char* text, const char* detail = "") {
  if (msg && cap > 0) std::snprintf(msg, (size_t)cap, "%s%s", text, detail);
  return code;
}

}  // namespace

extern "C" int bsnls_reprojection_cost_noalloc(const double* d_values, const int32_t* d_q_offset,
                                               const int32_t* d_t_offset, const int32_t* d_p_offset,
                                               const int32_t* d_camera_index, const double* d_pixel,
                                               const double* d_cameras, int32_t num_observations, double* d_cost,
                                               void* stream, char* error_message, int32_t error_capacity) {
  if (!d_values || !d_q_offset || !d_t_offset || !d_p_offset || !d_camera_index || !d_pixel || !d_cameras || !d_cost)
    return Fail(error_message, error_capacity, BSNLS_ERR_INVALID, "bsnls_reprojection_cost_noalloc: null pointer argument");
  if (num_observations <= 0)
    return Fail(error_message, error_capacity, BSNLS_ERR_INVALID, "bsnls_reprojection_cost_noalloc: num_observations must be positive");
  if (num_observations > BSNLS_NOALLOC_MAX_OBSERVATIONS) {
    char d[96]; std::snprintf(d, sizeof(d), " (%d requested, limit %d)", num_observations, (int)BSNLS_NOALLOC_MAX_OBSERVATIONS);
    return Fail(error_message, error_capacity, BSNLS_ERR_INVALID,
                "bsnls_reprojection_cost_noalloc: too many observations for the fixed scratch array", d);
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return Fail(error_message, error_capacity, BSNLS_ERR_CUDA, "cudaGetDevice: ", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return Fail(error_message, error_capacity, BSNLS_ERR_UNSUPPORTED, "device ordinal beyond the per-device table");
  double* partials = nullptr;
  {
    // once per device, with that device current: turn the symbol into the address kernels can use
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_partials_addr[dev]) {
      double* p = nullptr;
      e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_partials);
      if (e != cudaSuccess || !p) return Fail(error_message, error_capacity, BSNLS_ERR_CUDA, "cudaGetSymbolAddress(scratch): ", cudaGetErrorString(e));
      g_partials_addr[dev] = p;
    }
    partials = g_partials_addr[dev];
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int nblocks = (num_observations + kBlock - 1) / kBlock;
  noalloc_cost_kernel<<<nblocks, kBlock, 0, s>>>(d_values, d_q_offset, d_t_offset, d_p_offset, d_camera_index, d_pixel,
                                                 reinterpret_cast<const bsnls::CameraModel*>(d_cameras), num_observations, partials);
  noalloc_sum_kernel<<<1, kBlock, 0, s>>>(partials, nblocks, d_cost);
  e = cudaGetLastError();
  if (e != cudaSuccess) return Fail(error_message, error_capacity, BSNLS_ERR_CUDA, "kernel launch: ", cudaGetErrorString(e));
  return BSNLS_OK;
}
