// Micro-benchmark: latency and single-warp issue interval of DFMA, DMMA m8n8k4, rsqrt(double) and LDS.64 on this GPU
// (what bounds the pivot chain of the tile Cholesky panel). One CTA; cycles from clock64 of lane 0.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build_tmp/fp64_latency scripts/fp64_latency.cu
#include <cuda_runtime.h>
#include <cstdio>
template <int CHAINS>
__global__ void dfma_lat(double* out, long long* cyc, int iters) {
  double a[CHAINS];
  for (int i = 0; i < CHAINS; ++i) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) a[i] = fma(a[i], b, c);
  const long long t1 = clock64();
  double s = 0; for (int i = 0; i < CHAINS; ++i) s += a[i];
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int CHAINS>
__global__ void dmma_lat(double* out, long long* cyc, int iters) {
  double d[CHAINS][2];
  for (int i = 0; i < CHAINS; ++i) { d[i][0] = i; d[i][1] = -i; }
  const double a = 1.0 + threadIdx.x * 1e-6, b = 1e-3;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d[i][0]), "+d"(d[i][1]) : "d"(a), "d"(b));
  const long long t1 = clock64();
  double s = 0; for (int i = 0; i < CHAINS; ++i) s += d[i][0] + d[i][1];
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void rsqrt_lat(double* out, long long* cyc, int iters) {
  double a = 2.0 + threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) a = rsqrt(a) + 1.5;
  const long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lds_lat(double* out, long long* cyc, int iters) {
  __shared__ int nxt[256];
  nxt[threadIdx.x] = (threadIdx.x + 33) & 255;
  __syncthreads();
  int p = threadIdx.x;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) p = nxt[p];
  const long long t1 = clock64();
  out[threadIdx.x] = p;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void bar_lat(double* out, long long* cyc, int iters) {
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) __syncthreads();
  const long long t1 = clock64();
  out[threadIdx.x] = 0;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 8 * 1024); cudaMalloc(&cyc, 8);
  const int iters = 4096; long long c;
#define RUN(name, kern, threads, per_iter)                                                                     \
  for (int rep = 0; rep < 2; ++rep) { kern<<<1, threads>>>(out, cyc, iters); cudaDeviceSynchronize(); }         \
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);                                                               \
  printf("%-44s %3d threads: %7.2f cycles per instruction\n", name, threads, (double)c / iters / (per_iter));
  RUN("DFMA dependent chain (latency)", dfma_lat<1>, 32, 1)
  RUN("DFMA 2 chains", dfma_lat<2>, 32, 2)
  RUN("DFMA 4 chains", dfma_lat<4>, 32, 4)
  RUN("DFMA 8 chains", dfma_lat<8>, 32, 8)
  RUN("DFMA 16 chains, 1 warp", dfma_lat<16>, 32, 16)
  RUN("DFMA 16 chains, 4 warps (1 / SMSP)", dfma_lat<16>, 128, 16)
  RUN("DFMA 16 chains, 8 warps (2 / SMSP)", dfma_lat<16>, 256, 16)
  RUN("DFMA 16 chains, 16 warps (4 / SMSP)", dfma_lat<16>, 512, 16)
  RUN("DMMA dependent chain (latency)", dmma_lat<1>, 32, 1)
  RUN("DMMA 2 chains", dmma_lat<2>, 32, 2)
  RUN("DMMA 4 chains", dmma_lat<4>, 32, 4)
  RUN("DMMA 8 chains, 1 warp", dmma_lat<8>, 32, 8)
  RUN("DMMA 8 chains, 4 warps (1 / SMSP)", dmma_lat<8>, 128, 8)
  RUN("DMMA 8 chains, 8 warps (2 / SMSP)", dmma_lat<8>, 256, 8)
  RUN("DMMA 8 chains, 16 warps (4 / SMSP)", dmma_lat<8>, 512, 8)
  RUN("rsqrt(double) + add, dependent", rsqrt_lat, 32, 1)
  RUN("LDS.32 dependent", lds_lat, 32, 1)
  RUN("__syncthreads, 256 threads", bar_lat, 256, 1)
  printf("cuda: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
