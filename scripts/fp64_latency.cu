// Dependent-issue latencies that bound the sequential part of the Cholesky kernels (pivot chain):
// DFMA, DMUL, rsqrt(double), 1/x (double), sqrt(double), shared-memory round trip, __syncthreads with 8 warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu ; run on one B200.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP> __global__ void chain(double* out, long long* cyc, double seed, int iters) {
  __shared__ double sm[256];
  double x = seed + threadIdx.x * 1e-9, y = 1.0000001;
  sm[threadIdx.x] = x;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (OP == 0) x = fma(x, y, 1e-9);
    else if (OP == 1) x = x * y;
    else if (OP == 2) x = rsqrt(x) + 1.0;
    else if (OP == 3) x = 1.0 / x + 1.0;
    else if (OP == 4) x = sqrt(x) + 1.0;
    else if (OP == 5) { sm[threadIdx.x] = x; __syncwarp(); x = sm[threadIdx.x ^ 1] + 1e-9; __syncwarp(); }
    else if (OP == 6) { sm[threadIdx.x] = x; __syncthreads(); x = sm[(threadIdx.x + 32) & 255] + 1e-9; __syncthreads(); }
    else if (OP == 7) x = __shfl_xor_sync(0xffffffffu, x, 1) + 1e-9;
    else if (OP == 8) { float f = rsqrtf((float)x); double r = (double)f; const double hx = 0.5 * x; r = r * (1.5 - hx * r * r); r = r * (1.5 - hx * r * r); x = r + 1.0; }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int threads) {
  double* out; long long* cyc; cudaMalloc(&out, 8 * 1024); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  chain<OP><<<1, threads>>>(out, cyc, 1.5, iters); chain<OP><<<1, threads>>>(out, cyc, 1.5, iters);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-44s %3d threads: %7.1f cycles / iteration\n", name, threads, (double)h / iters);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  run<0>("DFMA dependent", 32); run<1>("DMUL dependent", 32); run<2>("rsqrt(double) + DADD", 32); run<3>("1/x (double) + DADD", 32);
  run<4>("sqrt(double) + DADD", 32); run<8>("rsqrtf + 2 Newton (double) + DADD", 32); run<7>("SHFL 64-bit + DADD", 32);
  run<5>("smem store, syncwarp, load, syncwarp + DADD", 32); run<6>("smem store, barrier, load, barrier + DADD", 256);
  run<0>("DFMA dependent, 8 warps", 256); run<2>("rsqrt(double) + DADD, 8 warps", 256);
  return 0;
}
