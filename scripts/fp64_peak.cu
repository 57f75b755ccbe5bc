// Micro-benchmark: FP64 peak of the vector pipe (DFMA) and of the tensor pipe (DMMA m8n8k4) on this GPU.
#include <cuda_runtime.h>
#include <cstdio>
__global__ void dfma_kernel(double* out, int iters) {
  double a[16]; for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fma(a[i], b, c);
  double s = 0; for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void dmma_kernel(double* out, int iters) {
  double d[8][2]; for (int i = 0; i < 8; ++i) { d[i][0] = i; d[i][1] = -i; }
  const double a = 1.0 + threadIdx.x * 1e-6, b = 1e-3;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d[i][0]), "+d"(d[i][1]) : "d"(a), "d"(b));
  double s = 0; for (int i = 0; i < 8; ++i) s += d[i][0] + d[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// larger FP64 tensor shapes (PTX ISA: sm_90+): fewer instructions and fragment loads per flop than m8n8k4
__global__ void dmma_16816_kernel(double* out, int iters) {
  double d[4][4]; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) d[i][j] = i - j;
  double a[8]; for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-6 + i * 1e-7;
  double b[4] = {1e-3, 2e-3, 3e-3, 4e-3};
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                   : "+d"(d[i][0]), "+d"(d[i][1]), "+d"(d[i][2]), "+d"(d[i][3])
                   : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
  double s = 0; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) s += d[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void dmma_1684_kernel(double* out, int iters) {
  double d[4][4]; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) d[i][j] = i - j;
  double a[2] = {1.0 + threadIdx.x * 1e-6, 1.0 - threadIdx.x * 1e-6};
  double b = 1e-3;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                   : "+d"(d[i][0]), "+d"(d[i][1]), "+d"(d[i][2]), "+d"(d[i][3]) : "d"(a[0]), "d"(a[1]), "d"(b));
  double s = 0; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) s += d[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0); dfma_kernel<<<sms * 8, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (rep) printf("DFMA: %.2f TFLOP/s (%d SMs, %.3f ms)\n", 2.0 * 16 * iters * sms * 8 * 256 / (ms * 1e-3) / 1e12, sms, ms);
    cudaEventRecord(e0); dmma_kernel<<<sms * 8, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep) printf("DMMA m8n8k4: %.2f TFLOP/s (%.3f ms)\n", 2.0 * 256 * 8 * iters * (double)sms * 8 * 8 / (ms * 1e-3) / 1e12, ms);
    cudaEventRecord(e0); dmma_16816_kernel<<<sms * 8, 256>>>(out, iters / 4); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep) printf("DMMA m16n8k16: %.2f TFLOP/s (%.3f ms)\n", 2.0 * 16 * 8 * 16 * 4 * (iters / 4) * (double)sms * 8 * 8 / (ms * 1e-3) / 1e12, ms);
    cudaEventRecord(e0); dmma_1684_kernel<<<sms * 8, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep) printf("DMMA m16n8k4: %.2f TFLOP/s (%.3f ms)\n", 2.0 * 16 * 8 * 4 * 4 * iters * (double)sms * 8 * 8 / (ms * 1e-3) / 1e12, ms);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
