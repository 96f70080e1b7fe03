// Microbenchmark: FP64 dependent-issue latency and throughput per SM sub-partition on this GPU.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void chain(double *out, int iters, int ilp_mode, long long *cycles) {
  double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
  double b = 1.0000001, c = 1e-9;
  long long t0 = clock64();
  if (ilp_mode == 1) {
    for (int i = 0; i < iters; ++i) { a0 = fma(a0, b, c); a0 = fma(a0, b, c); a0 = fma(a0, b, c); a0 = fma(a0, b, c); }
  } else if (ilp_mode == 2) {
    for (int i = 0; i < iters; ++i) { a0 = fma(a0, b, c); a1 = fma(a1, b, c); a0 = fma(a0, b, c); a1 = fma(a1, b, c); }
  } else {
    for (int i = 0; i < iters; ++i) { a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c); }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
__global__ void divchain(double *out, int iters, long long *cycles) {
  double a = 1.5 + threadIdx.x * 1e-3;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) { a = 1.0 / a + 0.5; }
  long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) *cycles = t1 - t0;
}
__global__ void sqrtchain(double *out, int iters, long long *cycles) {
  double a = 1.5 + threadIdx.x * 1e-3;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) { a = sqrt(a) + 0.5; }
  long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) *cycles = t1 - t0;
}
int main() {
  double *out; long long *cyc, h;
  cudaMalloc(&out, 1 << 24); cudaMalloc(&cyc, 8);
  int iters = 4096;
  for (int mode = 1; mode <= 3; ++mode)
    for (int warps = 1; warps <= 16; warps *= 2) {
      chain<<<1, warps * 32 * 4, 0>>>(out, iters, mode, cyc);  // warps per SMSP (4 SMSPs)
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      printf("dfma ilp_mode=%d warps/SMSP=%2d: %.2f cycles per DFMA per warp, %.3f DFMA/cycle/SMSP\n", mode, warps,
             (double)h / (iters * 4), warps * iters * 4.0 / h);
    }
  divchain<<<1, 32>>>(out, 1024, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("dependent (1/a + 0.5): %.1f cycles\n", (double)h / 1024);
  sqrtchain<<<1, 32>>>(out, 1024, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("dependent (sqrt(a) + 0.5): %.1f cycles\n", (double)h / 1024);
  for (int warps = 1; warps <= 8; warps *= 2) {
    divchain<<<1, warps * 128>>>(out, 1024, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("div chain with %d warps/SMSP: %.1f cycles per iteration per warp\n", warps, (double)h / 1024);
  }
  return 0;
}
