// Microbenchmark: FP64 tensor-core MMA (mma.sync.aligned.m8n8k4 .f64, the DMMA of sm_80+) against plain DFMA on this
// GPU, per SM sub-partition: dependent-issue latency and throughput in FMA per cycle.  Question behind it
// (BASELINE.json north_star): would phase B1 of the stage kernel — the contraction of the contravariant fluxes with
// the Q x Q derivative matrix, batched over cells — run faster on the tensor cores once the kernel is FP64 bound?
// One m8n8k4 is 8 x 8 x 4 = 256 FMA per warp instruction; one DFMA is 32.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void Dmma(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int ILP>
__global__ void DmmaChain(double *out, int iters, long long *cycles) {
  double c[ILP][2];
  for (int k = 0; k < ILP; ++k) { c[k][0] = threadIdx.x * 1e-3 + k; c[k][1] = 1 + k; }
  const double a = 1.0000001, b = 1e-9 * (threadIdx.x + 1);
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) Dmma(c[k][0], c[k][1], a, b);
  }
  long long t1 = clock64();
  double s = 0;
  for (int k = 0; k < ILP; ++k) s += c[k][0] + c[k][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int ILP>
__global__ void DfmaChain(double *out, int iters, long long *cycles) {
  double c[ILP];
  for (int k = 0; k < ILP; ++k) c[k] = threadIdx.x * 1e-3 + k;
  const double a = 1.0000001, b = 1e-9;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) c[k] = fma(c[k], a, b);
  }
  long long t1 = clock64();
  double s = 0;
  for (int k = 0; k < ILP; ++k) s += c[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main() {
  double *out; long long *cyc, h;
  cudaMalloc(&out, 1 << 24); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  for (int warps = 1; warps <= 8; warps *= 2) {   // warps per sub-partition (blocks of 4 x warps warps on one SM)
    DmmaChain<1><<<1, warps * 128>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double lat = (double)h / iters;
    DmmaChain<4><<<1, warps * 128>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double dmma_rate = warps * iters * 4.0 * 256 / h;
    DfmaChain<8><<<1, warps * 128>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double dfma_rate = warps * iters * 8.0 * 32 / h;
    printf("warps/SMSP=%d: DMMA m8n8k4 dependent latency %.1f cycles (1 warp chain), %.1f FMA/cycle/SMSP; DFMA %.1f FMA/cycle/SMSP\n",
           warps, lat, dmma_rate, dfma_rate);
  }
  // whole GPU, all SMs busy: FP64 TFLOP/s either way
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  float ms;
  DmmaChain<4><<<sms * 2, 512>>>(out, iters, cyc);
  cudaEventRecord(e0); DmmaChain<4><<<sms * 2, 512>>>(out, iters * 4, cyc); cudaEventRecord(e1); cudaEventSynchronize(e1);
  cudaEventElapsedTime(&ms, e0, e1);
  printf("DMMA, %d SMs x 32 warps: %.1f TFLOP/s FP64\n", sms, 2.0 * sms * 2 * 16 * (iters * 4.0) * 4 * 256 / (ms * 1e-3) / 1e12);
  DfmaChain<8><<<sms * 2, 512>>>(out, iters, cyc);
  cudaEventRecord(e0); DfmaChain<8><<<sms * 2, 512>>>(out, iters * 4, cyc); cudaEventRecord(e1); cudaEventSynchronize(e1);
  cudaEventElapsedTime(&ms, e0, e1);
  printf("DFMA, %d SMs x 32 warps: %.1f TFLOP/s FP64\n", sms, 2.0 * sms * 2 * 16 * (iters * 4.0) * 8 * 32 / (ms * 1e-3) / 1e12);
  return 0;
}
