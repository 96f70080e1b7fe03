// Does an FP64 warp instruction hold the dispatch port of its SM sub-partition for one cycle or for both cycles it
// occupies the FP64 pipe?  8 warps per sub-partition run independent DFMA chains mixed with K independent IMAD per
// DFMA (K = 0..3).  If dispatch is one cycle, cycles per DFMA stay at 2 until K + 1 > 2; if it is two, they grow as 2 + K.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_issue_mix fp64_issue_mix.cu && ./fp64_issue_mix
#include <cstdio>
#include <cuda_runtime.h>

template <int K>
__global__ void Mix(double *out, int *iout, long long *cycles, int iters) {
  double a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
  const double m = 1.0000001, c = 1e-9;
  int i0 = threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3, i4 = i0 + 4, i5 = i0 + 5, i6 = i0 + 6, i7 = i0 + 7, i8 = i0 + 8, i9 = i0 + 9, ia = i0 + 10, ib = i0 + 11;
  const int im = 3, ic = 7;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a0) : "d"(m), "d"(c));
      if (K >= 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i0) : "r"(im), "r"(ic));
      if (K >= 2) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i1) : "r"(im), "r"(ic));
      if (K >= 3) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i2) : "r"(im), "r"(ic));
      asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a1) : "d"(m), "d"(c));
      if (K >= 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i3) : "r"(im), "r"(ic));
      if (K >= 2) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i4) : "r"(im), "r"(ic));
      if (K >= 3) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i5) : "r"(im), "r"(ic));
      asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a2) : "d"(m), "d"(c));
      if (K >= 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i6) : "r"(im), "r"(ic));
      if (K >= 2) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i7) : "r"(im), "r"(ic));
      if (K >= 3) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i8) : "r"(im), "r"(ic));
      asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a3) : "d"(m), "d"(c));
      if (K >= 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(i9) : "r"(im), "r"(ic));
      if (K >= 2) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(ia) : "r"(im), "r"(ic));
      if (K >= 3) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(ib) : "r"(im), "r"(ic));
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
  iout[blockIdx.x * blockDim.x + threadIdx.x] = i0 + i1 + i2 + i3 + i4 + i5 + i6 + i7 + i8 + i9 + ia + ib;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int K>
void Run(double *out, int *iout, long long *cyc, int warps_per_partition) {
  const int iters = 4000, threads = warps_per_partition * 4 * 32;
  Mix<K><<<148, threads>>>(out, iout, cyc, iters);
  cudaDeviceSynchronize();
  Mix<K><<<148, threads>>>(out, iout, cyc, iters);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
  const double dfma_per_partition = (double)iters * 16 * warps_per_partition;
  std::printf("warps/SMSP %d, %d IMAD per DFMA: %.2f cycles per DFMA warp-instruction per sub-partition\n", warps_per_partition, K,
      c / dfma_per_partition);
}

int main() {
  double *out; int *iout; long long *cyc;
  cudaMalloc(&out, 148 * 1024 * sizeof(double)); cudaMalloc(&iout, 148 * 1024 * sizeof(int)); cudaMalloc(&cyc, sizeof(long long));
  for (int w : {2, 8}) { Run<0>(out, iout, cyc, w); Run<1>(out, iout, cyc, w); Run<2>(out, iout, cyc, w); Run<3>(out, iout, cyc, w); }
  return 0;
}
