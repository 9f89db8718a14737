// Microbenchmark: what the FP64 pipe of one B200 SM sub-partition sustains from FEW warps — the shape of the fused kernels
// (single-warp CTAs, 12 per SM = 3 warps per scheduler) — and how integer instructions issued beside it cost.
//   op mix  F: DFMA only   A: DADD only   M: DMUL only   X: DFMA, DMUL, DADD in turn   I: X with one IMAD per FP64 instruction
//   ILP     independent dependent-chains per thread (4 or 8)
// Prints cycles per FP64 warp instruction and scheduler (peak of the pipe: 2.0 = 16 lanes per scheduler and clock).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_mix fp64_mix.cu && ./fp64_mix
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double x_dfma(double a, double b, double c) { double r; asm volatile("fma.rn.f64 %0, %1, %2, %3;" : "=d"(r) : "d"(a), "d"(b), "d"(c)); return r; }
__device__ __forceinline__ double x_dadd(double a, double b) { double r; asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(r) : "d"(a), "d"(b)); return r; }
__device__ __forceinline__ double x_dmul(double a, double b) { double r; asm volatile("mul.rn.f64 %0, %1, %2;" : "=d"(r) : "d"(a), "d"(b)); return r; }
__device__ __forceinline__ int x_imad(int a, int b, int c) { int r; asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }

template <char MIX, int ILP>
__global__ void __launch_bounds__(32) k(double* out, int iters) {
  double a[ILP];
  int q[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) { a[i] = threadIdx.x * 1e-3 + i; q[i] = threadIdx.x + i; }
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        if (MIX == 'F') a[i] = x_dfma(a[i], b, c);
        else if (MIX == 'A') a[i] = x_dadd(a[i], c);
        else if (MIX == 'M') a[i] = x_dmul(a[i], b);
        else {
          a[i] = r == 0 ? x_dfma(a[i], b, c) : r == 1 ? x_dmul(a[i], b) : x_dadd(a[i], c);
          if (MIX == 'I') q[i] = x_imad(q[i], 3, it);
        }
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += a[i] + q[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <char MIX, int ILP> void run(double* out, int ctas_per_sm, int sms) {
  const int iters = 1 << 13, grid = sms * ctas_per_sm;
  k<MIX, ILP><<<grid, 32>>>(out, 64);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0); k<MIX, ILP><<<grid, 32>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  int mhz = 0; cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, 0); mhz /= 1000;
  const double fp64_per_sched = 3.0 * ILP * iters * (ctas_per_sm / 4.0);       // warp instructions per scheduler (4 per SM)
  const double cycles = best * 1e-3 * mhz * 1e6;
  printf("mix %c  ILP %d  %2d warps/SM (%4.1f per scheduler): %7.3f ms  %5.2f cycles per FP64 warp instruction and scheduler (at %d MHz)\n",
         MIX, ILP, ctas_per_sm, ctas_per_sm / 4.0, best, cycles / fp64_per_sched, mhz);
}

template <char MIX> void sweep(double* out, int sms) {
  for (int c : {4, 8, 12, 16, 32}) run<MIX, 4>(out, c, sms);
  for (int c : {4, 8, 12, 16}) run<MIX, 8>(out, c, sms);
}

int main() {
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  printf("%s, %d SMs, sm_%d%d\n", pr.name, pr.multiProcessorCount, pr.major, pr.minor);
  double* out; cudaMalloc(&out, sizeof(double) * 148 * 32 * 32 * 2);
  sweep<'F'>(out, pr.multiProcessorCount);
  sweep<'A'>(out, pr.multiProcessorCount);
  sweep<'M'>(out, pr.multiProcessorCount);
  sweep<'X'>(out, pr.multiProcessorCount);
  sweep<'I'>(out, pr.multiProcessorCount);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
