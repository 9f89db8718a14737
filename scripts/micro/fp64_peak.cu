// Microbenchmark: FP64 pipe peak of one B200 — the denominator SURVEY 8d asks for before any FP64 compute-roofline
// fraction is quoted (MEASURED_PEAKS.json holds HBM and bf16 only).
//   MODE 0: 16 independent scalar DFMA chains per thread (fma.rn.f64)
//   MODE 1: 8 independent DMMA accumulators per warp (mma.sync.m8n8k4.f64: 512 flop per warp instruction)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu && ./fp64_peak
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double dfma(double a, double b, double c) {
  double r; asm volatile("fma.rn.f64 %0, %1, %2, %3;" : "=d"(r) : "d"(a), "d"(b), "d"(c)); return r;
}
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void __launch_bounds__(256) k(double* out, int iters) {
  double a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = dfma(a[i], b, c);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) dmma(a[2 * i], a[2 * i + 1], b, c);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE> void run(const char* name, double* out, int ctas_per_sm, int sms, int mhz) {
  const int iters = 1 << 14, grid = sms * ctas_per_sm, tpb = 256;
  k<MODE><<<grid, tpb>>>(out, 64);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0); k<MODE><<<grid, tpb>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  // MODE 0: 16 DFMA x 2 flop per thread-step; MODE 1: 8 DMMA x 512 flop per warp-step
  const double flop = MODE == 0 ? 32.0 * iters * (double)grid * tpb : 8.0 * 512.0 * iters * (double)grid * (tpb / 32);
  const double tf = flop / (best * 1e-3) / 1e12;
  printf("%-34s %d CTAs/SM: %8.3f ms  %7.2f TFLOP/s  = %6.1f flop/clk/SM at the %d MHz max SM clock\n", name, ctas_per_sm, best, tf,
         tf * 1e12 / sms / (mhz * 1e6), mhz);
}

int main() {
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  int mhz = 0; cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, 0); mhz /= 1000;
  printf("%s, %d SMs, sm_%d%d\n", pr.name, pr.multiProcessorCount, pr.major, pr.minor);
  double* out; cudaMalloc(&out, sizeof(double) * 148 * 8 * 256 * 2);
  for (int c = 1; c <= 4; c *= 2) run<0>("scalar DFMA (16 chains/thread)", out, c, pr.multiProcessorCount, mhz);
  for (int c = 1; c <= 4; c *= 2) run<1>("DMMA m8n8k4 (8 accumulators/warp)", out, c, pr.multiProcessorCount, mhz);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
