// Microbenchmark: issue / pipe throughput of scalar FFMA against packed FFMA2 (fma.rn.f32x2) on sm_100a, with and
// without integer instructions interleaved (does FFMA2 free issue slots for the non-FP work of an issue-bound kernel?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_probe ffma2_probe.cu && ./ffma2_probe
// Measured on B200 (round 1): 16 FFMA 72.1 TFLOP/s, 8 FFMA2 73.5, 16 FFMA + 8 IADD 69.8, 8 FFMA2 + 8 IADD 72.6 — scalar
// FFMA already reaches the FP32 pipe peak (128 FMA/clk/SM) and integer adds ride along almost for free, so packing the
// fp32 model math into FFMA2 would not speed up the issue-bound cfg-3 kernels; not pursued.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ unsigned iadd(unsigned a, unsigned b) { unsigned r; asm volatile("add.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }

template <int MODE>   // 0: 16 scalar FFMA per step; 1: 8 FFMA2 per step (same flops); 2/3: the same plus 8 integer adds per step
__global__ void k(float* out, int iters) {
  float a[16]; u64 p[8]; unsigned q[8];
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
  for (int i = 0; i < 8; ++i) { float2 v = make_float2(a[2 * i], a[2 * i + 1]); p[i] = *reinterpret_cast<u64*>(&v); q[i] = threadIdx.x + i; }
  const float b = 1.0001f, c = 1e-6f;
  float2 bb = make_float2(b, b), cc = make_float2(c, c);
  const u64 b2 = *reinterpret_cast<u64*>(&bb), c2 = *reinterpret_cast<u64*>(&cc);
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fma1(a[i], b, c);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
    }
    if (MODE >= 2) {
#pragma unroll
      for (int i = 0; i < 8; ++i) q[i] = iadd(q[i], 3u);
    }
  }
  float s = 0; unsigned t = 0;
  for (int i = 0; i < 16; ++i) s += a[i];
  for (int i = 0; i < 8; ++i) { float2 v = *reinterpret_cast<float2*>(&p[i]); s += v.x + v.y; t += q[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
}
template <int MODE> void run(const char* name, float* out) {
  const int iters = 20000, blocks = 148 * 8, tpb = 256;
  k<MODE><<<blocks, tpb>>>(out, 100);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0); k<MODE><<<blocks, tpb>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double fma = (double)blocks * tpb * iters * 16;
  printf("%-28s %8.3f ms  %7.2f TFLOP/s fp32 (FMA = 2 flop)\n", name, ms, 2 * fma / ms / 1e9);
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
  run<0>("16 FFMA", out); run<1>("8 FFMA2", out); run<2>("16 FFMA + 8 IADD", out); run<3>("8 FFMA2 + 8 IADD", out);
  return cudaDeviceSynchronize() != cudaSuccess;
}
