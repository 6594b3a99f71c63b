// Issue-slot vs FMA-pipe microbenchmark for packed fp32 (FFMA2/FMUL2/FADD2) on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_f32x2 tools/ubench/f32x2.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
#define CHAINS 8
// mode 0: 2*CHAINS scalar FFMA per iteration; mode 1: CHAINS FFMA2; mode 2: scalar + CHAINS FMNMX(alu); mode 3: FFMA2 + CHAINS FMNMX
template <int MODE>
__global__ void __launch_bounds__(256) kern(float* out, int iters, float s) {
  float a[2 * CHAINS], m[CHAINS];
  for (int i = 0; i < 2 * CHAINS; i++) a[i] = threadIdx.x * 0.001f + i;
  for (int i = 0; i < CHAINS; i++) m[i] = (float)i;
  float b = s, c = 1.0f - s;
  u64 b2, c2;
  asm("mov.b64 %0, {%1, %1};" : "=l"(b2) : "f"(b));
  asm("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(c));
  u64 p[CHAINS];
  for (int i = 0; i < CHAINS; i++) asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 2 * CHAINS; i++) a[i] = fmaf(a[i], b, c);
    } else {
#pragma unroll
      for (int i = 0; i < CHAINS; i++) p[i] = ffma2(p[i], b2, c2);
    }
    if (MODE >= 2) {
#pragma unroll
      for (int i = 0; i < CHAINS; i++) asm volatile("max.f32 %0, %0, %1;" : "+f"(m[i]) : "f"(b));
    }
  }
  float r = 0;
  for (int i = 0; i < 2 * CHAINS; i++) r += a[i];
  for (int i = 0; i < CHAINS; i++) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(p[i])); r += x + y + m[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE>
void run(float* out, const char* name) {
  const int iters = 4096, blocks = 148 * 8;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  kern<MODE><<<blocks, 256>>>(out, iters, 0.999f);
  cudaEventRecord(e0);
  kern<MODE><<<blocks, 256>>>(out, iters, 0.999f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double fma = (double)blocks * 256 * iters * 2 * CHAINS;
  printf("%-28s %.3f ms  %.1f TFLOP/s fp32  (%.2f lane-FMA per clk per SM @1.965GHz)\n", name, ms, 2 * fma / ms / 1e9, fma / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  run<0>(out, "scalar FFMA");
  run<1>(out, "FFMA2");
  run<2>(out, "scalar FFMA + FMNMX(1:2)");
  run<3>(out, "FFMA2 + FMNMX(1:1)");
  return cudaGetLastError() != cudaSuccess;
}
