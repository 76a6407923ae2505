// Microbenchmark: issue throughput of scalar FFMA/FADD vs packed FFMA2/FADD2/FMUL2 on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_bench f32x2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0,{%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

template <int MODE> __global__ void k(float* out, int iters, float s) {
  constexpr int U = 8;
  float a[U], b[U];
  u64 A[U], B[U];
  for (int i = 0; i < U; ++i) { a[i] = threadIdx.x + i; b[i] = s + i; A[i] = pk(a[i], b[i]); B[i] = pk(b[i], a[i]); }
  const u64 S = pk(s, s * 0.5f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < U; ++i) {
      if (MODE == 0) { a[i] = fmaf(a[i], s, b[i]); b[i] = fmaf(b[i], s, a[i]); }        // 2 FFMA
      if (MODE == 1) { a[i] = a[i] + b[i]; b[i] = b[i] - a[i]; }                          // 2 FADD
      if (MODE == 2) { A[i] = fma2(A[i], S, B[i]); B[i] = fma2(B[i], S, A[i]); }          // 2 FFMA2
      if (MODE == 3) { A[i] = add2(A[i], B[i]); B[i] = add2(B[i], A[i]); }                // 2 FADD2
      if (MODE == 4) { A[i] = mul2(A[i], S); B[i] = mul2(B[i], S); }                      // 2 FMUL2
      if (MODE == 5) { A[i] = add2(A[i], B[i]); a[i] = a[i] + b[i]; }                     // FADD2 + FADD mix
    }
  }
  float r = 0;
  for (int i = 0; i < U; ++i) r += a[i] + b[i] + (float)(A[i] & 0xffff) + (float)(B[i] >> 40);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, float* d) {
  const int iters = 4096, blocks = 148 * 4, threads = 512;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<blocks, threads>>>(d, 16, 1.0001f);
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(d, iters, 1.0001f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double winstr = (double)blocks * threads / 32 * iters * 16;  // warp-instructions
  printf("%-14s %.3f ms  %.1f Gwarp-instr/s  (%.2f instr/clk/SM at 1.9 GHz)\n", name, ms, winstr / ms * 1e-6,
         winstr / ms * 1e-6 / 148 / 1.9);
}
int main() {
  float* d; cudaMalloc(&d, 148 * 4 * 512 * 4);
  run<0>("FFMA", d); run<1>("FADD", d); run<2>("FFMA2", d); run<3>("FADD2", d); run<4>("FMUL2", d); run<5>("FADD2+FADD", d);
  return 0;
}
