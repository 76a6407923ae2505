// Micro-benchmark: achievable HBM throughput of the column kernel's ACCESS MIX without any arithmetic.
// Per "environment" plane of R rows x W complex64: read 3 planes, write 2 planes (8 B elements) and 2 paired
// planes (16 B elements), in tiles of TC columns (runs of TC*8 / TC*16 bytes per row).  Variants: tile width.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

template <int TC, int RD, int WR8, int WR16>
__global__ void tile_mix(const float2* __restrict__ in, float2* __restrict__ out8, float4* __restrict__ out16,
                         int R, int W, int B, size_t plane) {
  // CTA = 256 threads; tile = R rows x TC columns of env b; thread (c, t): column c, rows t, t + 256/TC, ...
  const int tiles = W / TC;
  const int c = threadIdx.x % TC, t = threadIdx.x / TC, TPF = 256 / TC;
  for (int tile = blockIdx.x; tile < tiles * B; tile += gridDim.x) {
    const int b = tile / tiles, tx = tile % tiles;
    const size_t base = (size_t)b * plane + tx * TC + c;
    float2 acc = make_float2(0.f, 0.f);
    for (int r = t; r < R; r += TPF) {
#pragma unroll
      for (int k = 0; k < RD; ++k) {
        const float2 v = in[(size_t)k * B * plane + base + (size_t)r * W];
        acc.x += v.x; acc.y += v.y;
      }
    }
    for (int r = t; r < R; r += TPF) {
#pragma unroll
      for (int k = 0; k < WR8; ++k) out8[(size_t)k * B * plane + base + (size_t)r * W] = acc;
#pragma unroll
      for (int k = 0; k < WR16; ++k)
        out16[(size_t)k * B * plane + base + (size_t)r * W] = make_float4(acc.x, acc.y, acc.x, acc.y);
    }
  }
}

template <int TC, int RD, int WR8, int WR16> float run(const float2* in, float2* o8, float4* o16, int R, int W, int B, int grid) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const size_t plane = (size_t)R * W;
  for (int i = 0; i < 3; ++i) tile_mix<TC, RD, WR8, WR16><<<grid, 256>>>(in, o8, o16, R, W, B, plane);
  CK(cudaEventRecord(e0));
  const int reps = 20;
  for (int i = 0; i < reps; ++i) tile_mix<TC, RD, WR8, WR16><<<grid, 256>>>(in, o8, o16, R, W, B, plane);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  CK(cudaGetLastError());
  return ms / reps;
}

int main() {
  const int R = 256, W = 128, B = 256;
  const size_t plane = (size_t)R * W;
  float2 *in, *o8; float4* o16;
  CK(cudaMalloc(&in, sizeof(float2) * plane * B * 4));
  CK(cudaMalloc(&o8, sizeof(float2) * plane * B * 2));
  CK(cudaMalloc(&o16, sizeof(float4) * plane * B * 4));
  CK(cudaMemset(in, 0, sizeof(float2) * plane * B * 4));
  const double C = (double)plane * B * 8;
  for (int grid : {148 * 2, 148 * 4, 148 * 8}) {
#define RUN(TC, RD, W8, W16) { float ms = run<TC, RD, W8, W16>(in, o8, o16, R, W, B, grid); \
      printf("grid %4d TC=%2d rd %d wr8 %d wr16 %d: %.1f us  %.0f GB/s\n", grid, TC, RD, W8, W16, ms * 1e3, (RD + W8 + 2 * W16) * C / ms / 1e6); }
    RUN(8, 3, 2, 2) RUN(16, 3, 2, 2) RUN(32, 3, 2, 2)
    RUN(8, 4, 0, 0) RUN(16, 4, 0, 0)        // pure read, 4C
    RUN(8, 0, 2, 2) RUN(16, 0, 2, 2)        // pure write, 6C
    RUN(8, 4, 1, 0) RUN(16, 4, 1, 0)        // row-kernel-like mix (4C in as 2 paired.. approximated, 1C out)
  }
  return 0;
}
