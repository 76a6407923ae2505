// Ceiling of a 3-read : 6-write stream (the column kernel's mix), contiguous and row-tiled, no arithmetic.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

template <int RD, int WR>
__global__ void stream_mix(const float4* __restrict__ in, float4* __restrict__ out, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float4 a = make_float4(0, 0, 0, 0);
#pragma unroll
    for (int k = 0; k < RD; ++k) { const float4 v = in[k * n + i]; a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
#pragma unroll
    for (int k = 0; k < WR; ++k) out[k * n + i] = a;
  }
}
// same, but each CTA walks a [R rows][TCB bytes] tile of a [B][R][W bytes] array (rows W bytes apart)
template <int RD, int WR, int TCB>
__global__ void tile_mix(const float4* __restrict__ in, float4* __restrict__ out, int R, int Wb, int B, size_t n) {
  constexpr int LPR = TCB / 16;  // lanes (float4) per row run
  const int tiles = Wb / TCB;
  const int c = threadIdx.x % LPR, t = threadIdx.x / LPR, RPP = blockDim.x / LPR;
  for (int tile = blockIdx.x; tile < tiles * B; tile += gridDim.x) {
    const int b = tile / tiles, tx = tile % tiles;
    const size_t base = ((size_t)b * R * Wb + (size_t)tx * TCB) / 16 + c;
    for (int r = t; r < R; r += RPP) {
      const size_t i = base + (size_t)r * (Wb / 16);
      float4 a = make_float4(0, 0, 0, 0);
#pragma unroll
      for (int k = 0; k < RD; ++k) { const float4 v = in[k * n + i]; a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
#pragma unroll
      for (int k = 0; k < WR; ++k) out[k * n + i] = a;
    }
  }
}

int main() {
  const int R = 256, Wb = 1024, B = 256;  // plane = 256 rows x 128 complex64
  const size_t n = (size_t)R * Wb * B / 16;
  float4 *in, *out;
  CK(cudaMalloc(&in, 16 * n * 4));
  CK(cudaMalloc(&out, 16 * n * 6));
  CK(cudaMemset(in, 0, 16 * n * 4));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const double C = 16.0 * n;
  auto time = [&](auto launch, const char* name, int rd, int wr) {
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 20; ++i) launch();
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 20;
    CK(cudaGetLastError());
    printf("%-28s rd %d wr %d: %7.1f us %6.0f GB/s\n", name, rd, wr, ms * 1e3, (rd + wr) * C / ms / 1e6);
  };
  const int g = 148 * 8;
  time([&] { stream_mix<1, 1><<<g, 256>>>(in, out, n); }, "stream copy", 1, 1);
  time([&] { stream_mix<3, 6><<<g, 256>>>(in, out, n); }, "stream", 3, 6);
  time([&] { stream_mix<4, 1><<<g, 256>>>(in, out, n); }, "stream", 4, 1);
  time([&] { stream_mix<3, 3><<<g, 256>>>(in, out, n); }, "stream", 3, 3);
  time([&] { stream_mix<1, 6><<<g, 256>>>(in, out, n); }, "stream", 1, 6);
  time([&] { tile_mix<3, 6, 64><<<g, 256>>>(in, out, R, Wb, B, n); }, "tile 64B runs", 3, 6);
  time([&] { tile_mix<3, 6, 128><<<g, 256>>>(in, out, R, Wb, B, n); }, "tile 128B runs", 3, 6);
  time([&] { tile_mix<3, 6, 256><<<g, 256>>>(in, out, R, Wb, B, n); }, "tile 256B runs", 3, 6);
  time([&] { tile_mix<4, 1, 64><<<g, 256>>>(in, out, R, Wb, B, n); }, "tile 64B runs", 4, 1);
  time([&] { tile_mix<4, 1, 128><<<g, 256>>>(in, out, R, Wb, B, n); }, "tile 128B runs", 4, 1);
  time([&] { tile_mix<4, 1, 1024><<<g, 256>>>(in, out, R, Wb, B, n); }, "tile full rows", 4, 1);
  return 0;
}
