// Micro-benchmark: TMA feed rate of the column kernel's input tiles, no arithmetic.  296 persistent CTAs, each
// loops over (env, 8-column) tiles of 3 arrays [B][256][128] complex64 with a 3-deep ring:
//   mode 0: 2-D tensor-map boxes of 256 rows x 64 B (what col_tma_kernel issues today)
//   mode 1: 1-D bulk copies of 16 KB from a TILE-MAJOR copy of the same data ([B][16][256][8])
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile("{\n.reg .pred P1;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra LAB_WAIT;\nDONE:\n}\n"
               ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

constexpr int N = 256, MH = 128, TC = 8, TILES = MH / TC, NBUF = 3;
constexpr uint32_t kTile = N * TC * 8;

template <int MODE>
__global__ void __launch_bounds__(128) feed(const __grid_constant__ CUtensorMap m0, const __grid_constant__ CUtensorMap m1,
                                            const __grid_constant__ CUtensorMap m2, const float2* t0, const float2* t1,
                                            const float2* t2, int ntiles, float* sink) {
  extern __shared__ __align__(1024) unsigned char smem[];
  float2* bufs = reinterpret_cast<float2*>(smem);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NBUF * kTile);
  if (threadIdx.x == 0) {
    for (int i = 0; i < NBUF; ++i) mbar_init(bars + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int tile) {
    const int b = tile / TILES, tx = tile % TILES;
    const CUtensorMap* maps[3] = {&m0, &m1, &m2};
    const float2* tl[3] = {t0, t1, t2};
    for (int i = 0; i < NBUF; ++i) {
      mbar_expect_tx(bars + i, kTile);
      if (MODE == 0) tma_load_2d(bufs + (size_t)i * N * TC, maps[i], tx * TC * 2, b * N, bars + i);
      else bulk_load_1d(bufs + (size_t)i * N * TC, tl[i] + (size_t)tile * N * TC, kTile, bars + i);
    }
  };
  if (threadIdx.x == 0 && blockIdx.x < ntiles) issue(blockIdx.x);
  float acc = 0.f;
  uint32_t it = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    for (int i = 0; i < NBUF; ++i) mbar_wait(bars + i, it & 1u);
    for (int i = 0; i < NBUF; ++i) acc += bufs[(size_t)i * N * TC + threadIdx.x * 16].x;  // touch
    __syncthreads();
    if (threadIdx.x == 0 && tile + gridDim.x < ntiles) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(tile + gridDim.x);
    }
  }
  if (acc == 123.456f) *sink = acc;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int B = 256;
  const size_t plane = (size_t)N * MH;
  float2* a[3]; float* sink;
  for (int i = 0; i < 3; ++i) { CK(cudaMalloc(&a[i], 8 * plane * B)); CK(cudaMemset(a[i], 0, 8 * plane * B)); }
  CK(cudaMalloc(&sink, 4));
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  EncodeTiledFn enc = (EncodeTiledFn)p;
  CUtensorMap maps[3];
  const cuuint64_t dims[2] = {2 * MH, (cuuint64_t)B * N};
  const cuuint64_t strides[1] = {2 * MH * 4};
  const cuuint32_t box[2] = {2 * TC, 256}, estr[2] = {1, 1};
  for (int i = 0; i < 3; ++i)
    if (enc(&maps[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, a[i], dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
  const int ntiles = B * TILES;
  const size_t smem = NBUF * kTile + 64;
  CK(cudaFuncSetAttribute(feed<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(feed<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int grid : {148, 296, 592}) for (int mode = 0; mode < 2; ++mode) {
    auto launch = [&] {
      if (mode == 0) feed<0><<<grid, 128, smem>>>(maps[0], maps[1], maps[2], a[0], a[1], a[2], ntiles, sink);
      else feed<1><<<grid, 128, smem>>>(maps[0], maps[1], maps[2], a[0], a[1], a[2], ntiles, sink);
    };
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 20; ++i) launch();
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 20;
    CK(cudaGetLastError());
    printf("grid %3d %s: %.1f us  %.0f GB/s (3C read only)\n", grid, mode == 0 ? "2-D boxes 256 x 64 B" : "1-D bulk 16 KB      ", ms * 1e3,
           3.0 * 8 * plane * B / ms / 1e6);
  }
  return 0;
}
