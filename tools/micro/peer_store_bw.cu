// Remote-store bandwidth over NVLink from a kernel, all GPUs storing to all peers at once (the slab column pass's
// pattern: per unit, 8 KB runs to each destination).  Variants: 8-byte stores, 16-byte stores, TMA bulk copies.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o peer_store_bw peer_store_bw.cu ; ./peer_store_bw [ngpu]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)
constexpr int kMaxP = 8;
struct Args { char* dst[kMaxP]; int P; int units; int self; size_t run; };  // run = bytes per (unit, destination)

// 8-byte stores: thread t stores element t + nthreads*m of the unit's P*run bytes (256-byte runs per warp)
__global__ void __launch_bounds__(512) k_st8(Args a, int skip_self) {
  const int per = (int)(a.run / 8);
  for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
    for (int e = threadIdx.x; e < per * a.P; e += blockDim.x) {
      const int d = e / per, i = e - d * per;
      if (skip_self && d == a.self) continue;
      float2 v = make_float2((float)u, (float)e);
      reinterpret_cast<float2*>(a.dst[d] + (size_t)u * a.run)[i] = v;
    }
  }
}
__global__ void __launch_bounds__(512) k_st16(Args a, int skip_self) {
  const int per = (int)(a.run / 16);
  for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
    for (int e = threadIdx.x; e < per * a.P; e += blockDim.x) {
      const int d = e / per, i = e - d * per;
      if (skip_self && d == a.self) continue;
      float4 v = make_float4((float)u, (float)e, 0.f, 1.f);
      reinterpret_cast<float4*>(a.dst[d] + (size_t)u * a.run)[i] = v;
    }
  }
}
// TMA bulk copies shared -> (peer) global: one thread issues P copies of `run` bytes per unit
__global__ void __launch_bounds__(512) k_bulk(Args a, int skip_self) {
  extern __shared__ __align__(128) char sm[];
  const size_t tot = a.run * a.P;
  for (size_t i = threadIdx.x * 16; i < tot; i += blockDim.x * 16) *reinterpret_cast<float4*>(sm + i) = make_float4(1.f, 2.f, 3.f, 4.f);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
      for (int d = 0; d < a.P; ++d) {
        if (skip_self && d == a.self) continue;
        const uint32_t s = (uint32_t)__cvta_generic_to_shared(sm + d * a.run);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.dst[d] + (size_t)u * a.run), "r"(s), "r"((uint32_t)a.run) : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // at most one unit's reads outstanding
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main(int argc, char** argv) {
  int P = argc > 1 ? atoi(argv[1]) : 8;
  int have = 0; CK(cudaGetDeviceCount(&have)); if (have < P) P = have;
  const size_t run = 8192; const int units = 2048;  // 4 fields x 512 columns
  const size_t slot = run * units;                  // bytes per (source, destination)
  std::vector<char*> buf(P); std::vector<cudaStream_t> st(P); std::vector<cudaEvent_t> e0(P), e1(P);
  for (int d = 0; d < P; ++d) {
    CK(cudaSetDevice(d));
    for (int o = 0; o < P; ++o) if (o != d) { cudaError_t e = cudaDeviceEnablePeerAccess(o, 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { printf("peer %d->%d: %s\n", d, o, cudaGetErrorString(e)); return 1; } (void)cudaGetLastError(); }
    CK(cudaMalloc(&buf[d], slot * P)); CK(cudaMemset(buf[d], 0, slot * P));
    CK(cudaStreamCreate(&st[d])); CK(cudaEventCreate(&e0[d])); CK(cudaEventCreate(&e1[d]));
  }
  CK(cudaSetDevice(0));
  for (int d = 0; d < P; ++d) { CK(cudaSetDevice(d)); CK(cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(run * kMaxP))); }
  struct V { const char* name; int kind; int grid; int skip; };
  V vs[] = {{"st8  grid148", 0, 148, 1}, {"st8  grid296", 0, 296, 1}, {"st8  grid592", 0, 592, 1}, {"st16 grid148", 1, 148, 1}, {"st16 grid296", 1, 296, 1},
            {"st16 grid592", 1, 592, 1}, {"bulk grid148", 2, 148, 1}, {"bulk grid296", 2, 296, 1}, {"bulk grid444", 2, 444, 1},
            {"st8  grid296 +self", 0, 296, 0}, {"bulk grid296 +self", 2, 296, 0}};
  for (const V& v : vs) {
    float best = 1e9f;
    for (int rep = 0; rep < 4; ++rep) {
      for (int d = 0; d < P; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
      for (int d = 0; d < P; ++d) {
        CK(cudaSetDevice(d));
        Args a; a.P = P; a.units = units; a.self = d; a.run = run;
        for (int o = 0; o < P; ++o) a.dst[o] = buf[o] + slot * d;
        CK(cudaEventRecord(e0[d], st[d]));
        for (int it = 0; it < 4; ++it) {
          if (v.kind == 0) k_st8<<<v.grid, 512, 0, st[d]>>>(a, v.skip);
          else if (v.kind == 1) k_st16<<<v.grid, 512, 0, st[d]>>>(a, v.skip);
          else k_bulk<<<v.grid, 512, run * P, st[d]>>>(a, v.skip);
        }
        CK(cudaEventRecord(e1[d], st[d]));
      }
      float worst = 0;
      for (int d = 0; d < P; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); float ms; CK(cudaEventElapsedTime(&ms, e0[d], e1[d])); worst = ms > worst ? ms : worst; }
      if (rep > 0 && worst < best) best = worst;
    }
    const double remote = 4.0 * slot * (P - 1);
    printf("%-22s P=%d  %.3f ms per launch  remote stores %.1f GB/s per GPU\n", v.name, P, best / 4, remote / (best * 1e-3) / 1e9);
    fflush(stdout);
  }
  return 0;
}
