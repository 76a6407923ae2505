// Persistent, TMA-fed column kernel for the RK sub-stage (sm_100a).
//
// Same arithmetic as col_kernel<..., kColStage> (kernels.cuh) but the three input tiles of a
// (environment, 16-column) tile -- advection rows, RK register h, omega_hat -- arrive in shared
// memory through cp.async.bulk.tensor (TMA, SASS UTMALDG) one tile AHEAD of their use, so the
// FFT/pointwise phases never wait on a global load:
//
//   buffer   filled by TMA with     consumed by                      then reused as
//   buf_adv  adv rows of tile i     first FFT pass (registers)       exchange buffer of the forward FFT
//   buf_h    h of tile i            RK accumulate                    split exchange buffer of the two lane-paired
//                                                                    inverse FFTs (real, then imaginary parts)
//   buf_w    omega_hat of tile i    Crank-Nicolson update            the NEW omega_hat, read back by the prologue
//                                                                    (keeps 2*E registers free for the paired FFT)
//
// As soon as a buffer's last reader has passed a barrier, thread 0 re-arms its mbarrier and
// issues the load of the NEXT tile of this CTA (tile + gridDim.x): adv right after the forward
// FFT (it is needed first), h and omega_hat after the last inverse FFT (they are needed only
// after the next tile's forward FFT, which is longer than a TMA round trip).  Outputs leave with
// plain coalesced stores (128-bit for the lane-paired fields).  One CTA = TPF*TC threads;
// grid = (#SMs x resident CTAs), tiles round-robin.
#pragma once
#include <cuda.h>  // CUtensorMap (type only; the encoder is fetched at run time)

#include "kernels.cuh"

namespace kf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 2-D tiled TMA load: box at element coordinates (c0 = inner, c1 = row) -> dense [rows][inner] in smem
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

template <typename T, class P, int TC> struct ColTmaSmem {
  static constexpr int N = P::L;
  static constexpr size_t kTile = (size_t)N * TC * sizeof(Cx<T>);
  static_assert(kTile == col_exchange_bytes<T, N, TC>(), "a tile buffer doubles as the exchange buffer");
  static constexpr size_t bytes() { return 3 * kTile + 2 * (size_t)N * sizeof(Cx<T>) + 64; }
};

template <typename T, class P, int TC, int MINB>
__global__ void __launch_bounds__(P::TPF* TC, MINB)
    col_tma_kernel(StageArgs<T> a, const __grid_constant__ CUtensorMap m_adv, const __grid_constant__ CUtensorMap m_h,
                   const __grid_constant__ CUtensorMap m_w, int ntiles) {
  constexpr int N = P::L, E = P::E, TPF = P::TPF, MH = N / 2, TILES = MH / TC;
  constexpr int RB = N < 256 ? N : 256, NBOX = N / RB;  // TMA box rows (<= 256)
  constexpr uint32_t kBytes = (uint32_t)((size_t)N * TC * sizeof(Cx<T>));
  extern __shared__ __align__(1024) unsigned char smem_tma[];
  Cx<T>* buf_adv = reinterpret_cast<Cx<T>*>(smem_tma);
  Cx<T>* buf_h = buf_adv + (size_t)N * TC;
  Cx<T>* buf_w = buf_h + (size_t)N * TC;
  Cx<T>* s_tw = buf_w + (size_t)N * TC;
  Cx<T>* s_nyq = s_tw + N;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_nyq + N);  // [0] adv, [1] h, [2] w
  const int tid = threadIdx.x, c = tid % TC, t = tid / TC;
  const bool use_h = (a.beta != (T)0);
  const size_t plane = (size_t)N * MH;

  if (tid == 0) {
    mbar_init(bars + 0, 1);
    mbar_init(bars + 1, 1);
    mbar_init(bars + 2, 1);
    fence_mbar_init();
  }
  for (int i = tid; i < N; i += TPF * TC) s_tw[i] = ldc(a.tw + i);
  __syncthreads();

  // Work list: the nb tiles with tx == 0 first (they also carry the ky = M/2 column and cost more), then
  // the others.  CTA i takes positions i, i + gridDim.x, ...: every CTA gets at most one expensive tile
  // instead of the same few CTAs getting all of them (gridDim.x is often a multiple of TILES).
  const int nb = ntiles / TILES;
  auto decode = [&](int pos, int& b, int& tx) {
    if (pos < nb) { b = pos; tx = 0; }
    else { constexpr int TM1 = TILES > 1 ? TILES - 1 : 1; const int r = pos - nb; b = r / TM1; tx = 1 + (r - b * TM1); }
  };
  auto issue = [&](Cx<T>* buf, const CUtensorMap* map, uint64_t* bar, int tile) {
    int b, tx;
    decode(tile, b, tx);
    mbar_expect_tx(bar, kBytes);
#pragma unroll
    for (int r = 0; r < NBOX; ++r)
      tma_load_2d(buf + (size_t)r * RB * TC, map, tx * TC * 2, b * N + r * RB, bar);
  };

  const int first = blockIdx.x;
  if (tid == 0 && first < ntiles) {
    issue(buf_adv, &m_adv, bars + 0, first);
    issue(buf_w, &m_w, bars + 2, first);
    if (use_h) issue(buf_h, &m_h, bars + 1, first);
  }

  uint32_t it = 0;
  for (int tile = first; tile < ntiles; tile += gridDim.x, ++it) {
    const uint32_t par = it & 1u;
    int b, tx;
    decode(tile, b, tx);
    const int q = tx * TC + c;
    const bool special = (tx == 0);  // CTA-uniform
    const int next = tile + gridDim.x;
    const bool has_next = next < ntiles;
    Cx<T>* wmain = a.wmain + b * plane + q;
    Cx<T>* hmain = a.hmain + b * plane + q;
    Cx<T>* wnyq = a.wnyq + (size_t)b * N;
    Cx<T>* hnyq = a.hnyq + (size_t)b * N;
    const Cx<T>* G = a.G + (size_t)b * (MH + 1);
    Cx<T> g0 = mk<T>((T)0, (T)0);
    if (t == 0) g0 = ldc(G + q);  // forcing + control, row kx = 0 (needed after the FFT)

    Cx<T> v[E];
    mbar_wait(bars + 0, par);
#pragma unroll
    for (int m = 0; m < E; ++m) v[m] = ldc(buf_adv + (size_t)(t + TPF * m) * TC + c);
    __syncthreads();  // all rows are in registers before buf_adv becomes the exchange buffer
    ColEx<T, TC> exA{buf_adv, c};
    fft_regs<P, -1, T>(v, s_tw, t, exA);

    if (special) {
      // unpack column 0: A0[kx] = (W[kx] + conj(W[-kx]))/2, An[kx] = (W[kx] - conj(W[-kx]))/(2i)
      if (c == 0) {
#pragma unroll
        for (int m = 0; m < E; ++m) stc(buf_adv + t + TPF * m, v[m]);
      }
      __syncthreads();
      nyquist_column_update<T, N>(a, buf_adv, s_nyq, hnyq, wnyq, G, use_h, tid, TPF * TC);
      if (c == 0) {
#pragma unroll
        for (int m = 0; m < E; ++m)
          v[m] = (T)0.5 * add_conj(v[m], ldc(buf_adv + ((N - (t + TPF * m)) & (N - 1))));
      }
      __syncthreads();
    }
    if (tid == 0 && has_next) {  // buf_adv is free: prefetch the next tile's advection rows
      fence_proxy_async();
      issue(buf_adv, &m_adv, bars + 0, next);
    }

    if (use_h) mbar_wait(bars + 1, par);
    mbar_wait(bars + 2, par);
    Cx<T> w[E];
    rk_update_column<T, P>(a, v, w, g0, use_h, t, q, hmain, wmain,
                           [&](int, int kx) { return ldc(buf_h + kx * TC + c); },
                           [&](int, int kx) { return ldc(buf_w + kx * TC + c); });
    // the new omega_hat replaces the old one in buf_w (each thread rewrites exactly what it read)
#pragma unroll
    for (int m = 0; m < E; ++m) stc(buf_w + (t + TPF * m) * TC + c, w[m]);
    __syncthreads();  // buf_h has been consumed and becomes the exchange buffer

    auto load_w = [&](int, int kx) { return ldc(buf_w + kx * TC + c); };
    col_field_pair<0, T, P, TC>(a, load_w, s_tw, s_nyq, buf_h, c, t, q, b, special);
    col_field_pair<1, T, P, TC>(a, load_w, s_tw, s_nyq, buf_h, c, t, q, b, special);
    __syncthreads();  // every thread has read its omega_hat for the second pair
    if (tid == 0 && has_next) {
      fence_proxy_async();
      if (use_h) issue(buf_h, &m_h, bars + 1, next);
      issue(buf_w, &m_w, bars + 2, next);
    }
  }
}

}  // namespace kf
