// Persistent, TMA-fed column kernel for the RK sub-stage (sm_100a).
//
// Same arithmetic as col_kernel<..., kColStage> (kernels.cuh), but the three input tiles of a
// (environment, TC-column) tile -- advection rows, RK register h, omega_hat -- arrive in shared
// memory through cp.async.bulk.tensor (TMA, SASS UTMALDG) on one mbarrier each, one tile AHEAD of
// their use: a CTA owns ONE {adv, h, w} buffer set and re-arms each buffer as soon as the current
// tile has consumed it,
//
//   buffer  filled by TMA with     consumed by                  then reused as
//   adv     adv rows of tile i     first FFT pass (registers)   exchange buffer of the forward FFT; re-armed for
//                                                               tile i + gridDim.x right after that FFT
//   h       h of tile i            RK accumulate                split exchange buffer of the two lane-paired
//                                                               inverse FFTs; re-armed at the end of the tile
//   w       omega_hat of tile i    Crank-Nicolson update        the NEW omega_hat (read back by the prologue of
//                                                               the next stage); re-armed at the end of the tile
//
// so the advection load of the next tile overlaps the RK update and both inverse transforms of
// this one.  Results leave with per-thread stores (64-bit h / omega_hat, 128-bit lane-paired fields, runs
// of TC columns per row).  A two-stage ring with TMA stores (UTMASTG) was built and measured in round 1
// (131.7 -> 139.5 us per launch at 256 envs: slower) and has been removed.
// One CTA = TPF*TC threads; grid = #SMs x resident CTAs.
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
__device__ __forceinline__ void tma_load_2d_hint(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], "
      "[%4], %5;" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

// XBUF: a dedicated exchange buffer of N x TC lane-paired complex numbers for the inverse transforms (one exchange
// round, two barriers) instead of reusing the h tile for a split exchange (two rounds, four barriers).
template <typename T, class P, int TC, bool XBUF = false> struct ColTmaSmem {
  static constexpr int N = P::L;
  static constexpr size_t kTile = (size_t)N * TC * sizeof(Cx<T>);
  static_assert(kTile == col_exchange_bytes<T, N, TC>(), "a tile buffer doubles as the exchange buffer");
  static constexpr size_t bytes() { return (XBUF ? 5 : 3) * kTile + 2 * (size_t)N * sizeof(Cx<T>) + 64; }
};

template <typename T, class P, int TC, int MINB, bool XBUF = false>
__global__ void __launch_bounds__(P::TPF* TC, MINB)
    col_tma_kernel(StageArgs<T> a, const __grid_constant__ CUtensorMap m_adv, const __grid_constant__ CUtensorMap m_h,
                   const __grid_constant__ CUtensorMap m_w, int ntiles) {
  constexpr int N = P::L, E = P::E, TPF = P::TPF, MH = N / 2, TILES = MH / TC;
  constexpr int RB = N < 256 ? N : 256, NBOX = N / RB;  // TMA box rows (<= 256)
  constexpr uint32_t kBytes = (uint32_t)((size_t)N * TC * sizeof(Cx<T>));
  constexpr size_t kTileEl = (size_t)N * TC;
  extern __shared__ __align__(1024) unsigned char smem_tma[];
  Cx<T>* buf_adv = reinterpret_cast<Cx<T>*>(smem_tma);
  Cx<T>* buf_h = buf_adv + kTileEl;
  Cx<T>* buf_w = buf_h + kTileEl;
  Cx<T>* s_xbuf = buf_w + kTileEl;  // 2 tiles, only with XBUF
  Cx<T>* s_tw = s_xbuf + (XBUF ? 2 * kTileEl : 0);
  Cx<T>* s_nyq = s_tw + N;
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_nyq + N);  // [3]: adv, h, w
  const int tid = threadIdx.x, c = tid % TC, t = tid / TC;
  const bool use_h = (a.beta != (T)0);
  const size_t plane = (size_t)N * MH;

  pdl_launch_dependents();
  if (tid == 0) {
    for (int i = 0; i < 3; ++i) mbar_init(bar + i, 1);
    fence_mbar_init();
  }
  for (int i = tid; i < N; i += TPF * TC) s_tw[i] = ldc(a.tw + i);
  __syncthreads();
  pdl_wait();  // everything below reads the row kernel's advection rows or overwrites the fields it still reads

  // Work list: the nb tiles with tx == 0 first (they also carry the ky = M/2 column and cost more), then
  // the others.  CTA i takes positions i, i + gridDim.x, ...: every CTA gets at most one expensive tile
  // instead of the same few CTAs getting all of them (gridDim.x is often a multiple of TILES).
  const int nb = ntiles / TILES;
  auto decode = [&](int pos, int& b, int& tx) {
    if (pos < nb) { b = pos; tx = 0; }
    else { constexpr int TM1 = TILES > 1 ? TILES - 1 : 1; const int r = pos - nb; b = r / TM1; tx = 1 + (r - b * TM1); }
  };
  const uint64_t pol_state = l2_policy(a.l2_hints != 0);  // omega_hat / h: streamed when the hints are on
  auto issue = [&](Cx<T>* buf, const CUtensorMap* map, uint64_t* br, int tile) {
    int b, tx;
    decode(tile, b, tx);
    mbar_expect_tx(br, kBytes);
    const bool state = (map != &m_adv);
#pragma unroll
    for (int r = 0; r < NBOX; ++r) {
      if (state) tma_load_2d_hint(buf + (size_t)r * RB * TC, map, tx * TC * 2, b * N + r * RB, br, pol_state);
      else tma_load_2d(buf + (size_t)r * RB * TC, map, tx * TC * 2, b * N + r * RB, br);
    }
  };
  constexpr int km = ((MH + 1) * 2) / 3;  // 2/3 mask: columns ky >= km of the advection term are discarded

  if (tid == 0 && (int)blockIdx.x < ntiles) {
    int b0, tx0;
    decode(blockIdx.x, b0, tx0);
    if (!(a.dealias && tx0 * TC >= km)) issue(buf_adv, &m_adv, bar + 0, blockIdx.x);
    if (use_h) issue(buf_h, &m_h, bar + 1, blockIdx.x);
    issue(buf_w, &m_w, bar + 2, blockIdx.x);
  }

  uint32_t it = 0, adv_phase = 0;  // adv_phase: parity of the advection barrier (masked tiles skip a phase)
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const uint32_t par = it & 1u;
    int b, tx;
    decode(tile, b, tx);
    const int q = tx * TC + c;
    const bool special = (tx == 0);  // CTA-uniform
    const int next = tile + gridDim.x;  // the tile that will reuse the buffers
    const bool has_next = next < ntiles;
    // 2/3 mask on: tiles made only of columns ky >= km neither load nor transform the advection term (the row
    // kernel does not write those columns)
    const bool masked_tile = a.dealias && tx * TC >= km;  // CTA-uniform
    bool next_masked = false;
    if (has_next && a.dealias) {
      int nb2, ntx;
      decode(next, nb2, ntx);
      next_masked = ntx * TC >= km;
    }
    Cx<T>* wnyq = a.wnyq + (size_t)b * N;
    Cx<T>* hnyq = a.hnyq + (size_t)b * N;
    const Cx<T>* G = a.G + (size_t)b * (MH + 1);
    Cx<T> g0 = mk<T>((T)0, (T)0);
    if (t == 0) g0 = ldc(G + q);  // forcing + control, row kx = 0 (needed after the FFT)

    Cx<T> v[E];
    if (!masked_tile) {
      mbar_wait(bar + 0, adv_phase);
      adv_phase ^= 1u;  // one phase per advection tile actually loaded
#pragma unroll
      for (int m = 0; m < E; ++m) v[m] = ldc(buf_adv + (size_t)(t + TPF * m) * TC + c);
      __syncthreads();  // all rows are in registers before buf_adv becomes the exchange buffer
      ColEx<T, TC> exA{buf_adv, c};
      fft_regs<P, -1, T>(v, s_tw, t, exA);
    } else {
#pragma unroll
      for (int m = 0; m < E; ++m) v[m] = mk<T>((T)0, (T)0);
    }

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
    if (tid == 0 && has_next && !next_masked) {  // buf_adv is free: refill it for the next tile now
      fence_proxy_async();
      issue(buf_adv, &m_adv, bar + 0, next);
    }

    if (use_h) mbar_wait(bar + 1, par);
    mbar_wait(bar + 2, par);
    {
      Cx<T> w[E];
      Cx<T>* wmain = a.wmain + b * plane + q;
      Cx<T>* hmain = a.hmain + b * plane + q;
      rk_update_column<T, P>(a, v, w, g0, use_h, t, q,
                             [&](int, int kx, Cx<T> h) { if (a.store_h) stc_hint(hmain + (size_t)kx * MH, h, pol_state); },
                             [&](int, int kx, Cx<T> wn) {
                               stc_hint(wmain + (size_t)kx * MH, wn, pol_state);
                               stc(buf_w + kx * TC + c, wn);
                             },
                             [&](int, int kx) { return ldc(buf_h + kx * TC + c); },
                             [&](int, int kx) { return ldc(buf_w + kx * TC + c); },
                             [&](int, int kx) {
                               return a.ctrl_main ? ldc(a.ctrl_main + (size_t)kx * MH + q) : mk<T>((T)0, (T)0);
                             });
      __syncthreads();  // buf_h has been consumed and becomes the exchange buffer
    }

    auto load_w = [&](int, int kx) { return ldc(buf_w + kx * TC + c); };
    if constexpr (XBUF) {
      // buf_h is free as soon as the RK update is done (the barrier above): re-arm it a whole inverse-transform
      // phase earlier than the omega_hat tile
      if (tid == 0 && has_next && use_h) {
        fence_proxy_async();
        issue(buf_h, &m_h, bar + 1, next);
      }
      col_field_pair<0, T, P, TC, false>(a, load_w, s_tw, s_nyq, s_xbuf, c, t, q, b, special);
      col_field_pair<1, T, P, TC, false>(a, load_w, s_tw, s_nyq, s_xbuf, c, t, q, b, special);
    } else {
      col_field_pair<0, T, P, TC>(a, load_w, s_tw, s_nyq, buf_h, c, t, q, b, special);
      col_field_pair<1, T, P, TC>(a, load_w, s_tw, s_nyq, buf_h, c, t, q, b, special);
    }
    __syncthreads();  // every thread has read its omega_hat for the second pair and left the last exchange
    if (tid == 0 && has_next) {
      fence_proxy_async();
      if (use_h && !XBUF) issue(buf_h, &m_h, bar + 1, next);
      issue(buf_w, &m_w, bar + 2, next);
    }
  }
}

}  // namespace kf
