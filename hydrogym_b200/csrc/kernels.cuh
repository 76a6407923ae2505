// Fused kernels of one IMEX Runge-Kutta sub-stage of the Kolmogorov-flow solve.
//
// Reference path: RK4_CN.time_step_fn (hydrogym/jax/solvers/base.py:50-57) calling
// PseudoSpectralNavierStokes2D.nonlinear_terms / linear_terms / implicit_timestep
// (hydrogym/jax/envs/kolmogorov.py:209-257).  One sub-stage is two launches:
//
//   col_kernel : forward x-FFT of the advection rows -> + forcing/control on row
//                kx=0 -> h, omega_hat update (RK accumulate + Crank-Nicolson solve)
//                -> u_hat, v_hat, d/dx, d/dy of the NEW omega_hat -> inverse x-FFT
//   row_kernel : c2r y-FFT of the 4 fields -> -(gx*u + gy*v) -> r2c y-FFT
//
// Layouts (per environment):
//   spectral  omega_hat, h : main[N][M/2] (ky = 0..M/2-1) + nyq[N] (ky = M/2)
//   x-space   adv          : [N][M/2] complex with column 0 PACKED as
//             Re(column ky=0) + i*Re(column ky=M/2): irfftn drops the imaginary
//             part of those two columns after the x-transform, and rfftn's y-pass
//             makes them real, so one complex column carries both.
//   x-space   fields       : two arrays FA, FB of [N][M/2] LANE-PAIRED complex
//             numbers (re0, re1, im0, im1): FA carries (u_hat, gx_hat), FB carries
//             (v_hat, gy_hat), column 0 packed as above.  The pairs share every
//             twiddle, so the inverse x- and y-FFTs run two transforms per
//             instruction (F2 lanes: FFMA2 / FADD2 / FMUL2) with 128-bit accesses.
// Wavenumbers are the integers p = fftfreq index, q = rfftfreq index (2*pi*kx = p).
#pragma once
#include <cstdint>
#include <type_traits>

#include "fft.cuh"

namespace kf {

template <typename T> struct StageArgs {
  using L2 = typename PairOf<T>::type;
  Cx<T>* wmain;        // [Bc][N][M/2]
  Cx<T>* wnyq;         // [Bc][N]
  Cx<T>* hmain;        // [Bc][N][M/2]
  Cx<T>* hnyq;         // [Bc][N]
  const Cx<T>* G;      // [Bc][M/2+1]  forcing + control row (kx = 0)
  const Cx<T>* ctrl_main;  // [N][M/2] dense control spectrum shared by all environments (general control_field of
  const Cx<T>* ctrl_nyq;   // [N]      solve(), kolmogorov.py:259-272); NULL = none (the env path: row kx = 0 only)
  Cx<T>* advx;         // [Bc][N][M/2] packed x-space advection
  Cx<L2>* fa;          // [Bc][N][M/2] lanes (u_hat, gx_hat) in x-space
  Cx<L2>* fb;          // [Bc][N][M/2] lanes (v_hat, gy_hat) in x-space
  const Cx<T>* tw;     // [N] exp(-2*pi*i*m/N)
  T beta, gdt, munu;   // beta_s, gamma_s*dt, mu_s*nu
  T scale;             // 1/(N*M), folded into the prologue
  int dealias;
  int store_h;         // 0 in the last sub-stage: the next RK step starts with beta = 0 and never reads h
  int l2_hints;        // 1: omega_hat / h tiles are loaded and stored with the L2 evict_first policy (they are touched
                       // once per sub-stage and only by the column kernel) so that the x-space fields and the
                       // advection rows -- produced by one kernel and consumed by the next -- keep the L2
};

// L2 cache policies (createpolicy): evict_first for streamed data, evict_normal otherwise
__device__ __forceinline__ uint64_t l2_policy(bool evict_first) {
  uint64_t p;
  if (evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 64-bit / 128-bit global stores with an L2 policy
__device__ __forceinline__ void stc_hint(Cx<float>* p, Cx<float> a, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v2.f32 [%0], {%1, %2}, %3;" ::"l"(p), "f"(a.x), "f"(a.y), "l"(pol) : "memory");
}
__device__ __forceinline__ void stc_hint(Cx<double>* p, Cx<double> a, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(a.x), "d"(a.y), "l"(pol) : "memory");
}

// exchange policy of the column kernel: TC columns interleaved, element (pos, c) at pos * TC + c.  Reads (pos = t +
// TPF * m) of a warp's 32 / TC consecutive t are side by side; the pass-0 scatters (pos = 16 t + r) of adjacent t are
// 16 * TC elements apart, a 2-way conflict for 8-byte elements at TC = 8.  An XOR swizzle of the position (bit 0 with
// bit 4) removes it within the same footprint but costs more address arithmetic than it saves: 256^2 x 256 fp32
// 0.881 -> 0.909 ms per RK step, 512^2 1.235 -> 1.348 (profiles/r2_exchange_layout_ab.txt).  Left as it is.
template <typename X, int TC> struct ColEx {
  static constexpr bool kSplit = false;
  Cx<X>* sm; int c;
  __device__ __forceinline__ void store(int pos, Cx<X> a) { stc(sm + pos * TC + c, a); }
  __device__ __forceinline__ Cx<X> load(int pos) { return ldc(sm + pos * TC + c); }
  __device__ __forceinline__ void sync() { __syncthreads(); }
};

// split variant for lane-paired data: real parts, then imaginary parts, through N x TC lane values
template <typename X, int TC> struct ColExSplit {
  static constexpr bool kSplit = true;
  X* sm; int c;
  __device__ __forceinline__ void store_part(int pos, X a) { stl(sm + pos * TC + c, a); }
  __device__ __forceinline__ X load_part(int pos) { return ldl(sm + pos * TC + c); }
  __device__ __forceinline__ void sync() { __syncthreads(); }
};

// exchange policy of the row kernels: one padded buffer per FFT group.
// Lane-paired complex128 elements (Cx<D2>, 32 bytes) are kept as TWO PLANES of 16-byte lane pairs (x parts in
// [0, n), y parts in [n, 2n), n = elements in the buffer) when n > 0: 128-bit shared accesses are served per
// quarter-warp, and 8 threads on consecutive 32-byte elements hit every bank group twice (ncu on the fp64 row kernel:
// 40 % of the shared wavefronts were conflicts, LSU pipe at 81 %); on consecutive 16-byte slots they are conflict-free,
// and so are the pass-0 scatters (element stride 17 after padding).  Same footprint as the interleaved layout.
template <typename X, int TPF> struct RowEx {
  static constexpr bool kSplit = false;
  static constexpr bool kPlanar = std::is_same<X, D2>::value;
  Cx<X>* sm;
  int bar = 0;  // > 0: named barrier of this group's TPF threads (groups of one CTA run out of phase); 0: whole CTA
  int n = 0;    // Cx<D2> only: > 0 selects the two-plane layout over n elements
  __device__ __forceinline__ static int pad(int pos) { return pos + (pos >> 4); }
  __device__ __forceinline__ void store(int pos, Cx<X> a) {
    if constexpr (kPlanar) {
      if (n > 0) {
        X* pl = reinterpret_cast<X*>(sm);
        stl(pl + pad(pos), a.x);
        stl(pl + n + pad(pos), a.y);
        return;
      }
    }
    stc(sm + pad(pos), a);
  }
  __device__ __forceinline__ Cx<X> load(int pos) {
    if constexpr (kPlanar) {
      if (n > 0) {
        const X* pl = reinterpret_cast<const X*>(sm);
        return mk<X>(ldl(pl + pad(pos)), ldl(pl + n + pad(pos)));
      }
    }
    return ldc(sm + pad(pos));
  }
  __device__ __forceinline__ void sync() {
    if (TPF <= 32) __syncwarp();
    else if (bar == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(TPF) : "memory");
  }
};

// 1/x: MUFU.RCP + one Newton step for fp32 (branch-free, ~1 ulp); IEEE division for fp64
__device__ __forceinline__ float fast_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return fmaf(r, fmaf(-x, r, 1.0f), r);
}
__device__ __forceinline__ double fast_rcp(double x) { return 1.0 / x; }

// integer wavenumber p of row kx = t + TPF*m as a real: the wrap at N/2 depends on m only
template <typename T, class P> __device__ __forceinline__ T wave_p(T tf, int m) {
  return tf + (T)(P::TPF * m - (m >= P::E / 2 ? P::L : 0));
}

template <typename T> __device__ __forceinline__ T inv_k2(T p, T q) {
  // k2 is an integer >= 1 except at p = q = 0, where the coefficient it scales is 0 anyway
  return fast_rcp(fmax(p * p + q * q, (T)1));
}

// Lane-paired coefficients c with field_hat = c * (i * omega_hat), given the integer wavenumbers p, q
// (kolmogorov.py:236-246 with 2*pi*kx = p, 2*pi*ky = q; `scale` = 1/(N*M) of the inverse transform):
//   PAIR 0: (u_hat = i q w / k^2,  gx_hat = i p w)      PAIR 1: (v_hat = -i p w / k^2,  gy_hat = i q w)
template <int PAIR, typename T> __device__ __forceinline__ typename PairOf<T>::type pair_coef(T p, T q, T scale) {
  using L2 = typename PairOf<T>::type;
  const T sp = p * scale, sq = q * scale, ik2 = inv_k2<T>(p, q);
  return PAIR == 0 ? L2(sq * ik2, sp) : L2(-(sp * ik2), sq);
}
// cf * (i * w) over both lanes
template <typename T> __device__ __forceinline__ Cx<typename PairOf<T>::type> coef_times_iw(typename PairOf<T>::type cf,
                                                                                          Cx<T> w) {
  using L2 = typename PairOf<T>::type;
  return mk<L2>(cf * bc<L2>(-w.y), cf * bc<L2>(w.x));
}

// ky = M/2 column of a tile-0 CTA, all threads cooperating (one kx per thread): unpack
// An[kx] = (W[kx] - conj(W[-kx]))/(2i) from the packed forward transform W (in s_w[0..N)),
// then the same RK accumulate + Crank-Nicolson update as the main columns.  New omega_hat
// values go to global memory and to s_nyq for the prologue.
template <typename T, int N>
__device__ __forceinline__ void nyquist_column_update(const StageArgs<T>& a, const Cx<T>* s_w, Cx<T>* s_nyq,
                                                      Cx<T>* hnyq, Cx<T>* wnyq, const Cx<T>* G, bool use_h,
                                                      int tid, int nthreads) {
  constexpr int MH = N / 2;
  for (int kx = tid; kx < N; kx += nthreads) {
    const Cx<T> d = ldc(s_w + kx) - conj(ldc(s_w + ((N - kx) & (N - 1))));
    Cx<T> h = mk<T>((T)0.5 * d.y, (T)-0.5 * d.x);
    if (a.dealias) h = mk<T>((T)0, (T)0);  // ky = M/2 >= km always
    if (kx == 0) h = h + ldc(G + MH);
    if (a.ctrl_nyq) h = h + ldc(a.ctrl_nyq + kx);
    if (use_h) h = h + a.beta * ldc(hnyq + kx);
    if (a.store_h) stc(hnyq + kx, h);
    const T p = (T)(kx < N / 2 ? kx : kx - N);
    const T lk = a.munu * (p * p + (T)MH * (T)MH);
    Cx<T> wv = ldc(wnyq + kx);
    wv = fast_rcp((T)1 + lk) * (wv + a.gdt * h - lk * wv);
    stc(wnyq + kx, wv);
    stc(s_nyq + kx, wv);
  }
}

enum ColMode { kColStage = 0, kColInit = 1 };

// Shared-memory bytes of the column kernels' exchange buffer: N x TC scalar complex numbers, which is
// also N x TC lane PAIRS (the lane-paired FFTs exchange real and imaginary parts in two rounds).
template <typename T, int N, int TC> __host__ __device__ constexpr size_t col_exchange_bytes() {
  static_assert(TC >= 4, "the packed-column scratch (2 x N paired complex) must fit in the exchange buffer");
  return (size_t)N * TC * sizeof(Cx<T>);
}

// ---------------------------------------------------------------------------
// Prologue of the NEXT nonlinear term for one column: lane pair PAIR of the four
// spectral fields of omega_hat (w[] in registers), inverse x-FFT, store to FA/FB.
// Thread (c, t) owns column q = tile*TC + c and rows {t + TPF*m}.
// ---------------------------------------------------------------------------
// SPLIT: the exchange buffer holds N x TC lane values (real and imaginary parts go through it in two rounds, four
// barriers per exchange); !SPLIT: it holds N x TC lane-paired complex numbers (one round, two barriers).
template <int PAIR, typename T, class P, int TC, bool SPLIT = true, class LoadW>
__device__ __forceinline__ void col_field_pair(const StageArgs<T>& a, LoadW load_w, const Cx<T>* s_tw,
                                               const Cx<T>* s_nyq, void* s_exchange, int c, int t, int q, int b,
                                               bool special) {
  using L2 = typename PairOf<T>::type;
  constexpr int N = P::L, E = P::E, TPF = P::TPF, MH = N / 2;
  const T tf = (T)t, qf = (T)q;
  Cx<L2> v[E];
#pragma unroll
  for (int m = 0; m < E; ++m)
    v[m] = coef_times_iw<T>(pair_coef<PAIR, T>(wave_p<T, P>(tf, m), qf, a.scale), load_w(m, t + TPF * m));
  if (special) {
    // column 0 <- herm(field[:,0]) + i*herm(field[:,M/2]); herm(a)[kx] = (a[kx]+conj(a[-kx]))/2
    Cx<L2>* sA = reinterpret_cast<Cx<L2>*>(s_exchange);
    Cx<L2>* sB = sA + N;
    if (c == 0) {
#pragma unroll
      for (int m = 0; m < E; ++m) {
        const int kx = t + TPF * m;
        stc(sA + kx, v[m]);
        stc(sB + kx, coef_times_iw<T>(pair_coef<PAIR, T>(wave_p<T, P>(tf, m), (T)MH, a.scale), ldc(s_nyq + kx)));
      }
    }
    __syncthreads();
    if (c == 0) {
#pragma unroll
      for (int m = 0; m < E; ++m) {
        const int kx = t + TPF * m;
        const int mkx = (N - kx) & (N - 1);
        const Cx<L2> ah = add_conj(v[m], ldc(sA + mkx));           // 2 * herm(A)
        const Cx<L2> bh = add_conj(ldc(sB + kx), ldc(sB + mkx));   // 2 * herm(B)
        v[m] = mk<L2>((ah.x - bh.y) * L2(0.5), (ah.y + bh.x) * L2(0.5));
      }
    }
    __syncthreads();
  }
  if constexpr (SPLIT) {
    ColExSplit<L2, TC> ex{reinterpret_cast<L2*>(s_exchange), c};
    fft_regs<P, +1, L2>(v, s_tw, t, ex);
  } else {
    ColEx<L2, TC> ex{reinterpret_cast<Cx<L2>*>(s_exchange), c};
    fft_regs<P, +1, L2>(v, s_tw, t, ex);
  }
  Cx<L2>* out = (PAIR == 0 ? a.fa : a.fb) + (size_t)b * N * MH + q;
#pragma unroll
  for (int m = 0; m < E; ++m) stc(out + (size_t)(t + TPF * m) * MH, v[m]);
}

// RK accumulate + Crank-Nicolson update of one column (S:53-56, K:209-220): v[] holds the forward
// x-transform of the advection column (column 0 already unpacked); load_h / load_w fetch the old
// h / omega_hat of row kx, store_h / store_w dispose of the new ones, extra(m, kx) is a further source term added
// after the mask (the dense control spectrum; callers without one pass zero_extra).  Returns the new omega_hat in w[].
template <typename T> struct ZeroExtra {
  __device__ __forceinline__ Cx<T> operator()(int, int) const { return mk<T>((T)0, (T)0); }
};
template <typename T, class P, class Args, class StoreH, class StoreW, class LoadH, class LoadW, class Extra>
__device__ __forceinline__ void rk_update_column(const Args& a, const Cx<T>* v, Cx<T>* w, Cx<T> g0, bool use_h,
                                                 int t, int q, StoreH store_h, StoreW store_w, LoadH load_h,
                                                 LoadW load_w, Extra extra) {
  constexpr int N = P::L, E = P::E, TPF = P::TPF, MH = N / 2;
  constexpr int kn = (N / 2) * 2 / 3, km = ((MH + 1) * 2) / 3;
  const T tf = (T)t, qf = (T)q;
#pragma unroll
  for (int m = 0; m < E; ++m) {
    const int kx = t + TPF * m;
    Cx<T> h = v[m];
    if (a.dealias && ((kx >= kn && kx < 2 * kn) || q >= km)) h = mk<T>((T)0, (T)0);
    if (m == 0) h = h + g0;  // g0 is zero except on row kx = 0 (t == 0)
    h = h + extra(m, kx);
    if (use_h) h = h + a.beta * load_h(m, kx);
    store_h(m, kx, h);
    const T p = wave_p<T, P>(tf, m);
    const T lk = a.munu * (p * p + qf * qf);
    const Cx<T> wv = load_w(m, kx);
    w[m] = fast_rcp((T)1 + lk) * (wv + a.gdt * h - lk * wv);
    store_w(m, kx, w[m]);
  }
}

// ---------------------------------------------------------------------------
// Column kernel (direct-load variant).  grid = (M/2 / TC, Bc), block = TPF*TC threads;
// every global access is a run of TC consecutive elements per row.  Tile 0 also carries
// the ky = M/2 column, packed with ky = 0 in x-space.
//   MODE kColStage : epilogue of a sub-stage + prologue (4 fields) for the next one
//   MODE kColInit  : prologue only from the stored omega_hat (first stage of a rollout,
//                    observation of a stored trajectory)
// The RK sub-stages normally run through col_tma_kernel (col_tma.cuh); this kernel is the
// KF_NO_TMA fallback and the prologue-only path.
// ---------------------------------------------------------------------------
template <typename T, class P, int TC, int MODE, int MINB>
__global__ void __launch_bounds__(P::TPF* TC, MINB) col_kernel(StageArgs<T> a) {
  constexpr int N = P::L, E = P::E, TPF = P::TPF, MH = N / 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Cx<T>* s_sc = reinterpret_cast<Cx<T>*>(smem_raw);  // exchange buffer, scalar-complex view
  Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(smem_raw + col_exchange_bytes<T, N, TC>());
  Cx<T>* s_nyq = s_tw + N;
  const int c = threadIdx.x % TC, t = threadIdx.x / TC;
  const int b = blockIdx.y;
  const int q = blockIdx.x * TC + c;
  const bool special = (blockIdx.x == 0);  // CTA-uniform
  const size_t plane = (size_t)N * MH;
  Cx<T>* wmain = a.wmain + b * plane + q;
  Cx<T>* wnyq = a.wnyq + (size_t)b * N;

  for (int i = threadIdx.x; i < N; i += TPF * TC) s_tw[i] = ldc(a.tw + i);
  __syncthreads();

  Cx<T> w[E];
  if constexpr (MODE == kColStage) {
    Cx<T>* hmain = a.hmain + b * plane + q;
    Cx<T>* hnyq = a.hnyq + (size_t)b * N;
    const Cx<T>* G = a.G + (size_t)b * (MH + 1);
    const Cx<T>* advx = a.advx + b * plane + q;
    const bool use_h = (a.beta != (T)0);
    Cx<T> v[E];
    // with the 2/3 mask on, the advection term of every column ky >= km is discarded: such tiles neither read
    // it nor transform it (the row kernel does not write those columns either)
    constexpr int km = ((MH + 1) * 2) / 3;
    const bool masked_tile = a.dealias && (int)blockIdx.x * TC >= km;  // CTA-uniform
    if (!masked_tile) {
#pragma unroll
      for (int m = 0; m < E; ++m) v[m] = ldc(advx + (size_t)(t + TPF * m) * MH);
      ColEx<T, TC> ex{s_sc, c};
      fft_regs<P, -1, T>(v, s_tw, t, ex);
    } else {
#pragma unroll
      for (int m = 0; m < E; ++m) v[m] = mk<T>((T)0, (T)0);
    }

    if (special) {
      // unpack column 0: A0[kx] = (W[kx] + conj(W[-kx]))/2, An[kx] = (W[kx] - conj(W[-kx]))/(2i)
      if (c == 0) {
#pragma unroll
        for (int m = 0; m < E; ++m) stc(s_sc + t + TPF * m, v[m]);
      }
      __syncthreads();
      nyquist_column_update<T, N>(a, s_sc, s_nyq, hnyq, wnyq, G, use_h, threadIdx.x, TPF * TC);
      if (c == 0) {
#pragma unroll
        for (int m = 0; m < E; ++m)
          v[m] = (T)0.5 * add_conj(v[m], ldc(s_sc + ((N - (t + TPF * m)) & (N - 1))));
      }
      __syncthreads();
    }
    Cx<T> g0 = mk<T>((T)0, (T)0);
    if (t == 0) g0 = ldc(G + q);
    rk_update_column<T, P>(a, v, w, g0, use_h, t, q,
                           [&](int, int kx, Cx<T> h) { if (a.store_h) stc(hmain + (size_t)kx * MH, h); },
                           [&](int, int kx, Cx<T> wn) { stc(wmain + (size_t)kx * MH, wn); },
                           [&](int, int kx) { return ldc(hmain + (size_t)kx * MH); },
                           [&](int, int kx) { return ldc(wmain + (size_t)kx * MH); },
                           [&](int, int kx) {
                             return a.ctrl_main ? ldc(a.ctrl_main + (size_t)kx * MH + q) : mk<T>((T)0, (T)0);
                           });
  } else {
#pragma unroll
    for (int m = 0; m < E; ++m) w[m] = ldc(wmain + (size_t)(t + TPF * m) * MH);
    if (special) {
      for (int i = threadIdx.x; i < N; i += TPF * TC) s_nyq[i] = ldc(wnyq + i);
      __syncthreads();
    }
  }

  auto load_w = [&](int m, int) { return w[m]; };
  col_field_pair<0, T, P, TC>(a, load_w, s_tw, s_nyq, s_sc, c, t, q, b, special);
  col_field_pair<1, T, P, TC>(a, load_w, s_tw, s_nyq, s_sc, c, t, q, b, special);
}

// ---------------------------------------------------------------------------
// Row kernels.  A group of TPF threads owns two x-rows (x0, x0+1).  Per row ONE
// lane-paired inverse FFT of length M gives (u + i v) in lane 0 and (gx + i gy) in
// lane 1 -- four real c2r transforms per paired complex FFT -- and the two rows'
// advection values go through ONE forward complex FFT, W = FFT(adv[x0] + i adv[x0+1]),
// split afterwards by Hermitian symmetry.  Global loads are direct (coalesced both
// for k and for the mirrored index M-k); the only shared-memory traffic is the FFT
// exchange.
// ---------------------------------------------------------------------------

// Z[k] = A[k] + i B[k] for the Hermitian extensions of the packed half-rows A, B (any lane type);
// element k of a half-row lives at offset at(k) (identity for the batched layout, blocked for slabs)
template <typename X, class P, class At>
__device__ __forceinline__ void load_pair_hermitian_at(Cx<X>* z, const Cx<X>* __restrict__ A,
                                                       const Cx<X>* __restrict__ B, int t, At at) {
  constexpr int M = P::L, MH = M / 2, E = P::E, TPF = P::TPF;
#pragma unroll
  for (int m = 0; m < E; ++m) {
    const int k = t + TPF * m;
    if (m < E / 2) {
      const Cx<X> a = ldc(A + at(k)), b = ldc(B + at(k));
      const Cx<X> z0 = mk<X>(a.x, b.x), zk = mk<X>(a.x - b.y, a.y + b.x);
      z[m] = (k == 0) ? z0 : zk;
    } else {
      const int src = (k == MH) ? 0 : M - k;
      const Cx<X> a = ldc(A + at(src)), b = ldc(B + at(src));
      const Cx<X> zh = mk<X>(a.y, b.y), zk = mk<X>(a.x + b.y, b.x - a.y);
      z[m] = (k == MH) ? zh : zk;
    }
  }
}

template <typename X, class P>
__device__ __forceinline__ void load_pair_hermitian(Cx<X>* z, const Cx<X>* __restrict__ A, const Cx<X>* __restrict__ B,
                                                    int t) {
  load_pair_hermitian_at<X, P>(z, A, B, t, [](int k) { return k; });
}

// Programmatic dependent launch (the host sets cudaLaunchAttributeProgrammaticStreamSerialization when KF_PDL=1): a
// kernel lets its successor in the stream start early (launch_dependents, first instruction) and the successor runs
// its own prologue -- twiddle table to shared memory, mbarrier setup -- while this grid drains; it must not touch
// anything the predecessor reads or writes before pdl_wait(), which returns once the predecessor grid has completed
// and its writes are visible.  Both are no-ops in a launch without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// L2 prefetch of `bytes` (multiple of 16) starting at p: no destination, no completion to wait for
__device__ __forceinline__ void prefetch_l2_bulk(const void* p, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

#ifndef KF_ROW_AHEAD
// Experiment knob: L2 bulk prefetch of the rows of the CTA `KF_ROW_AHEAD` launches ahead.  Measured on B200 at
// 256^2 x 256 envs fp32: row kernel 73.9 us without, 82.1 / 88.4 / 92.8 us with 300 / 592 / 1184 ahead -- the
// prefetches displace the column kernel's outputs that the row kernel would otherwise still find in L2.  Off.
#define KF_ROW_AHEAD 0
#endif

template <typename T, int M> __host__ __device__ constexpr size_t row_exchange_elems() { return (size_t)M + M / 16; }

template <typename T, class P, int GROUPS, int MINB>
__global__ void __launch_bounds__(P::TPF* GROUPS, MINB) row_kernel(StageArgs<T> a) {
  using L2 = typename PairOf<T>::type;
  constexpr int M = P::L, MH = M / 2, E = P::E, TPF = P::TPF, N = M;
  constexpr int EXN = (int)row_exchange_elems<T, M>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t = threadIdx.x % TPF;
  Cx<L2>* s_ex = reinterpret_cast<Cx<L2>*>(smem_raw) + (size_t)g * EXN;
  Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(smem_raw + (size_t)GROUPS * EXN * sizeof(Cx<L2>));
  RowEx<L2, TPF> ex2{s_ex, 0, EXN};
  RowEx<T, TPF> ex1{reinterpret_cast<Cx<T>*>(s_ex)};
  // the column kernel walks the environments in ascending order, so its LAST outputs are the ones
  // still in L2: start from the last environment
  const int b = gridDim.y - 1 - blockIdx.y;
  const int x0 = 2 * (blockIdx.x * GROUPS + g);
  const size_t plane = (size_t)N * MH;
  pdl_launch_dependents();

  if (KF_ROW_AHEAD > 0 && threadIdx.x == 0) {
    // the rows of the CTA that will run KF_ROW_AHEAD launches later (2*GROUPS contiguous rows of fa and of fb):
    // pull them into L2 now so that its loads see L2 latency instead of HBM latency
    const int lin = blockIdx.y * gridDim.x + blockIdx.x + KF_ROW_AHEAD;
    const int by = lin / gridDim.x, bx = lin - by * gridDim.x;
    if (by < gridDim.y) {
      const size_t off = (size_t)(gridDim.y - 1 - by) * plane + (size_t)(2 * GROUPS * bx) * MH;
      constexpr unsigned kBytes = 2 * GROUPS * MH * sizeof(Cx<L2>);
      prefetch_l2_bulk(a.fa + off, kBytes);
      prefetch_l2_bulk(a.fb + off, kBytes);
    }
  }
  for (int i = threadIdx.x; i < M; i += TPF * GROUPS) s_tw[i] = ldc(a.tw + i);
  __syncthreads();
  pdl_wait();  // the fields below are the column kernel's output; the advection rows we write are its input

  if (a.l2_hints & 2) {
    // experiment: pull the SECOND row of every group into L1 now, so that its loads (issued after the first row's
    // transform) see L1 latency; 2 * GROUPS rows x 2 arrays of MH lane-paired complex numbers, 128-byte lines
    constexpr int kLinesPerRow = (int)(MH * sizeof(Cx<L2>) / 128);
    for (int i = threadIdx.x; i < GROUPS * 2 * kLinesPerRow; i += TPF * GROUPS) {
      const int gg = i / (2 * kLinesPerRow), rest = i - gg * 2 * kLinesPerRow;
      const int arr = rest / kLinesPerRow, line = rest - arr * kLinesPerRow;
      const size_t off = b * plane + (size_t)(2 * (blockIdx.x * GROUPS + gg) + 1) * MH;
      const char* src = reinterpret_cast<const char*>((arr ? a.fb : a.fa) + off) + (size_t)line * 128;
      asm volatile("prefetch.global.L1 [%0];" ::"l"(src));
    }
  }

  Cx<T> v[E];
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const size_t roff = b * plane + (size_t)(x0 + rr) * MH;
    Cx<L2> z[E];
    load_pair_hermitian<L2, P>(z, a.fa + roff, a.fb + roff, t);
    fft_regs<P, +1, L2>(z, s_tw, t, ex2);
    // lanes: z.x = (u, gx), z.y = (v, gy)  ->  adv = -(gx*u + gy*v)   (kolmogorov.py:249)
#pragma unroll
    for (int m = 0; m < E; ++m) {
      const T adv = -(z[m].x.lo() * z[m].x.hi() + z[m].y.lo() * z[m].y.hi());
      if (rr == 0) v[m].x = adv; else v[m].y = adv;
    }
  }

  fft_regs<P, -1, T>(v, s_tw, t, ex1);
#pragma unroll
  for (int m = 0; m < E; ++m) ex1.store(t + TPF * m, v[m]);
  ex1.sync();
  Cx<T>* o0 = a.advx + b * plane + (size_t)x0 * MH;
  Cx<T>* o1 = o0 + MH;
  constexpr int km = ((MH + 1) * 2) / 3;
#pragma unroll
  for (int m = 0; m < E / 2; ++m) {
    const int k = t + TPF * m;
    if (a.dealias && k >= km) continue;  // masked by the column kernel anyway: never read, so never written
    Cx<T> r0, r1;
    if (k == 0) {
      const Cx<T> wn = ex1.load(MH);
      r0 = mk<T>(v[m].x, wn.x);
      r1 = mk<T>(v[m].y, wn.y);
    } else {
      const Cx<T> wm = ex1.load(M - k);
      const Cx<T> s = add_conj(v[m], wm), d = sub_conj(v[m], wm);
      r0 = (T)0.5 * s;
      r1 = mk<T>((T)0.5 * d.y, (T)-0.5 * d.x);
    }
    stc(o0 + k, r0);
    stc(o1 + k, r1);
  }
}

// ---------------------------------------------------------------------------
// Reward / observation row kernel (reference: compute_tke utils.py:175-194 and
// _trajectory_mean_obs kolmogorov.py:426-446).  Consumes lane 0 (u_hat, v_hat) of the
// x-space fields of one snapshot; per environment accumulates sum u, sum v, sum u^2,
// sum v^2 (fp64, warp-shuffle reduction + one atomic per warp) and adds
// |velocity| at the obs_size x obs_size strided probes into the running obs mean.
// ---------------------------------------------------------------------------
template <typename T> struct ObsArgs {
  using L2 = typename PairOf<T>::type;
  const Cx<L2>* fa;    // [Bc][N][M/2]
  const Cx<L2>* fb;
  const Cx<T>* tw;
  double* acc;         // [Bc][4] for this snapshot (stride acc_stride between envs)
  size_t acc_stride;
  T* obs;              // [Bc][nobs*nobs], accumulated with weight obs_w
  int nobs, sx, sy;
  T obs_w;
};

template <typename T, class P, int GROUPS>
__global__ void __launch_bounds__(P::TPF* GROUPS) row_obs_kernel(ObsArgs<T> a) {
  using L2 = typename PairOf<T>::type;
  constexpr int M = P::L, MH = M / 2, E = P::E, TPF = P::TPF, N = M;
  constexpr int EXN = (int)row_exchange_elems<T, M>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t = threadIdx.x % TPF;
  Cx<L2>* s_ex = reinterpret_cast<Cx<L2>*>(smem_raw) + (size_t)g * EXN;
  Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(smem_raw + (size_t)GROUPS * EXN * sizeof(Cx<L2>));
  RowEx<L2, TPF> ex{s_ex, 0, EXN};
  const int b = blockIdx.y;
  const int x = blockIdx.x * GROUPS + g;
  const size_t roff = ((size_t)b * N + x) * MH;

  for (int i = threadIdx.x; i < M; i += TPF * GROUPS) s_tw[i] = ldc(a.tw + i);
  __syncthreads();

  Cx<L2> z[E];
  load_pair_hermitian<L2, P>(z, a.fa + roff, a.fb + roff, t);
  fft_regs<P, +1, L2>(z, s_tw, t, ex);

  T su = 0, sv = 0, suu = 0, svv = 0;
#pragma unroll
  for (int m = 0; m < E; ++m) {
    const T u = z[m].x.lo(), vv = z[m].y.lo();
    su += u; sv += vv;
    suu += u * u; svv += vv * vv;
  }
  if (x % a.sx == 0 && x / a.sx < a.nobs) {
#pragma unroll
    for (int m = 0; m < E; ++m) {
      const int y = t + TPF * m;
      if (y % a.sy == 0 && y / a.sy < a.nobs) {
        const T u = z[m].x.lo(), vv = z[m].y.lo();
        T* o = a.obs + (size_t)b * a.nobs * a.nobs + (x / a.sx) * a.nobs + y / a.sy;
        *o += a.obs_w * sqrt(u * u + vv * vv);  // single writer per launch
      }
    }
  }
  double r[4] = {(double)su, (double)sv, (double)suu, (double)svv};
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
    for (int i = 0; i < 4; ++i) r[i] += __shfl_xor_sync(0xffffffffu, r[i], off);
  }
  if ((threadIdx.x & 31) == 0) {
    double* acc = a.acc + (size_t)b * a.acc_stride;
#pragma unroll
    for (int i = 0; i < 4; ++i) atomicAdd(acc + i, r[i]);
  }
}
// Physical-space fields of one snapshot (diagnostics: compute_energy_dissipation / compute_divergence /
// compute_real_velocity_point, utils.py:129-172, 213-229, take irfftn of u_hat, v_hat): the c2r y-transform of the
// x-space fields written out as real arrays out[f][b][x][y], f = 0..3 -> u, v, d(omega)/dx, d(omega)/dy.
template <typename T, class P, int GROUPS>
__global__ void __launch_bounds__(P::TPF* GROUPS) row_fields_kernel(ObsArgs<T> a, T* out, int nfields, size_t fstride) {
  using L2 = typename PairOf<T>::type;
  constexpr int M = P::L, MH = M / 2, E = P::E, TPF = P::TPF, N = M;
  constexpr int EXN = (int)row_exchange_elems<T, M>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t = threadIdx.x % TPF;
  Cx<L2>* s_ex = reinterpret_cast<Cx<L2>*>(smem_raw) + (size_t)g * EXN;
  Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(smem_raw + (size_t)GROUPS * EXN * sizeof(Cx<L2>));
  RowEx<L2, TPF> ex{s_ex, 0, EXN};
  const int b = blockIdx.y;
  const int x = blockIdx.x * GROUPS + g;
  const size_t roff = ((size_t)b * N + x) * MH;
  for (int i = threadIdx.x; i < M; i += TPF * GROUPS) s_tw[i] = ldc(a.tw + i);
  __syncthreads();
  Cx<L2> z[E];
  load_pair_hermitian<L2, P>(z, a.fa + roff, a.fb + roff, t);
  fft_regs<P, +1, L2>(z, s_tw, t, ex);
  T* o = out + ((size_t)b * N + x) * M;
#pragma unroll
  for (int m = 0; m < E; ++m) {
    const int y = t + TPF * m;
    o[y] = z[m].x.lo();
    if (nfields > 1) o[fstride + y] = z[m].y.lo();
    if (nfields > 2) o[2 * fstride + y] = z[m].x.hi();
    if (nfields > 3) o[3 * fstride + y] = z[m].y.hi();
  }
}

// mean_tke[b] = mean_j 0.5*(sum((u-U)^2) + sum((v-V)^2))/n^2 ; nonfinite flag
template <typename T>
__global__ void tke_finalize_kernel(const double* acc, int B, int nsnap, int N, int M, T* mean_tke, int* nonfinite) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double cnt = (double)N * (double)M;
  double s = 0.0;
  for (int j = 0; j < nsnap; ++j) {
    const double* r = acc + ((size_t)b * nsnap + j) * 4;
    s += 0.5 * ((r[2] - r[0] * r[0] / cnt) + (r[3] - r[1] * r[1] / cnt)) / ((double)N * (double)N);
  }
  s /= (double)nsnap;
  if (mean_tke) mean_tke[b] = (T)s;
  if (nonfinite) nonfinite[b] = isfinite(s) ? 0 : 1;
}

// ---------------------------------------------------------------------------
// Layout conversion and the per-step forcing row.
// ---------------------------------------------------------------------------
// user omega_hat [B][N][M/2+1] (row stride = user_row, env stride = user_env) <-> internal
template <typename T>
__global__ void import_kernel(const Cx<T>* __restrict__ user, size_t user_env, Cx<T>* main, Cx<T>* nyq, int N, int MH) {
  const int b = blockIdx.y;
  const int total = N * (MH + 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int kx = i / (MH + 1), q = i - kx * (MH + 1);
    const Cx<T> v = ldc(user + (size_t)b * user_env + i);
    if (q == MH) stc(nyq + (size_t)b * N + kx, v);
    else stc(main + ((size_t)b * N + kx) * MH + q, v);
  }
}
template <typename T>
__global__ void export_kernel(Cx<T>* __restrict__ user, size_t user_env, const Cx<T>* main, const Cx<T>* nyq, int N, int MH) {
  const int b = blockIdx.y;
  const int total = N * (MH + 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int kx = i / (MH + 1), q = i - kx * (MH + 1);
    const Cx<T> v = (q == MH) ? ldc(nyq + (size_t)b * N + kx) : ldc(main + ((size_t)b * N + kx) * MH + q);
    stc(user + (size_t)b * user_env + i, v);
  }
}

// G[b][q] = F[q] + sum_m clip(a[b][m]) * C[m][q]   (kolmogorov.py:386-395, 259-272, 274-292, env_core.py:419-420)
template <typename T>
__global__ void forcing_row_kernel(Cx<T>* G, const Cx<T>* F, const Cx<T>* C, const T* actions, T lo, T hi, int MH1) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= MH1) return;
  Cx<T> g = ldc(F + q);
  if (actions) {
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      T am = actions[(size_t)b * 4 + m];
      am = am < lo ? lo : (am > hi ? hi : am);
      g = g + am * ldc(C + (size_t)m * MH1 + q);
    }
  }
  stc(G + (size_t)b * MH1 + q, g);
}

}  // namespace kf
