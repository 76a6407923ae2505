// On-chip-resident RK solver steps for small grids (sm_100a): ONE thread-block cluster per environment keeps
// omega_hat, the RK register h, one lane pair of x-space fields and the advection rows in (distributed) shared
// memory and runs ALL sub-stages of ALL requested RK steps without touching global memory -- no launch per
// sub-stage, no L2 round trip between the x- and y-transforms.  Cluster size CS = 1 (64^2: one CTA per environment),
// 2 (128^2 fp32), ...; with CS > 1 the column <-> row transposition goes through distributed shared memory
// (mapa + st.shared::cluster) and the phases are separated by barrier.cluster.
//
// Same arithmetic as the two-kernel path (kernels.cuh; reference RK4_CN.time_step_fn hydrogym/jax/solvers/base.py:50-57
// over nonlinear_terms / linear_terms / implicit_timestep hydrogym/jax/envs/kolmogorov.py:209-257), cut differently:
//
//   K layout  rank r owns MHc = (N/2)/CS packed ky columns, all kx, stored column-major [c][kx] (+pad):
//             W (omega_hat), H (RK register), ADV (advection after the y-transform); ky = N/2 lives beside
//             packed column 0 on rank 0 (WN, HN).
//   X layout  rank r owns Nc = N/CS rows x, all packed ky: FX [row][ky] lane-paired complex (one PAIR at a time:
//             (u_hat, gx_hat), then (v_hat, gy_hat)).
//
//   per sub-stage:  K1  pair 0 of the 4 spectral fields of omega_hat -> inverse x-FFT -> scatter into FX
//                   X1  rows (x0, x0+1): ONE lane-paired inverse y-FFT of FX[x0] + i FX[x0+1] -> u*gx kept in registers
//                   K2  pair 1 -> inverse x-FFT -> FX
//                   X2  inverse y-FFT -> adv = -(u*gx + v*gy) -> forward y-FFT of adv[x0] + i adv[x0+1] -> split by
//                       Hermitian symmetry -> scatter into ADV (K layout)
//                   K3  forward x-FFT of ADV -> unpack ky = 0 / N/2 -> + forcing/control on row kx = 0 -> RK
//                       accumulate + Crank-Nicolson update of H, W
//   with one cluster barrier between consecutive phases (4 per sub-stage).  Every transform belongs to one group of TPF
//   threads inside a warp: exchanges are private padded buffers + __syncwarp, never a block barrier.
#pragma once
#include "kernels.cuh"
#include "slab.cuh"  // RowExSplit

namespace kf {

template <typename T> struct ResidentArgs {
  Cx<T>* omega;     // [B][N][N/2+1] user layout, advanced in place
  const Cx<T>* G;   // [B][N/2+1] forcing + control row (kx = 0)
  const Cx<T>* tw;  // [N] transposed twiddle table of the plan (fft.cuh)
  T beta[5], gdt[5], munu[5];
  T scale;          // 1/(N*N)
  int dealias;
  int n_steps;
};

template <int CS> __device__ __forceinline__ unsigned cluster_ctarank() {
  if constexpr (CS == 1) return 0u;
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// all threads of all CTAs of the cluster; orders shared-memory (incl. remote) writes before the reads that follow
template <int CS> __device__ __forceinline__ void cluster_barrier() {
  if constexpr (CS == 1) {
    __syncthreads();
  } else {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
}
// store `val` at the shared-memory location `p` of CTA `rank` of the cluster (p is this CTA's own address of it)
template <int CS, typename V> __device__ __forceinline__ void st_rank(V* p, unsigned rank, V val) {
  static_assert(sizeof(V) % 8 == 0 && sizeof(V) <= 32, "8-, 16- or 32-byte values");
  const unsigned long long* w = reinterpret_cast<const unsigned long long*>(&val);
  if constexpr (CS == 1) {
    (void)rank;
    if constexpr (sizeof(V) == 8) {
      *reinterpret_cast<unsigned long long*>(p) = w[0];
    } else {
      reinterpret_cast<ulonglong2*>(p)[0] = make_ulonglong2(w[0], w[1]);
      if constexpr (sizeof(V) == 32) reinterpret_cast<ulonglong2*>(p)[1] = make_ulonglong2(w[2], w[3]);
    }
  } else {
    uint32_t la = (uint32_t)__cvta_generic_to_shared(p), ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(rank));
    if constexpr (sizeof(V) == 8) {
      asm volatile("st.shared::cluster.b64 [%0], %1;" ::"r"(ra), "l"(w[0]) : "memory");
    } else if constexpr (sizeof(V) == 16) {
      asm volatile("st.shared::cluster.v2.b64 [%0], {%1, %2};" ::"r"(ra), "l"(w[0]), "l"(w[1]) : "memory");
    } else {
      asm volatile("st.shared::cluster.v2.b64 [%0], {%1, %2};" ::"r"(ra), "l"(w[0]), "l"(w[1]) : "memory");
      asm volatile("st.shared::cluster.v2.b64 [%0], {%1, %2};" ::"r"(ra + 16), "l"(w[2]), "l"(w[3]) : "memory");
    }
  }
}

template <typename T, class P, int CS> struct ResidentSmem {
  using L2 = typename PairOf<T>::type;
  static constexpr int N = P::L, MH = N / 2, MHc = MH / CS, Nc = N / CS;
  // Strides in elements.  64-bit shared accesses are served per HALF-warp, and with TPF = 8 a half-warp is two groups
  // (two columns / two private exchange buffers): consecutive groups must sit 64 bytes apart modulo 128, i.e. the
  // stride must be 8 modulo 16 elements (ncu: with N + 4 / N + N/16 every LDS.64 / STS.64 of the K phases and of the
  // exchanges took 2x its ideal wavefronts and the kernel was bound by the shared-memory pipe at 80 %; fixing it took
  // 64^2 x 1024 fp32 from 183 to 175 us per RK step).  fp64 elements are 16 bytes, one group per quarter-warp: the
  // original strides are the better ones there (438 vs 456 us), so the padding depends on the element size.
  static constexpr bool kF32 = sizeof(T) == 4;
  static constexpr int KS = kF32 ? N + 8 : N + 4;  // column stride of the K-layout buffers
  static constexpr int XS = MH + 1;   // row stride of FX (16-byte elements: one quarter-warp = one group, 8 distinct slots)
  static constexpr int kExMin = N + N / 16;  // padded private exchange buffer of one group (RowEx::pad)
  static constexpr int EXN = kF32 ? kExMin + ((8 - kExMin % 16) + 16) % 16 : kExMin;
  static constexpr int GROUPS = 256 / P::TPF;
  static_assert(GROUPS == MHc && 2 * GROUPS == Nc, "256 threads: one group per owned column and per owned row pair");
  static constexpr size_t kK = (size_t)MHc * KS * sizeof(Cx<T>);
  static constexpr size_t kFX = (size_t)Nc * XS * sizeof(Cx<L2>);
  static constexpr size_t kEX = (size_t)GROUPS * EXN * sizeof(Cx<T>);  // = lane values of the split paired exchange
  static constexpr size_t kSC = 2 * (size_t)N * sizeof(Cx<L2>);        // packed-column scratch (group 0 of rank 0)
  static constexpr size_t kSmall = (size_t)(3 * N + MH + 2) * sizeof(Cx<T>);  // twiddles, WN, HN, G row
  static constexpr size_t bytes() { return 3 * kK + kFX + kEX + kSC + kSmall + 16; }
};

template <typename T, class P, int CS, int MINB>
__global__ void __launch_bounds__(256, MINB) resident_kernel(ResidentArgs<T> a) {
  using L2 = typename PairOf<T>::type;
  using S = ResidentSmem<T, P, CS>;
  constexpr int N = P::L, MH = N / 2, E = P::E, TPF = P::TPF, MHc = S::MHc, Nc = S::Nc, KS = S::KS, XS = S::XS;
  constexpr int kn = (N / 2) * 2 / 3, km = ((MH + 1) * 2) / 3;
  extern __shared__ __align__(16) unsigned char smem_res[];
  unsigned char* sp = smem_res;
  Cx<T>* W = reinterpret_cast<Cx<T>*>(sp); sp += S::kK;
  Cx<T>* H = reinterpret_cast<Cx<T>*>(sp); sp += S::kK;
  Cx<T>* ADV = reinterpret_cast<Cx<T>*>(sp); sp += S::kK;
  Cx<L2>* FX = reinterpret_cast<Cx<L2>*>(sp); sp += S::kFX;
  Cx<T>* EXb = reinterpret_cast<Cx<T>*>(sp); sp += S::kEX;
  Cx<L2>* SC = reinterpret_cast<Cx<L2>*>(sp); sp += S::kSC;
  Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(sp);
  Cx<T>* WN = s_tw + N;
  Cx<T>* HN = WN + N;
  Cx<T>* GS = HN + N;

  const int tid = threadIdx.x, g = tid / TPF, t = tid % TPF;
  const unsigned rank = cluster_ctarank<CS>();
  const int b = blockIdx.x / CS;
  const int q = (int)rank * MHc + g;        // packed ky column of this group in the K phases
  const bool special = (q == 0);            // carries ky = 0 and ky = N/2 (group 0 of rank 0: lanes 0..TPF-1 of warp 0)
  constexpr unsigned kGroupMask = TPF >= 32 ? 0xffffffffu : ((1u << TPF) - 1u);
  const int xg0 = (int)rank * Nc + 2 * g;   // global x of this group's first row in the X phases
  const T tf = (T)t, qf = (T)q;
  Cx<T>* Wc = W + (size_t)g * KS;
  Cx<T>* Hc = H + (size_t)g * KS;
  Cx<T>* Ac = ADV + (size_t)g * KS;
  Cx<T>* ex_group = EXb + (size_t)g * S::EXN;
  RowExSplit<L2, TPF> ex2{reinterpret_cast<L2*>(ex_group)};
  RowEx<T, TPF> ex1{ex_group};

  // ---- prologue: tables and this cluster's environment -> shared memory --------------------------------------
  Cx<T>* user = a.omega + (size_t)b * N * (MH + 1);
  for (int i = tid; i < N; i += 256) s_tw[i] = ldc(a.tw + i);
  for (int i = tid; i <= MH; i += 256) GS[i] = ldc(a.G + (size_t)b * (MH + 1) + i);
  for (int i = tid; i < N * MHc; i += 256) {  // consecutive threads -> consecutive ky of one row (coalesced)
    const int kx = i / MHc, c = i - kx * MHc;
    W[(size_t)c * KS + kx] = ldc(user + (size_t)kx * (MH + 1) + (int)rank * MHc + c);
  }
  if (rank == 0)
    for (int i = tid; i < N; i += 256) WN[i] = ldc(user + (size_t)i * (MH + 1) + MH);
  cluster_barrier<CS>();

  T p0[E], p1[E];  // u*gx of rows x0, x0+1, carried from X1 to X2
  for (int step = 0; step < a.n_steps; ++step) {
#pragma unroll 1
    for (int s = 0; s < 5; ++s) {
      const T beta = a.beta[s], gdt = a.gdt[s], munu = a.munu[s];
      const bool use_h = (s != 0), store_h = (s != 4);
#pragma unroll
      for (int pair = 0; pair < 2; ++pair) {
        // ---- K1 / K2: lane pair `pair` of the spectral fields of column q, inverse x-FFT, scatter into FX -------
        {
          Cx<L2> v[E];
#pragma unroll
          for (int m = 0; m < E; ++m) {
            const L2 cf = pair == 0 ? pair_coef<0, T>(wave_p<T, P>(tf, m), qf, a.scale)
                                    : pair_coef<1, T>(wave_p<T, P>(tf, m), qf, a.scale);
            v[m] = coef_times_iw<T>(cf, Wc[t + TPF * m]);
          }
          if (special) {
            // column 0 <- herm(field[:,0]) + i*herm(field[:,N/2]); herm(f)[kx] = (f[kx] + conj(f[-kx]))/2
            Cx<L2>* sA = SC;
            Cx<L2>* sB = SC + N;
#pragma unroll
            for (int m = 0; m < E; ++m) {
              const int kx = t + TPF * m;
              const L2 cf = pair == 0 ? pair_coef<0, T>(wave_p<T, P>(tf, m), (T)MH, a.scale)
                                      : pair_coef<1, T>(wave_p<T, P>(tf, m), (T)MH, a.scale);
              stc(sA + kx, v[m]);
              stc(sB + kx, coef_times_iw<T>(cf, WN[kx]));
            }
            __syncwarp(kGroupMask);
#pragma unroll
            for (int m = 0; m < E; ++m) {
              const int kx = t + TPF * m, mkx = (N - kx) & (N - 1);
              const Cx<L2> ah = add_conj(v[m], ldc(sA + mkx));
              const Cx<L2> bh = add_conj(ldc(sB + kx), ldc(sB + mkx));
              v[m] = mk<L2>((ah.x - bh.y) * L2(0.5), (ah.y + bh.x) * L2(0.5));
            }
            __syncwarp(kGroupMask);
          }
          fft_regs<P, +1, L2>(v, s_tw, t, ex2);
#pragma unroll
          for (int m = 0; m < E; ++m) {
            const int x = t + TPF * m;
            st_rank<CS>(FX + (size_t)(x % Nc) * XS + q, (unsigned)(x / Nc), v[m]);
          }
        }
        cluster_barrier<CS>();
        // ---- X1 / X2: rows (x0, x0+1) of this pair: ONE lane-paired inverse y-FFT of FX[x0] + i FX[x0+1] ----------
        {
          Cx<L2> z[E];
          load_pair_hermitian<L2, P>(z, FX + (size_t)(2 * g) * XS, FX + (size_t)(2 * g + 1) * XS, t);
          fft_regs<P, +1, L2>(z, s_tw, t, ex2);
          if (pair == 0) {
#pragma unroll
            for (int m = 0; m < E; ++m) {
              p0[m] = z[m].x.lo() * z[m].x.hi();  // u * gx, row x0
              p1[m] = z[m].y.lo() * z[m].y.hi();  // row x0 + 1
            }
          } else {
            // adv = -(gx*u + gy*v) (kolmogorov.py:249); forward y-FFT of adv[x0] + i adv[x0+1]
            Cx<T> v[E];
#pragma unroll
            for (int m = 0; m < E; ++m)
              v[m] = mk<T>(-fmaL(z[m].x.lo(), z[m].x.hi(), p0[m]), -fmaL(z[m].y.lo(), z[m].y.hi(), p1[m]));
            fft_regs<P, -1, T>(v, s_tw, t, ex1);
#pragma unroll
            for (int m = 0; m < E; ++m) ex1.store(t + TPF * m, v[m]);
            ex1.sync();
#pragma unroll
            for (int m = 0; m < E / 2; ++m) {
              const int k = t + TPF * m;
              if (a.dealias && k >= km) continue;  // masked in K3: never read
              Cx<T> o0, o1;
              if (k == 0) {
                const Cx<T> wn = ex1.load(MH);
                o0 = mk<T>(v[m].x, wn.x);
                o1 = mk<T>(v[m].y, wn.y);
              } else {
                const Cx<T> wm = ex1.load(N - k);
                const Cx<T> sm = add_conj(v[m], wm), d = sub_conj(v[m], wm);
                o0 = (T)0.5 * sm;
                o1 = mk<T>((T)0.5 * d.y, (T)-0.5 * d.x);
              }
              // rows x0 (even) and x0+1 are adjacent in the K layout: one store of two complex numbers
              struct alignas(16) Two { Cx<T> a, b; };
              Two two{o0, o1};
              st_rank<CS>(reinterpret_cast<Two*>(ADV + (size_t)(k % MHc) * KS + xg0), (unsigned)(k / MHc), two);
            }
          }
        }
        cluster_barrier<CS>();
      }
      // ---- K3: forward x-FFT of the advection column, forcing, RK accumulate + Crank-Nicolson update ------------
      {
        Cx<T> v[E];
        const bool masked_col = a.dealias && q >= km;
#pragma unroll
        for (int m = 0; m < E; ++m) v[m] = masked_col ? mk<T>((T)0, (T)0) : Ac[t + TPF * m];
        fft_regs<P, -1, T>(v, s_tw, t, ex1);
        if (special) {
          // unpack column 0: A0[kx] = (V[kx] + conj(V[-kx]))/2 (ky = 0), An[kx] = (V[kx] - conj(V[-kx]))/(2i) (ky = N/2)
          Cx<T>* sV = reinterpret_cast<Cx<T>*>(SC);
#pragma unroll
          for (int m = 0; m < E; ++m) stc(sV + t + TPF * m, v[m]);
          __syncwarp(kGroupMask);
#pragma unroll
          for (int m = 0; m < E; ++m) {
            const int kx = t + TPF * m, mkx = (N - kx) & (N - 1);
            const Cx<T> vm = ldc(sV + mkx);
            const Cx<T> d = v[m] - conj(vm);
            Cx<T> hn = mk<T>((T)0.5 * d.y, (T)-0.5 * d.x);
            if (a.dealias) hn = mk<T>((T)0, (T)0);  // ky = N/2 >= km always
            if (kx == 0) hn = hn + GS[MH];
            if (use_h) hn = hn + beta * HN[kx];
            if (store_h) HN[kx] = hn;
            const T p = wave_p<T, P>(tf, m);
            const T lk = munu * (p * p + (T)MH * (T)MH);
            Cx<T> wv = WN[kx];
            wv = fast_rcp((T)1 + lk) * (wv + gdt * hn - lk * wv);
            WN[kx] = wv;
            v[m] = (T)0.5 * add_conj(v[m], vm);
          }
          __syncwarp(kGroupMask);
        }
        Cx<T> g0 = mk<T>((T)0, (T)0);
        if (t == 0) g0 = GS[q];
#pragma unroll
        for (int m = 0; m < E; ++m) {
          const int kx = t + TPF * m;
          Cx<T> h = v[m];
          if (a.dealias && ((kx >= kn && kx < 2 * kn) || q >= km)) h = mk<T>((T)0, (T)0);
          if (m == 0) h = h + g0;  // zero except on row kx = 0 (t == 0)
          if (use_h) h = h + beta * Hc[kx];
          if (store_h) Hc[kx] = h;
          const T p = wave_p<T, P>(tf, m);
          const T lk = munu * (p * p + qf * qf);
          const Cx<T> wv = Wc[kx];
          Wc[kx] = fast_rcp((T)1 + lk) * (wv + gdt * h - lk * wv);
        }
        // the next K1 reads W / WN of this group's own column only (same threads): a warp-level fence is enough
        __syncwarp();
      }
    }
  }

  // ---- epilogue: omega_hat back to the user layout -------------------------------------------------------------
  cluster_barrier<CS>();
  for (int i = tid; i < N * MHc; i += 256) {
    const int kx = i / MHc, c = i - kx * MHc;
    stc(user + (size_t)kx * (MH + 1) + (int)rank * MHc + c, W[(size_t)c * KS + kx]);
  }
  if (rank == 0)
    for (int i = tid; i < N; i += 256) stc(user + (size_t)i * (MH + 1) + MH, WN[i]);
}

}  // namespace kf
