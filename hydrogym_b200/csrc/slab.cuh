// Slab-decomposed single-grid kernels (BASELINE.json configs[4]: one N x N Kolmogorov flow over P GPUs).
//
// The reference has no multi-device path for this solve; the arithmetic is the same RK sub-stage as
// kernels.cuh (hydrogym/jax/solvers/base.py:50-57, hydrogym/jax/envs/kolmogorov.py:209-257), split at
// the two places where a 2-D FFT changes direction:
//
//   K layout (rank r owns Q = M/2/P packed ky columns, ALL kx, kx contiguous: [Q][N]):
//       slab_col_kernel : forward x-FFT of the advection columns -> forcing/control -> RK accumulate +
//                         Crank-Nicolson update of h / omega_hat -> 4 spectral fields -> inverse x-FFT
//   X layout (rank r owns X = N/P x rows, ALL ky, blocked by the peer that owns the columns: [P][X][Q]):
//       slab_row_kernel : c2r y-FFT of the 4 fields -> -(gx*u + gy*v) -> r2c y-FFT
//
// Between them the data crosses ranks with an all-to-all whose blocks are exactly the [X][Q] tiles of the
// blocked X layout; transpose_kernel turns K-layout rows into those tiles (and back) on the local GPU.
// Column 0 of the packed x-space layout carries ky = 0 and ky = M/2 together (see kernels.cuh); it lives on
// rank 0 and is handled by the SPECIAL instance of slab_col_kernel (one CTA).
#pragma once
#include <type_traits>

#include "kernels.cuh"

namespace kf {

#ifndef KF_MAX_PEERS
#define KF_MAX_PEERS 16
#endif

template <typename T> struct SlabArgs {
  using L2 = typename PairOf<T>::type;
  // K layout
  Cx<T>* wk;            // [Q][N] omega_hat
  Cx<T>* hk;            // [Q][N] RK register
  Cx<T>* wnyq;          // [N] ky = M/2 column (rank 0)
  Cx<T>* hnyq;          // [N]
  const Cx<T>* G;       // [M/2+1] forcing + control row (kx = 0)
  const Cx<T>* advk;    // [Q][N] packed x-space advection, transposed
  Cx<T>* fk;            // [4][Q][N] planes u_hat, gx_hat, v_hat, gy_hat (x-space after the column kernel)
  // X layout, blocked: element (xl, k) of a half-row of plane f at (k / Q) * fblk + f * X*Q + xl * Q + k % Q
  const Cx<T>* fx;      // fields received from the peers
  size_t fblk;          // block stride of fx (4*X*Q: the four planes of one peer are adjacent)
  // advection rows leave in blocks of Q columns: block d (the columns rank d owns) starts at adv_dst[d] and is
  // [X][Q] -- a slice of the local send buffer (NCCL all-to-all follows) or of rank d's receive buffer mapped
  // over NVLink (peer-to-peer stores straight from this kernel, no separate collective)
  Cx<T>* adv_dst[KF_MAX_PEERS];
  const Cx<T>* tw;      // [N] twiddle table of the length-N plan (fft.cuh)
  const Cx<T>* tw2;     // [N/2] twiddle table of the length-N/2 plan (slab_row_half_kernel)
  const Cx<T>* twh;     // [N/2] exp(-2*pi*i*k/N)
  T beta, gdt, munu, scale;
  int dealias;
  int q0;               // global packed-column index of local column 0 (rank * Q)
  int x0;               // global x of local row 0 (rank * X)
  int Q, X, lgQ;
  int row_prefetch;     // slab_row_half_kernel: L2 bulk prefetch of the next pair of field rows (KF_SLAB_ROW_PREFETCH)
  int cb0, cb1;         // slab_col_scalar_kernel: range of column blocks (of GROUPS columns) this launch covers
  // direct mode: the column kernels store their x-space field columns straight into the destination ranks' receive
  // buffers (K-layout pieces [4][Q][X] in this rank's slot: runs of X contiguous elements per column), over NVLink
  // and under their own butterflies; the receiver transposes the pieces locally.  0 = fields go to fk (sender transposes)
  int direct, lgX;
  Cx<T>* fld_dst[KF_MAX_PEERS];
};

// where element x of local column ql of field plane fld goes
template <typename T, int N> __device__ __forceinline__ Cx<T>* direct_dst(const SlabArgs<T>& a, int fld, int ql, int x) {
  return a.fld_dst[x >> a.lgX] + ((size_t)fld * a.Q + ql) * a.X + (x & (a.X - 1));
}

// split exchange of one FFT group over a padded buffer of lane values (real parts, then imaginary parts)
template <typename X, int TPF> struct RowExSplit {
  static constexpr bool kSplit = true;
  X* sm;
  __device__ __forceinline__ static int pad(int pos) { return pos + (pos >> 4); }
  __device__ __forceinline__ void store_part(int pos, X a) { stl(sm + pad(pos), a); }
  __device__ __forceinline__ X load_part(int pos) { return ldl(sm + pad(pos)); }
  __device__ __forceinline__ void sync() {
    if (TPF <= 32) __syncwarp(); else __syncthreads();
  }
};

template <int N> __host__ __device__ constexpr size_t slab_exchange_elems() { return (size_t)N + N / 16; }
// per-CTA shared memory of the slab FFT kernels: one exchange buffer per group + the twiddle table
// (whole lane-paired complex numbers: the non-split exchange needs half the barriers and 2*E fewer registers
// than the split one, and 8192 x 16 B + twiddles still fits the 227 KB of an SM)
template <typename T, int N, int GROUPS> __host__ __device__ constexpr size_t slab_smem_bytes() {
  return sizeof(Cx<typename PairOf<T>::type>) * GROUPS * slab_exchange_elems<N>() + sizeof(Cx<T>) * (size_t)N;
}

enum SlabColMode { kSlabStage = 0, kSlabInit = 1 };

// Reload of a value this thread stored earlier in the kernel.  An opaque, ordered load: the compiler can neither
// forward the stored registers (that is the point -- keeping omega_hat live across the field transforms is what
// made the 8192-point kernels spill) nor hoist the load above the store.
__device__ __forceinline__ Cx<float> reload(const Cx<float>* p) {
  Cx<float> r;
  asm volatile("ld.global.cg.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ Cx<double> reload(const Cx<double>* p) {
  Cx<double> r;
  asm volatile("ld.global.cg.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p) : "memory");
  return r;
}


// integer wavenumber of row index kx as a real
template <typename T, int N> __device__ __forceinline__ T wave_of(int kx) { return (T)(kx < N / 2 ? kx : kx - N); }

// ---------------------------------------------------------------------------
// K-layout column kernel.  A group of TPF threads owns local column ql (a contiguous run of N kx values);
// thread t holds rows {t + TPF*m}.  grid = Q / GROUPS (normal) or 1 CTA with GROUPS = 1 (SPECIAL, rank 0,
// packed column 0 = ky 0 and ky M/2).  The normal instance computes global column 0 too but discards it.
// ---------------------------------------------------------------------------
// LEAN (fp64 at 8192 points, where a whole lane-paired transform plus the twiddle table would need 409 KB): the
// lane-paired exchange goes through lane values in two rounds (RowExSplit, N x 16 B) and the twiddles are read from
// global memory -- 139 KB of shared memory instead.
template <typename T, class P, int GROUPS, int MODE, bool SPECIAL, int MINB, bool LEAN = false>
__global__ void __launch_bounds__(P::TPF* GROUPS, MINB) slab_col_kernel(SlabArgs<T> a) {
  using L2 = typename PairOf<T>::type;
  constexpr int N = P::L, E = P::E, TPF = P::TPF, MH = N / 2;
  constexpr size_t EXN = slab_exchange_elems<N>();
  static_assert(!SPECIAL || GROUPS == 1, "the packed column is one CTA");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t0 = threadIdx.x % TPF;
  unsigned char* s_g = smem_raw + (size_t)g * EXN * (LEAN ? sizeof(Cx<T>) : sizeof(Cx<L2>));
  const Cx<T>* s_tw = LEAN ? a.tw : reinterpret_cast<const Cx<T>*>(smem_raw + (size_t)GROUPS * EXN * sizeof(Cx<L2>));
  RowEx<T, TPF> ex1{reinterpret_cast<Cx<T>*>(s_g)};
  typename std::conditional<LEAN, RowExSplit<L2, TPF>, RowEx<L2, TPF>>::type ex2{
      reinterpret_cast<typename std::conditional<LEAN, L2, Cx<L2>>::type*>(s_g)};
  if constexpr (!LEAN) {
    Cx<T>* tw_w = reinterpret_cast<Cx<T>*>(smem_raw + (size_t)GROUPS * EXN * sizeof(Cx<L2>));
    for (int i = threadIdx.x; i < N; i += TPF * GROUPS) tw_w[i] = ldc(a.tw + i);
    __syncthreads();
  }

  // persistent CTAs (the host launches at most one wave): the twiddle table is staged once per CTA
  const int ncb = SPECIAL ? 1 : a.Q / GROUPS;
#pragma unroll 1
  for (int cb = blockIdx.x; cb < ncb; cb += gridDim.x) {
  // an opaque copy of the thread index per column (and per field pair below): otherwise every address, wavenumber
  // and exchange offset is loop-invariant and the compiler keeps them all in registers across the transforms
  int t = t0;
  asm volatile("" : "+r"(t));
  const int ql = SPECIAL ? 0 : cb * GROUPS + g;
  const int q = a.q0 + ql;
  const bool live = SPECIAL || q != 0;
  const T tf = (T)t, qf = (T)q;
  Cx<T>* wrow = a.wk + (size_t)ql * N;
  Cx<T>* hrow = a.hk + (size_t)ql * N;

  Cx<T> w[E];
  Cx<T> wn[SPECIAL ? E : 1];  // ky = M/2 column (SPECIAL only)
  if constexpr (MODE == kSlabStage) {
    const bool use_h = (a.beta != (T)0);
    Cx<T> v[E];
    const Cx<T>* arow = a.advk + (size_t)ql * N;
#pragma unroll
    for (int m = 0; m < E; ++m) v[m] = ldc(arow + t + TPF * m);
    fft_regs<P, -1, T>(v, s_tw, t, ex1);
    if constexpr (!SPECIAL) asm volatile("" : "+r"(t));
    if constexpr (SPECIAL) {
      // unpack: A0[kx] = (W[kx] + conj(W[-kx]))/2 (ky = 0), An[kx] = (W[kx] - conj(W[-kx]))/(2i) (ky = M/2)
#pragma unroll
      for (int m = 0; m < E; ++m) ex1.store(t + TPF * m, v[m]);
      __syncthreads();
#pragma unroll
      for (int m = 0; m < E; ++m) {
        const int kx = t + TPF * m;
        const Cx<T> wm = ex1.load((N - kx) & (N - 1));
        const Cx<T> d = v[m] - conj(wm);
        v[m] = (T)0.5 * add_conj(v[m], wm);
        Cx<T> h = mk<T>((T)0.5 * d.y, (T)-0.5 * d.x);
        if (a.dealias) h = mk<T>((T)0, (T)0);  // ky = M/2 >= km always
        if (kx == 0) h = h + ldc(a.G + MH);
        if (use_h) h = h + a.beta * ldc(a.hnyq + kx);
        stc(a.hnyq + kx, h);
        const T p = wave_of<T, N>(kx);
        const T lk = a.munu * (p * p + (T)MH * (T)MH);
        const Cx<T> wv = ldc(a.wnyq + kx);
        wn[m] = fast_rcp((T)1 + lk) * (wv + a.gdt * h - lk * wv);
        stc(a.wnyq + kx, wn[m]);
      }
      __syncthreads();
    }
    Cx<T> g0 = mk<T>((T)0, (T)0);
    if (t == 0) g0 = ldc(a.G + q);
    rk_update_column<T, P>(a, v, w, g0, use_h, t, q,
                           [&](int, int kx, Cx<T> h) { if (live) stc(hrow + kx, h); },
                           [&](int, int kx, Cx<T> wv) { if (live) stc(wrow + kx, wv); },
                           [&](int, int kx) { return ldc(hrow + kx); },
                           [&](int, int kx) { return ldc(wrow + kx); }, ZeroExtra<T>());
  } else if constexpr (SPECIAL) {
#pragma unroll
    for (int m = 0; m < E; ++m) {
      w[m] = ldc(wrow + t + TPF * m);
      wn[m] = ldc(a.wnyq + t + TPF * m);
    }
  }

  // mirrored omega_hat values of the packed column, read back from global memory (this CTA wrote them)
  Cx<T> w0m[SPECIAL ? E : 1], wnm[SPECIAL ? E : 1];
  if constexpr (SPECIAL) {
    __syncthreads();
#pragma unroll
    for (int m = 0; m < E; ++m) {
      const int mkx = (N - (t + TPF * m)) & (N - 1);
      w0m[m] = ldc(wrow + mkx);
      wnm[m] = ldc(a.wnyq + mkx);
    }
  }

#pragma unroll
  for (int pair = 0; pair < 2; ++pair) {
    if constexpr (!SPECIAL) asm volatile("" : "+r"(t));
    Cx<L2> f[E];
#pragma unroll
    for (int m = 0; m < E; ++m) {
      const T p = wave_p<T, P>(tf, m);
      const L2 cf = pair == 0 ? pair_coef<0, T>(p, qf, a.scale) : pair_coef<1, T>(p, qf, a.scale);
      // the ordinary columns read omega_hat back (L2-hot) instead of holding it across both transforms
      Cx<T> wv;
      if constexpr (SPECIAL) wv = w[m]; else wv = reload(wrow + t + TPF * m);
      f[m] = coef_times_iw<T>(cf, wv);
      if constexpr (SPECIAL) {
        // column 0 <- herm(field[:,0]) + i*herm(field[:,M/2]); herm(c)[kx] = (c[kx] + conj(c[-kx]))/2
        const int mkx = (N - (t + TPF * m)) & (N - 1);
        const T pm = wave_of<T, N>(mkx);
        const L2 cfm = pair == 0 ? pair_coef<0, T>(pm, (T)0, a.scale) : pair_coef<1, T>(pm, (T)0, a.scale);
        const L2 cn = pair == 0 ? pair_coef<0, T>(p, (T)MH, a.scale) : pair_coef<1, T>(p, (T)MH, a.scale);
        const L2 cnm = pair == 0 ? pair_coef<0, T>(pm, (T)MH, a.scale) : pair_coef<1, T>(pm, (T)MH, a.scale);
        const Cx<L2> ah = add_conj(f[m], coef_times_iw<T>(cfm, w0m[m]));
        const Cx<L2> bh = add_conj(coef_times_iw<T>(cn, wn[m]), coef_times_iw<T>(cnm, wnm[m]));
        f[m] = mk<L2>((ah.x - bh.y) * L2(0.5), (ah.y + bh.x) * L2(0.5));
      }
    }
    fft_regs<P, +1, L2>(f, s_tw, t, ex2);
    Cx<T>* o0 = a.fk + ((size_t)(2 * pair) * a.Q + ql) * N;  // planes 2 pair (lane 0) and 2 pair + 1 (lane 1)
    Cx<T>* o1 = o0 + (size_t)a.Q * N;
    if (live && a.direct) {
#pragma unroll
      for (int m = 0; m < E; ++m) {
        stc(direct_dst<T, N>(a, 2 * pair, ql, t + TPF * m), mk<T>(f[m].x.lo(), f[m].y.lo()));
        stc(direct_dst<T, N>(a, 2 * pair + 1, ql, t + TPF * m), mk<T>(f[m].x.hi(), f[m].y.hi()));
      }
    } else if (live) {
#pragma unroll
      for (int m = 0; m < E; ++m) {
        stc(o0 + t + TPF * m, mk<T>(f[m].x.lo(), f[m].y.lo()));
        stc(o1 + t + TPF * m, mk<T>(f[m].x.hi(), f[m].y.hi()));
      }
    }
  }
  }  // column blocks
}

// ---------------------------------------------------------------------------
// K-layout column kernel of the ordinary columns, scalar lanes.  The same arithmetic as slab_col_kernel (which now
// only serves the packed column), but every transform is a scalar one: 32 data registers per thread instead of
// 64, so the kernel fits 64 registers and TWO 512-thread CTAs share an SM at 8192 points -- one CTA's loads
// overlap the other's butterflies, which is what the lane-paired form (one CTA per SM, all warps in the same
// phase) could not do.  The four fields are four transforms written to four planes.  Twiddles come from global
// memory through L1 (the table is shared by both CTAs; the streaming operands bypass or leave L1 early).
// grid: persistent over the Q local columns (global column 0 is computed but not stored).
// ---------------------------------------------------------------------------
template <typename T, int N, int GROUPS, int TWS = 0> __host__ __device__ constexpr size_t slab_scalar_smem_bytes() {
  return sizeof(Cx<T>) * (GROUPS * slab_exchange_elems<N>() + (TWS ? (size_t)N : 0));
}
__device__ __forceinline__ Cx<float> ld_stream(const Cx<float>* p) {
  const float2 v = __ldcs(reinterpret_cast<const float2*>(p));
  return mk<float>(v.x, v.y);
}
__device__ __forceinline__ Cx<double> ld_stream(const Cx<double>* p) {
  const double2 v = __ldcs(reinterpret_cast<const double2*>(p));
  return mk<double>(v.x, v.y);
}

// TWS = 1: the twiddle table is copied to shared memory once per CTA and the groups of the CTA synchronise on their own
// named barriers, so ONE CTA of GROUPS = 2 transforms per SM (1024 threads at 8192 points) replaces the two CTAs: the
// same two out-of-phase transforms per SM, but the 15 twiddles a thread needs per pass come from shared memory instead
// of an L1 that the streaming operands keep evicting (ncu: long_scoreboard on the first use of each twiddle was 37 % of
// the stall samples of the two-CTA form).
template <typename T, class P, int GROUPS, int MODE, int MINB, int TWS = 0>
__global__ void __launch_bounds__(P::TPF* GROUPS, MINB) slab_col_scalar_kernel(SlabArgs<T> a) {
  constexpr int N = P::L, E = P::E, TPF = P::TPF;
  constexpr size_t EXN = slab_exchange_elems<N>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t0 = threadIdx.x % TPF;
  RowEx<T, TPF> ex{reinterpret_cast<Cx<T>*>(smem_raw) + (size_t)g * EXN};
  const Cx<T>* __restrict__ tw = a.tw;
  if constexpr (TWS != 0) {
    Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(smem_raw) + (size_t)GROUPS * EXN;
    for (int i = threadIdx.x; i < N; i += TPF * GROUPS) s_tw[i] = ldc(a.tw + i);
    __syncthreads();
    tw = s_tw;
    if (GROUPS > 1 && TPF > 32) ex.bar = 1 + g;
  }

  const int ncb = a.cb1;
#pragma unroll 1
  for (int cb = a.cb0 + blockIdx.x; cb < ncb; cb += gridDim.x) {
    int t = t0;
    asm volatile("" : "+r"(t));  // see slab_col_kernel
    const int ql = cb * GROUPS + g;
    const int q = a.q0 + ql;
    const bool live = q != 0;
    const T qf = (T)q;
    Cx<T>* wrow = a.wk + (size_t)ql * N;
    Cx<T>* hrow = a.hk + (size_t)ql * N;

    if constexpr (MODE == kSlabStage) {
      const bool use_h = (a.beta != (T)0);
      constexpr unsigned kColBytes = (unsigned)(N * sizeof(Cx<T>));
      if (t0 == 0 && g == 0) {
        // the update below reads this column of h and omega_hat right after the transform, and the next
        // iteration starts with its advection column: have L2 fetch them meanwhile
#pragma unroll 1
        for (int gg = 0; gg < GROUPS; ++gg) {
          if (use_h) prefetch_l2_bulk(hrow + (size_t)gg * N, kColBytes);
          prefetch_l2_bulk(wrow + (size_t)gg * N, kColBytes);
          if (cb + (int)gridDim.x < ncb)
            prefetch_l2_bulk(a.advk + ((size_t)(cb + gridDim.x) * GROUPS + gg) * N, kColBytes);
        }
      }
      Cx<T> v[E];
      const Cx<T>* arow = a.advk + (size_t)ql * N;
#pragma unroll
      for (int m = 0; m < E; ++m) v[m] = ld_stream(arow + t + TPF * m);
      fft_regs<P, -1, T>(v, tw, t, ex);
      asm volatile("" : "+r"(t));
      Cx<T> g0 = mk<T>((T)0, (T)0);
      if (t == 0) g0 = ldc(a.G + q);
      // rk_update_column (kernels.cuh), with the loads of h and omega_hat issued four rows at a time: the stores
      // of one row may alias the loads of the next as far as the compiler knows, so row-by-row every row costs a
      // full memory round trip
      constexpr int MH = N / 2, kn = (N / 2) * 2 / 3, km = ((MH + 1) * 2) / 3, RB = 4;
      const T tf = (T)t;
#pragma unroll
      for (int mb = 0; mb < E; mb += RB) {
        Cx<T> hh[RB], ww[RB];
#pragma unroll
        for (int j = 0; j < RB; ++j) {
          const int kx = t + TPF * (mb + j);
          hh[j] = use_h ? ld_stream(hrow + kx) : mk<T>((T)0, (T)0);
          ww[j] = ld_stream(wrow + kx);
        }
#pragma unroll
        for (int j = 0; j < RB; ++j) {
          const int m = mb + j, kx = t + TPF * m;
          Cx<T> h = v[m];
          if (a.dealias && ((kx >= kn && kx < 2 * kn) || q >= km)) h = mk<T>((T)0, (T)0);
          if (m == 0) h = h + g0;  // g0 is zero except on row kx = 0 (t == 0)
          if (use_h) h = h + a.beta * hh[j];
          if (live) stc(hrow + kx, h);
          const T p = wave_p<T, P>(tf, m);
          const T lk = a.munu * (p * p + qf * qf);
          const Cx<T> wn = fast_rcp((T)1 + lk) * (ww[j] + a.gdt * h - lk * ww[j]);
          if (live) stc(wrow + kx, wn);
        }
      }
    }

#pragma unroll 1
    for (int fld = 0; fld < 4; ++fld) {  // u_hat, gx_hat, v_hat, gy_hat = c * (i omega_hat), kernels.cuh pair_coef
      asm volatile("" : "+r"(t));
      const T tf = (T)t;
      // omega_hat is read back (L2-hot: this thread stored it above) through a pointer the compiler cannot see
      // through: no forwarding of the stored registers, no hoisting above the stores, but all E loads in flight
      const Cx<T>* wr = wrow;
      asm volatile("" : "+l"(wr) : : "memory");
      Cx<T> f[E];
#pragma unroll
      for (int m = 0; m < E; ++m) {
#ifdef KF_SLAB_WCG
        const Cx<T> wv = reload(wr + t + TPF * m);  // L2 only: the 64 KB row must not evict the twiddle table from L1
#else
        const Cx<T> wv = ldc(wr + t + TPF * m);
#endif
        const T p = wave_p<T, P>(tf, m);
        const T sp = p * a.scale, sq = qf * a.scale, ik2 = inv_k2<T>(p, qf);
        const T c = fld == 0 ? sq * ik2 : fld == 1 ? sp : fld == 2 ? -(sp * ik2) : sq;
        f[m] = mk<T>(c * -wv.y, c * wv.x);
      }
      fft_regs<P, +1, T>(f, tw, t, ex);
      Cx<T>* out = a.fk + ((size_t)fld * a.Q + ql) * N;
      if (live && a.direct) {
#pragma unroll
        for (int m = 0; m < E; ++m) stc(direct_dst<T, N>(a, fld, ql, t + TPF * m), f[m]);
      } else if (live) {
#pragma unroll
        for (int m = 0; m < E; ++m) stc(out + t + TPF * m, f[m]);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// X-layout row kernel, half-length form.  A real row of length M is one complex transform of length H = M/2 of
// its even / odd samples:
//     z[n] = x[2n] + i x[2n+1] = IDFT_H( Ev + i Od ),  Ev[k] = X[k] + conj X[H-k],  Od[k] = (X[k] - conj X[H-k]) e^{+2 pi i k / M}
//     X[k] = (C[k] + conj C[H-k]) / 2 - (i/2) e^{-2 pi i k / M} (C[k] - conj C[H-k]),  C = DFT_H(z)
// so a group of TPF = H/E threads owns ONE row: two lane-paired inverse transforms ((u, gx) and (v, gy)) and one
// scalar forward transform of length H.  Against slab_row_kernel (two rows through length-M transforms) the
// register file holds half as much per CTA, so two CTAs fit on an SM and one's loads overlap the other's
// butterflies.  PH is the plan of length H; a.tw2 its twiddle table, a.twh[k] = e^{-2 pi i k / M}, k < H.
// grid: persistent, row blocks of GROUPS rows.
// ---------------------------------------------------------------------------
template <typename T, int H, int GROUPS> __host__ __device__ constexpr size_t slab_half_smem_bytes() {
  return sizeof(Cx<typename PairOf<T>::type>) * GROUPS * slab_exchange_elems<H>() + sizeof(Cx<T>) * (size_t)H;
}

template <typename T, class PH, int GROUPS, int MINB>
__global__ void __launch_bounds__(PH::TPF* GROUPS, MINB) slab_row_half_kernel(SlabArgs<T> a) {
  using L2 = typename PairOf<T>::type;
  constexpr int H = PH::L, E = PH::E, TPF = PH::TPF;
  constexpr size_t EXN = slab_exchange_elems<H>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t = threadIdx.x % TPF;
  unsigned char* s_g = smem_raw + (size_t)g * EXN * sizeof(Cx<L2>);
  Cx<T>* s_tw = reinterpret_cast<Cx<T>*>(smem_raw + (size_t)GROUPS * EXN * sizeof(Cx<L2>));
  RowEx<T, TPF> ex1{reinterpret_cast<Cx<T>*>(s_g)};
  RowEx<L2, TPF> ex2{reinterpret_cast<Cx<L2>*>(s_g), 0, (int)EXN};  // fp64: two-plane layout (kernels.cuh)
  const int Q = a.Q, lgQ = a.lgQ;
  const unsigned blk = (unsigned)a.fblk;  // element offsets of a slab fit 32 bits (N * N complex at most)

  for (int i = threadIdx.x; i < H; i += TPF * GROUPS) s_tw[i] = ldc(a.tw2 + i);
  __syncthreads();

  const int nrb = a.X / GROUPS;
#pragma unroll 1
  for (int rb = blockIdx.x; rb < nrb; rb += gridDim.x) {
    const int xl = rb * GROUPS + g;
    const size_t roff = (size_t)xl * Q;
    T acc_e[E], acc_o[E];  // sum over the two lane pairs of (velocity * gradient) at samples 2n, 2n+1
#pragma unroll
    for (int m = 0; m < E; ++m) acc_e[m] = (T)0, acc_o[m] = (T)0;
#pragma unroll 1
    for (int pair = 0; pair < 2; ++pair) {
      // planes (u, gx) for pair 0, (v, gy) for pair 1, packed into the two lanes as they are loaded
      const Cx<T>* F0 = a.fx + (size_t)(2 * pair) * a.X * Q + roff;
      const Cx<T>* F1 = F0 + (size_t)a.X * Q;
      if (a.row_prefetch && t < 2 * (H >> lgQ)) {
        // the two plane rows this group transforms NEXT (pair 1 of this row, or pair 0 of the CTA's next row block):
        // one bulk L2 prefetch per plane and per peer block, issued before this pair's own loads
        const int nrb_next = rb + (int)gridDim.x;
        if (pair == 0 || nrb_next < nrb) {
          const size_t nroff = pair == 0 ? roff : (size_t)(nrb_next * GROUPS + g) * Q;
          const Cx<T>* nx = a.fx + (size_t)(2 * (1 - pair) + (t & 1)) * a.X * Q + nroff + (size_t)(t >> 1) * blk;
          prefetch_l2_bulk(nx, (unsigned)(Q * sizeof(Cx<T>)));
        }
      }
      auto ldp = [&](unsigned off) {
        const Cx<T> p0 = ldc(F0 + off), p1 = ldc(F1 + off);
        return mk<L2>(L2(p0.x, p1.x), L2(p0.y, p1.y));
      };
      // nothing below depends on `pair` or `rb` except through F: an opaque copy of the thread index keeps the
      // compiler from hoisting the 2 E element offsets, the twiddle addresses and the exchange addresses out of
      // the loops (they would occupy the registers the transform needs)
      int tp = t;
      asm volatile("" : "+r"(tp));
      auto at = [&](int k) { return (unsigned)(k >> lgQ) * blk + (unsigned)(k & (Q - 1)); };
      Cx<L2> z[E];
#pragma unroll
      for (int m = 0; m < E; ++m) {
        const int k = tp + TPF * m;
        const Cx<L2> xk = ldp(at(k));
        const Cx<L2> xm = ldp(at(k == 0 ? 0 : H - k));
        const Cx<T> th = ldc(a.twh + k);  // (cos, -sin) of 2 pi k / M
        const Cx<L2> ev = add_conj(xk, xm), d = sub_conj(xk, xm);
        // od = d * conj(th);  z = ev + i * od
        const L2 c = bc<L2>(th.x), sn = bc<L2>(-th.y);
        const Cx<L2> od = mk<L2>(d.x * c - d.y * sn, d.x * sn + d.y * c);
        const Cx<L2> zk = mk<L2>(ev.x - od.y, ev.y + od.x);
        // packed entry 0 holds the two real values X[0], X[H]:  z[0] = (X0 + XH) + i (X0 - XH)
        const Cx<L2> z0 = mk<L2>(xk.x + xk.y, xk.x - xk.y);
        z[m] = (k == 0) ? z0 : zk;
      }
      fft_regs<PH, +1, L2>(z, s_tw, tp, ex2);
#pragma unroll
      for (int m = 0; m < E; ++m) {
        acc_e[m] = fmaL(z[m].x.lo(), z[m].x.hi(), acc_e[m]);  // kolmogorov.py:249
        acc_o[m] = fmaL(z[m].y.lo(), z[m].y.hi(), acc_o[m]);
      }
    }
    Cx<T> v[E];
#pragma unroll
    for (int m = 0; m < E; ++m) v[m] = mk<T>(-acc_e[m], -acc_o[m]);
    int tq = t;
    asm volatile("" : "+r"(tq));
    fft_regs<PH, -1, T>(v, s_tw, tq, ex1);
#pragma unroll
    for (int m = 0; m < E; ++m) ex1.store(tq + TPF * m, v[m]);
    ex1.sync();
#pragma unroll
    for (int m = 0; m < E; ++m) {
      const int k = tq + TPF * m;
      Cx<T> r;
      if (k == 0) {
        r = mk<T>(v[m].x + v[m].y, v[m].x - v[m].y);  // (X[0], X[H]) packed
      } else {
        const Cx<T> cm = ex1.load(H - k);
        const Cx<T> th = ldc(a.twh + k);
        const Cx<T> sm = add_conj(v[m], cm), d = sub_conj(v[m], cm);
        const T pa = th.x * d.x - th.y * d.y, pb = th.x * d.y + th.y * d.x;  // th * d
        r = mk<T>((T)0.5 * (sm.x + pb), (T)0.5 * (sm.y - pa));
      }
      stc(a.adv_dst[k >> lgQ] + (size_t)xl * Q + (k & (Q - 1)), r);
    }
    ex1.sync();
  }
}

// Reward / observation partial sums of this rank's rows (compute_tke utils.py:175-194,
// _trajectory_mean_obs kolmogorov.py:426-446): acc[4] += (sum u, sum v, sum u^2, sum v^2) over the local
// rows; |velocity| at the probes that fall on local rows is added to obs with weight obs_w.  The host sums
// acc and obs over the ranks.  One row per group; grid = X / GROUPS.
template <typename T, class P, int GROUPS, bool LEAN = false>
__global__ void __launch_bounds__(P::TPF* GROUPS) slab_row_obs_kernel(SlabArgs<T> a, double* acc, T* obs, int nobs,
                                                                      int sx, int sy, T obs_w) {
  constexpr int M = P::L, E = P::E, TPF = P::TPF;
  constexpr size_t EXN = slab_exchange_elems<M>();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int g = threadIdx.x / TPF, t = threadIdx.x % TPF;
  const Cx<T>* s_tw = LEAN ? a.tw : reinterpret_cast<const Cx<T>*>(smem_raw) + (size_t)GROUPS * EXN;
  RowEx<T, TPF> ex{reinterpret_cast<Cx<T>*>(smem_raw) + (size_t)g * EXN};
  const int xl = blockIdx.x * GROUPS + g;
  const int x = a.x0 + xl;
  const int Q = a.Q, lgQ = a.lgQ;

  if constexpr (!LEAN) {  // LEAN: twiddles straight from global memory (no room for the table beside the exchange)
    Cx<T>* tw_w = reinterpret_cast<Cx<T>*>(smem_raw) + (size_t)GROUPS * EXN;
    for (int i = threadIdx.x; i < M; i += TPF * GROUPS) tw_w[i] = ldc(a.tw + i);
    __syncthreads();
  }

  const size_t roff = (size_t)xl * Q, blk = a.fblk;
  // one scalar transform of u_hat + i v_hat (planes 0 and 2): real part u, imaginary part v
  Cx<T> z[E];
  load_pair_hermitian_at<T, P>(z, a.fx + roff, a.fx + 2 * (size_t)a.X * Q + roff, t,
                               [&](int k) { return (size_t)(k >> lgQ) * blk + (k & (Q - 1)); });
  fft_regs<P, +1, T>(z, s_tw, t, ex);

  T su = 0, sv = 0, suu = 0, svv = 0;
#pragma unroll
  for (int m = 0; m < E; ++m) {
    const T u = z[m].x, vv = z[m].y;
    su += u; sv += vv;
    suu += u * u; svv += vv * vv;
  }
  if (obs && x % sx == 0 && x / sx < nobs) {
#pragma unroll
    for (int m = 0; m < E; ++m) {
      const int y = t + TPF * m;
      if (y % sy == 0 && y / sy < nobs) {
        const T u = z[m].x, vv = z[m].y;
        obs[(x / sx) * nobs + y / sy] += obs_w * sqrt(u * u + vv * vv);  // single writer per launch
      }
    }
  }
  double r[4] = {(double)su, (double)sv, (double)suu, (double)svv};
  // threads of a warp may belong to different rows of the same slab: all of them add into the same sums
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
    for (int i = 0; i < 4; ++i) r[i] += __shfl_xor_sync(0xffffffffu, r[i], off);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) atomicAdd(acc + i, r[i]);
  }
}

// ---------------------------------------------------------------------------
// Tiled transpose with blocked row addressing on both sides: logical in[R][Cn] -> out[Cn][R], where
// logical input row r starts at (r / in_rpb) * in_blk + (r % in_rpb) * in_ld and logical output row c at
// (c / out_rpb) * out_blk + (c % out_rpb) * out_ld.  blockIdx.z selects one of several arrays
// (in + z * in_z, out + z * out_z).  V is any 8 / 16 / 32-byte element.
// ---------------------------------------------------------------------------
struct TransposeArgs {
  int R, Cn;
  int in_rpb, out_rpb;
  size_t in_blk, in_ld, in_z;
  size_t out_blk, out_ld, out_z;
  // if out_tab[0] != nullptr, output block j (logical rows j*out_rpb ...) starts at out_tab[j] instead of
  // out + j*out_blk: the blocks may live on different GPUs (peer-mapped receive buffers)
  void* out_tab[KF_MAX_PEERS];
  int in_lg, out_lg;  // log2 of in_rpb / out_rpb (both are powers of two; filled in by the launcher)
};

template <typename V> __global__ void __launch_bounds__(256) transpose_kernel(const V* __restrict__ in, V* __restrict__ out,
                                                                              TransposeArgs p) {
  __shared__ V tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  in += (size_t)blockIdx.z * p.in_z;
  out += (size_t)blockIdx.z * p.out_z;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = r0 + ty + i, c = c0 + tx;
    if (r < p.R && c < p.Cn)
      tile[ty + i][tx] = in[(size_t)(r >> p.in_lg) * p.in_blk + (size_t)(r & (p.in_rpb - 1)) * p.in_ld + c];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, r = r0 + tx;
    if (r < p.R && c < p.Cn) {
      const int blk = c >> p.out_lg;
      V* dst = p.out_tab[0] ? reinterpret_cast<V*>(p.out_tab[blk]) + (size_t)blockIdx.z * p.out_z
                            : out + (size_t)blk * p.out_blk;
      dst[(size_t)(c & (p.out_rpb - 1)) * p.out_ld + r] = tile[tx][ty + i];
    }
  }
}

// strided column copy: dst[i * dst_ld] = src[i * src_ld], i < n (the ky = M/2 column of the user layout)
template <typename V> __global__ void strided_copy_kernel(const V* __restrict__ src, size_t src_ld, V* __restrict__ dst,
                                                          size_t dst_ld, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[(size_t)i * dst_ld] = src[(size_t)i * src_ld];
}

}  // namespace kf
