// Register-resident mixed-radix Stockham FFT for sm_100a.
//
// One FFT of length L is computed by TPF cooperating threads; thread t holds the
// E = L/TPF elements {t + TPF*m}.  A pass of radix R does E/R butterflies per
// thread entirely in registers; between passes the values are exchanged through
// shared memory (one write + one read per element).  The index formulas are the
// ones checked on the CPU by tests/kernel_model.py::stockham_fft:
//   butterfly j reads in[j + r*L/R], twiddles by w^(r*(j%Ns)), DFT-R,
//   writes out[(j/Ns)*Ns*R + (j%Ns) + r*Ns].
// The last pass writes back to {t + TPF*m}, so output and input register layouts
// coincide and FFTs chain without touching memory.
//
// Replaces jnp.fft.{rfftn,irfftn} on the reference's hot path
// (hydrogym/jax/envs/kolmogorov.py:243,247,250); no cuFFT.
#pragma once
#include <cuda_runtime.h>

namespace kf {

template <typename T> struct Cx { T x, y; };
template <typename T> struct VecOf;
template <> struct VecOf<float> { using type = float2; };
template <> struct VecOf<double> { using type = double2; };

template <typename T> __device__ __forceinline__ Cx<T> mk(T a, T b) { Cx<T> r; r.x = a; r.y = b; return r; }
template <typename T> __device__ __forceinline__ Cx<T> operator+(Cx<T> a, Cx<T> b) { return mk<T>(a.x + b.x, a.y + b.y); }
template <typename T> __device__ __forceinline__ Cx<T> operator-(Cx<T> a, Cx<T> b) { return mk<T>(a.x - b.x, a.y - b.y); }
template <typename T> __device__ __forceinline__ Cx<T> operator*(Cx<T> a, Cx<T> b) {
  return mk<T>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
template <typename T> __device__ __forceinline__ Cx<T> operator*(T s, Cx<T> a) { return mk<T>(s * a.x, s * a.y); }
template <typename T> __device__ __forceinline__ Cx<T> conj(Cx<T> a) { return mk<T>(a.x, -a.y); }
// multiply by +i / -i
template <typename T> __device__ __forceinline__ Cx<T> mul_i(Cx<T> a) { return mk<T>(-a.y, a.x); }
template <typename T> __device__ __forceinline__ Cx<T> mul_mi(Cx<T> a) { return mk<T>(a.y, -a.x); }

// vectorised (64-bit for complex64, 128-bit for complex128) global/shared access
template <typename T> __device__ __forceinline__ Cx<T> ldc(const Cx<T>* p) {
  using V = typename VecOf<T>::type;
  V v = *reinterpret_cast<const V*>(p);
  return mk<T>(v.x, v.y);
}
template <typename T> __device__ __forceinline__ void stc(Cx<T>* p, Cx<T> a) {
  using V = typename VecOf<T>::type;
  V v; v.x = a.x; v.y = a.y;
  *reinterpret_cast<V*>(p) = v;
}

// cos/sin(2*pi*k/32), k = 0..8, folded to immediates (k is always a constant expression)
__host__ __device__ constexpr double cos32(int k) {
  return k == 0 ? 1.0 : k == 1 ? 0.98078528040323044913 : k == 2 ? 0.92387953251128675613
       : k == 3 ? 0.83146961230254523708 : k == 4 ? 0.70710678118654752440
       : k == 5 ? 0.55557023301960222474 : k == 6 ? 0.38268343236508977173
       : k == 7 ? 0.19509032201612826785 : 0.0;
}
// exp(-2*pi*i*k/R) for 0 <= k < R/2, R in {2,4,8,16,32}
__host__ __device__ constexpr double tw_cos(int k, int R) {
  int a = k * (32 / R);  // in 32nds of a turn, 0..15
  return a <= 8 ? cos32(a) : -cos32(16 - a);
}
__host__ __device__ constexpr double tw_sin(int k, int R) {  // sin(2*pi*k/R) >= 0 on this range
  int a = k * (32 / R);
  return a <= 8 ? cos32(8 - a) : cos32(a - 8);
}

// Natural-order in, natural-order out DFT of compile-time size R (power of two),
// recursive decimation in time; DIR = -1 forward (exp(-i..)), +1 inverse.
template <int R, int S, int DIR, typename T> struct Dft {
  __device__ __forceinline__ static void run(const Cx<T>* in, Cx<T>* out) {
    Cx<T> e[R / 2], o[R / 2];
    Dft<R / 2, 2 * S, DIR, T>::run(in, e);
    Dft<R / 2, 2 * S, DIR, T>::run(in + S, o);
#pragma unroll
    for (int k = 0; k < R / 2; ++k) {
      Cx<T> w;
      if (k == 0) w = o[0];
      else if (4 * k == R) w = (DIR < 0) ? mul_mi(o[k]) : mul_i(o[k]);
      else {
        const T c = (T)tw_cos(k, R), s = (T)(DIR < 0 ? -tw_sin(k, R) : tw_sin(k, R));
        w = mk<T>(o[k].x * c - o[k].y * s, o[k].x * s + o[k].y * c);
      }
      out[k] = e[k] + w;
      out[k + R / 2] = e[k] - w;
    }
  }
};
template <int S, int DIR, typename T> struct Dft<1, S, DIR, T> {
  __device__ __forceinline__ static void run(const Cx<T>* in, Cx<T>* out) { out[0] = in[0]; }
};

// Compile-time FFT plan: length L, TPF threads per FFT, up to 4 radix passes.
template <int L_, int TPF_, int R0_, int R1_, int R2_ = 1, int R3_ = 1> struct Plan {
  static constexpr int L = L_, TPF = TPF_, E = L_ / TPF_;
  __host__ __device__ static constexpr int radix(int p) { return p == 0 ? R0_ : p == 1 ? R1_ : p == 2 ? R2_ : R3_; }
  static_assert(R0_ * R1_ * R2_ * R3_ == L_, "radices must multiply to L");
  static_assert(E % R0_ == 0 && E % R1_ == 0 && E % R2_ == 0 && E % R3_ == 0, "radix must divide E");
};

// One radix pass (and, recursively, the following ones).  `Ex` is the exchange
// policy: store(pos, value), load(pos), sync() over the TPF threads of this FFT.
template <class P, int PASS, int NS, int DIR, typename T, class Ex> struct PassRunner {
  __device__ __forceinline__ static void run(Cx<T>* v, const Cx<T>* __restrict__ tw, int t, Ex& ex) {
    constexpr int R = P::radix(PASS);
    constexpr int NB = P::E / R;
    constexpr bool kLast = (NS * R == P::L);
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const int j = t + P::TPF * b;
      const int k = j & (NS - 1);
      Cx<T> a[R], o[R];
#pragma unroll
      for (int r = 0; r < R; ++r) a[r] = v[b + r * NB];
      if (NS > 1) {
#pragma unroll
        for (int r = 1; r < R; ++r) {
          Cx<T> w = tw[r * k * (P::L / (NS * R))];
          if (DIR > 0) w.y = -w.y;
          a[r] = a[r] * w;
        }
      }
      Dft<R, 1, DIR, T>::run(a, o);
      if (kLast) {
#pragma unroll
        for (int r = 0; r < R; ++r) v[b + r * NB] = o[r];
      } else {
        const int base = (j / NS) * NS * R + k;
#pragma unroll
        for (int r = 0; r < R; ++r) ex.store(base + r * NS, o[r]);
      }
    }
    if constexpr (!kLast) {
      ex.sync();
#pragma unroll
      for (int m = 0; m < P::E; ++m) v[m] = ex.load(t + P::TPF * m);
      ex.sync();
      PassRunner<P, PASS + 1, NS * R, DIR, T, Ex>::run(v, tw, t, ex);
    }
  }
};

// In-register FFT over the TPF threads that share `ex`.  v[m] <-> element t + TPF*m.
// Unnormalised in both directions.
template <class P, int DIR, typename T, class Ex>
__device__ __forceinline__ void fft_regs(Cx<T>* v, const Cx<T>* __restrict__ tw, int t, Ex& ex) {
  PassRunner<P, 0, 1, DIR, T, Ex>::run(v, tw, t, ex);
}

}  // namespace kf
