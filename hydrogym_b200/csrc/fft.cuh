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

template <typename T> __host__ __device__ __forceinline__ Cx<T> mk(T a, T b) { Cx<T> r; r.x = a; r.y = b; return r; }
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator+(Cx<T> a, Cx<T> b) { return mk<T>(a.x + b.x, a.y + b.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator-(Cx<T> a, Cx<T> b) { return mk<T>(a.x - b.x, a.y - b.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator*(Cx<T> a, Cx<T> b) {
  return mk<T>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator*(T s, Cx<T> a) { return mk<T>(s * a.x, s * a.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> conj(Cx<T> a) { return mk<T>(a.x, -a.y); }
// multiply by +i / -i
template <typename T> __host__ __device__ __forceinline__ Cx<T> mul_i(Cx<T> a) { return mk<T>(-a.y, a.x); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> mul_mi(Cx<T> a) { return mk<T>(a.y, -a.x); }

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

// cos(2*pi*a/32) for any integer a, folded to an immediate (a is always a constant expression)
__host__ __device__ constexpr double cos32q(int k) {  // first quadrant, k = 0..8
  return k == 0 ? 1.0 : k == 1 ? 0.98078528040323044913 : k == 2 ? 0.92387953251128675613
       : k == 3 ? 0.83146961230254523708 : k == 4 ? 0.70710678118654752440
       : k == 5 ? 0.55557023301960222474 : k == 6 ? 0.38268343236508977173
       : k == 7 ? 0.19509032201612826785 : 0.0;
}
__host__ __device__ constexpr double cos32(int a) {
  a = ((a % 32) + 32) % 32;
  return a <= 8 ? cos32q(a) : a <= 16 ? -cos32q(16 - a) : a <= 24 ? -cos32q(a - 16) : cos32q(32 - a);
}
__host__ __device__ constexpr double sin32(int a) { return cos32(a - 8); }

// a * exp(DIR * 2*pi*i * E32/32) with the constant folded; trivial rotations cost no multiplies
template <int E32, int DIR, typename T> __host__ __device__ __forceinline__ Cx<T> rot32(Cx<T> a) {
  constexpr int e = (((DIR < 0 ? -E32 : E32) % 32) + 32) % 32;  // counter-clockwise 32nds of a turn
  if constexpr (e == 0) return a;
  else if constexpr (e == 8) return mul_i(a);
  else if constexpr (e == 16) return mk<T>(-a.x, -a.y);
  else if constexpr (e == 24) return mul_mi(a);
  else if constexpr (e % 8 == 4) {
    constexpr T h = (T)0.70710678118654752440;
    constexpr int sc = (e == 4 || e == 28) ? 1 : -1, ss = (e == 4 || e == 12) ? 1 : -1;
    // (x + i y)(sc + i ss) h
    return mk<T>(h * ((T)sc * a.x - (T)ss * a.y), h * ((T)ss * a.x + (T)sc * a.y));
  } else {
    constexpr T c = (T)cos32(e), sn = (T)sin32(e);
    return mk<T>(a.x * c - a.y * sn, a.x * sn + a.y * c);
  }
}

// In-place DFT of compile-time size R on registers a[0..R).  Output X[k] is left at
// position Radix<R>::pos(k) (digit-reversed); callers index with pos(), which costs
// nothing once unrolled.  DIR = -1 forward, +1 inverse (unnormalised).
template <int R> struct Radix;
template <> struct Radix<1> {
  __host__ __device__ static constexpr int pos(int k) { return k; }
  template <int DIR, typename T> __host__ __device__ __forceinline__ static void run(Cx<T>*) {}
};
template <> struct Radix<2> {
  __host__ __device__ static constexpr int pos(int k) { return k; }
  template <int DIR, typename T> __host__ __device__ __forceinline__ static void run(Cx<T>* a) {
    const Cx<T> t = a[0] - a[1];
    a[0] = a[0] + a[1];
    a[1] = t;
  }
};
template <> struct Radix<4> {
  __host__ __device__ static constexpr int pos(int k) { return k; }
  template <int DIR, typename T> __host__ __device__ __forceinline__ static void run(Cx<T>* a) {
    const Cx<T> t0 = a[0] + a[2], t1 = a[0] - a[2], t2 = a[1] + a[3];
    const Cx<T> t3 = (DIR < 0) ? mul_mi(a[1] - a[3]) : mul_i(a[1] - a[3]);
    a[0] = t0 + t2; a[2] = t0 - t2; a[1] = t1 + t3; a[3] = t1 - t3;
  }
};
// R = 4 * R2: radix-4 over stride R2, twiddle, then R2-point DFTs on contiguous blocks.
// X[k1 + 4*k2] ends up at position R2*k1 + Radix<R2>::pos(k2).
template <int R> struct Radix {
  static constexpr int R2 = R / 4;
  static_assert(R == 8 || R == 16 || R == 32, "supported radices: 2, 4, 8, 16, 32");
  __host__ __device__ static constexpr int pos(int k) { return R2 * (k % 4) + Radix<R2>::pos(k / 4); }
  template <int DIR, typename T, int N2> __host__ __device__ __forceinline__ static void column(Cx<T>* a) {
    if constexpr (N2 < R2) {
      Cx<T> b[4] = {a[N2], a[R2 + N2], a[2 * R2 + N2], a[3 * R2 + N2]};
      Radix<4>::run<DIR, T>(b);
      a[N2] = b[0];
      a[R2 + N2] = rot32<N2 * 1 * (32 / R), DIR, T>(b[1]);
      a[2 * R2 + N2] = rot32<N2 * 2 * (32 / R), DIR, T>(b[2]);
      a[3 * R2 + N2] = rot32<N2 * 3 * (32 / R), DIR, T>(b[3]);
      column<DIR, T, N2 + 1>(a);
    }
  }
  template <int DIR, typename T> __host__ __device__ __forceinline__ static void run(Cx<T>* a) {
    column<DIR, T, 0>(a);
#pragma unroll
    for (int k1 = 0; k1 < 4; ++k1) Radix<R2>::template run<DIR, T>(a + R2 * k1);
  }
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
      Cx<T> a[R];
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
      Radix<R>::template run<DIR, T>(a);
      if (kLast) {
#pragma unroll
        for (int r = 0; r < R; ++r) v[b + r * NB] = a[Radix<R>::pos(r)];
      } else {
        const int base = (j / NS) * NS * R + k;
#pragma unroll
        for (int r = 0; r < R; ++r) ex.store(base + r * NS, a[Radix<R>::pos(r)]);
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
