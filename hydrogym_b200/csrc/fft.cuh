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
// Everything is generic over a LANE type: float / double run one transform per
// thread group; F2 (two floats in one 64-bit register, sm_100a FFMA2/FADD2/FMUL2)
// and D2 (two doubles) run TWO independent transforms with the same twiddles in
// the two lanes -- for fp32 that halves the issue slots of the butterfly math and
// makes every exchange / global access 128 bits wide.
//
// Replaces jnp.fft.{rfftn,irfftn} on the reference's hot path
// (hydrogym/jax/envs/kolmogorov.py:243,247,250); no cuFFT.
#pragma once
#include <cuda_runtime.h>

#include <cmath>

namespace kf {

// ---------------------------------------------------------------------------
// Lane types
// ---------------------------------------------------------------------------
struct F2 {  // (lo, hi) floats in one 64-bit register; arithmetic is device-only (host stubs never run)
  unsigned long long v;
  F2() = default;
  __host__ __device__ __forceinline__ F2(float lo, float hi) {
#ifdef __CUDA_ARCH__
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi));
#else
    v = 0; (void)lo; (void)hi;
#endif
  }
  __host__ __device__ __forceinline__ explicit F2(double c) : F2((float)c, (float)c) {}
  __device__ __forceinline__ float lo() const { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); (void)b; return a; }
  __device__ __forceinline__ float hi() const { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); (void)a; return b; }
};
__device__ __forceinline__ F2 operator+(F2 a, F2 b) { F2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ F2 operator-(F2 a, F2 b) { F2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ F2 operator*(F2 a, F2 b) { F2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ F2 fmaL(F2 a, F2 b, F2 c) {
  F2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r;
}

struct D2 {  // two doubles; plain scalar arithmetic per lane (no packed fp64 instructions exist)
  double a, b;
  D2() = default;
  __host__ __device__ __forceinline__ D2(double lo, double hi) : a(lo), b(hi) {}
  __host__ __device__ __forceinline__ explicit D2(double c) : a(c), b(c) {}
  __host__ __device__ __forceinline__ double lo() const { return a; }
  __host__ __device__ __forceinline__ double hi() const { return b; }
};
__host__ __device__ __forceinline__ D2 operator+(D2 x, D2 y) { return D2(x.a + y.a, x.b + y.b); }
__host__ __device__ __forceinline__ D2 operator-(D2 x, D2 y) { return D2(x.a - y.a, x.b - y.b); }
__host__ __device__ __forceinline__ D2 operator*(D2 x, D2 y) { return D2(x.a * y.a, x.b * y.b); }
__host__ __device__ __forceinline__ D2 fmaL(D2 x, D2 y, D2 z) { return D2(fma(x.a, y.a, z.a), fma(x.b, y.b, z.b)); }

__host__ __device__ __forceinline__ float fmaL(float a, float b, float c) { return fmaf(a, b, c); }
__host__ __device__ __forceinline__ double fmaL(double a, double b, double c) { return fma(a, b, c); }

// scalar type of a lane type, the two-transform lane type of a scalar, and broadcast
template <typename L> struct LaneOf { using S = L; static constexpr int kLanes = 1; };
template <> struct LaneOf<F2> { using S = float; static constexpr int kLanes = 2; };
template <> struct LaneOf<D2> { using S = double; static constexpr int kLanes = 2; };
template <typename S> struct PairOf;
template <> struct PairOf<float> { using type = F2; };
template <> struct PairOf<double> { using type = D2; };
template <typename L> __host__ __device__ __forceinline__ L bc(typename LaneOf<L>::S s) { return L(s); }
template <> __host__ __device__ __forceinline__ F2 bc<F2>(float s) { return F2(s, s); }
template <> __host__ __device__ __forceinline__ D2 bc<D2>(double s) { return D2(s, s); }

// ---------------------------------------------------------------------------
// Complex numbers over a lane type
// ---------------------------------------------------------------------------
template <typename T> struct Cx { T x, y; };

template <typename T> __host__ __device__ __forceinline__ Cx<T> mk(T a, T b) { Cx<T> r; r.x = a; r.y = b; return r; }
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator+(Cx<T> a, Cx<T> b) { return mk<T>(a.x + b.x, a.y + b.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator-(Cx<T> a, Cx<T> b) { return mk<T>(a.x - b.x, a.y - b.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator*(Cx<T> a, Cx<T> b) {
  return mk<T>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
template <typename T> __host__ __device__ __forceinline__ Cx<T> operator*(T s, Cx<T> a) { return mk<T>(s * a.x, s * a.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> conj(Cx<T> a) { return mk<T>(a.x, -a.y); }
// multiply by +i / -i (scalar lanes only: the negation is free there)
template <typename T> __host__ __device__ __forceinline__ Cx<T> mul_i(Cx<T> a) { return mk<T>(-a.y, a.x); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> mul_mi(Cx<T> a) { return mk<T>(a.y, -a.x); }
// a + conj(b), a - conj(b) without a negation instruction
template <typename T> __host__ __device__ __forceinline__ Cx<T> add_conj(Cx<T> a, Cx<T> b) { return mk<T>(a.x + b.x, a.y - b.y); }
template <typename T> __host__ __device__ __forceinline__ Cx<T> sub_conj(Cx<T> a, Cx<T> b) { return mk<T>(a.x - b.x, a.y + b.y); }
// a * w for a scalar twiddle w broadcast over the lanes: 2 mul + 2 fma
template <typename L> __host__ __device__ __forceinline__ Cx<L> cmul_tw(Cx<L> a, Cx<typename LaneOf<L>::S> w) {
  const L wx = bc<L>(w.x), wy = bc<L>(w.y), nwy = bc<L>(-w.y);
  return mk<L>(fmaL(a.y, nwy, a.x * wx), fmaL(a.y, wx, a.x * wy));
}

// vectorised global/shared access: 64 bits for complex64, 128 for complex128 and for two packed complex64
template <typename T> struct VecOf;
template <> struct VecOf<float> { using type = float2; };
template <> struct VecOf<double> { using type = double2; };
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
template <> __device__ __forceinline__ Cx<F2> ldc<F2>(const Cx<F2>* p) {
  const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(p);
  Cx<F2> r; r.x.v = v.x; r.y.v = v.y;
  return r;
}
template <> __device__ __forceinline__ void stc<F2>(Cx<F2>* p, Cx<F2> a) {
  ulonglong2 v; v.x = a.x.v; v.y = a.y.v;
  *reinterpret_cast<ulonglong2*>(p) = v;
}
template <> __device__ __forceinline__ Cx<D2> ldc<D2>(const Cx<D2>* p) {
  const double2 u = reinterpret_cast<const double2*>(p)[0], w = reinterpret_cast<const double2*>(p)[1];
  return mk<D2>(D2(u.x, u.y), D2(w.x, w.y));
}
template <> __device__ __forceinline__ void stc<D2>(Cx<D2>* p, Cx<D2> a) {
  reinterpret_cast<double2*>(p)[0] = make_double2(a.x.a, a.x.b);
  reinterpret_cast<double2*>(p)[1] = make_double2(a.y.a, a.y.b);
}
// lane values alone (the halves of a split exchange)
__device__ __forceinline__ float ldl(const float* p) { return *p; }
__device__ __forceinline__ double ldl(const double* p) { return *p; }
__device__ __forceinline__ F2 ldl(const F2* p) { F2 r; r.v = *reinterpret_cast<const unsigned long long*>(p); return r; }
__device__ __forceinline__ D2 ldl(const D2* p) { const double2 u = *reinterpret_cast<const double2*>(p); return D2(u.x, u.y); }
__device__ __forceinline__ void stl(float* p, float a) { *p = a; }
__device__ __forceinline__ void stl(double* p, double a) { *p = a; }
__device__ __forceinline__ void stl(F2* p, F2 a) { *reinterpret_cast<unsigned long long*>(p) = a.v; }
__device__ __forceinline__ void stl(D2* p, D2 a) { *reinterpret_cast<double2*>(p) = make_double2(a.a, a.b); }

__device__ __forceinline__ F2 operator-(F2 a) { return F2(0.0) - a; }
__host__ __device__ __forceinline__ D2 operator-(D2 x) { return D2(-x.a, -x.b); }

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

// a * exp(DIR * 2*pi*i * E32/32) with the constant folded; trivial rotations cost no multiplies.
// Written without unary negation of lane values (packed lanes have no free negate in PTX).
template <int E32, int DIR, typename T> __host__ __device__ __forceinline__ Cx<T> rot32(Cx<T> a) {
  constexpr int e = (((DIR < 0 ? -E32 : E32) % 32) + 32) % 32;  // counter-clockwise 32nds of a turn
  if constexpr (e == 0) return a;
  else if constexpr (e == 8) return mk<T>(-a.y, a.x);
  else if constexpr (e == 16) return mk<T>(-a.x, -a.y);
  else if constexpr (e == 24) return mk<T>(a.y, -a.x);
  else if constexpr (e % 8 == 4) {
    // (x + i y)(sc + i ss) h with sc, ss = +-1: x' = h (sc x - ss y), y' = h (ss x + sc y)
    constexpr double h = 0.70710678118654752440;
    constexpr bool cpos = (e == 4 || e == 28), spos = (e == 4 || e == 12);
    if constexpr (cpos && spos) return mk<T>((a.x - a.y) * T(h), (a.x + a.y) * T(h));
    else if constexpr (cpos && !spos) return mk<T>((a.x + a.y) * T(h), (a.y - a.x) * T(h));
    else if constexpr (!cpos && spos) return mk<T>((a.x + a.y) * T(-h), (a.x - a.y) * T(h));
    else return mk<T>((a.y - a.x) * T(h), (a.x + a.y) * T(-h));
  } else {
    constexpr double c = cos32(e), sn = sin32(e);
    return mk<T>(fmaL(a.y, T(-sn), a.x * T(c)), fmaL(a.y, T(c), a.x * T(sn)));
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
    const Cx<T> t0 = a[0] + a[2], t1 = a[0] - a[2], t2 = a[1] + a[3], d = a[1] - a[3];
    a[0] = t0 + t2;
    a[2] = t0 - t2;
    if (DIR < 0) {  // t3 = -i d = (d.y, -d.x)
      a[1] = mk<T>(t1.x + d.y, t1.y - d.x);
      a[3] = mk<T>(t1.x - d.y, t1.y + d.x);
    } else {        // t3 = +i d = (-d.y, d.x)
      a[1] = mk<T>(t1.x - d.y, t1.y + d.x);
      a[3] = mk<T>(t1.x + d.y, t1.y - d.x);
    }
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
  // Twiddles live in a TRANSPOSED table: pass p >= 1 (sub-transform length NS, radix R) owns the block
  // [r - 1][k] = exp(-2*pi*i * r*k / (NS*R)), r = 1..R-1, k = 0..NS-1, at offset tw_off(p).  Threads of one FFT
  // have consecutive k, so every twiddle load is a run of consecutive words (no bank conflicts, unlike
  // tw[r*k*stride] in a linear table).  Total size sum (R-1)*NS = L - R0 <= L.
  __host__ __device__ static constexpr int ns(int p) { return p == 0 ? 1 : ns(p - 1) * radix(p - 1); }
  __host__ __device__ static constexpr int tw_off(int p) { return p <= 1 ? 0 : tw_off(p - 1) + (radix(p - 1) - 1) * ns(p - 1); }
  static_assert(R0_ * R1_ * R2_ * R3_ == L_, "radices must multiply to L");
  static_assert(E % R0_ == 0 && E % R1_ == 0 && E % R2_ == 0 && E % R3_ == 0, "radix must divide E");
};

// One radix pass (and, recursively, the following ones).  `Ex` is the exchange
// policy over the TPF threads of this FFT; T is the lane type of the data, twiddles are
// scalars of LaneOf<T>::S broadcast over the lanes.  Two kinds of policy:
//   Ex::kSplit == false : store(pos, Cx<T>), load(pos), sync()  -- whole complex numbers
//   Ex::kSplit == true  : store_part(pos, T), load_part(pos), sync() -- real parts, then
//                         imaginary parts, through a buffer of HALF the size (two rounds)
template <class P, int PASS, int NS, int DIR, typename T, class Ex> struct PassRunner {
  using S = typename LaneOf<T>::S;
  __device__ __forceinline__ static void run(Cx<T>* v, const Cx<S>* __restrict__ tw, int t, Ex& ex) {
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
          Cx<S> w = ldc(tw + P::tw_off(PASS) + (r - 1) * NS + k);
          if (DIR > 0) w.y = -w.y;
          a[r] = cmul_tw<T>(a[r], w);
        }
      }
      Radix<R>::template run<DIR, T>(a);
      if constexpr (kLast || Ex::kSplit) {
#pragma unroll
        for (int r = 0; r < R; ++r) v[b + r * NB] = a[Radix<R>::pos(r)];
      } else {
        const int base = (j / NS) * NS * R + k;
#pragma unroll
        for (int r = 0; r < R; ++r) ex.store(base + r * NS, a[Radix<R>::pos(r)]);
      }
    }
    if constexpr (!kLast) {
      if constexpr (Ex::kSplit) {
        T nx[P::E];
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          const int j = t + P::TPF * b;
          const int base = (j / NS) * NS * R + (j & (NS - 1));
#pragma unroll
          for (int r = 0; r < R; ++r) ex.store_part(base + r * NS, v[b + r * NB].x);
        }
        ex.sync();
#pragma unroll
        for (int m = 0; m < P::E; ++m) nx[m] = ex.load_part(t + P::TPF * m);
        ex.sync();
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          const int j = t + P::TPF * b;
          const int base = (j / NS) * NS * R + (j & (NS - 1));
#pragma unroll
          for (int r = 0; r < R; ++r) ex.store_part(base + r * NS, v[b + r * NB].y);
        }
        ex.sync();
#pragma unroll
        for (int m = 0; m < P::E; ++m) v[m] = mk<T>(nx[m], ex.load_part(t + P::TPF * m));
        ex.sync();
      } else {
        ex.sync();
#pragma unroll
        for (int m = 0; m < P::E; ++m) v[m] = ex.load(t + P::TPF * m);
        ex.sync();
      }
      PassRunner<P, PASS + 1, NS * R, DIR, T, Ex>::run(v, tw, t, ex);
    }
  }
};

// Host side: fill the transposed twiddle table of plan P (L complex numbers as interleaved doubles).
template <class P> inline void fill_twiddles(double* out) {
  const double kPi = 3.14159265358979323846264338327950288;
  for (int i = 0; i < 2 * P::L; ++i) out[i] = 0.0;
  int off = 0, ns = P::radix(0);
  for (int p = 1; p < 4 && ns < P::L; ++p) {
    const int R = P::radix(p);
    for (int r = 1; r < R; ++r)
      for (int k = 0; k < ns; ++k) {
        const double ang = -2.0 * kPi * (double)r * (double)k / ((double)ns * R);
        out[2 * (off + (r - 1) * ns + k)] = cos(ang);
        out[2 * (off + (r - 1) * ns + k) + 1] = sin(ang);
      }
    off += (R - 1) * ns;
    ns *= R;
  }
}

// In-register FFT over the TPF threads that share `ex`.  v[m] <-> element t + TPF*m.
// Unnormalised in both directions.
template <class P, int DIR, typename T, class Ex>
__device__ __forceinline__ void fft_regs(Cx<T>* v, const Cx<typename LaneOf<T>::S>* __restrict__ tw, int t, Ex& ex) {
  PassRunner<P, 0, 1, DIR, T, Ex>::run(v, tw, t, ex);
}

}  // namespace kf
