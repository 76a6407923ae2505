// 3-D channel-flow step: tile kernels.  Reference: C = hydrogym/jax/envs/channel.py,
// S = hydrogym/jax/solvers/base.py.  See DESIGN.md §7b for the derivation; in short:
//
// * the state is FFT'd over (x, y) and physical in z (fastest axis, 72 points), so every operator of the reference acts
//   on ONE axis: x-transforms on tiles (x, z) at fixed y / ky ("X tiles", rows of 72 contiguous z), y-transforms on tiles
//   (y, z) at fixed x ("Y tiles", one contiguous 72 x 72 block), Chebyshev matrices and the per-mode pressure inverses
//   along z inside either tile.  A CTA owns one tile of one field in shared memory; z is the batch axis of the
//   length-72 transforms (72 = 8 x 9 Cooley-Tukey, two register-resident butterfly stages), so shared-memory accesses
//   are conflict-free and global accesses are runs of 72 contiguous elements.
// * wall conditions replace whole z-planes of a field (C:227-241), and an (x, y) FFT acts plane by plane, so the
//   reference's "to physical, fix the walls, back to spectral" is a plane replacement in EITHER space; everything the
//   reference computes from the wall-corrected field that is linear in it (the viscous term C:286-300, the whole
//   projection C:311-372, the mass-flux shift S:166-183) is evaluated spectrally per (kx, ky) mode, with the Hermitian
//   projection of `.real` (C:154) applied where it matters: the x- (y-) Nyquist row of an x- (y-) derivative is zero.
// * only the advective products need physical space: per RK stage one X pass (finish the previous stage's right-hand
//   side, RK update, inverse x-transforms of the state and of its x-derivative), one Y pass (inverse y-transforms,
//   two real fields per complex transform, walls, z-derivative) and one Y pass for the products + forward y-transform.
// * a thread owns two adjacent z-columns (16-byte accesses); in fp32 the 72 x 72 z-matrix products run on the tensor
//   cores (mma.sync TF32, 3-product hi / lo split), in fp64 on FMAs; the matrices arrive by cp.async and are waited for
//   only before the product; the epilogue's operands are prefetched into L2 at kernel start.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <type_traits>

namespace kc {

constexpr int N = 72;
constexpr int NN = N * N;
constexpr int V = N * N * N;
constexpr int ZP = N / 2;       // each thread owns two adjacent z-columns (16-byte accesses in fp32)
constexpr int THREADS = ZP * 9;  // 36 column pairs x 9 butterfly lanes
constexpr int NACT = 24;
constexpr int NAB = 37;
template <typename T> constexpr int ctas_per_sm() { return sizeof(T) == 4 ? 2 : 1; }

template <typename T> struct __align__(2 * sizeof(T)) C2 { T x, y; };
template <typename T> __device__ __forceinline__ C2<T> mk(T a, T b) { C2<T> r; r.x = a; r.y = b; return r; }
template <typename T> __device__ __forceinline__ C2<T> operator+(C2<T> a, C2<T> b) { return mk<T>(a.x + b.x, a.y + b.y); }
template <typename T> __device__ __forceinline__ C2<T> operator-(C2<T> a, C2<T> b) { return mk<T>(a.x - b.x, a.y - b.y); }
template <typename T> __device__ __forceinline__ C2<T> operator*(T s, C2<T> a) { return mk<T>(s * a.x, s * a.y); }
template <typename T> __device__ __forceinline__ C2<T> cmul(C2<T> a, C2<T> b) { return mk<T>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
template <typename T> __device__ __forceinline__ void cfma(C2<T>& acc, C2<T> a, C2<T> b) {
  acc.x = fma(a.x, b.x, acc.x); acc.x = fma(-a.y, b.y, acc.x);
  acc.y = fma(a.x, b.y, acc.y); acc.y = fma(a.y, b.x, acc.y);
}
template <typename T> __device__ __forceinline__ C2<T> mul_i(C2<T> a) { return mk<T>(-a.y, a.x); }  // i * a

// integer wavenumber of index i (fftfreq order; index 36 is the Nyquist line -36)
__device__ __forceinline__ int wave(int i) { return i < N / 2 ? i : i - N; }
__device__ __forceinline__ bool kept(int i) { int p = wave(i); return (p < 0 ? -p : p) <= N / 3; }  // U:8-19

// In-register 8-point DFT (radix-2, three stages) and 9-point DFT (3 x 3), exponent sign - (forward) or + (INV).
template <typename T, bool INV> __device__ __forceinline__ C2<T> rot(C2<T> a) {  // (-i) * a forward, (+i) * a inverse
  return INV ? mk<T>(-a.y, a.x) : mk<T>(a.y, -a.x);
}
template <typename T, bool INV> __device__ __forceinline__ void dft8(C2<T> (&x)[8]) {
  constexpr T h = T(0.70710678118654752440);
  const C2<T> a0 = x[0] + x[4], a1 = x[0] - x[4], a2 = x[2] + x[6], a3 = rot<T, INV>(x[2] - x[6]);
  const C2<T> a4 = x[1] + x[5], a5 = x[1] - x[5], a6 = x[3] + x[7], a7 = rot<T, INV>(x[3] - x[7]);
  const C2<T> b0 = a0 + a2, b2 = a0 - a2, b1 = a1 + a3, b3 = a1 - a3;
  const C2<T> b4 = a4 + a6, b6 = rot<T, INV>(a4 - a6), c5 = a5 + a7, c7 = a5 - a7;
  // W8 * c5 and W8^3 * c7: W8 = (1 -+ i) / sqrt(2), W8^3 = (-1 -+ i) / sqrt(2)
  const C2<T> b5 = INV ? mk<T>(h * (c5.x - c5.y), h * (c5.x + c5.y)) : mk<T>(h * (c5.x + c5.y), h * (c5.y - c5.x));
  const C2<T> b7 = INV ? mk<T>(-h * (c7.x + c7.y), h * (c7.x - c7.y)) : mk<T>(h * (c7.y - c7.x), -h * (c7.x + c7.y));
  x[0] = b0 + b4; x[4] = b0 - b4; x[1] = b1 + b5; x[5] = b1 - b5;
  x[2] = b2 + b6; x[6] = b2 - b6; x[3] = b3 + b7; x[7] = b3 - b7;
}
template <typename T, bool INV> __device__ __forceinline__ void dft3(C2<T>& a, C2<T>& b, C2<T>& c) {
  constexpr T s3 = T(0.86602540378443864676);
  const C2<T> s = b + c, d = b - c;
  const C2<T> t = mk<T>(a.x - T(0.5) * s.x, a.y - T(0.5) * s.y);
  const C2<T> u = INV ? mk<T>(-s3 * d.y, s3 * d.x) : mk<T>(s3 * d.y, -s3 * d.x);   // (-+ i sqrt(3)/2) * d
  a = a + s; b = t + u; c = t - u;
}
template <typename T, bool INV> __device__ __forceinline__ void dft9(C2<T> (&x)[9]) {
  constexpr T c1 = T(0.76604444311897803520), s1 = T(0.64278760968653932632);   // cos / sin of 2 pi / 9
  constexpr T c2 = T(0.17364817766693034885), s2 = T(0.98480775301220805937);   // 4 pi / 9
  constexpr T c4 = T(-0.93969262078590838405), s4 = T(0.34202014332566873304);  // 8 pi / 9
  auto tw = [](T c, T s) { return INV ? mk<T>(c, s) : mk<T>(c, -s); };
  // n = 3 n1 + n2, k = k1 + 3 k2: DFT3 over n1 for each n2, twiddle W9^(n2 k1), DFT3 over n2 for each k1
  dft3<T, INV>(x[0], x[3], x[6]);
  dft3<T, INV>(x[1], x[4], x[7]);
  dft3<T, INV>(x[2], x[5], x[8]);
  x[4] = cmul(x[4], tw(c1, s1)); x[7] = cmul(x[7], tw(c2, s2));   // n2 = 1: k1 = 1, 2
  x[5] = cmul(x[5], tw(c2, s2)); x[8] = cmul(x[8], tw(c4, s4));   // n2 = 2: k1 = 1, 2
  dft3<T, INV>(x[0], x[1], x[2]);   // k1 = 0 -> X[0], X[3], X[6]
  dft3<T, INV>(x[3], x[4], x[5]);   // k1 = 1 -> X[1], X[4], X[7]
  dft3<T, INV>(x[6], x[7], x[8]);   // k1 = 2 -> X[2], X[5], X[8]
  const C2<T> y1 = x[3], y2 = x[6], y3 = x[1], y5 = x[7], y6 = x[2], y7 = x[5];
  x[1] = y1; x[2] = y2; x[3] = y3; x[5] = y5; x[6] = y6; x[7] = y7;
}

// asynchronous global -> shared copy of n bytes (multiple of 16, both 16-byte aligned) by the whole CTA, and its wait
__device__ __forceinline__ void async_copy(void* smem_dst, const void* gmem_src, int nbytes, int tid, int nthreads) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  const char* g = reinterpret_cast<const char*>(gmem_src);
  for (int o = tid * 16; o < nbytes; o += nthreads * 16)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d + o), "l"(g + o) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void async_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// programmatic dependent launch: let the next kernel of the stream start its prologue now; wait for the previous
// kernel's results (no-ops when the launch does not carry the attribute)
__device__ __forceinline__ void grid_dependency_wait() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// two adjacent z-columns of a tile row / global row, moved as one vector
template <typename E> struct __align__(2 * sizeof(E)) Two { E a, b; };
template <typename E> __device__ __forceinline__ Two<E> ld2(const E* p) { return *reinterpret_cast<const Two<E>*>(p); }
template <typename E> __device__ __forceinline__ void st2(E* p, E a, E b) { Two<E> v; v.a = a; v.b = b; *reinterpret_cast<Two<E>*>(p) = v; }

// Length-72 DFT along the slow axis of a [72][72] shared-memory tile, for the two columns z0, z0 + 1, butterfly lane r
// (0..8).  72 = 8 x 9: lane r does the 8-point DFT over n1 of x[9 n1 + r] and the twiddle W72^(r k1) (in place: it owns
// those 8 slots), then lanes 0..7 do the 9-point DFT over n2 for k1 = r.  tw[j] = exp(-2 pi i j / 72); INV conjugates.
// Unnormalised.  Ends with a barrier.
template <typename T, bool INV> __device__ __forceinline__ void fft72(C2<T>* t, const C2<T>* tw, int z0, int r) {
  {
    C2<T> a[2][8];
#pragma unroll
    for (int n1 = 0; n1 < 8; ++n1) {
      const Two<C2<T>> v = ld2(t + (9 * n1 + r) * N + z0);
      a[0][n1] = v.a; a[1][n1] = v.b;
    }
    dft8<T, INV>(a[0]);
    dft8<T, INV>(a[1]);
    st2(t + r * N + z0, a[0][0], a[1][0]);
#pragma unroll
    for (int k1 = 1; k1 < 8; ++k1) {
      C2<T> w = tw[r * k1];   // r k1 <= 56
      if (INV) w.y = -w.y;
      st2(t + (r + 9 * k1) * N + z0, cmul(a[0][k1], w), cmul(a[1][k1], w));
    }
  }
  __syncthreads();
  C2<T> b[2][9];
  if (r < 8) {
#pragma unroll
    for (int n2 = 0; n2 < 9; ++n2) {
      const Two<C2<T>> v = ld2(t + (n2 + 9 * r) * N + z0);
      b[0][n2] = v.a; b[1][n2] = v.b;
    }
    dft9<T, INV>(b[0]);
    dft9<T, INV>(b[1]);
  }
  __syncthreads();
  if (r < 8) {
#pragma unroll
    for (int k2 = 0; k2 < 9; ++k2) st2(t + (r + 8 * k2) * N + z0, b[0][k2], b[1][k2]);
  }
  __syncthreads();
}

// The same transform on two tiles at once (stage by stage: 3 barriers instead of 6).
template <typename T, bool INV> __device__ __forceinline__ void fft72_pair(C2<T>* ta, C2<T>* tb, const C2<T>* tw, int z0, int r) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    C2<T>* t = h ? tb : ta;
    C2<T> a[2][8];
#pragma unroll
    for (int n1 = 0; n1 < 8; ++n1) {
      const Two<C2<T>> v = ld2(t + (9 * n1 + r) * N + z0);
      a[0][n1] = v.a; a[1][n1] = v.b;
    }
    dft8<T, INV>(a[0]);
    dft8<T, INV>(a[1]);
    st2(t + r * N + z0, a[0][0], a[1][0]);
#pragma unroll
    for (int k1 = 1; k1 < 8; ++k1) {
      C2<T> w = tw[r * k1];
      if (INV) w.y = -w.y;
      st2(t + (r + 9 * k1) * N + z0, cmul(a[0][k1], w), cmul(a[1][k1], w));
    }
  }
  __syncthreads();
  C2<T> b[2][2][9];
  if (r < 8) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      C2<T>* t = h ? tb : ta;
#pragma unroll
      for (int n2 = 0; n2 < 9; ++n2) {
        const Two<C2<T>> v = ld2(t + (n2 + 9 * r) * N + z0);
        b[h][0][n2] = v.a; b[h][1][n2] = v.b;
      }
    }
  }
  __syncthreads();
  if (r < 8) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      C2<T>* t = h ? tb : ta;
      dft9<T, INV>(b[h][0]);
      dft9<T, INV>(b[h][1]);
#pragma unroll
      for (int k2 = 0; k2 < 9; ++k2) st2(t + (r + 8 * k2) * N + z0, b[h][0][k2], b[h][1][k2]);
    }
  }
  __syncthreads();
}

// out[c][m] = sum_j Mt[j][i0 + c] * in[(r + 9 m)][j] for the 8 rows of lane r and the two columns i0, i0 + 1: a 72 x 72
// real matrix (stored transposed, so consecutive i are consecutive words) applied along the fast axis of a [72][72]
// tile.  The operand is read with 16-byte broadcast loads (4 floats / 2 doubles / 2 complex floats / 1 complex double),
// the matrix as pairs: 10 shared-memory instructions per 64 (fp32 complex) multiply-adds.
template <typename T> __device__ __forceinline__ void zmat_real(const T* Mt, int i0, int r, const T* in, T (&out)[2][8]) {
  constexpr int JV = 16 / sizeof(T);
  struct __align__(16) Vec { T v[JV]; };
#pragma unroll
  for (int m = 0; m < 8; ++m) { out[0][m] = T(0); out[1][m] = T(0); }
  for (int j = 0; j < N; j += JV) {
    Two<T> mj[JV];
#pragma unroll
    for (int q = 0; q < JV; ++q) mj[q] = ld2(Mt + (j + q) * N + i0);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const Vec v = *reinterpret_cast<const Vec*>(in + (r + 9 * m) * N + j);
#pragma unroll
      for (int q = 0; q < JV; ++q) {
        out[0][m] = fma(mj[q].a, v.v[q], out[0][m]);
        out[1][m] = fma(mj[q].b, v.v[q], out[1][m]);
      }
    }
  }
}
template <typename T> __device__ __forceinline__ void zmat_cplx(const T* Mt, int i0, int r, const C2<T>* in, C2<T> (&out)[2][8]) {
  constexpr int JV = 16 / sizeof(C2<T>);
  struct __align__(16) Vec { C2<T> v[JV]; };
#pragma unroll
  for (int m = 0; m < 8; ++m) { out[0][m] = mk<T>(T(0), T(0)); out[1][m] = out[0][m]; }
  for (int j = 0; j < N; j += JV) {
    Two<T> mj[JV];
#pragma unroll
    for (int q = 0; q < JV; ++q) mj[q] = ld2(Mt + (j + q) * N + i0);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const Vec v = *reinterpret_cast<const Vec*>(in + (r + 9 * m) * N + j);
#pragma unroll
      for (int q = 0; q < JV; ++q) {
        out[0][m].x = fma(mj[q].a, v.v[q].x, out[0][m].x);
        out[0][m].y = fma(mj[q].a, v.v[q].y, out[0][m].y);
        out[1][m].x = fma(mj[q].b, v.v[q].x, out[1][m].x);
        out[1][m].y = fma(mj[q].b, v.v[q].y, out[1][m].y);
      }
    }
  }
}

// ---- fp32 z-matrix products on the tensor cores: mma.sync m16n8k8 TF32 with the 3-product split ------------------
// x = hi + lo with hi = tf32(x), lo = tf32(x - hi); D += a_lo b_hi + a_hi b_lo + a_hi b_hi keeps fp32-level accuracy
// (the dropped a_lo b_lo term is 2^-22 relative).  A 72 x 72 real matrix times the 144 (re / im rows) x 72 operand of
// an X tile is 9 x 9 x 9 MMA tiles x 3 products; the FMA version spends half of the kernel's instructions on it.
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
  const float rest = x - __uint_as_float(hi);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(rest));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// D[rho][i] = sum_j A[rho][j] Mt[j][i] for NT consecutive 8-column tiles starting at tile n_first, one 16-row tile:
// `aoff(rho)` gives the float offset of A[rho][0] in `A`, `astep` the float stride between consecutive j.
template <int NT, typename AOff>
__device__ __forceinline__ void mma_rows(const float* A, AOff aoff, int astep, const float* Mt, int m0, int n_first, int lane,
                                         float (&d)[NT][4]) {
  const int g = lane >> 2, t = lane & 3;
  const int oa = aoff(m0 + g), ob = aoff(m0 + g + 8);
#pragma unroll
  for (int n = 0; n < NT; ++n) { d[n][0] = 0.f; d[n][1] = 0.f; d[n][2] = 0.f; d[n][3] = 0.f; }
  for (int k0 = 0; k0 < N; k0 += 8) {
    uint32_t ahi[4], alo[4];
    tf32_split(A[oa + astep * (k0 + t)], ahi[0], alo[0]);
    tf32_split(A[ob + astep * (k0 + t)], ahi[1], alo[1]);
    tf32_split(A[oa + astep * (k0 + t + 4)], ahi[2], alo[2]);
    tf32_split(A[ob + astep * (k0 + t + 4)], ahi[3], alo[3]);
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      const int col = (n_first + n) * 8 + g;
      uint32_t bhi[2], blo[2];
      tf32_split(Mt[(k0 + t) * N + col], bhi[0], blo[0]);
      tf32_split(Mt[(k0 + t + 4) * N + col], bhi[1], blo[1]);
      mma_tf32(d[n], alo, bhi);
      mma_tf32(d[n], ahi, blo);
      mma_tf32(d[n], ahi, bhi);
    }
  }
}
// complex X-tile operand in place: t (72 x 72 complex floats) := t * Mt along z.  Two barriers inside; the caller has
// synchronised the operand.  Rows rho = 2 * row + part (re / im interleaved: 2-way bank conflicts on the A loads only).
__device__ __forceinline__ void zmat_mma_cplx_inplace(const float* Mt, C2<float>* tile, int tid) {
  float* A = reinterpret_cast<float*>(tile);
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  auto aoff = [](int rho) { return (rho >> 1) * (2 * N) + (rho & 1); };
  float d[9][4];
  if (warp < 9) mma_rows<9>(A, aoff, 2, Mt, warp * 16, 0, lane, d);
  __syncthreads();
  if (warp < 9) {
    const int oa = aoff(warp * 16 + g), ob = aoff(warp * 16 + g + 8);
#pragma unroll
    for (int n = 0; n < 9; ++n) {
      const int c = n * 8 + 2 * t;
      A[oa + 2 * c] = d[n][0]; A[oa + 2 * (c + 1)] = d[n][1];
      A[ob + 2 * c] = d[n][2]; A[ob + 2 * (c + 1)] = d[n][3];
    }
  }
  __syncthreads();
}
// real Y-tile operand: out[row][i] = sum_j in[row][j] Mt[j][i], rows 0..71 (in / out: float [>= 80][72], distinct).
// Ten warps: 5 row tiles (the last one half empty) x 2 column halves.  One barrier at the end.
__device__ __forceinline__ void zmat_mma_real(const float* Mt, const float* in, float* out, int tid) {
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  auto aoff = [](int row) { return row * N; };
  if (warp < 10) {
    const int m0 = (warp % 5) * 16, nf = (warp / 5) * 5;
    float d[5][4];
    if (nf == 0) {
      mma_rows<5>(in, aoff, 1, Mt, m0, 0, lane, d);
    } else {
      float d4[4][4];
      mma_rows<4>(in, aoff, 1, Mt, m0, 5, lane, d4);
#pragma unroll
      for (int n = 0; n < 4; ++n) { d[n][0] = d4[n][0]; d[n][1] = d4[n][1]; d[n][2] = d4[n][2]; d[n][3] = d4[n][3]; }
    }
    const int nt = nf == 0 ? 5 : 4;
    const int ra = m0 + g, rb = ra + 8;
#pragma unroll
    for (int n = 0; n < 5; ++n) {
      if (n < nt) {
        const int c = (nf + n) * 8 + 2 * t;
        if (ra < N) { out[ra * N + c] = d[n][0]; out[ra * N + c + 1] = d[n][1]; }
        if (rb < N) { out[rb * N + c] = d[n][2]; out[rb * N + c + 1] = d[n][3]; }
      }
    }
  }
  __syncthreads();
}

template <typename T> struct Consts {
  const C2<T>* tw;     // [72]
  const T* kx;         // [72] C:114-118
  const T* ky;
  const T* dzt;        // [72][72] transposed Dz
  const T* dzzt;
  const T* ainv_t;     // [37*37][72][72] transposed inverses
  const T* jets;       // [24][72*72]
  const T* jets_dx;
  const T* jets_dy;
  const C2<T>* jets_hat;
};

enum XMode { kXFirst = 0, kXMid = 1, kXLast = 2, kXForward = 3 };

template <typename T> struct XArgs {
  Consts<T> c;
  int mode, stage, B, jets_on;
  const C2<T>* BN;   // [3][B][V] forward-y-transformed products (x, ky, z)
  C2<T>* S0;         // base state of the RK4 step
  C2<T>* S;          // stage state
  C2<T>* ACC;        // k1 + 2 k2 + 2 k3
  C2<T>* A;          // inverse-x-transformed stage state (x, ky, z)
  C2<T>* AX;         // ... of its x-derivative
  C2<T>* RHS;        // optional export of this stage's k
  const C2<T>* M0hat;  // [B][72*72] fft2 of the jet pattern of each environment's action
  const T* gain;     // [B] -gain(t) of this stage's time (C:231-239)
  T nu, force;       // force = fx * 72 * 72 on mode (0, 0) of field 0 (C:267-275)
  T a_dt;            // stage state = S0 + a_dt * k
  T acc_w;           // ACC (+)= acc_w * k
  T dt6;             // kXLast: S = S0 + dt6 * (ACC + k)
};

// copy a tile between global memory (row stride `gs` elements) and shared memory, two columns per access
template <typename E> __device__ __forceinline__ void tile_load(E* t, const E* g, size_t gs, int z0, int r) {
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const Two<E> v = ld2(g + (size_t)(r + 9 * m) * gs + z0);
    st2(t + (r + 9 * m) * N + z0, v.a, v.b);
  }
}
template <typename E> __device__ __forceinline__ void tile_store(E* g, const E* t, size_t gs, int z0, int r) {
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const Two<E> v = ld2(t + (r + 9 * m) * N + z0);
    st2(g + (size_t)(r + 9 * m) * gs + z0, v.a, v.b);
  }
}

// grid (3 fields, 72 ky, B), THREADS threads, dynamic smem = 2 tiles + dzzt + twiddles
template <typename T> __global__ void __launch_bounds__(THREADS, ctas_per_sm<T>()) x_kernel(const XArgs<T> a) {
  extern __shared__ __align__(32) unsigned char smem_raw[];
  C2<T>* t1 = reinterpret_cast<C2<T>*>(smem_raw);
  C2<T>* t2 = t1 + NN;
  T* mt = reinterpret_cast<T*>(t2 + NN);
  C2<T>* tw = reinterpret_cast<C2<T>*>(mt + NN);
  const int f = blockIdx.x, kyi = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, z0 = 2 * (tid % ZP), r = tid / ZP;
  const size_t base = ((size_t)f * a.B + b) * V + (size_t)kyi * N;  // + x * NN + z
  if (tid < N) tw[tid] = a.c.tw[tid];
  if (a.mode == kXMid || a.mode == kXLast)
    async_copy(mt, a.c.dzzt, NN * (int)sizeof(T), tid, THREADS);   // needed only by the matrix product
  grid_dependency_wait();
  if (a.mode == kXMid || a.mode == kXLast) {
    // the epilogue's operands (base state, accumulator) on their way into L2 while the transforms run
    constexpr int SEGS = (N * (int)sizeof(C2<T>) + 127) / 128;   // 72 rows x 576 (1152) bytes in 128-byte lines
    for (int l = tid; l < 2 * N * SEGS; l += THREADS) {
      const int arr = l / (N * SEGS), row = (l / SEGS) % N, seg = l % SEGS;
      const C2<T>* p = (arr ? a.ACC : a.S0) + base + (size_t)row * NN;
      prefetch_l2(reinterpret_cast<const char*>(p) + seg * 128);
    }
  }
  const bool my = kept(kyi);

  if (a.mode == kXForward) {  // plain forward x-transform of BN into S0 (C:156-161)
    tile_load(t1, a.BN + base, NN, z0, r);
    __syncthreads();
    fft72<T, false>(t1, tw, z0, r);
    tile_store(a.S0 + base, t1, NN, z0, r);
    return;
  }

  if (a.mode == kXFirst) {
    tile_load(t1, a.S0 + base, NN, z0, r);
    __syncthreads();
  } else {
    // ---- this stage's right-hand side k = N_hat + L (C:243-309) on the tile (kx, z) of mode column ky ----
    const C2<T>* state = a.stage == 0 ? a.S0 : a.S;
    if (my) tile_load(t1, a.BN + base, NN, z0, r);
    tile_load(t2, state + base, NN, z0, r);
    __syncthreads();
    if (my) fft72<T, false>(t1, tw, z0, r);
    // t1 := -fft(N) * dealias + nu * (-k2 * state) on the UNcorrected state (C:279-283, 296-299; own slots), then the
    // wall planes of the state tile are replaced (C:227-241) and Dzz is applied along z
    const T kyv = a.c.ky[kyi];
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const int kxi = r + 9 * m;
      const T kxv = a.c.kx[kxi];
      const T k2 = kxv * kxv + kyv * kyv;
      const Two<C2<T>> v = ld2(t2 + kxi * N + z0);
      C2<T> p0 = (-(a.nu * k2)) * v.a, p1 = (-(a.nu * k2)) * v.b;
      if (my && kept(kxi)) {
        const Two<C2<T>> nh = ld2(t1 + kxi * N + z0);
        p0 = p0 - nh.a; p1 = p1 - nh.b;
      }
      if (f == 0 && kxi == 0 && kyi == 0) { p0.x += a.force; p1.x += a.force; }   // + fft of the uniform body force
      st2(t1 + kxi * N + z0, p0, p1);
    }
    {
      const T g = a.gain[b];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int z = z0 + c;
        const bool wall = z == 0 || z == N - 1;
        const bool jet = f == 2 && a.jets_on && z >= 1 && z <= 5;
        if (wall || jet) {
#pragma unroll
          for (int m = 0; m < 8; ++m) {
            const int kxi = r + 9 * m;
            t2[kxi * N + z] = wall ? mk<T>(T(0), T(0)) : g * a.M0hat[(size_t)b * NN + kxi * N + kyi];
          }
        }
      }
    }
    async_wait();
    __syncthreads();
    C2<T> visc[2][8];
    if constexpr (std::is_same<T, float>::value) {
      zmat_mma_cplx_inplace(mt, t2, tid);   // tensor cores; t2 := Dzz * state_bc
#pragma unroll
      for (int m = 0; m < 8; ++m) {
        const Two<C2<T>> v = ld2(t2 + (r + 9 * m) * N + z0);
        visc[0][m] = v.a; visc[1][m] = v.b;
      }
    } else {
      zmat_cplx<T>(mt, z0, r, t2, visc);
    }
    __syncthreads();  // t2 is overwritten below
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const int kxi = r + 9 * m;
      const size_t g = base + (size_t)kxi * NN + z0;
      const Two<C2<T>> pre = ld2(t1 + kxi * N + z0);
      const Two<C2<T>> s0 = ld2(a.S0 + g);
      Two<C2<T>> acc;
      acc.a = mk<T>(T(0), T(0)); acc.b = acc.a;
      if (a.mode == kXLast || a.stage != 0) acc = ld2(a.ACC + g);
      C2<T> nxt[2], kk[2];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const C2<T> k = (c ? pre.b : pre.a) + a.nu * visc[c][m];
        kk[c] = k;
        const C2<T> s0c = c ? s0.b : s0.a;
        const C2<T> ac = c ? acc.b : acc.a;
        if (a.mode == kXLast) {
          nxt[c] = s0c + a.dt6 * (ac + k);                         // S:148-152
        } else {
          nxt[c] = s0c + a.a_dt * k;                               // S:139-146
        }
      }
      if (a.RHS) st2(a.RHS + g, kk[0], kk[1]);
      if (a.mode != kXLast) {
        if (a.stage == 0) st2(a.ACC + g, a.acc_w * kk[0], a.acc_w * kk[1]);
        else st2(a.ACC + g, acc.a + a.acc_w * kk[0], acc.b + a.acc_w * kk[1]);
      }
      st2(a.S + g, nxt[0], nxt[1]);
      st2(t1 + kxi * N + z0, nxt[0], nxt[1]);
    }
    if (a.mode == kXLast) return;
  }
  // ---- inverse x-transforms of the stage state and of its x-derivative (C:153-154, 170-171) ----
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int kxi = r + 9 * m;
    const T kxv = kxi == N / 2 ? T(0) : a.c.kx[kxi];   // Hermitian projection of `.real`: Nyquist row of d/dx is zero
    const Two<C2<T>> v = ld2(t1 + kxi * N + z0);
    st2(t2 + kxi * N + z0, kxv * mul_i(v.a), kxv * mul_i(v.b));
  }
  __syncthreads();
  fft72_pair<T, true>(t1, t2, tw, z0, r);
  tile_store(a.A + base, t1, NN, z0, r);
  tile_store(a.AX + base, t2, NN, z0, r);
}

template <typename T> struct YArgs {
  Consts<T> c;
  int B, bc, jets_on, mode;   // mode 0: products (Y2) / derivatives (Y1); 1: plain
  const C2<T>* A;
  const C2<T>* AX;
  T* F;                       // [3 fields][4: f, fx, fy, fz][B][V]
  const T* M0; const T* M0dx; const T* M0dy;   // [B][72*72]
  const T* gain;              // [B]
  const T* phys;              // [3][B][V] (plain forward transform of the stored state)
  T* phys_out;                // Y1 with bc == 0: write the physical field here instead of F
  C2<T>* BN;
};

// Y1: grid (3 fields, 72 x, B).  Inverse y-transforms (two real fields per complex transform), walls, d/dz.
template <typename T> __global__ void __launch_bounds__(THREADS, ctas_per_sm<T>()) y1_kernel(const YArgs<T> a) {
  extern __shared__ __align__(32) unsigned char smem_raw[];
  C2<T>* t1 = reinterpret_cast<C2<T>*>(smem_raw);
  C2<T>* t2 = t1 + NN;
  T* mt = reinterpret_cast<T*>(t2 + NN);
  C2<T>* tw = reinterpret_cast<C2<T>*>(mt + NN);
  const int f = blockIdx.x, xi = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, z0 = 2 * (tid % ZP), r = tid / ZP;
  const size_t base = ((size_t)f * a.B + b) * V + (size_t)xi * NN;  // + y * N + z: one contiguous block
  if (tid < N) tw[tid] = a.c.tw[tid];
  if (a.bc) async_copy(mt, a.c.dzt, NN * (int)sizeof(T), tid, THREADS);   // needed only by the z-derivative
  grid_dependency_wait();
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int kyi = r + 9 * m;
    const Two<C2<T>> va = ld2(a.A + base + kyi * N + z0);
    if (a.bc) {
      const Two<C2<T>> vx = ld2(a.AX + base + kyi * N + z0);
      st2(t1 + kyi * N + z0, va.a + mul_i(vx.a), va.b + mul_i(vx.b));   // f + i * df/dx: both real after the transform
      const T kyv = kyi == N / 2 ? T(0) : a.c.ky[kyi];
      st2(t2 + kyi * N + z0, kyv * mul_i(va.a), kyv * mul_i(va.b));
    } else {
      st2(t1 + kyi * N + z0, va.a, va.b);
    }
  }
  __syncthreads();
  const T sc = T(1) / T(NN);
  if (!a.bc) {  // plain `to_physical` (C:163-168)
    fft72<T, true>(t1, tw, z0, r);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const Two<C2<T>> v = ld2(t1 + (r + 9 * m) * N + z0);
      st2(a.phys_out + base + (r + 9 * m) * N + z0, sc * v.a.x, sc * v.b.x);
    }
    return;
  }
  fft72_pair<T, true>(t1, t2, tw, z0, r);
  T fv[2][8], fxv[2][8], fyv[2][8];
  const T g = a.gain[b];
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int yi = r + 9 * m;
    const Two<C2<T>> c1 = ld2(t1 + yi * N + z0), c2 = ld2(t2 + yi * N + z0);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int z = z0 + c;
      const bool wall = z == 0 || z == N - 1;
      const bool jet = f == 2 && a.jets_on && z >= 1 && z <= 5;
      const C2<T> v1 = c ? c1.b : c1.a, v2 = c ? c2.b : c2.a;
      fv[c][m] = sc * v1.x; fxv[c][m] = sc * v1.y; fyv[c][m] = sc * v2.x;
      if (wall) { fv[c][m] = T(0); fxv[c][m] = T(0); fyv[c][m] = T(0); }
      if (jet) {
        const size_t q = (size_t)b * NN + xi * N + yi;
        fv[c][m] = g * a.M0[q]; fxv[c][m] = g * a.M0dx[q]; fyv[c][m] = g * a.M0dy[q];
      }
    }
  }
  __syncthreads();
  T* fr = reinterpret_cast<T*>(t1);  // wall-corrected field as a real tile for the z-derivative
#pragma unroll
  for (int m = 0; m < 8; ++m) st2(fr + (r + 9 * m) * N + z0, fv[0][m], fv[1][m]);
  async_wait();
  __syncthreads();
  T fz[2][8];
  if constexpr (std::is_same<T, float>::value) {
    float* dz_out = reinterpret_cast<float*>(t2);   // t2's values are in registers (fyv) by now
    zmat_mma_real(mt, fr, dz_out, tid);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const Two<T> v = ld2(dz_out + (r + 9 * m) * N + z0);
      fz[0][m] = v.a; fz[1][m] = v.b;
    }
  } else {
    zmat_real<T>(mt, z0, r, fr, fz);
  }
  T* out = a.F + ((size_t)f * 4 * a.B + b) * V + (size_t)xi * NN;
  const size_t fs = (size_t)a.B * V;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int o = (r + 9 * m) * N + z0;
    st2(out + o, fv[0][m], fv[1][m]);
    st2(out + fs + o, fxv[0][m], fxv[1][m]);
    st2(out + 2 * fs + o, fyv[0][m], fyv[1][m]);
    st2(out + 3 * fs + o, fz[0][m], fz[1][m]);
  }
}

// Y2: grid (3 fields, 72 x, B).  Advective product of field f (C:262-264) or the stored physical field, forward in y.
template <typename T> __global__ void __launch_bounds__(THREADS, ctas_per_sm<T>()) y2_kernel(const YArgs<T> a) {
  extern __shared__ __align__(32) unsigned char smem_raw[];
  C2<T>* t1 = reinterpret_cast<C2<T>*>(smem_raw);
  C2<T>* tw = t1 + NN;
  const int f = blockIdx.x, xi = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, z0 = 2 * (tid % ZP), r = tid / ZP;
  if (tid < N) tw[tid] = a.c.tw[tid];
  grid_dependency_wait();
  const size_t fs = (size_t)a.B * V;
  const size_t e0 = (size_t)b * V + (size_t)xi * NN;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int o = (r + 9 * m) * N + z0;
    T n0, n1;
    if (a.mode == 1) {
      const Two<T> p = ld2(a.phys + (size_t)f * fs + e0 + o);
      n0 = p.a; n1 = p.b;
    } else {
      const T* d = a.F + (size_t)f * 4 * fs + e0 + o;
      const Two<T> u = ld2(a.F + e0 + o), v = ld2(a.F + 4 * fs + e0 + o), w = ld2(a.F + 8 * fs + e0 + o);
      const Two<T> dx = ld2(d + fs), dy = ld2(d + 2 * fs), dz = ld2(d + 3 * fs);
      n0 = u.a * dx.a + v.a * dy.a + w.a * dz.a;
      n1 = u.b * dx.b + v.b * dy.b + w.b * dz.b;
    }
    st2(t1 + o, mk<T>(n0, T(0)), mk<T>(n1, T(0)));
  }
  __syncthreads();
  fft72<T, false>(t1, tw, z0, r);
  tile_store(a.BN + (size_t)f * fs + e0, t1, N, z0, r);
}

}  // namespace kc
