// Host-side check of the in-register DFT butterflies in hydrogym_b200/csrc/fft.cuh
// (they are __host__ __device__, so the very code the kernels inline runs here on the CPU).
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "../../hydrogym_b200/csrc/fft.cuh"

template <int R, int DIR, typename T> double check() {
  kf::Cx<T> a[R], in[R];
  for (int i = 0; i < R; ++i) {
    in[i].x = (T)(std::rand() / (double)RAND_MAX - 0.5);
    in[i].y = (T)(std::rand() / (double)RAND_MAX - 0.5);
    a[i] = in[i];
  }
  kf::Radix<R>::template run<DIR, T>(a);
  double err = 0;
  for (int k = 0; k < R; ++k) {
    long double re = 0, im = 0;
    for (int n = 0; n < R; ++n) {
      const long double ang = DIR * 2.0L * 3.14159265358979323846264338327950288L * ((n * k) % R) / R;
      re += in[n].x * cosl(ang) - in[n].y * sinl(ang);
      im += in[n].x * sinl(ang) + in[n].y * cosl(ang);
    }
    const kf::Cx<T> got = a[kf::Radix<R>::pos(k)];
    err = std::fmax(err, std::fmax((double)fabsl(got.x - re), (double)fabsl(got.y - im)));
  }
  return err;
}

int main() {
  double e64 = 0, e32 = 0;
#define RUN(R)                                                                              \
  e64 = std::fmax(e64, std::fmax(check<R, -1, double>(), check<R, +1, double>()));         \
  e32 = std::fmax(e32, std::fmax(check<R, -1, float>(), check<R, +1, float>()));
  RUN(2) RUN(4) RUN(8) RUN(16) RUN(32)
  std::printf("max_err_f64 %.3e max_err_f32 %.3e\n", e64, e32);
  return (e64 < 1e-13 && e32 < 1e-5) ? 0 : 1;
}
