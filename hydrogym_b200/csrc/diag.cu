// Spectral diagnostics as one-pass CUDA reductions (used by the configs[3] long-horizon
// energy-spectrum check).  Per-mode kinetic energy follows compute_energy_mode
// (hydrogym/jax/utils/utils.py:89-101) with u_hat, v_hat from compute_velocity_fft (utils.py:53-69):
//     e(kx, ky) = 0.5 * (|u_hat|^2 + |v_hat|^2) / (n*m)^2 = 0.5 * |omega_hat|^2 / kappa^2 / (n*m)^2,   kappa != 0.
// The radial binning (bin = round(kappa), half-spectrum weights 2 for 0 < ky < m/2) is ours: the reference
// ships the per-mode quantity only.  No handle: works on the user layout [B][N][M/2+1].  sm_100a only.
#include "../../include/kolmo_b200.h"

#include <cuda_runtime.h>

#include <string>

int kf_set_error(const char* msg);  // kolmo_b200.cu

namespace {

template <typename T>
__global__ void energy_spectrum_kernel(const T* __restrict__ omega /* interleaved complex */, int N, int MH1, int nbins,
                                       double* __restrict__ spec) {
  extern __shared__ double s_bins[];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < nbins; i += blockDim.x) s_bins[i] = 0.0;
  __syncthreads();
  const size_t env = (size_t)N * MH1;
  const double norm = 0.5 / (((double)N * N) * ((double)N * N));
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < env; i += (size_t)gridDim.x * blockDim.x) {
    const int kx = (int)(i / MH1), q = (int)(i - (size_t)kx * MH1);
    const int p = kx < N / 2 ? kx : kx - N;
    const double k2 = (double)p * p + (double)q * q;
    if (k2 == 0.0) continue;
    const double re = (double)omega[2 * ((size_t)b * env + i)], im = (double)omega[2 * ((size_t)b * env + i) + 1];
    const double wgt = (q == 0 || q == MH1 - 1) ? 1.0 : 2.0;
    const int bin = (int)(sqrt(k2) + 0.5);
    if (bin < nbins) atomicAdd(&s_bins[bin], wgt * norm * (re * re + im * im) / k2);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nbins; i += blockDim.x)
    if (s_bins[i] != 0.0) atomicAdd(spec + (size_t)b * nbins + i, s_bins[i]);
}


// linear_terms (kolmogorov.py:209-211): nu * (2 pi i)^2 (kx^2 + ky^2) * w = -nu * kappa^2 * w, and
// implicit_timestep (kolmogorov.py:213-220): w / (1 - dt * nu * (2 pi i)^2 (kx^2 + ky^2)) = w / (1 + dt nu kappa^2),
// on the integer wavenumbers kappa^2 = p^2 + q^2 (2 pi kx = p).  mode 0 = linear, 1 = implicit.
template <typename T>
__global__ void spectral_pointwise_kernel(const T* __restrict__ in, T* __restrict__ out, int N, int MH1, size_t total,
                                          T nu, T ts, int mode) {
  const size_t env = (size_t)N * MH1;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i % env;
    const int kx = (int)(r / MH1), q = (int)(r - (size_t)kx * MH1);
    const T p = (T)(kx < N / 2 ? kx : kx - N);
    const T k2 = p * p + (T)q * (T)q;
    const T f = mode == 0 ? -nu * k2 : (T)1 / ((T)1 + ts * nu * k2);
    out[2 * i] = f * in[2 * i];
    out[2 * i + 1] = f * in[2 * i + 1];
  }
}

// numpy / jax `gradient` along one axis with unit spacing: central differences inside, one-sided at the two ends
template <typename T> __device__ __forceinline__ T grad_at(const T* __restrict__ f, int i, int n, size_t stride) {
  if (i == 0) return f[stride] - f[0];
  if (i == n - 1) return f[(size_t)(n - 1) * stride] - f[(size_t)(n - 2) * stride];
  return (T)0.5 * (f[(size_t)(i + 1) * stride] - f[(size_t)(i - 1) * stride]);
}

// compute_energy_dissipation (utils.py:145-172) and compute_divergence (utils.py:213-229) on the physical velocity
// fields u, v [B][N][M]: eps[b] = sum 2 nu (du_dx^2 + dv_dy^2 + (du_dy + dv_dx)^2) / (4 pi^2) (fp64 block reduction +
// one atomic per block); div[b][x][y] = du_dx + dv_dy.  Either output may be NULL.
template <typename T>
__global__ void strain_kernel(const T* __restrict__ u, const T* __restrict__ v, int N, int M, double nu,
                              double* __restrict__ eps, T* __restrict__ div) {
  const int b = blockIdx.y;
  const size_t env = (size_t)N * M;
  const T* ub = u + (size_t)b * env;
  const T* vb = v + (size_t)b * env;
  double acc = 0.0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < env; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i / M), y = (int)(i - (size_t)x * M);
    const T du_dx = grad_at(ub + y, x, N, (size_t)M), dv_dy = grad_at(vb + (size_t)x * M, y, M, 1);
    if (div) div[(size_t)b * env + i] = du_dx + dv_dy;
    if (eps) {
      const T du_dy = grad_at(ub + (size_t)x * M, y, M, 1), dv_dx = grad_at(vb + y, x, N, (size_t)M);
      const T sh = du_dy + dv_dx;
      acc += 2.0 * nu * ((double)(du_dx * du_dx) + (double)(dv_dy * dv_dy) + (double)(sh * sh));
    }
  }
  if (eps) {
    __shared__ double s_red[32];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
      double r = threadIdx.x < (blockDim.x >> 5) ? s_red[threadIdx.x] : 0.0;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) r += __shfl_xor_sync(0xffffffffu, r, off);
      if (threadIdx.x == 0) atomicAdd(eps + b, r / (4.0 * 3.14159265358979323846 * 3.14159265358979323846));
    }
  }
}

int check_launch(const char* what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return kf_set_error((std::string(what) + ": " + cudaGetErrorString(e)).c_str());
  return 0;
}

template <typename T>
int spectral_pointwise(int N, int B, double nu, double ts, const void* in, void* out, int mode, cudaStream_t s) {
  const size_t total = (size_t)B * N * (N / 2 + 1);
  size_t blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  spectral_pointwise_kernel<T><<<(unsigned)blocks, 256, 0, s>>>((const T*)in, (T*)out, N, N / 2 + 1, total, (T)nu, (T)ts, mode);
  return check_launch("spectral_pointwise_kernel");
}

}  // namespace

extern "C" int kf_energy_spectrum(int dtype, int N, int B, const void* omega_hat_dev, double* spectrum_dev, int nbins,
                                  void* stream) {
  if (!omega_hat_dev || !spectrum_dev) return kf_set_error("kf_energy_spectrum: null argument");
  if (N < 2 || B < 1 || nbins < 1 || nbins > 4096) return kf_set_error("kf_energy_spectrum: bad size (nbins <= 4096)");
  if (dtype != KF_F32 && dtype != KF_F64) return kf_set_error("kf_energy_spectrum: dtype must be KF_F32 or KF_F64");
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(spectrum_dev, 0, sizeof(double) * (size_t)B * nbins, s);
  if (e != cudaSuccess) return kf_set_error(cudaGetErrorString(e));
  const int MH1 = N / 2 + 1;
  const size_t env = (size_t)N * MH1;
  int gx = (int)((env + 256 * 8 - 1) / (256 * 8));
  if (gx > 148 * 4) gx = 148 * 4;
  const dim3 grid(gx, B);
  const size_t smem = sizeof(double) * nbins;
  if (dtype == KF_F32)
    energy_spectrum_kernel<float><<<grid, 256, smem, s>>>((const float*)omega_hat_dev, N, MH1, nbins, spectrum_dev);
  else
    energy_spectrum_kernel<double><<<grid, 256, smem, s>>>((const double*)omega_hat_dev, N, MH1, nbins, spectrum_dev);
  e = cudaGetLastError();
  if (e != cudaSuccess) return kf_set_error(cudaGetErrorString(e));
  return 0;
}

extern "C" int kf_linear_terms(int dtype, int N, int B, double nu, const void* omega_hat_dev, void* out_dev, void* stream) {
  if (!omega_hat_dev || !out_dev) return kf_set_error("kf_linear_terms: null argument");
  if (N < 2 || B < 1) return kf_set_error("kf_linear_terms: bad size");
  if (dtype == KF_F32) return spectral_pointwise<float>(N, B, nu, 0.0, omega_hat_dev, out_dev, 0, (cudaStream_t)stream);
  if (dtype == KF_F64) return spectral_pointwise<double>(N, B, nu, 0.0, omega_hat_dev, out_dev, 0, (cudaStream_t)stream);
  return kf_set_error("kf_linear_terms: dtype must be KF_F32 or KF_F64");
}

extern "C" int kf_implicit_timestep(int dtype, int N, int B, double nu, double time_step, const void* omega_hat_dev,
                                    void* out_dev, void* stream) {
  if (!omega_hat_dev || !out_dev) return kf_set_error("kf_implicit_timestep: null argument");
  if (N < 2 || B < 1) return kf_set_error("kf_implicit_timestep: bad size");
  if (dtype == KF_F32)
    return spectral_pointwise<float>(N, B, nu, time_step, omega_hat_dev, out_dev, 1, (cudaStream_t)stream);
  if (dtype == KF_F64)
    return spectral_pointwise<double>(N, B, nu, time_step, omega_hat_dev, out_dev, 1, (cudaStream_t)stream);
  return kf_set_error("kf_implicit_timestep: dtype must be KF_F32 or KF_F64");
}

extern "C" int kf_strain_diagnostics(int dtype, int N, int B, double nu, const void* u_dev, const void* v_dev,
                                     double* dissipation_dev, void* divergence_dev, void* stream) {
  if (!u_dev || !v_dev || (!dissipation_dev && !divergence_dev)) return kf_set_error("kf_strain_diagnostics: null argument");
  if (N < 2 || B < 1) return kf_set_error("kf_strain_diagnostics: bad size");
  if (dtype != KF_F32 && dtype != KF_F64) return kf_set_error("kf_strain_diagnostics: dtype must be KF_F32 or KF_F64");
  cudaStream_t s = (cudaStream_t)stream;
  if (dissipation_dev) {
    const cudaError_t e = cudaMemsetAsync(dissipation_dev, 0, sizeof(double) * (size_t)B, s);
    if (e != cudaSuccess) return kf_set_error(cudaGetErrorString(e));
  }
  const size_t env = (size_t)N * N;
  int gx = (int)((env + 256 * 4 - 1) / (256 * 4));
  if (gx > 148 * 4) gx = 148 * 4;
  const dim3 grid(gx, B);
  if (dtype == KF_F32)
    strain_kernel<float><<<grid, 256, 0, s>>>((const float*)u_dev, (const float*)v_dev, N, N, nu, dissipation_dev,
                                              (float*)divergence_dev);
  else
    strain_kernel<double><<<grid, 256, 0, s>>>((const double*)u_dev, (const double*)v_dev, N, N, nu, dissipation_dev,
                                               (double*)divergence_dev);
  return check_launch("strain_kernel");
}
