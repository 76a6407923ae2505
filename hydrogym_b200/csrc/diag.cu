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
