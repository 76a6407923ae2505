"""Spectral diagnostics on device: CUDA reductions behind ``kf_energy_spectrum`` plus the
per-mode helpers of hydrogym/jax/utils/utils.py:89-126 as thin indexing ops over torch tensors."""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def energy_spectrum(omega_hat: torch.Tensor, nbins: int | None = None) -> torch.Tensor:
    """Radial kinetic-energy spectrum E[b, k] (float64) of ``omega_hat`` (B, N, N//2+1) or (N, N//2+1).

    Per-mode energy as in ``compute_energy_mode`` (utils.py:89-101): 0.5 (|u_hat|^2 + |v_hat|^2) / (n m)^2 with
    the velocities of ``compute_velocity_fft`` (utils.py:53-69); bin = round(|k|); the conjugate half-plane is
    counted by weights 2 on 0 < ky < N/2.
    """
    squeeze = omega_hat.dim() == 2
    w = omega_hat[None] if squeeze else omega_hat
    if w.dim() != 3 or w.shape[2] != w.shape[1] // 2 + 1:
        raise ValueError("omega_hat must be (B, N, N//2+1) complex")
    if w.dtype not in (torch.complex64, torch.complex128) or not w.is_cuda:
        raise ValueError("omega_hat must be a complex CUDA tensor")
    w = w.contiguous()
    B, N = int(w.shape[0]), int(w.shape[1])
    nbins = int(nbins) if nbins is not None else N
    out = torch.empty((B, nbins), dtype=torch.float64, device=w.device)
    lib = _lib.load()
    with torch.cuda.device(w.device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(w.device).cuda_stream)
        _lib.check(lib.kf_energy_spectrum(_lib.KF_F32 if w.dtype == torch.complex64 else _lib.KF_F64, N, B,
                                          ctypes.c_void_p(w.data_ptr()), ctypes.c_void_p(out.data_ptr()), nbins,
                                          stream), "kf_energy_spectrum")
    return out[0] if squeeze else out


def compute_velocity_fft(omega_hat: torch.Tensor):
    """utils.py:53-69 on integer wavenumbers (2 pi kx = p, 2 pi ky = q): u_hat = i q psi_hat, v_hat = -i p psi_hat."""
    N = omega_hat.shape[-2]
    rd = torch.float32 if omega_hat.dtype == torch.complex64 else torch.float64
    p = torch.fft.fftfreq(N, 1.0 / N, device=omega_hat.device, dtype=rd)[:, None]
    q = torch.arange(N // 2 + 1, device=omega_hat.device, dtype=rd)[None, :]
    k2 = p * p + q * q
    k2[0, 0] = 1.0
    psi = omega_hat / k2
    return 1j * q * psi, -1j * p * psi


def compute_energy_mode(uhat, vhat, kx, ky, n, m):
    """utils.py:89-101 (kx, ky integer wavenumbers)."""
    return 0.5 * (uhat[..., kx % n, ky % m].abs() ** 2 + vhat[..., kx % n, ky % m].abs() ** 2) / float(n * m) ** 2


def compute_velocity_mode(uhat, vhat, kx, ky, n, m):
    """utils.py:104-121."""
    return torch.sqrt((uhat[..., kx % n, ky % m].abs() ** 2 + vhat[..., kx % n, ky % m].abs() ** 2) / float(n * m) ** 2)


_STEPPERS = {}


def _stepper(N, B, dtype, device):
    """One cached ``KolmogorovStepper`` per (N, B, dtype, device): the diagnostics only use its x-space-field, row
    and reduction kernels, so the workspace, twiddle tables and TMA descriptors are built once, not per call."""
    from .solver import KolmogorovStepper
    key = (int(N), int(B), dtype, str(device))
    st = _STEPPERS.get(key)
    if st is None:
        if len(_STEPPERS) >= 8:  # bounded: drop the oldest workspace
            _STEPPERS.pop(next(iter(_STEPPERS))).close()
        st = _STEPPERS[key] = KolmogorovStepper(grid_size=(N, N), num_envs=B, dtype=dtype, device=device,
                                                n_inner=1, n_outer=1)
    return st


def clear_cache():
    """Free the cached diagnostic workspaces."""
    while _STEPPERS:
        _STEPPERS.popitem()[1].close()


def _velocity(omega_hat):
    """(u, v) physical fields (B, N, N) of omega_hat (B, N, Mh) or (N, Mh) through the row-FFT kernels."""
    w = omega_hat[None] if omega_hat.dim() == 2 else omega_hat
    rd = torch.float32 if w.dtype == torch.complex64 else torch.float64
    uv = _stepper(w.shape[1], w.shape[0], rd, w.device).physical_fields(w.contiguous(), 2)
    return uv[0], uv[1]


def _strain(omega_hat, nu, want_eps, want_div):
    u, v = _velocity(omega_hat)
    B, N = int(u.shape[0]), int(u.shape[1])
    eps = torch.empty((B,), dtype=torch.float64, device=u.device) if want_eps else None
    div = torch.empty_like(u) if want_div else None
    lib = _lib.load()
    with torch.cuda.device(u.device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(u.device).cuda_stream)
        _lib.check(lib.kf_strain_diagnostics(
            _lib.KF_F32 if u.dtype == torch.float32 else _lib.KF_F64, N, B, float(nu), ctypes.c_void_p(u.data_ptr()),
            ctypes.c_void_p(v.data_ptr()), ctypes.c_void_p(eps.data_ptr() if want_eps else 0),
            ctypes.c_void_p(div.data_ptr() if want_div else 0), stream), "kf_strain_diagnostics")
    return eps, div


def compute_real_velocity_point(omega_hat, x_idx, y_idx):
    """utils.py:129-142: |velocity| at grid point (x_idx, y_idx)."""
    u, v = _velocity(omega_hat)
    out = torch.sqrt(u[:, x_idx, y_idx].abs() ** 2 + v[:, x_idx, y_idx].abs() ** 2)
    return out[0] if omega_hat.dim() == 2 else out


def compute_energy_dissipation(omega_hat, nu):
    """utils.py:145-172: 2 nu <S_ij S_ij> with the reference's finite differences in INDEX units
    (jnp.gradient, unit spacing, one-sided at the edges), summed and divided by 4 pi^2 -- one CUDA reduction
    (``kf_strain_diagnostics``) over the physical velocity fields.  Returns float64."""
    eps, _ = _strain(omega_hat, nu, True, False)
    return eps[0] if omega_hat.dim() == 2 else eps


def compute_divergence(omega_hat):
    """utils.py:213-229: du/dx + dv/dy with the same index-unit finite differences (not spectrally exact)."""
    _, div = _strain(omega_hat, 0.0, False, True)
    return div[0] if omega_hat.dim() == 2 else div


def compute_tke(omega_hat):
    """utils.py:175-194: turbulent kinetic energy sum(0.5 (|u - mean u|^2 + |v - mean v|^2)) / n^2 of one state (N, Mh)
    or a batch (B, N, Mh), through the fused reward / observation kernel (``kf_reward_obs`` on a one-snapshot
    trajectory: c2r of u_hat, v_hat, fp64 warp-shuffle sums)."""
    w = omega_hat[None] if omega_hat.dim() == 2 else omega_hat
    rd = torch.float32 if w.dtype == torch.complex64 else torch.float64
    _, tke, _ = _stepper(w.shape[1], w.shape[0], rd, w.device).reward_obs(w[:, None].contiguous())
    return tke[0] if omega_hat.dim() == 2 else tke


def compute_reward(omega_hat, actions):
    """utils.py:197-210 as written: TKE minus the plain sum of the four actions (not the env's reward, K:462-470)."""
    a = torch.as_tensor(actions, device=omega_hat.device)
    return compute_tke(omega_hat) - (a[..., 0] + a[..., 1] + a[..., 2] + a[..., 3])
