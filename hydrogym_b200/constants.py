"""Host-side (init-time) constants of the Kolmogorov flow, sampled the way the reference samples them.

Reference (paths relative to the reference repository):
  hydrogym/jax/envs/kolmogorov.py:46-57    load_mesh  -- ENDPOINT-INCLUSIVE linspace
  hydrogym/jax/envs/kolmogorov.py:89-101   load_fft_mesh -- cycles-per-unit wavenumbers
  hydrogym/jax/envs/kolmogorov.py:103-129  initialize_state
  hydrogym/jax/envs/kolmogorov.py:274-292  forcing_term
  hydrogym/jax/envs/kolmogorov.py:259-272  control_term  (no 2*pi*i factor)
  hydrogym/jax/solvers/base.py:100-103     step counts (float floor division)

Because the real-space grid includes the endpoint 2*pi while the FFT assumes
period 2*pi, sin(k*y) leaks over every ky bin; the spectra are therefore computed
numerically here (fp64, cast once), not replaced by analytic single modes.  Both
the forcing and the control field are x-independent, so their 2-D spectra are
non-zero on row kx = 0 only and one row of M/2+1 complex numbers carries them.
"""
from __future__ import annotations

import numpy as np


def check_domain(domain_x, domain_y):
    """The kernels work on integer wavenumbers (2*pi*kx = p), i.e. on the reference's default 2*pi-periodic box
    (kolmogorov.py:33-34 DEFAULT_DOMAIN_X / _Y).  Any other domain would make the host-side forcing / control rows
    and the device-side derivative and viscous operators disagree, so it is refused instead of silently mixed."""
    for name, d in (("domain_x", domain_x), ("domain_y", domain_y)):
        lo, hi = float(d[0]), float(d[1])
        if lo != 0.0 or abs(hi - 2 * np.pi) > 1e-12:
            raise ValueError(f"hydrogym_b200 supports the reference's default domain (0, 2*pi) only; got {name}={tuple(d)}")


def step_counts(dt: float, end_time: float, save_n: int = 1):
    """(n_inner, n_outer) with the reference's Python float floor-division (base.py:100-103)."""
    if end_time < 1:
        raise ValueError(
            "This flow configuration requires the end time to be at least 1. Please adjust t_span and run again.")
    step_to_save = int(save_n // dt)
    total_steps = int(end_time // dt)
    outer_steps = int(total_steps // step_to_save)
    return step_to_save, outer_steps


def real_mesh(grid_size, domain_x=(0, 2 * np.pi), domain_y=(0, 2 * np.pi)):
    nx, ny = grid_size[1], grid_size[0]
    x = np.linspace(domain_x[0], domain_x[1], nx)
    y = np.linspace(domain_y[0], domain_y[1], ny)
    return np.meshgrid(x, y, indexing="ij")


def fft_mesh(grid_size, domain_x=(0, 2 * np.pi), domain_y=(0, 2 * np.pi)):
    N, M = grid_size
    kx = np.fft.fftfreq(N, domain_x[1] / N)
    ky = np.fft.rfftfreq(M, domain_y[1] / M)
    return np.meshgrid(kx, ky, indexing="ij")


def _row0_spectrum_of_sin(k, grid_size, domain_y):
    """Row kx=0 of rfft2(sin(k*Y)) for Y[i, j] = y[j]: N * rfft(sin(k*y))."""
    N, M = grid_size
    y = np.linspace(domain_y[0], domain_y[1], N)  # load_mesh: y has grid_size[0] points
    assert N == M, "square grids only"
    return N * np.fft.rfft(np.sin(k * y))


def forcing_row(k, grid_size, domain_y=(0, 2 * np.pi)):
    """Row kx=0 of forcing_term(): 2*pi*i*(fy_hat*kx - fx_hat*ky) with fx = sin(k*Y), fy = 0."""
    M = grid_size[1]
    ky = np.fft.rfftfreq(M, domain_y[1] / M)
    return 2j * np.pi * (-_row0_spectrum_of_sin(k, grid_size, domain_y) * ky)


def control_rows(ks, grid_size, domain_y=(0, 2 * np.pi)):
    """Rows kx=0 of control_term() for unit action on each wavenumber in ``ks``: -ky * rfft2(sin(k_i*Y))."""
    M = grid_size[1]
    ky = np.fft.rfftfreq(M, domain_y[1] / M)
    return np.stack([-ky * _row0_spectrum_of_sin(k, grid_size, domain_y) for k in ks])


def initial_vorticity_hat(grid_size, domain_x=(0, 2 * np.pi), domain_y=(0, 2 * np.pi)):
    """initialize_state(): rfftn(-sin X + cos Y) on the endpoint-inclusive mesh."""
    X, Y = real_mesh(grid_size, domain_x, domain_y)
    return np.fft.rfftn(-np.sin(X) + np.cos(Y))
