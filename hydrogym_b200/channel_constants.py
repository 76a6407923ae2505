"""Operator constants of the 3-D channel-flow solver, computed once on the host in fp64 and uploaded by
``kc_set_operators`` (``include/channel_b200.h``).  They are constants of the solve -- the reference builds the same
arrays in its constructor -- not part of the step.

Reference: C = hydrogym/jax/envs/channel.py, U = hydrogym/jax/utils/utils.py.
"""
from __future__ import annotations

import functools

import numpy as np

N = 72                                   # C:420-422
LX, LY, LZ = 2.0 * np.pi, np.pi, 2.0     # C:416-418
NU = 1.9e-3                              # C:419
N_JETS_X, N_JETS_Y = 6, 4                # C:191-192
NAB = N // 2 + 1


def chebyshev_operators(n=N, lz=LZ):
    """(z, Dz, Dzz) of U:22-50: Gauss-Lobatto points ascending in [-1, 1] scaled to z = lz/2 * x, collocation
    derivative with the negative-row-sum diagonal, second derivative as the matrix square."""
    j = np.arange(n)
    x = np.cos(np.pi * j / (n - 1))[::-1]
    c = np.where((j == 0) | (j == n - 1), 2.0, 1.0) * (-1.0) ** j
    d = np.outer(c, 1.0 / c) / ((x[:, None] - x[None, :]) + np.eye(n))
    d -= np.diag(d.sum(axis=1))
    s = 2.0 / lz
    return 0.5 * lz * x, s * d, s * s * (d @ d)


def wavenumbers(n, length):
    """C:114-118: ``fftfreq(n, d=length/n) * 2 * pi``."""
    return np.fft.fftfreq(n, d=length / n) * 2.0 * np.pi


def pressure_inverses(dz, dzz, n=N, lx=LX, ly=LY):
    """C:128-148 for the (n/2+1)^2 distinct |kx|, |ky| pairs (k2 is even in both wavenumbers, so the modes
    (+-kx, +-ky) share one operator): rows 0 and n-1 of ``Dzz - k2 I`` are replaced by the Neumann rows of Dz; the
    mean mode's middle row pins the pressure level.  Returns (n/2+1, n/2+1, n, n)."""
    kx, ky = wavenumbers(n, lx), wavenumbers(n, ly)
    h = n // 2 + 1
    # index n/2 is the Nyquist line (-n/2): same k2 as +n/2 would have
    k2 = kx[:h, None] ** 2 + ky[None, :h] ** 2
    a = dzz[None, None] - k2[:, :, None, None] * np.eye(n)[None, None]
    a[:, :, 0, :] = dz[0]
    a[:, :, -1, :] = dz[-1]
    a[0, 0, n // 2, :] = 0.0
    a[0, 0, n // 2, n // 2] = 1.0
    return np.linalg.inv(a)


def jet_basis(n=N, slit_length_x=3.0, slit_width_y=2.0, x_span_frac=2 / 3):
    """The zero-mean wall pattern of each of the 24 jets (C:194-217 is linear in the amplitudes ``Vjets``):
    Gaussian slits centred on a 6 x 4 lattice of grid indices.  Returns (24, n, n) in the reference's
    ``reshape(nx_jets, ny_jets)`` order."""
    xc = np.linspace(3.0, float(int(n * x_span_frac) - 4), N_JETS_X)
    yc = np.linspace(4.0, float(n - 5), N_JETS_Y)
    i = np.arange(n, dtype=np.float64)
    gx = np.exp(-0.5 * (i[None, :] - xc[:, None]) ** 2 / slit_length_x ** 2)   # (6, n)
    gy = np.exp(-0.5 * (i[None, :] - yc[:, None]) ** 2 / slit_width_y ** 2)    # (4, n)
    pat = gx[:, None, :, None] * gy[None, :, None, :]                          # (6, 4, n, n)
    pat = pat - pat.mean(axis=(2, 3), keepdims=True)
    return pat.reshape(N_JETS_X * N_JETS_Y, n, n)


def jet_basis_spectral(jets, lx=LX, ly=LY):
    """(fft2(jets), Re ifft2(i kx fft2(jets)), Re ifft2(i ky fft2(jets))): what the reference's ``fft_xy`` /
    ``dx_hat`` / ``ifft_xy(...).real`` make of a jet plane (C:250-259)."""
    n = jets.shape[-1]
    kx, ky = wavenumbers(n, lx), wavenumbers(n, ly)
    h = np.fft.fftn(jets, axes=(1, 2))
    dx = np.fft.ifftn(1j * kx[None, :, None] * h, axes=(1, 2)).real
    dy = np.fft.ifftn(1j * ky[None, None, :] * h, axes=(1, 2)).real
    return h, dx, dy


@functools.lru_cache(maxsize=1)
def operators():
    """Everything ``kc_set_operators`` takes, as contiguous fp64 arrays."""
    z, dz, dzz = chebyshev_operators()
    a_inv = pressure_inverses(dz, dzz)
    jets = jet_basis()
    jh, jdx, jdy = jet_basis_spectral(jets)
    jh2 = np.stack([jh.real, jh.imag], axis=-1)
    return tuple(np.ascontiguousarray(a, dtype=np.float64) for a in (z, dz, dzz, a_inv, jets, jdx, jdy, jh2))


def probe_indices(n=N, m=8):
    """C:69-73."""
    return np.linspace(0, n - 1, m).astype(np.int64)


def synthetic_initial_fields(dtype=np.float64):
    """Deterministic stand-in for the HF-hosted ``Channel_3D_Retau180/initial_field/{U,V,W}.npy`` (C:446-455; not
    reachable offline): a flat turbulent-like mean profile with grid mean 8 plus smooth O(0.5) three-dimensional
    perturbations that vanish at the walls.  The bench and the examples use it; real runs pass
    ``env_config["initial_field_dir"]``."""
    z = chebyshev_operators()[0]
    x = np.arange(N) * LX / N
    y = np.arange(N) * LY / N
    X, Y, Z = np.meshgrid(x, y, z, indexing="ij")
    env = 1.0 - Z ** 2
    prof = 1.0 - Z ** 8
    U = 8.0 * prof / prof.mean() + 0.6 * env * np.sin(2 * X) * np.cos(2 * Y) * np.cos(0.5 * np.pi * Z)
    V = 0.5 * env * np.cos(X + 2 * Y) * (1.0 + Z)
    W = 0.3 * env ** 2 * np.sin(3 * X) * np.sin(2 * Y)
    return U.astype(dtype), V.astype(dtype), W.astype(dtype)
