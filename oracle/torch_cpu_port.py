"""torch-CPU (MKL FFT, multi-threaded elementwise) variant of the oracle's RK4-CN solver step.

TEST / BASELINE INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs time this beside the NumPy oracle so that the CPU gets its best configuration
(BASELINE.md section 4.2): one process, the whole batch of environments in one tensor, ``torch.fft`` (MKL / pocketfft
with intra-op threads) and threaded pointwise kernels.  The arithmetic is ``kolmogorov_oracle``'s statement for
statement -- constants (grids, forcing, control spectrum) are taken from the NumPy oracle object itself --
and ``tests/test_oracle.py`` holds it to the NumPy oracle at round-off.

Reference lines followed: hydrogym/jax/envs/kolmogorov.py:209-257 (linear / implicit / nonlinear terms),
hydrogym/jax/solvers/base.py:50-57 (RK4_CN.time_step_fn), hydrogym/jax/utils/utils.py:72-86 (no-op de-aliasing).
"""
from __future__ import annotations

import numpy as np
import torch

from .kolmogorov_oracle import ALPHA_RK4, BETA_RK4, GAMMA_RK4


class TorchCpuSolver:
    def __init__(self, env, threads=None):
        """``env``: an ``oracle.kolmogorov_oracle.KolmogorovFlow`` (supplies the constants and the dtype)."""
        if threads:
            torch.set_num_threads(int(threads))
        eq = env.equation
        cd = torch.complex64 if eq.cdtype == np.complex64 else torch.complex128
        self.cd = cd
        self.dt, self.nu, self.dealias = env.integrator.dt, env.flow.nu, eq.dealias
        kx, ky = torch.from_numpy(np.asarray(eq.kx)), torch.from_numpy(np.asarray(eq.ky))
        dd = ((2j * np.pi) ** 2 * (kx ** 2 + ky ** 2)).to(cd)          # K:216, K:234
        self.lin = (self.nu * dd).to(cd)                                # K:209-211
        self.dd = dd
        dd1 = dd.clone()
        dd1[0, 0] = 1                                                   # K:235
        self.inv_dd1 = (-1.0 / dd1).to(cd)                              # psi_hat = -omega_hat / dd (K:237)
        self.iky = ((2j * np.pi) * ky).to(cd)                           # K:238, K:246
        self.ikx = ((2j * np.pi) * kx).to(cd)                           # K:239 (sign applied below), K:245
        self.forcing = torch.from_numpy(np.asarray(eq._forcing)).to(cd)  # K:274-292
        n, m = env.flow.grid_size
        self.kn, self.km = int(n // 2 * 2 // 3), int((m // 2 + 1) * 2 // 3)

    def nonlinear_terms(self, w, control_hat=None):
        """K:222-257."""
        psi = w * self.inv_dd1
        u = torch.fft.irfft2(self.iky * psi)
        v = torch.fft.irfft2(-(self.ikx * psi))
        gx = torch.fft.irfft2(self.ikx * w)
        gy = torch.fft.irfft2(self.iky * w)
        adv = torch.fft.rfft2(-(gx * u + gy * v)).to(self.cd)
        if self.dealias:  # opt-in: what utils.py:72-86 intended (the reference's version is the identity)
            adv[..., self.kn:2 * self.kn, :] = 0
            adv[..., :, self.km:] = 0
        terms = adv + self.forcing
        return terms if control_hat is None else terms + control_hat

    def time_step(self, u, control_hat=None):
        """S:50-57."""
        h = 0
        for k in range(5):
            h = self.nonlinear_terms(u, control_hat) + BETA_RK4[k] * h
            mu = 0.5 * self.dt * (ALPHA_RK4[k + 1] - ALPHA_RK4[k])
            yn = u + GAMMA_RK4[k] * self.dt * h + mu * (self.lin * u)
            u = (1 / (1 - mu * self.nu * self.dd) * yn).to(self.cd)     # K:213-220
        return u

    def rk_steps(self, u, n, control_hat=None):
        u = torch.from_numpy(np.ascontiguousarray(u)).to(self.cd)
        ch = None if control_hat is None else torch.from_numpy(np.ascontiguousarray(control_hat)).to(self.cd)
        for _ in range(int(n)):
            u = self.time_step(u, ch)
        return u.numpy()
