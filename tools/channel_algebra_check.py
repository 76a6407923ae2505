"""NumPy emulation of the channel kernels' restructured algebra (plane replacement in spectral
space, spectral viscous term / projection / mass-flux shift, two real fields per complex transform, Nyquist rows of
derivatives zeroed) against the oracle.  Run from the repo root: python tools/channel_algebra_check.py"""
import sys, numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import channel_oracle as orc
from hydrogym_b200 import channel_constants as cc
N = 72
z, dz, dzz, a_inv, jets, jdx, jdy, jh = cc.operators()
jh = jh[..., 0] + 1j * jh[..., 1]
kx = cc.wavenumbers(N, cc.LX); ky = cc.wavenumbers(N, cc.LY)
kxn = kx.copy(); kxn[36] = 0; kyn = ky.copy(); kyn[36] = 0
p = np.where(np.arange(N) < 36, np.arange(N), np.arange(N) - N)
keep = np.abs(p) <= 24
def gain(t):
    g = 0.3
    if t < 10: g = g * (t * 0.1)
    if t >= 40: g = g * ((50 - t) * 0.1)
    return g
def planes(act):
    return (np.tensordot(act, jets, 1), np.tensordot(act, jdx, 1), np.tensordot(act, jdy, 1), np.tensordot(act, jh, 1))
def rhs(S, act, t, nu=cc.NU, fx=2.0):
    g = -gain(t); M0, M0dx, M0dy, M0h = planes(act)
    F = []
    for f in range(3):
        A = np.fft.ifft(S[f], axis=0) * N
        AX = np.fft.ifft(1j * kxn[:, None, None] * S[f], axis=0) * N
        c1 = np.fft.ifft(A + 1j * AX, axis=1) * N / (N * N)
        c2 = np.fft.ifft(1j * kyn[None, :, None] * A, axis=1) * N / (N * N)
        fv, fxv, fyv = c1.real.copy(), c1.imag.copy(), c2.real.copy()
        for a_ in (fv, fxv, fyv):
            a_[:, :, 0] = 0; a_[:, :, -1] = 0
        if f == 2:
            for k in range(1, 6):
                fv[:, :, k] = g * M0; fxv[:, :, k] = g * M0dx; fyv[:, :, k] = g * M0dy
        fz = np.einsum("ij,xyj->xyi", dz, fv)
        F.append((fv, fxv, fyv, fz))
    u, v, w = F[0][0], F[1][0], F[2][0]
    out = []
    for f in range(3):
        n = u * F[f][1] + v * F[f][2] + w * F[f][3]
        nh = np.fft.fftn(n, axes=(0, 1)) * (keep[:, None, None] & keep[None, :, None])
        sbc = S[f].copy(); sbc[:, :, 0] = 0; sbc[:, :, -1] = 0
        if f == 2:
            for k in range(1, 6): sbc[:, :, k] = g * M0h
        k2 = (kx[:, None] ** 2 + ky[None, :] ** 2)[:, :, None]
        k = nu * (-k2 * S[f] + np.einsum("ij,xyj->xyi", dzz, sbc)) - nh
        if f == 0: k[0, 0, :] += fx * N * N
        out.append(k)
    return out
def project(S, act, t, dt, target=8.0):
    g = -gain(t); M0h = planes(act)[3]
    u, v, w = S
    ia = np.minimum(np.arange(N), N - np.arange(N))
    Ai = a_inv[ia[:, None], ia[None, :]]          # (72,72,72,72)
    r = (1j * kxn[:, None, None] * u + 1j * kyn[None, :, None] * v + np.einsum("ij,xyj->xyi", dz, w)) / dt
    r[:, :, 0] = w[:, :, 0] / dt; r[:, :, -1] = w[:, :, -1] / dt; r[0, 0, 36] = 0
    ph = np.einsum("xyij,xyj->xyi", Ai, r)
    u = u - dt * 1j * kxn[:, None, None] * ph; v = v - dt * 1j * kyn[None, :, None] * ph
    w = w - dt * np.einsum("ij,xyj->xyi", dz, ph)
    for a_ in (u, v, w): a_[:, :, 0] = 0; a_[:, :, -1] = 0
    for k in range(1, 6): w[:, :, k] = g * M0h
    mean = u[0, 0, :].real.sum() / (N * N * N)
    u[0, 0, 1:-1] += (target - mean) * N * N
    return [u, v, w]
def rk4(S, act, t, dt):
    k1 = rhs(S, act, t); k2 = rhs([s + 0.5 * dt * k for s, k in zip(S, k1)], act, t + 0.5 * dt)
    k3 = rhs([s + 0.5 * dt * k for s, k in zip(S, k2)], act, t + 0.5 * dt); k4 = rhs([s + dt * k for s, k in zip(S, k3)], act, t + dt)
    star = [s + (dt / 6) * ((a + 2 * b + 2 * c) + d) for s, a, b, c, d in zip(S, k1, k2, k3, k4)]
    return project(star, act, t + dt, dt)
fields = orc.synthetic_initial_fields()
env = orc.ChannelFlowSpectralEnv(initial_fields=fields)
eq = env.equation
act = np.clip(0.9 * np.sin(1.3 * np.arange(24) + 0.4), -1, 1)
s0 = eq.to_spectral(orc.VelocityState(*fields))
rel = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)
for t in (3.0, 44.5, 60.0):
    want = eq.rhs(s0, action=act, t=t, fx=2.0); got = rhs([s0.u, s0.v, s0.w], act, t)
    print("rhs t", t, [rel(g, w) for g, w in zip(got, (want.u, want.v, want.w))])
S = [s0.u, s0.v, s0.w]; so = s0
for n in range(3):
    S = rk4(S, act, 38 + n, orc.DT); so = env.integrator.rk4_step(so, dt=orc.DT, action=act, t=38 + n, fx=2.0, target_bulk_u=8.0)
    print("rk4 step", n, [rel(g, w) for g, w in zip(S, (so.u, so.v, so.w))])
