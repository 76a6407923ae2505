"""Golden summary of ONE RK4-CN solver step at 8192^2 from the NumPy oracle (BASELINE.json configs[4]).

Run in the build container (minutes of CPU, ~10 GB of RAM):

    python tests/golden/make_slab8192_golden.py            # writes tests/golden/slab8192_rk1_fp64.npz

The oracle (oracle/kolmogorov_oracle.py, pinned against the reference's own source at 64^2) advances a BROADBAND
random state -- content on every wavenumber including the Nyquist row kx = -N/2 and column ky = N/2 -- by one solver
step in fp64 with the reference's forcing and a non-zero control action.  The full result (537 MB) cannot be
committed; the fixture keeps what localises an error: the norm and the sum of the new state, the |kx|, ky <= 32 block,
the norms of 64 row bands and 64 column bands, and 512 scattered entries (128 of them on the Nyquist row / column).
``slab_state(N, seed)`` regenerates the input on the GPU box from its seed (tests/test_gpu_slab.py).
Reference lines exercised: hydrogym/jax/envs/kolmogorov.py:209-292, hydrogym/jax/solvers/base.py:50-57.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

N = 8192
RE = 1000.0
DT = 1e-4
SEED = 8192
ACTION = np.array([0.3, -0.2, 0.1, -0.4])
KLOW = 32
NBANDS = 64


def slab_state(n=N, seed=SEED):
    """Broadband random omega_hat (n, n/2+1), rounded to complex64 (the fp32 kernels' input) and returned as
    complex128.  Envelope ~ (1 + |k|/48)^-2: rms vorticity about 4, Nyquist lines populated."""
    rng = np.random.default_rng(seed)
    mh = n // 2 + 1
    p = np.abs(np.fft.fftfreq(n, 1.0 / n))[:, None]
    q = np.arange(mh)[None, :]
    env = 1.0 / (1.0 + np.sqrt(p * p + q * q) / 48.0) ** 2
    w = rng.standard_normal((n, mh)) + 1j * rng.standard_normal((n, mh))
    w *= env * (n * 480.0)
    w[0, 0] = 0.0
    return w.astype(np.complex64).astype(np.complex128)


def scattered_indices(n=N, seed=SEED):
    """512 (kx, ky) index pairs: 384 anywhere, 64 on the Nyquist row kx = n/2, 64 on the Nyquist column ky = n/2."""
    rng = np.random.default_rng(seed + 1)
    mh = n // 2 + 1
    kx = np.concatenate([rng.integers(0, n, 384), np.full(64, n // 2), rng.integers(0, n, 64)])
    ky = np.concatenate([rng.integers(0, mh, 384), rng.integers(0, mh, 64), np.full(64, n // 2)])
    return kx, ky


def summarise(w, n=N):
    """What the fixture stores about a state (n, n/2+1)."""
    mh = n // 2 + 1
    rows = np.r_[0:KLOW + 1, n - KLOW:n]
    kx, ky = scattered_indices(n)
    rb = np.sqrt((np.abs(w) ** 2).reshape(NBANDS, n // NBANDS, mh).sum(axis=(1, 2)))
    cb = np.sqrt((np.abs(w[:, :mh - 1]) ** 2).reshape(n, NBANDS, (mh - 1) // NBANDS).sum(axis=(0, 2)))
    return dict(norm=np.sqrt((np.abs(w) ** 2).sum()), total=w.sum(), low=w[np.ix_(rows, np.arange(KLOW + 1))].copy(),
                row_bands=rb, col_bands=cb, nyq_col_norm=np.sqrt((np.abs(w[:, mh - 1]) ** 2).sum()),
                scattered=w[kx, ky].copy())


def main():
    sys.path.insert(0, ROOT)
    from oracle import kolmogorov_oracle as orc
    t0 = time.time()
    env = orc.KolmogorovFlow(dict(dt=DT), dict(Re=RE, grid_size=(N, N)), dtype=np.float64)
    w0 = slab_state()
    print(f"[8192 golden] state built in {time.time() - t0:.1f}s, |w0| = {np.linalg.norm(w0):.6e}", flush=True)
    clipped = env._clip_action(ACTION[None], env.default_params)
    ch = env.equation.control_term(w0[None], env._control_field(clipped, env.default_params))[0]
    w1 = env.integrator.rk_steps(w0, 1, ch)
    print(f"[8192 golden] one RK step in {time.time() - t0:.1f}s", flush=True)
    s0, s1 = summarise(w0), summarise(w1)
    out = {f"in_{k}": v for k, v in s0.items()}
    out.update({f"out_{k}": v for k, v in s1.items()})
    out.update(N=N, Re=RE, dt=DT, seed=SEED, action=ACTION, klow=KLOW)
    np.savez_compressed(os.path.join(HERE, "slab8192_rk1_fp64.npz"), **out)
    print(f"[8192 golden] wrote fixture; |w1| = {s1['norm']:.12e}, |w1 - w0| / |w0| = "
          f"{np.linalg.norm(w1 - w0) / np.linalg.norm(w0):.3e}", flush=True)


if __name__ == "__main__":
    main()
