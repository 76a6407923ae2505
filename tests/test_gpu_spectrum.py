"""Energy-spectrum diagnostics: the CUDA reduction against the oracle, and the BASELINE.json configs[3] check
(1024^2, Re=1000, 16 envs, long-horizon rollout: energy-spectrum statistics of the fp32 and fp64 kernels agree)."""
import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc
from test_gpu_parity import random_states

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N", [64, 256])
@pytest.mark.parametrize("dtype", [np.complex64, np.complex128])
def test_energy_spectrum_matches_oracle(N, dtype):
    from hydrogym_b200.diagnostics import energy_spectrum
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float64)
    w = random_states(ref, 3, seed=N).astype(dtype)
    want = orc.energy_spectrum(w.astype(np.complex128), ref.kx, ref.ky)
    got = energy_spectrum(torch.from_numpy(w).cuda()).cpu().numpy()
    assert got.shape == want.shape == (3, N)
    assert np.abs(got - want).max() <= 1e-12 * want.max()


def _synthetic_states(N, B, seed0=1234):
    """SURVEY 8d synthetic inputs: band-limited Gaussian random vorticity, rms(omega) = 4."""
    out = np.empty((B, N, N // 2 + 1), dtype=np.complex128)
    p = np.fft.fftfreq(N, 1.0 / N)[:, None]
    q = np.arange(N // 2 + 1)[None, :]
    k = np.sqrt(p * p + q * q)
    for b in range(B):
        rng = np.random.default_rng(seed0 + b)
        w = (rng.normal(size=k.shape) + 1j * rng.normal(size=k.shape)) * np.exp(-(k / 8.0) ** 2)
        w[(k < 1) | (k > N / 3)] = 0
        real = np.fft.irfft2(w, s=(N, N))
        real *= 4.0 / np.sqrt(np.mean(real ** 2))
        out[b] = np.fft.rfft2(real)
        out[b, 0, 0] = 0
    return out


def _mean_spectrum(N, B, dtype, w0, actions, n_blocks, block):
    from hydrogym_b200 import KolmogorovStepper
    from hydrogym_b200.diagnostics import energy_spectrum
    cd = torch.complex64 if dtype == torch.float32 else torch.complex128
    st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=dtype, device="cuda:0", dealias=True)
    w = torch.from_numpy(w0).to(cd).cuda()
    act = torch.from_numpy(actions).to(dtype).cuda()
    acc = torch.zeros(N, dtype=torch.float64, device="cuda")
    cnt = 0
    for j in range(n_blocks):
        st.rk_steps(w, act, block)
        if j >= n_blocks // 2:  # statistics over the second half of the horizon
            acc += energy_spectrum(w).mean(0)
            cnt += 1
    st.close()
    assert torch.isfinite(torch.view_as_real(w)).all()
    return (acc / cnt).cpu().numpy()


def test_long_horizon_energy_spectrum_statistics_1024():
    """configs[3]: 1024^2, Re=1000, 16 envs.  Over t = 4 (4000 RK steps, ~16 eddy turnovers at rms(omega) = 4)
    fp32 and fp64 trajectories decorrelate; their ensemble- and time-averaged radial energy spectra must
    still agree.  Tolerances are statistical (16 envs x 10 snapshots): total energy 5 %, energetic bins
    (k <= 32) 25 %, inertial range slope through log-spectrum distance."""
    N, B = 1024, 16
    w0 = _synthetic_states(N, B)
    actions = np.random.default_rng(4321).uniform(-0.5, 0.5, size=(B, 4))
    e32 = _mean_spectrum(N, B, torch.float32, w0, actions, n_blocks=20, block=200)
    e64 = _mean_spectrum(N, B, torch.float64, w0, actions, n_blocks=20, block=200)
    assert abs(e32.sum() - e64.sum()) < 0.05 * e64.sum(), (e32.sum(), e64.sum())
    ks = np.arange(1, 33)
    rel = np.abs(e32[ks] - e64[ks]) / e64[ks]
    assert rel.max() < 0.25, (int(ks[rel.argmax()]), float(rel.max()))
    kk = np.arange(33, N // 3)
    live = e64[kk] > 1e-14 * e64.max()
    d = np.abs(np.log10(e32[kk][live] / e64[kk][live]))
    assert np.median(d) < 0.05 and d.max() < 0.5, (float(np.median(d)), float(d.max()))


def test_energy_spectrum_statistics_match_oracle_64():
    """Same statistic against the NumPy oracle where it is affordable: 64^2, Re=100, zero action from the
    reference initial state; 1500 RK steps, spectra averaged over the last 500 (every 100)."""
    from hydrogym_b200 import KolmogorovStepper
    from hydrogym_b200.diagnostics import energy_spectrum
    N = 64
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float64)
    w_ref = ref.flow.initialize_state()[None].astype(np.complex128)
    st = KolmogorovStepper(grid_size=(N, N), Re=100, num_envs=1, dtype=torch.float32, device="cuda:0")
    w = torch.from_numpy(w_ref.astype(np.complex64)).cuda()
    acc_ref, acc = np.zeros(N), np.zeros(N)
    for j in range(15):
        w_ref = ref.integrator.rk_steps(w_ref, 100, None)
        st.rk_steps(w, None, 100)
        if j >= 10:
            acc_ref += orc.energy_spectrum(w_ref[0], ref.kx, ref.ky)
            acc += energy_spectrum(w[0]).cpu().numpy()
    st.close()
    live = acc_ref > 1e-12 * acc_ref.max()
    assert np.abs(acc[live] - acc_ref[live]).max() < 1e-3 * acc_ref.max()
    assert abs(acc.sum() - acc_ref.sum()) < 1e-4 * acc_ref.sum()


@pytest.mark.parametrize("dtype,tol", [(np.complex128, 1e-11), (np.complex64, 2e-5)])
def test_physical_space_diagnostics_match_oracle(dtype, tol):
    """compute_energy_dissipation / compute_divergence / compute_real_velocity_point (utils.py:129-172, 213-229)
    on the physical velocity fields the row-FFT kernels export."""
    from hydrogym_b200 import KolmogorovStepper, diagnostics
    N = 128
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float64)
    w = random_states(ref, 2, seed=8).astype(dtype)
    wt = torch.from_numpy(w).cuda()
    st = KolmogorovStepper(grid_size=(N, N), Re=100, num_envs=2, device="cuda:0",
                           dtype=torch.float64 if dtype == np.complex128 else torch.float32)
    f = st.physical_fields(wt, 4).cpu().numpy()
    st.close()
    for b in range(2):
        w64 = w[b].astype(np.complex128)
        uh, vh = orc.compute_velocity_fft(w64, ref.kx, ref.ky)
        p = np.fft.fftfreq(N, 1.0 / N)[:, None]
        q = np.arange(N // 2 + 1)[None, :]
        want = [np.fft.irfft2(uh), np.fft.irfft2(vh), np.fft.irfft2(1j * p * w64), np.fft.irfft2(1j * q * w64)]
        for i in range(4):
            assert np.abs(f[i, b] - want[i]).max() < tol * np.abs(want[i]).max(), (b, i)
        eps = float(diagnostics.compute_energy_dissipation(wt[b], 0.01))
        assert abs(eps - orc.compute_energy_dissipation(w64, ref.kx, ref.ky, 0.01, N)) < 10 * tol * abs(eps)
        div = diagnostics.compute_divergence(wt[b]).cpu().numpy()
        wdiv = orc.compute_divergence(w64, ref.kx, ref.ky)
        assert np.abs(div - wdiv).max() < 10 * tol * max(np.abs(want[0]).max(), np.abs(wdiv).max())
        pt = float(diagnostics.compute_real_velocity_point(wt[b], 5, 77))
        assert abs(pt - orc.compute_real_velocity_point(uh, vh, 5, 77)) < tol * max(1.0, pt)
