"""CPU tests of the oracle: golden vectors (oracle-generated, see tests/golden/make_golden.py), the
reference's quirks and analytic known answers."""
import os

import numpy as np
import pytest

from oracle import kolmogorov_oracle as orc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "kolmogorov_64_re100.npz"))
FLOW = dict(Re=100, grid_size=(64, 64))


def rel(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def test_step_counts_use_float_floor_division():
    assert orc.step_counts(1e-3, 10.0) == (999, 10)
    assert orc.step_counts(1e-2, 10.0) == (99, 10)
    assert orc.step_counts(5e-3, 10.0) == (199, 10)
    assert orc.step_counts(0.1, 10.0) == (9, 11)


def test_dealiasing_is_identity_as_written():
    a = np.random.default_rng(0).normal(size=(64, 33)) + 0j
    assert orc.dealiasing(a) is a
    masked = orc.dealiasing(a, apply_intended=True)
    assert np.all(masked[21:42] == 0) and np.all(masked[:, 22:] == 0) and np.all(masked[:21, :22] == a[:21, :22])


def test_forcing_and_control_live_on_row_zero_and_leak_over_ky():
    env = orc.KolmogorovFlow(flow_config=FLOW)
    f = env.equation._forcing
    assert np.abs(f[1:]).max() == 0.0
    assert (np.abs(f[0]) > 1e-6).sum() >= 30          # endpoint-inclusive grid -> leakage over all ky (F2)
    cf = env._control_field(np.array([1.0, 0, 0, 0]), env.default_params)
    c = env.equation.control_term(f, cf)
    assert np.abs(c[1:]).max() == 0.0
    ky = np.arange(33) / (2 * np.pi)
    assert np.allclose(c[0], -ky * np.fft.rfft2(np.sin(4 * env.y))[0])       # no 2*pi*i factor (F4)
    assert np.allclose(f[0], -1j * np.arange(33) * np.fft.rfft2(np.sin(4 * env.y))[0])


def test_control_is_linear_in_action():
    env = orc.KolmogorovFlow(flow_config=FLOW)
    a = np.array([0.3, -0.2, 0.1, 0.4])
    full = env.equation.control_term(None or env.equation._forcing, env._control_field(a, env.default_params))
    parts = sum(a[i] * env.equation.control_term(full, env._control_field(np.eye(4)[i], env.default_params))
                for i in range(4))
    assert np.allclose(full, parts, atol=1e-10)


def test_linear_decay_known_answer():
    """Without advection/forcing each mode decays by prod_s (1 - mu_s nu k^2)/(1 + mu_s nu k^2) per RK step."""
    env = orc.KolmogorovFlow(flow_config=FLOW)
    eq = env.equation
    eq.nonlinear_terms = lambda w, c=None: np.zeros_like(w)
    w0 = np.random.default_rng(1).normal(size=(64, 33)) + 1j * np.random.default_rng(2).normal(size=(64, 33))
    w1 = env.integrator.time_step(w0)
    p = np.fft.fftfreq(64, 1 / 64)[:, None]
    q = np.arange(33)[None, :]
    k2 = p * p + q * q
    fac = np.ones_like(k2)
    for s in range(5):
        mu = 0.5 * 1e-3 * (orc.ALPHA_RK4[s + 1] - orc.ALPHA_RK4[s])
        fac *= (1 - mu * 0.01 * k2) / (1 + mu * 0.01 * k2)
    assert np.allclose(w1, w0 * fac, rtol=1e-12)


def test_golden_vectors_fp64_and_fp32():
    for name, dt, tol in (("f64", np.float64, 1e-12), ("f32", np.float32, 2e-4)):
        env = orc.KolmogorovFlow(flow_config=FLOW, dtype=dt)
        w = env.flow.initialize_state()
        assert rel(w, GOLD[f"omega0_{name}"]) < 1e-6
        w = env.integrator.rk_steps(w, 10)
        assert rel(w, GOLD[f"zero_action_rk10_{name}"]) < tol
    # cross precision: fp32 vs fp64 trajectories agree to ~1e-5 over 100 RK steps
    assert rel(GOLD["zero_action_rk100_f32"], GOLD["zero_action_rk100_f64"]) < 1e-3
    assert rel(GOLD["action_rk100_f32"], GOLD["action_rk100_f64"]) < 1e-3


def test_survey_anchor_values():
    # TKE = 0.50622 @200, 0.60933 @1000, 0.64840 @2000 RK steps (64^2, Re=100, zero action)
    assert np.allclose(GOLD["tke_200_1000_2000"], [0.50622, 0.60933, 0.64840], atol=2e-5)
    env = orc.KolmogorovFlow(flow_config=FLOW)
    w = env.integrator.rk_steps(env.flow.initialize_state(), 200)
    assert abs(float(orc.compute_tke(w, env.kx, env.ky, 64)) - GOLD["tke_200_1000_2000"][0]) < 1e-10


def test_env_trace_matches_golden_and_reward_ignores_flow_by_default():
    env = orc.KolmogorovFlow(dict(dt=1e-2, action_time=2.0), FLOW)
    obs, state = env.reset_env()
    assert np.allclose(obs, GOLD["env_reset_obs"], rtol=1e-10)
    assert state["trajectory"].shape == (2, 64, 33)
    assert np.array_equal(state["trajectory"][-1], state["omega_hat"])      # F10
    a = GOLD["env_actions"][0]
    obs, state, reward, done, info = env.step_env(None, state, a)
    assert np.allclose(obs, GOLD["env_obs"][0], rtol=1e-9)
    assert reward == pytest.approx(-np.abs(a).sum())                        # reward_alpha = 0 (F8)
    assert info["mean_tke"] == pytest.approx(GOLD["env_mean_tke"][0], rel=1e-10)
    assert not done and info["discount"] == 1.0 and state["time"] == 1


def test_solve_rejects_short_horizons():
    env = orc.KolmogorovFlow(flow_config=FLOW)
    with pytest.raises(ValueError):
        env.integrator.solve(1e-3, (0.0, 0.5))


def test_batched_equals_looped():
    env = orc.KolmogorovFlow(flow_config=FLOW)
    rng = np.random.default_rng(3)
    w0 = env.flow.initialize_state()[None] + rng.normal(size=(3, 64, 33)) * 50
    acts = rng.uniform(-0.5, 0.5, (3, 4))
    ch = env.equation.control_term(w0, env._control_field(acts, env.default_params))
    both = env.integrator.rk_steps(w0, 3, ch)
    for b in range(3):
        chb = env.equation.control_term(w0[b], env._control_field(acts[b], env.default_params))
        assert np.allclose(both[b], env.integrator.rk_steps(w0[b], 3, chb), rtol=1e-12)


def test_energy_spectrum_sums_to_kinetic_energy():
    """The radial spectrum (oracle binning of utils.py:89-101 per-mode energies) sums to the physical-space
    kinetic energy 0.5 <u^2 + v^2> when the Nyquist lines are empty (Parseval)."""
    N = 32
    env = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float64)
    rng = np.random.default_rng(0)
    w = rng.normal(size=(N, N))
    wh = np.fft.rfft2(w)
    wh[N // 2, :] = 0
    wh[:, N // 2] = 0
    wh[0, 0] = 0
    spec = orc.energy_spectrum(wh, env.kx, env.ky)
    uh, vh = orc.compute_velocity_fft(wh, env.kx, env.ky)
    u, v = np.fft.irfft2(uh), np.fft.irfft2(vh)
    assert spec.shape == (N,)
    assert abs(spec.sum() - 0.5 * np.mean(u * u + v * v)) < 1e-12 * spec.sum()
    # one mode: matches compute_energy_mode on that mode and its conjugate
    one = np.zeros_like(wh)
    one[3, 4] = 2.0 - 1.0j
    s1 = orc.energy_spectrum(one, env.kx, env.ky)
    uh, vh = orc.compute_velocity_fft(one, env.kx, env.ky)
    assert abs(s1[5] - 2 * orc.compute_energy_mode(uh, vh, 3, 4, N, N)) < 1e-18
    assert s1.sum() == s1[5]


def test_torch_cpu_port_matches_numpy_oracle():
    """oracle/torch_cpu_port.py (the CPU baseline's torch / MKL variant) restates the same RK step."""
    from oracle.torch_cpu_port import TorchCpuSolver
    for dt_, tol in ((np.float64, 1e-13), (np.float32, 1e-6)):
        for dealias in (False, True):
            env = orc.KolmogorovFlow(dict(dt=1e-3), dict(Re=100, grid_size=(64, 64)), dtype=dt_, dealias=dealias)
            rng = np.random.default_rng(0)
            p = np.abs(np.fft.fftfreq(64, 1 / 64))[:, None]
            q = np.arange(33)[None, :]
            w0 = env.flow.initialize_state()[None] + (rng.normal(size=(3, 64, 33)) + 1j * rng.normal(size=(3, 64, 33))) \
                * np.exp(-(p * p + q * q) / 64) * 64 * 64 * 0.1
            w0 = w0.astype(orc._ctype(dt_))
            acts = rng.uniform(-0.5, 0.5, (3, 4)).astype(dt_)
            ch = env.equation.control_term(w0, env._control_field(acts, env.default_params))
            a = env.integrator.rk_steps(w0.copy(), 5, ch)
            b = TorchCpuSolver(env).rk_steps(w0, 5, ch)
            assert np.linalg.norm(a - b) / np.linalg.norm(a) < tol
