"""The channel-flow oracle (oracle/channel_oracle.py) and the product's host constants against the reference's own
source (tests/golden/reference_shim_channel.npz, made by tests/golden/make_reference_channel.py: channel.py,
solvers/base.py and utils/utils.py of the reference imported unmodified over NumPy stand-ins, fp64)."""
import os

import numpy as np
import pytest

from oracle import channel_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
SUB = (slice(None, None, 6), slice(None, None, 6), slice(None))
LOW = np.r_[0:6, 67:72]


def sub(a):
    """What the fixture keeps of a 72^3 array (make_reference_channel.py: pack)."""
    a = np.asarray(a)
    return a[np.ix_(LOW, LOW, np.arange(72))] if np.iscomplexobj(a) else a[SUB]


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "reference_shim_channel.npz"))


@pytest.fixture(scope="module")
def env():
    return orc.ChannelFlowSpectralEnv(dtype=np.float64)


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(np.asarray(b)), 1e-300)


def check_field(gold, name, a, tol):
    assert rel(sub(a), gold[name + "_sub"]) < tol, name
    assert abs(np.sqrt((np.abs(a) ** 2).sum()) - gold[name + "_norm"]) < tol * gold[name + "_norm"], name


def test_operator_constants_match_reference(gold, env):
    eq = env.equation
    assert np.array_equal(eq.z, gold["z"])
    assert rel(eq.Dz, gold["Dz"]) < 1e-15 and rel(eq.Dzz, gold["Dzz"]) < 1e-15
    assert np.array_equal(eq.dealias[:, :, 0], gold["dealias"])
    for m, ref in zip(gold["modes"], gold["A_inv_modes"]):
        assert rel(eq.A_inv[m], ref) < 1e-12
    clipped = np.clip(gold["action"], -1, 1)
    assert rel(eq.apply_jets_v(np.zeros((72, 72, 72)), clipped)[:, :, 2], gold["jets_plane"]) < 1e-14
    assert np.array_equal(env.get_obs(env.reset_env()[1]), gold["obs0"])


def test_product_host_constants_match_reference(gold):
    """hydrogym_b200/channel_constants.py (what kc_set_operators uploads) against the reference's constructor."""
    from hydrogym_b200 import channel_constants as cc
    z, dz, dzz, a_inv, jets, jdx, jdy, jh = cc.operators()
    assert np.array_equal(z, gold["z"])
    assert rel(dz, gold["Dz"]) < 1e-15 and rel(dzz, gold["Dzz"]) < 1e-15
    for m, ref in zip(gold["modes"], gold["A_inv_modes"]):
        kxi, kyi = divmod(int(m), 72)
        ia, ib = min(kxi, 72 - kxi), min(kyi, 72 - kyi)
        assert rel(a_inv[ia, ib], ref) < 1e-12
    clipped = np.clip(gold["action"], -1, 1)
    plane = np.tensordot(clipped, jets, axes=(0, 0))
    assert rel(plane, gold["jets_plane"]) < 1e-13
    # the spectral companions are what fft_xy / dx_hat / ifft_xy(...).real make of that plane
    kx = np.fft.fftfreq(72, d=cc.LX / 72) * 2 * np.pi
    h = np.fft.fftn(plane)
    assert rel(np.tensordot(clipped, jh[..., 0] + 1j * jh[..., 1], axes=(0, 0)), h) < 1e-13
    assert rel(np.tensordot(clipped, jdx, axes=(0, 0)), np.fft.ifftn(1j * kx[:, None] * h).real) < 1e-12
    assert list(cc.probe_indices()) == [0, 10, 20, 30, 40, 50, 60, 71]


def test_product_synthetic_fields_are_the_fixture_fields():
    from hydrogym_b200 import channel_constants as cc
    for a, b in zip(cc.synthetic_initial_fields(), orc.synthetic_initial_fields()):
        assert np.array_equal(a, b)


def test_rhs_terms_match_reference(gold, env):
    eq = env.equation
    s0 = eq.to_spectral(orc.VelocityState(env.U0, env.V0, env.W0))
    clipped = np.clip(gold["action"], -1, 1)
    n = eq.nonlinear_terms(s0, action=clipped, t=3.0, fx=2.0)
    l = eq.linear_terms(s0, action=clipped, t=3.0)
    for nm, a in zip(("Nu", "Nv", "Nw", "Lu", "Lv", "Lw"), (n.u, n.v, n.w, l.u, l.v, l.w)):
        check_field(gold, nm, a, 1e-12)


@pytest.mark.parametrize("tag", ["act", "noact"])
def test_rk4_steps_match_reference(gold, env, tag):
    eq = env.equation
    s = eq.to_spectral(orc.VelocityState(env.U0, env.V0, env.W0))
    a = np.clip(gold["action"], -1, 1) if tag == "act" else None
    for n in range(2):
        s = env.integrator.rk4_step(s, dt=orc.DT, action=a, t=n, fx=2.0, target_bulk_u=8.0)
        for c, f in zip("uvw", (s.u, s.v, s.w)):
            check_field(gold, f"rk{n + 1}_{tag}_{c}", f, 1e-11)


def test_jet_gain_ramp():
    g = orc.PseudoSpectralNavierStokes3D.jet_gain
    assert g(0.0) == 0.0 and abs(g(5.0) - 0.15) < 1e-15 and g(20.0) == 0.3
    assert abs(g(45.0) - 0.15) < 1e-15 and abs(g(60.0) + 0.3) < 1e-15   # negative beyond action_time, as the reference


def test_pressure_inverses_are_shared_by_the_four_sign_combinations(env):
    """kc_set_operators uploads 37 x 37 inverses (|kx|, |ky|): k2 is even in both wavenumbers, so the oracle's own
    72 x 72 inverses (C:128-148) must coincide for (+-kx, +-ky) -- including the Nyquist lines."""
    from hydrogym_b200 import channel_constants as cc
    a_inv = cc.operators()[3]
    A = env.equation.A_inv.reshape(72, 72, 72, 72)
    rng = np.random.default_rng(0)
    for kxi, kyi in [(0, 0), (0, 36), (36, 36), (71, 1), (5, 67)] + [tuple(rng.integers(0, 72, 2)) for _ in range(12)]:
        ia, ib = min(kxi, 72 - kxi), min(kyi, 72 - kyi)
        assert rel(a_inv[ia, ib], A[kxi, kyi]) < 1e-12, (kxi, kyi)


def test_jet_basis_is_linear_and_zero_mean(env):
    """The kernels combine 24 precomputed patterns linearly (C:194-217 is linear in Vjets): any amplitude vector
    gives the oracle's plane; each pattern has zero mean and lives on the 6 x 4 lattice of grid indices."""
    from hydrogym_b200 import channel_constants as cc
    jets = cc.jet_basis()
    assert jets.shape == (24, 72, 72) and np.abs(jets.mean(axis=(1, 2))).max() < 1e-15
    rng = np.random.default_rng(3)
    for _ in range(3):
        v = rng.uniform(-2, 2, 24)
        assert rel(np.tensordot(v, jets, 1), env.equation.jet_pattern(v)) < 1e-13
    peak = np.unravel_index(np.argmax(jets[0]), (72, 72))
    assert peak == (3, 4)                       # first slit centre: x index 3, y index 4 (C:199-200)
    peak = np.unravel_index(np.argmax(jets[23]), (72, 72))
    assert peak == (44, 67)                     # last: x = int(72 * 2/3) - 4, y = 72 - 5


def test_oracle_env_step_bookkeeping(env):
    """time += 50 regardless of the oracle's shortened rollout, terminal at the horizon, reward = -wss (C:502-505)."""
    short = orc.ChannelFlowSpectralEnv(nsteps=1)
    obs, st = short.reset_env()
    obs, st, reward, done, info = short.step_env(st, np.zeros(24))
    assert st["time"] == 50 and not done and info["discount"] == 1.0 and obs.shape == (128,)
    assert reward == -orc.wss_compute(4, st["U"], z=short.equation.z)
    st["time"] = orc.MAX_STEPS - 50
    _, st, _, done, info = short.step_env(st, np.zeros(24))
    assert done and info["discount"] == 0.0
