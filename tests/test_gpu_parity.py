"""GPU parity: the sm_100a kernels (through the C ABI / Python boundary) vs the NumPy oracle."""
import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc

pytestmark = pytest.mark.gpu

ACTION = np.array([-0.25, -0.03, 0.02, 0.01])  # the reference script's minimize_tke action


def rel_l2(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def random_states(ref_env, B, seed, amp=2.0):
    """reference initial state + smooth random perturbation, generally NOT Hermitian in columns 0 / M/2."""
    rng = np.random.default_rng(seed)
    N, Mh = ref_env.n, ref_env.m // 2 + 1
    p = np.abs(np.fft.fftfreq(N, 1.0 / N))[:, None]
    q = np.arange(Mh)[None, :]
    env = np.exp(-(p * p + q * q) / 64.0) * N * N
    w0 = ref_env.flow.initialize_state().astype(np.complex128)
    noise = (rng.normal(size=(B, N, Mh)) + 1j * rng.normal(size=(B, N, Mh))) * env * amp * 0.05
    return w0[None] + noise


def make_stepper(N, Re, B, dtype, dt=1e-3, **kw):
    from hydrogym_b200 import KolmogorovStepper
    return KolmogorovStepper(grid_size=(N, N), Re=Re, dt=dt, num_envs=B, dtype=dtype, device="cuda:0", **kw)


@pytest.mark.parametrize("N,steps", [(64, 20), (128, 6), (256, 3), (512, 2), (1024, 1)])
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-11), (torch.float32, 2e-5)])
def test_rk_steps_match_oracle(N, steps, dtype, tol):
    B = 3 if N <= 256 else 2
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=npdt)
    w0 = random_states(ref, B, seed=N).astype(orc._ctype(npdt))
    rng = np.random.default_rng(7)
    actions = rng.uniform(-0.7, 0.7, size=(B, 4)).astype(npdt)  # some outside the clip range
    clipped = ref._clip_action(actions, ref.default_params)
    ch = ref.equation.control_term(w0, ref._control_field(clipped, ref.default_params))
    want = ref.integrator.rk_steps(w0.copy(), steps, ch)
    st = make_stepper(N, 100, B, dtype)
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, torch.from_numpy(actions).cuda(), steps)
    got = w.cpu().numpy()
    assert np.isfinite(got).all()
    for b in range(B):
        assert rel_l2(got[b], want[b]) < tol, (b, rel_l2(got[b], want[b]))


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-5), (torch.float32, 1e-3)])
def test_100_rk_steps_tolerance_of_north_star(dtype, tol):
    """BASELINE.json: rel-L2 on vorticity <= 1e-5 (fp64) / 1e-3 (fp32) after 100 steps."""
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(64, 64)), dtype=npdt)
    w0 = ref.flow.initialize_state()[None]
    ch = ref.equation.control_term(w0, ref._control_field(ACTION[None].astype(npdt), ref.default_params))
    want = ref.integrator.rk_steps(w0.copy(), 100, ch)
    st = make_stepper(64, 100, 1, dtype)
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, torch.from_numpy(ACTION[None].astype(npdt)).cuda(), 100)
    got = w.cpu().numpy()
    vort_got, vort_want = np.fft.irfftn(got[0]), np.fft.irfftn(want[0])
    assert rel_l2(vort_got, vort_want) < tol
    assert rel_l2(got, want) < tol


def test_graph_and_chunked_paths_agree_with_direct_launches():
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(64, 64)))
    w0 = random_states(ref, 5, seed=3)
    acts = torch.from_numpy(np.random.default_rng(1).uniform(-0.5, 0.5, (5, 4))).cuda()
    outs = []
    for kw in (dict(use_graph=False), dict(use_graph=True), dict(use_graph=True, chunk=2), dict(use_graph=False, chunk=3)):
        st = make_stepper(64, 100, 5, torch.float64, **kw)
        w = torch.from_numpy(w0).cuda()
        with torch.cuda.stream(torch.cuda.Stream()):
            st.rk_steps(w, acts, 19)
        torch.cuda.synchronize()
        outs.append(w.cpu().numpy())
    for o in outs[1:]:
        assert np.array_equal(o, outs[0])


@pytest.mark.parametrize("N,B,steps", [(64, 2, 5), (128, 5, 3), (256, 40, 3), (1024, 2, 1)])
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-11), (torch.float32, 2e-5)])
def test_dealias_flag_matches_intended_mask(N, B, steps, dtype, tol):
    """2/3 mask ON (the bench configuration): masked column tiles skip the advection load / transform and the row
    kernel skips the masked columns; B = 40 at 256^2 gives every persistent CTA several tiles of both kinds."""
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=npdt, dealias=True)
    w0 = random_states(ref, B, seed=5).astype(orc._ctype(npdt))
    nref = min(B, 3)  # the oracle is slow: check the first, middle and last environments
    pick = sorted({0, B // 2, B - 1})[:nref]
    actions = np.random.default_rng(3).uniform(-0.5, 0.5, size=(B, 4)).astype(npdt)
    ch = ref.equation.control_term(w0[pick], ref._control_field(actions[pick], ref.default_params))
    want = ref.integrator.rk_steps(w0[pick].copy(), steps, ch)
    st = make_stepper(N, 100, B, dtype, dealias=True)
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, torch.from_numpy(actions).cuda(), steps)
    got = w.cpu().numpy()
    for i, b in enumerate(pick):
        assert rel_l2(got[b], want[i]) < tol, (b, rel_l2(got[b], want[i]))


@pytest.mark.parametrize("N", [64, 128, 256])
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-10), (torch.float32, 3e-5)])
def test_reward_obs_match_oracle(N, dtype, tol):
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=npdt)
    traj = random_states(ref, 2 * 3, seed=11).reshape(2, 3, N, N // 2 + 1).astype(orc._ctype(npdt))
    want_obs = ref._trajectory_mean_obs(traj)
    want_tke = ref._avg_tke(traj)
    st = make_stepper(N, 100, 2, dtype)
    obs, tke, nonf = st.reward_obs(torch.from_numpy(traj).cuda())
    assert np.abs(obs.cpu().numpy() - want_obs).max() < tol * np.abs(want_obs).max()
    assert np.abs(tke.cpu().numpy() - want_tke).max() < tol * np.abs(want_tke).max()
    assert int(nonf.sum()) == 0


def test_grid_beyond_1024_runs_on_the_slab_kernels_and_matches_oracle():
    """2048^2 (and 4096^2, 8192^2) have no batched plans: KolmogorovStepper serves them environment by environment
    through the single-grid slab kernels on one rank.  One RK step + reward / observation against the oracle."""
    N, B = 2048, 2
    ref = orc.KolmogorovFlow(dict(dt=2e-4), dict(Re=1000, grid_size=(N, N)), dtype=np.float32)
    w0 = random_states(ref, B, seed=77).astype(np.complex64)
    actions = np.array([[0.3, -0.2, 0.1, -0.4], [0.9, 0.0, -0.9, 0.2]], dtype=np.float32)  # second one gets clipped
    clipped = ref._clip_action(actions, ref.default_params)
    ch = ref.equation.control_term(w0, ref._control_field(clipped, ref.default_params))
    want = ref.integrator.rk_steps(w0.copy(), 1, ch)
    st = make_stepper(N, 1000, B, torch.float32, dt=2e-4, n_inner=1, n_outer=1)
    assert st.lanes == 1 and st.chunk == 1
    final, traj, obs, tke, nonf = st.step(torch.from_numpy(w0).cuda(), torch.from_numpy(actions).cuda())
    got = final.cpu().numpy()
    for b in range(B):
        assert rel_l2(got[b], want[b]) < 2e-5, (b, rel_l2(got[b], want[b]))
    assert torch.equal(traj[:, 0], final) and (nonf == 0).all()
    want_obs = ref._trajectory_mean_obs(want[:, None])
    want_tke = ref._avg_tke(want[:, None])
    assert np.abs(obs.cpu().numpy() - want_obs).max() < 1e-4 * max(1.0, np.abs(want_obs).max())
    assert np.abs(tke.cpu().numpy() / want_tke - 1).max() < 1e-4
    with pytest.raises(NotImplementedError):
        st.physical_fields(final)
    st.close()
    with pytest.raises(ValueError):
        make_stepper(3072, 1000, 1, torch.float32)
