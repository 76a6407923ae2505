"""GPU parity of the env boundary (reset_env / step_env / step / VectorEnv façade) vs the oracle."""
import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc

pytestmark = pytest.mark.gpu

ENV_CFG = dict(dt=1e-2, action_time=2.0)   # n_inner = 99, n_outer = 2: short horizon, still pre-turbulent
FLOW_CFG = dict(Re=100, grid_size=(64, 64))
ACTIONS = [np.array([-0.25, -0.03, 0.02, 0.01]), np.array([0.9, -0.9, 0.1, 0.0]), np.zeros(4)]


def rel_l2(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


@pytest.mark.parametrize("dtype,tol", [("float64", 1e-9), ("float32", 5e-4)])
def test_reset_and_three_steps_match_oracle(dtype, tol):
    from hydrogym_b200 import KolmogorovFlow
    npdt = np.float64 if dtype == "float64" else np.float32
    env = KolmogorovFlow(dict(ENV_CFG, dtype=dtype, device="cuda:0"), FLOW_CFG)
    ref = orc.KolmogorovFlow(ENV_CFG, FLOW_CFG, dtype=npdt)
    assert (env.n_inner, env.n_outer) == (99, 2)
    params = env.default_params
    obs, state = env.reset_env(None, params)
    robs, rstate = ref.reset_env()
    assert obs.shape == (64,) and state.omega_hat.shape == (64, 33) and state.trajectory.shape == (2, 64, 33)
    assert rel_l2(state.omega_hat.cpu().numpy(), rstate["omega_hat"]) < tol
    assert np.abs(obs.cpu().numpy() - robs).max() < tol * max(1.0, np.abs(robs).max())
    for a in ACTIONS:
        obs, state, reward, done, info = env.step_env(None, state, torch.tensor(a), params)
        robs, rstate, rreward, rdone, rinfo = ref.step_env(None, rstate, a)
        assert rel_l2(state.omega_hat.cpu().numpy(), rstate["omega_hat"]) < tol
        assert rel_l2(state.trajectory.cpu().numpy(), rstate["trajectory"]) < tol
        assert np.abs(obs.cpu().numpy() - robs).max() < tol * max(1.0, np.abs(robs).max())
        assert abs(float(reward) - float(rreward)) < 1e-6
        assert abs(float(info["mean_tke"]) - float(rinfo["mean_tke"])) < tol * float(rinfo["mean_tke"])
        assert bool(done) == bool(rdone) and float(info["discount"]) == float(rinfo["discount"])
    assert int(state.time) == 3


def test_reward_alpha_and_time_limit_and_autoreset():
    from hydrogym_b200 import KolmogorovFlow
    env = KolmogorovFlow(dict(ENV_CFG, dtype="float64", device="cuda:0"), FLOW_CFG)
    ref = orc.KolmogorovFlow(ENV_CFG, FLOW_CFG)
    params = env.default_params.replace(reward_alpha=0.5, max_episode_steps=2)
    rparams = ref.default_params.replace(reward_alpha=0.5, max_episode_steps=2)
    obs0, state = env.reset(None, params, num_envs=3)
    assert obs0.shape == (3, 64) and state.omega_hat.shape == (3, 64, 33)
    _, rstate = ref.reset_env()
    acts = np.stack(ACTIONS)
    obs, state, reward, done, info = env.step(None, state, torch.from_numpy(acts), params)
    for b in range(3):
        _, rs, rr, rd, ri = ref.step_env(None, rstate, acts[b], rparams)
        assert abs(float(reward[b]) - float(rr)) < 1e-9
        assert rel_l2(state.omega_hat[b].cpu().numpy(), rs["omega_hat"]) < 1e-9
    assert not bool(done.any()) and state.time.tolist() == [1, 1, 1]
    obs, state, reward, done, info = env.step(None, state, torch.from_numpy(acts), params)
    assert bool(done.all()) and info["discount"].tolist() == [0.0, 0.0, 0.0]
    # auto-reset: state and obs are the post-reset ones again
    assert state.time.tolist() == [0, 0, 0]
    assert np.allclose(obs.cpu().numpy(), obs0.cpu().numpy())
    assert rel_l2(state.omega_hat[1].cpu().numpy(), rstate["omega_hat"]) < 1e-9


def test_vector_env_facade():
    from hydrogym_b200 import KolmogorovVectorEnv
    venv = KolmogorovVectorEnv(4, dict(ENV_CFG), FLOW_CFG, dtype="float64", device="cuda:0")
    ref = orc.KolmogorovFlow(ENV_CFG, FLOW_CFG)
    obs, info = venv.reset(seed=0)
    assert obs.shape == (4, 64) and venv.single_action_space.shape == (4,)
    _, rstate = ref.reset_env()
    acts = np.random.default_rng(0).uniform(-0.5, 0.5, (4, 4))
    obs, reward, terminated, truncated, info = venv.step(torch.from_numpy(acts))
    for b in range(4):
        robs, rs, rr, rd, ri = ref.step_env(None, rstate, acts[b])
        assert np.abs(obs[b].cpu().numpy() - robs).max() < 1e-9
        assert abs(float(reward[b]) - float(rr)) < 1e-12
        assert abs(float(info["mean_tke"][b]) - float(ri["mean_tke"])) < 1e-9
    assert not bool(terminated.any()) and not bool(truncated.any())
    venv.close()


def test_solver_surface_and_end_time_guard():
    from hydrogym_b200 import FlowConfig
    from hydrogym_b200.solvers import RungeKuttaCrankNicolson
    flow = FlowConfig(**FLOW_CFG)
    integ = RungeKuttaCrankNicolson(flow, dt=1e-2, save_n=1, dtype=torch.float64, device="cuda:0")
    with pytest.raises(ValueError):
        integ.solve(1e-2, flow, (0.0, 0.5))
    final, traj = integ.solve(1e-2, flow, (0.0, 3.0), save_n=1)
    ref = orc.KolmogorovFlow(dict(dt=1e-2, action_time=3.0), FLOW_CFG)
    rfinal, rtraj = ref.integrator.solve(1e-2, (0.0, 3.0), 1)
    assert traj.shape == (3, 64, 33)
    assert rel_l2(traj.cpu().numpy(), rtraj) < 1e-9
    # x-independent control field through the documented tuple interface
    cf = ref._control_field(ACTIONS[0], ref.default_params)
    final, traj = integ.solve(1e-2, flow, (0.0, 1.0), save_n=1, control_field=cf)
    rfinal, rtraj = ref.integrator.solve(1e-2, (0.0, 1.0), 1, control_field=cf)
    assert rel_l2(final.cpu().numpy(), rfinal) < 1e-9


def test_wrapper_stack_over_the_cuda_env_matches_oracle():
    """The reference PPO nesting (run_ppo.py:165-172) over the B200 env: raw quantities from the oracle env, wrapper
    arithmetic from the NumPy restatement of env_core.py:449-742."""
    from hydrogym_b200 import KolmogorovFlow
    from hydrogym_b200 import wrappers as W
    from oracle import wrappers_oracle as wo
    B, gamma = 3, 0.99
    env = KolmogorovFlow(dict(ENV_CFG, dtype="float64", device="cuda:0"), FLOW_CFG)
    ref = orc.KolmogorovFlow(ENV_CFG, FLOW_CFG, dtype=np.float64)
    params = env.default_params
    w = W.NormalizeVecReward(W.NormalizeVecObservation(W.VecEnv(W.ClipAction(W.LogWrapper(env)), num_envs=B)), gamma)
    assert w.name == "KolmogorovFlow" and w.num_actions == 4          # attribute proxying through 5 wrappers
    obs, st = w.reset(None, params)
    robs0, rstate = ref.reset_env()
    nobs, rno = wo.normalize_obs_reset(np.repeat(robs0[None], B, 0))
    assert np.abs(obs.cpu().numpy() - nobs).max() < 1e-7  # identical envs: variance 0 -> eps dominates
    rrw = wo.normalize_reward_reset(B)
    rlog = dict(episode_returns=np.zeros(B), episode_lengths=np.zeros(B, dtype=np.int64),
                returned_episode_returns=np.zeros(B), returned_episode_lengths=np.zeros(B, dtype=np.int64),
                timestep=np.zeros(B, dtype=np.int64))
    acts = np.array([[-0.25, -0.03, 0.02, 0.01], [1.7, -0.9, 0.1, 0.0], [0.0, 0.0, 0.0, 0.0]])
    rstates = [dict(rstate) for _ in range(B)]
    for _ in range(2):
        obs, st, r, d, info = w.step(None, st, torch.tensor(acts, device="cuda"), params)
        raw = [ref.step_env(None, rstates[b], np.clip(acts[b], -1.0, 1.0)) for b in range(B)]
        rstates = [x[1] for x in raw]
        raw_obs = np.stack([x[0] for x in raw])
        raw_rew = np.array([float(x[2]) for x in raw])
        raw_done = np.array([bool(x[3]) for x in raw])
        nobs, rno = wo.normalize_obs_step(rno, raw_obs)
        nrew, rrw = wo.normalize_reward_step(rrw, raw_rew, raw_done.astype(np.float64), gamma)
        linfo, rlog = wo.log_wrapper_step(rlog, raw_rew, raw_done.astype(np.int64))
        assert np.abs(obs.cpu().numpy() - nobs).max() < 1e-6 * max(1.0, np.abs(nobs).max())
        assert np.abs(r.cpu().numpy() - nrew).max() < 1e-8 * max(1.0, np.abs(nrew).max())
        assert np.array_equal(info["timestep"].cpu().numpy(), linfo["timestep"])
        assert np.allclose(st.env_state.env_state.episode_returns.cpu().numpy(), rlog["episode_returns"], atol=1e-9)
        assert not bool(d.any())


def test_solver_surface_callbacks_observations_and_npy(tmp_path):
    """Direct-solver surface (docs kolmogorov.md:56-102): solve() with a LogCallback, FlowConfig.get_observations,
    trajectory .npy round trip -- values against the oracle's solve."""
    from hydrogym_b200 import KolmogorovFlow, io
    env = KolmogorovFlow(dict(ENV_CFG, dtype="float64", device="cuda:0"), FLOW_CFG)
    ref = orc.KolmogorovFlow(ENV_CFG, FLOW_CFG, dtype=np.float64)
    log = tmp_path / "obs.txt"
    cb = io.LogCallback(postprocess=lambda f: f.get_observations().mean(dim=1), filename=str(log))
    final, traj = env.integrator.solve(dt=1e-2, flow=env.flow, t_span=(0.0, 2.0), callbacks=[cb], save_n=1)
    rfinal, rtraj = ref.integrator.solve(dt=1e-2, t_span=(0.0, 2.0), save_n=1)
    assert traj.shape == (2, 64, 33) and rel_l2(traj.cpu().numpy(), rtraj) < 1e-9
    want = np.stack([ref._trajectory_mean_obs(rtraj[t][None]) for t in range(2)])   # n = obs_size^2: same 8x8 points
    got = env.flow.get_observations()
    assert got.shape == (2, 64) and np.abs(got.cpu().numpy() - want).max() < 1e-9
    assert np.allclose(np.loadtxt(log), np.round(want.mean(axis=1), 5), atol=1e-5)
    io.save_trajectory(str(tmp_path / "traj.npy"), traj)
    back = io.load_trajectory(str(tmp_path / "traj.npy"), device="cuda")
    assert torch.equal(back, traj)


def test_vector_env_truncation_resets_state_obs_and_trajectory_without_reset_call():
    """Auto-reset after ``set_state`` (no ``reset()`` before): environments that hit ``max_episode_steps`` take the
    cached post-reset state, observation and trajectory (gymnax ``Environment.step`` semantics)."""
    from hydrogym_b200 import KolmogorovVectorEnv
    venv = KolmogorovVectorEnv(3, dict(ENV_CFG), FLOW_CFG, dtype="float64", device="cuda:0", keep_trajectory=True)
    venv.params = venv.params.replace(max_episode_steps=2)
    ref_obs, ref_state = venv.env.reset_env(None, venv.params)
    w0 = ref_state.omega_hat.unsqueeze(0).repeat(3, 1, 1) * torch.tensor([1.0, 0.9, 1.1], device="cuda")[:, None, None]
    venv.set_state(w0.contiguous())
    acts = torch.zeros(3, 4, dtype=torch.float64)
    _, _, term, trunc, _ = venv.step(acts)
    assert not bool(trunc.any()) and not bool(term.any()) and venv.time.tolist() == [1, 1, 1]
    obs, reward, term, trunc, info = venv.step(acts)
    assert bool(trunc.all()) and venv.time.tolist() == [0, 0, 0]
    for b in range(3):
        assert torch.equal(venv.omega_hat[b], ref_state.omega_hat)
        assert torch.equal(venv.trajectory[b], ref_state.trajectory)
        assert torch.equal(obs[b], ref_obs)
    assert not torch.equal(info["final_observation"][1], ref_obs)  # the step's own observation is kept in info
    venv.close()
