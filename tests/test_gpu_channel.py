"""3-D channel-flow step on the GPU (csrc/channel.cu[h] through the C ABI of include/channel_b200.h) against

* the reference's own source: tests/golden/reference_shim_channel.npz (fp64 run of the unmodified channel.py /
  RungeKutta4 over NumPy stand-ins; see tests/golden/make_reference_channel.py), and
* the NumPy oracle (oracle/channel_oracle.py) on cases the fixture does not hold (other times, batches).

Tolerances: fp64 1e-9 relative on fields after up to 100 solver steps (the kernels evaluate the linear term, the
projection and the wall conditions spectrally -- same algebra, different rounding order); fp32 2e-4."""
import os

import numpy as np
import pytest
import torch

from oracle import channel_oracle as orc

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
SUB = (slice(None, None, 6), slice(None, None, 6), slice(None))
LOW = np.r_[0:6, 67:72]


def sub(a):
    a = np.asarray(a)
    return a[np.ix_(LOW, LOW, np.arange(72))] if np.iscomplexobj(a) else a[SUB]


def rel(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(np.asarray(b)), 1e-300))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "reference_shim_channel.npz"))


@pytest.fixture(scope="module")
def fields():
    return orc.synthetic_initial_fields(np.float64)


def make_stepper(B=1, dtype=torch.float64, jets_on=True):
    from hydrogym_b200.envs.channel import ChannelStepper
    return ChannelStepper(num_envs=B, dtype=dtype, device="cuda:0", jets_on=jets_on)


def loaded(fields, B=1, dtype=torch.float64, jets_on=True):
    st = make_stepper(B, dtype, jets_on)
    U, V, W = (torch.from_numpy(np.broadcast_to(f, (B,) + f.shape).copy()) for f in fields)
    st.set_state(U, V, W)
    st.to_spectral()
    return st


def test_transforms_round_trip(fields):
    """to_spectral == fft2 over (x, y) of each z-plane (C:150-161); to_physical inverts it (C:153-168)."""
    st = loaded(fields)
    s = st.spectral().cpu().numpy()[:, 0]
    for got, f in zip(s, fields):
        assert rel(got, np.fft.fftn(f, axes=(0, 1))) < 1e-13
    st.to_physical()
    for got, f in zip(st.get_state(), fields):
        assert rel(got[0].cpu().numpy(), f) < 1e-13
    st.close()


def test_rhs_matches_reference_source(gold, fields):
    """rhs = nonlinear_terms + linear_terms (C:243-309) of the initial state with the test action at t = 3."""
    st = loaded(fields)
    k = st.rhs(torch.from_numpy(gold["action"]), 3.0).cpu().numpy()[:, 0]
    for c, got in zip("uvw", k):
        want_sub = gold[f"N{c}_sub"] + gold[f"L{c}_sub"]
        assert rel(sub(got), want_sub) < 1e-10, c
    st.close()


def test_rhs_matches_oracle_everywhere(fields):
    """Every mode of the right-hand side, at a time inside the closing ramp of the jets (t = 44.5)."""
    env = orc.ChannelFlowSpectralEnv(initial_fields=fields)
    eq = env.equation
    act = np.clip(0.8 * np.cos(0.7 * np.arange(24)), -1, 1)
    s0 = eq.to_spectral(orc.VelocityState(*fields))
    want = eq.rhs(s0, action=act, t=44.5, fx=2.0)
    st = loaded(fields)
    k = st.rhs(torch.from_numpy(act), 44.5).cpu().numpy()[:, 0]
    for got, w in zip(k, (want.u, want.v, want.w)):
        assert rel(got, w) < 1e-11
    st.close()


@pytest.mark.parametrize("tag", ["act", "noact"])
def test_rk4_steps_match_reference_source(gold, fields, tag):
    """RungeKutta4.rk4_step x 2 (S:126-185: 4 rhs, projection C:311-372, mass-flux shift), with and without jets."""
    st = loaded(fields, jets_on=(tag == "act"))
    act = torch.from_numpy(np.clip(gold["action"], -1, 1)) if tag == "act" else None
    for n in range(2):
        st.rk4_steps(act, float(n), n=1, dt=orc.DT)
        s = st.spectral().cpu().numpy()[:, 0]
        for c, got in zip("uvw", s):
            assert rel(sub(got), gold[f"rk{n + 1}_{tag}_{c}_sub"]) < 1e-10, (n, c)
            nrm = np.sqrt((np.abs(got) ** 2).sum())
            assert abs(nrm - gold[f"rk{n + 1}_{tag}_{c}_norm"]) < 1e-10 * gold[f"rk{n + 1}_{tag}_{c}_norm"]
    st.close()


def test_step_env_matches_reference_source(gold, fields):
    """Two consecutive `step_env` calls (2 x 50 solver steps; the second at t >= 50, negative gain ramp) through the
    env class: observation, reward, time, done and the fields against the reference's own run."""
    from hydrogym_b200.envs.channel import ChannelFlowSpectralEnv
    env = ChannelFlowSpectralEnv({"initial_fields": fields, "dtype": "float64", "device": "cuda:0"})
    obs, state = env.reset_env()
    assert np.array_equal(obs.cpu().numpy(), gold["obs0"])
    act = torch.from_numpy(gold["action"])   # holds two out-of-range entries: clipped by step_env (C:493)
    for i in (1, 2):
        obs, state, reward, done, info = env.step_env(None, state, act)
        assert int(state.time) == int(gold[f"step{i}_time"]) and bool(done) == bool(gold[f"step{i}_done"])
        assert float(info["discount"]) == float(gold[f"step{i}_discount"])
        assert abs(float(reward) - float(gold[f"step{i}_reward"])) < 1e-9 * abs(float(gold[f"step{i}_reward"]))
        assert rel(obs.cpu().numpy(), gold[f"step{i}_obs"]) < 1e-9
        for c, f in zip("UVW", (state.U, state.V, state.W)):
            f = f.cpu().numpy()
            assert rel(sub(f), gold[f"step{i}_{c}_sub"]) < 1e-9, (i, c)
            assert abs(np.linalg.norm(f) - gold[f"step{i}_{c}_norm"]) < 1e-9 * gold[f"step{i}_{c}_norm"]
        assert rel(state.U.cpu().numpy()[:, :, 4], gold[f"step{i}_U_plane4"]) < 1e-9
        # properties of the scheme that hold at any size: walls exactly at rest, grid-mean streamwise velocity on the
        # target (S:166-183), the five jet planes of w carry one (x, y) pattern with zero mean (C:194-225)
        U, V, W = (f.cpu().numpy() for f in (state.U, state.V, state.W))
        for f in (U, V, W):
            assert not f[:, :, 0].any() and not f[:, :, -1].any()
        # the shift is computed before the walls are zeroed again, so 2/72 of each step's (small) shift is lost
        assert abs(U.mean() - 8.0) < 1e-3
        for k in range(2, 6):
            assert np.abs(W[:, :, k] - W[:, :, 1]).max() < 1e-12 * max(1.0, np.abs(W[:, :, 1]).max())
        assert abs(W[:, :, 1].mean()) < 1e-12
    env.close()


def test_fp32_step_close_to_fp64(gold, fields):
    """The reference's default precision (env_config["dtype"] = "float32"): one env step against the fp64 record."""
    from hydrogym_b200.envs.channel import ChannelFlowSpectralEnv
    env = ChannelFlowSpectralEnv({"initial_fields": fields, "device": "cuda:0"})
    obs, state = env.reset_env()
    obs, state, reward, done, info = env.step_env(None, state, torch.from_numpy(gold["action"]))
    assert obs.dtype == torch.float32 and state.U.dtype == torch.float32
    assert rel(obs.cpu().numpy(), gold["step1_obs"]) < 2e-4
    assert abs(float(reward) - float(gold["step1_reward"])) < 2e-3 * abs(float(gold["step1_reward"]))
    assert rel(sub(state.U.cpu().numpy()), gold["step1_U_sub"]) < 2e-4
    assert rel(sub(state.W.cpu().numpy()), gold["step1_W_sub"]) < 5e-3   # |W| << |U|: absolute error ~ eps * |U|
    env.close()


def test_batched_envs_are_independent(fields):
    """B = 3 environments with different actions and clocks in one launch == three single runs (C: jax.vmap)."""
    rng = np.random.default_rng(5)
    acts = rng.uniform(-1.3, 1.3, (3, 24))
    times = np.array([0.0, 7.0, 43.0])
    stB = make_stepper(3)
    U, V, W = (torch.from_numpy(np.stack([f * (1 + 0.01 * b) for b in range(3)])) for f in fields)
    stB.set_state(U, V, W)
    obsB, rewB = stB.step(torch.from_numpy(acts), times, nsteps=3, dt=orc.DT)
    UB = stB.get_state()[0].cpu().numpy()
    for b in range(3):
        st1 = make_stepper(1)
        st1.set_state(U[b], V[b], W[b])
        obs1, rew1 = st1.step(torch.from_numpy(acts[b]), times[b], nsteps=3, dt=orc.DT)
        assert np.array_equal(obs1[0].cpu().numpy(), obsB[b].cpu().numpy())
        assert float(rew1[0]) == float(rewB[b])
        assert np.array_equal(st1.get_state()[0][0].cpu().numpy(), UB[b])
        st1.close()
    stB.close()


def test_short_rollout_matches_oracle_at_late_times(fields):
    """4 solver steps starting at t = 38 (the ramp-down branch switches on inside the rollout) against the oracle."""
    env = orc.ChannelFlowSpectralEnv(initial_fields=fields, nsteps=4)
    act = 0.9 * np.sin(0.9 * np.arange(24) + 1.0)
    state = dict(time=38, U=fields[0], V=fields[1], W=fields[2], dt=orc.DT, terminal=False)
    wss, (U1, V1, W1) = env.rollout(state, act)
    st = make_stepper(1)
    st.set_state(*[torch.from_numpy(f) for f in fields])
    obs, rew = st.step(torch.from_numpy(act), 38.0, nsteps=4, dt=orc.DT)
    got = [f[0].cpu().numpy() for f in st.get_state()]
    for g, w in zip(got, (U1, V1, W1)):
        assert np.linalg.norm(g - w) < 1e-11 * np.linalg.norm(U1)
    assert abs(float(rew[0]) + wss) < 1e-11 * abs(wss)
    nxt = dict(U=U1, W=W1)
    assert rel(obs[0].cpu().numpy(), env.get_obs(nxt)) < 1e-11
    st.close()


def test_env_api_and_auto_reset(fields):
    """Spaces, names, batched reset, terminal at the horizon and gymnax auto-reset (C:459-533)."""
    from hydrogym_b200.envs.channel import ChannelEnvParams, ChannelFlowSpectralEnv
    env = ChannelFlowSpectralEnv({"initial_fields": fields, "dtype": "float64", "device": "cuda:0"})
    assert env.name == "ChannelFlowSpectralEnv" and env.num_actions == 24
    assert env.action_space().shape == (24,) and env.observation_space().shape == (128,)
    params = ChannelEnvParams(max_steps_in_episode=50)
    obs, state = env.reset(None, params, num_envs=2)
    assert obs.shape == (2, 128) and state.U.shape == (2, 72, 72, 72)
    obs1, state1, reward, done, info = env.step(None, state, torch.zeros(2, 24, dtype=torch.float64), params)
    assert bool(done.all()) and float(info["discount"].sum()) == 0.0
    assert int(state1.time.sum()) == 0 and torch.equal(state1.U[0], env.U0)   # auto-reset state
    assert torch.equal(obs1, obs)
    assert reward.shape == (2,) and float(reward[0]) == float(reward[1]) and float(reward[0]) < 0
    with pytest.raises(FileNotFoundError):
        ChannelFlowSpectralEnv({"device": "cuda:0"})
    env.close()


def test_graph_replay_is_bit_identical_to_direct_launches(fields):
    """kc_step replays one CUDA graph of its launches (on an internal stream ordered by events) when the step count
    repeats; KC_GRAPH=0 launches directly.  Same kernels, same arguments: identical bits, also on the second call
    (graph reuse with new actions and clocks) and after a change of the step count (re-capture)."""
    rng = np.random.default_rng(11)
    acts = [torch.from_numpy(rng.uniform(-1, 1, (2, 24))) for _ in range(3)]
    out = {}
    old = os.environ.get("KC_GRAPH")
    try:
        for mode in ("0", "1"):
            os.environ["KC_GRAPH"] = mode
            st = make_stepper(2)
            U, V, W = (torch.from_numpy(np.stack([f, 1.02 * f])) for f in fields)
            st.set_state(U, V, W)
            res = []
            for a, t, n in zip(acts, ([0.0, 3.0], [5.0, 42.0], [47.0, 9.0]), (3, 3, 2)):
                obs, rew = st.step(a, np.array(t), nsteps=n, dt=orc.DT)
                res.append((obs.cpu().numpy(), rew.cpu().numpy(), st.get_state()[2].cpu().numpy()))
            out[mode] = res
            st.close()
    finally:
        if old is None:
            os.environ.pop("KC_GRAPH", None)
        else:
            os.environ["KC_GRAPH"] = old
    for a, b in zip(out["0"], out["1"]):
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert np.isfinite(out["1"][-1][2]).all()


def test_run_channel_pseudospectral_final_and_trajectory(fields):
    """C:375-403: the direct-solve helper returns the final fields, or the per-step trajectory whose last entry is
    that final state; both equal the stepper's own rollout."""
    from hydrogym_b200 import run_channel_pseudospectral
    act = torch.from_numpy(0.5 * np.cos(np.arange(24.0)))
    U, V, W = (torch.from_numpy(f) for f in fields)
    fin = run_channel_pseudospectral(U, V, W, act, dt=orc.DT, nsteps=3, device="cuda:0")
    traj = run_channel_pseudospectral(U, V, W, act, dt=orc.DT, nsteps=3, return_trajectory=True, device="cuda:0")
    assert traj[0].shape == (3, 72, 72, 72)
    for f, t in zip(fin, traj):
        assert torch.equal(f, t[-1])
    st = make_stepper(1)
    st.set_state(U, V, W)
    st.step(act, 0.0, nsteps=3, dt=orc.DT)
    for f, g in zip(fin, st.get_state()):
        assert torch.equal(f, g[0])
    st.close()


def test_reference_ppo_wrapper_stack_over_the_channel_env(fields):
    """run_ppo.py:165-172's nesting (LogWrapper, ClipAction, VecEnv, NormalizeVecObservation, NormalizeVecReward) over
    the channel env: the wrappers' arithmetic (NumPy restatement of env_core.py:449-742) applied to this env's own raw
    outputs."""
    from hydrogym_b200 import wrappers as W
    from hydrogym_b200.envs.channel import ChannelFlowSpectralEnv
    from oracle import wrappers_oracle as wo
    B, gamma = 2, 0.99
    cfg = {"initial_fields": fields, "dtype": "float64", "device": "cuda:0"}
    env, raw_env = ChannelFlowSpectralEnv(cfg), ChannelFlowSpectralEnv(cfg)
    params = env.default_params
    w = W.NormalizeVecReward(W.NormalizeVecObservation(W.VecEnv(W.ClipAction(W.LogWrapper(env)), num_envs=B)), gamma)
    assert w.name == "ChannelFlowSpectralEnv" and w.num_actions == 24
    obs, st = w.reset(None, params)
    robs, rstate = raw_env.reset(None, params, num_envs=B)
    nobs, rno = wo.normalize_obs_reset(robs.cpu().numpy())
    assert np.abs(obs.cpu().numpy() - nobs).max() < 1e-7
    rrw = wo.normalize_reward_reset(B)
    acts = torch.from_numpy(np.stack([1.4 * np.sin(np.arange(24.0)), -0.3 * np.ones(24)])).cuda()
    obs, st, r, d, info = w.step(None, st, acts, params)
    raw_obs, rstate, raw_rew, raw_done, _ = raw_env.step(None, rstate, torch.clamp(acts, -1.0, 1.0), params)
    nobs, rno = wo.normalize_obs_step(rno, raw_obs.cpu().numpy())
    nrew, rrw = wo.normalize_reward_step(rrw, raw_rew.cpu().numpy(), raw_done.cpu().numpy().astype(np.float64), gamma)
    assert np.abs(obs.cpu().numpy() - nobs).max() < 1e-6 * max(1.0, np.abs(nobs).max())
    assert np.abs(r.cpu().numpy() - nrew).max() < 1e-8 * max(1.0, np.abs(nrew).max())
    assert int(info["timestep"][0]) == 1 and not bool(d.any())
    env.close()
    raw_env.close()
