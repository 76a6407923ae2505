"""Wrapper stack (hydrogym_b200/wrappers.py) vs the NumPy restatement of env_core.py:449-742, driven by a scripted
dummy env on CPU tensors (the wrappers only see reset/step tuples, so no GPU is needed)."""
import numpy as np
import torch

from hydrogym_b200 import wrappers as W
from hydrogym_b200.spaces import Box
from oracle import wrappers_oracle as wo


class ScriptedEnv:
    """Batched env that replays pre-drawn obs / reward / done sequences; state = step counter."""

    def __init__(self, B, T, seed=0, obs_shape=(2, 3)):
        rng = np.random.default_rng(seed)
        self.obs = rng.normal(size=(T + 1, B) + obs_shape) * 3 + 1
        self.rew = rng.normal(size=(T + 1, B))
        self.done = rng.uniform(size=(T + 1, B)) < 0.3
        self.obs_shape = obs_shape
        self.last_action = None
        self.extra = "proxied"

    def observation_space(self, params=None):
        return Box(-np.inf, np.inf, self.obs_shape)

    def reset(self, key=None, params=None, num_envs=None):
        assert num_envs in (None, self.obs.shape[1])
        return torch.from_numpy(self.obs[0]), 0

    def step(self, key, state, action, params=None):
        t = state + 1
        self.last_action = action
        return (torch.from_numpy(self.obs[t]), t, torch.from_numpy(self.rew[t]), torch.from_numpy(self.done[t]),
                {"mean_tke": torch.zeros(self.obs.shape[1])})


def test_attribute_proxy_and_clip_and_transforms():
    env = ScriptedEnv(4, 3)
    w = W.TransformReward(W.TransformObservation(W.ClipAction(env, -0.25, 0.5), lambda o: o * 2), lambda r: r - 1)
    assert w.extra == "proxied" and w.obs_shape == (2, 3)
    obs, st = w.reset(None, None)
    assert torch.equal(obs, torch.from_numpy(env.obs[0]) * 2)
    obs, st, r, d, info = w.step(None, st, torch.tensor([[-1.0, 0.3, 0.7, 0.0]] * 4), None)
    assert torch.equal(env.last_action, torch.tensor([[-0.25, 0.3, 0.5, 0.0]] * 4))
    assert torch.equal(obs, torch.from_numpy(env.obs[1]) * 2) and torch.equal(r, torch.from_numpy(env.rew[1]) - 1)


def test_flatten_observation():
    env = ScriptedEnv(5, 2)
    w = W.FlattenObservationWrapper(env)
    assert w.observation_space(None).shape == (6,)
    obs, st = w.reset(None, None)
    assert obs.shape == (5, 6)
    obs, st, *_ = w.step(None, st, None, None)
    assert torch.equal(obs, torch.from_numpy(env.obs[1]).reshape(5, 6))


def test_vec_env_sizes_batch_from_keys_or_argument():
    env = ScriptedEnv(7, 1)
    assert W.VecEnv(env, num_envs=7).reset(None, None)[0].shape[0] == 7
    assert W.VecEnv(env).reset(torch.zeros(7, 2), None)[0].shape[0] == 7


def test_log_wrapper_matches_reference_arithmetic():
    B, T = 6, 25
    env = ScriptedEnv(B, T, seed=3)
    w = W.LogWrapper(env)
    obs, st = w.reset(None, None)
    ref = dict(episode_returns=np.zeros(B), episode_lengths=np.zeros(B, dtype=np.int64),
               returned_episode_returns=np.zeros(B), returned_episode_lengths=np.zeros(B, dtype=np.int64),
               timestep=np.zeros(B, dtype=np.int64))
    for t in range(1, T + 1):
        obs, st, r, d, info = w.step(None, st, None, None)
        rinfo, ref = wo.log_wrapper_step(ref, env.rew[t], env.done[t].astype(np.int64))
        for k in ("returned_episode_returns", "returned_episode_lengths", "timestep"):
            assert np.allclose(info[k].numpy(), rinfo[k], rtol=0, atol=1e-12), (t, k)
        assert np.array_equal(info["returned_episode"].numpy(), env.done[t])
        assert np.allclose(st.episode_returns.numpy(), ref["episode_returns"], atol=1e-12)
        assert np.array_equal(st.episode_lengths.numpy(), ref["episode_lengths"])
        assert "mean_tke" in info and st.env_state == t


def test_normalize_observation_and_reward_match_reference_arithmetic():
    B, T, gamma = 8, 20, 0.99
    env = ScriptedEnv(B, T, seed=5, obs_shape=(4,))
    w = W.NormalizeVecReward(W.NormalizeVecObservation(W.VecEnv(env, num_envs=B)), gamma)
    obs, st = w.reset(None, None)
    robs, rst = wo.normalize_obs_reset(env.obs[0])
    rrew = wo.normalize_reward_reset(B)
    assert np.allclose(obs.numpy(), robs, rtol=1e-12, atol=1e-12)
    for t in range(1, T + 1):
        obs, st, r, d, info = w.step(None, st, None, None)
        robs, rst = wo.normalize_obs_step(rst, env.obs[t])
        rr, rrew = wo.normalize_reward_step(rrew, env.rew[t], env.done[t].astype(np.float64), gamma)
        assert np.allclose(obs.numpy(), robs, rtol=1e-11, atol=1e-12), t
        assert np.allclose(r.numpy(), rr, rtol=1e-11, atol=1e-12), t
        assert np.allclose(st.env_state.mean.numpy(), rst["mean"], rtol=1e-11, atol=1e-12)
        assert np.allclose(st.return_val.numpy(), rrew["return_val"], rtol=1e-11, atol=1e-12)
        assert abs(float(st.count) - rrew["count"]) < 1e-9


def test_log_callback_and_trajectory_roundtrip(tmp_path):
    from hydrogym_b200 import io

    class Flow:
        vorticity = torch.arange(24, dtype=torch.float64).reshape(2, 3, 4).to(torch.complex128) * (1 + 2j)

    path = tmp_path / "log.txt"
    cb = io.LogCallback(postprocess=lambda f: f.vorticity.abs().sum(dim=(1, 2)), filename=str(path))
    vals = cb(Flow())
    assert np.allclose(np.loadtxt(path), np.round(vals, 5))
    npy = tmp_path / "traj.npy"
    io.save_trajectory(str(npy), Flow.vorticity)
    back = io.load_trajectory(str(npy))
    assert back.dtype == torch.complex128 and torch.equal(back, Flow.vorticity)
    assert io.CallbackBase(interval=3)(6, 0.0, None) and not io.CallbackBase(interval=3)(7, 0.0, None)
