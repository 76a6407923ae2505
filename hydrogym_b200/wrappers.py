"""Gymnax-style wrapper stack over torch tensors (reference: hydrogym/jax/env_core.py:449-742, itself adapted
from purejaxrl).  The wrappers compose purely through ``reset(key, params)`` / ``step(key, state, action, params)``
and proxy every other attribute to the wrapped env (E:456-457), so the reference's PPO loop
(examples/jax/getting_started/3_ppo/run_ppo.py:165-172) drives the B200 env with the same nesting:

    env = LogWrapper(KolmogorovFlow(...)); env = ClipAction(env); env = VecEnv(env, num_envs=B)
    env = NormalizeVecObservation(env); env = NormalizeVecReward(env, gamma)

Everything here is a handful of elementwise / reduction torch ops per step on (B, 64) observations and (B,)
rewards that stay on the GPU next to the env state; the running moments use the parallel-merge form of Welford's
update written once in ``merge_moments``.
"""
from __future__ import annotations

import dataclasses
from typing import Any, Callable, Optional

import numpy as np
import torch

from .spaces import Box


class GymnaxWrapper:
    """E:449-457: holds ``_env`` and forwards unknown attributes to it."""

    def __init__(self, env):
        self._env = env

    def __getattr__(self, name):
        if name == "_env":  # not set yet (copy / pickle): avoid infinite recursion
            raise AttributeError(name)
        return getattr(self._env, name)


# -- observation / reward / action transforms ------------------------------------------------------------------

class FlattenObservationWrapper(GymnaxWrapper):
    """E:460-495.  Observations keep their leading batch axes (if the state is batched) and lose the rest."""

    def observation_space(self, params=None):
        inner = self._env.observation_space(params)
        if not isinstance(inner, Box):
            raise TypeError("only Box observation spaces are supported")
        return Box(low=inner.low, high=inner.high, shape=(int(np.prod(inner.shape)),), dtype=inner.dtype)

    def _flat(self, obs, params):
        keep = obs.dim() - len(self._env.observation_space(params).shape)
        return obs.reshape(tuple(obs.shape[:keep]) + (-1,))

    def reset(self, key=None, params=None, **kw):
        obs, state = self._env.reset(key, params, **kw)
        return self._flat(obs, params), state

    def step(self, key, state, action, params=None):
        obs, state, reward, done, info = self._env.step(key, state, action, params)
        return self._flat(obs, params), state, reward, done, info


class ClipAction(GymnaxWrapper):
    """E:583-593: clip to a FIXED [low, high] (default [-1, 1]) before the env's own clip."""

    def __init__(self, env, low=-1.0, high=1.0):
        super().__init__(env)
        self.low, self.high = low, high

    def step(self, key, state, action, params=None):
        return self._env.step(key, state, torch.clamp(torch.as_tensor(action), self.low, self.high), params)


class TransformObservation(GymnaxWrapper):
    """E:596-607."""

    def __init__(self, env, transform_obs: Callable):
        super().__init__(env)
        self.transform_obs = transform_obs

    def reset(self, key=None, params=None, **kw):
        obs, state = self._env.reset(key, params, **kw)
        return self.transform_obs(obs), state

    def step(self, key, state, action, params=None):
        obs, state, reward, done, info = self._env.step(key, state, action, params)
        return self.transform_obs(obs), state, reward, done, info


class TransformReward(GymnaxWrapper):
    """E:610-617."""

    def __init__(self, env, transform_reward: Callable):
        super().__init__(env)
        self.transform_reward = transform_reward

    def step(self, key, state, action, params=None):
        obs, state, reward, done, info = self._env.step(key, state, action, params)
        return obs, state, self.transform_reward(reward), done, info


# -- episode statistics ---------------------------------------------------------------------------------------

@dataclasses.dataclass
class LogEnvState:
    """E:498-505."""
    env_state: Any
    episode_returns: Any
    episode_lengths: Any
    returned_episode_returns: Any
    returned_episode_lengths: Any
    timestep: Any


class LogWrapper(GymnaxWrapper):
    """E:508-546: running episode return / length, latched into ``returned_*`` when an episode ends."""

    def reset(self, key=None, params=None, **kw):
        obs, env_state = self._env.reset(key, params, **kw)
        feat = len(self._env.observation_space(params).shape)
        lead = obs.shape[:obs.dim() - feat]  # () for one env, (B,) for a batch
        zf = torch.zeros(lead, dtype=obs.dtype, device=obs.device)
        zi = torch.zeros(lead, dtype=torch.int64, device=obs.device)
        return obs, LogEnvState(env_state, zf, zi, zf.clone(), zi.clone(), zi.clone())

    def step(self, key, state: LogEnvState, action, params=None):
        obs, env_state, reward, done, info = self._env.step(key, state.env_state, action, params)
        ret = state.episode_returns + reward
        length = state.episode_lengths + 1
        alive_f, alive_i = (~done).to(ret.dtype), (~done).to(length.dtype)
        ended_f, ended_i = done.to(ret.dtype), done.to(length.dtype)
        new = LogEnvState(
            env_state=env_state,
            episode_returns=ret * alive_f,
            episode_lengths=length * alive_i,
            returned_episode_returns=state.returned_episode_returns * alive_f + ret * ended_f,
            returned_episode_lengths=state.returned_episode_lengths * alive_i + length * ended_i,
            timestep=state.timestep + 1)
        info = dict(info)
        info["returned_episode_returns"] = new.returned_episode_returns
        info["returned_episode_lengths"] = new.returned_episode_lengths
        info["timestep"] = new.timestep
        info["returned_episode"] = done
        return obs, new, reward, done, info


# -- batching -------------------------------------------------------------------------------------------------

class VecEnv(GymnaxWrapper):
    """E:609-613.  The reference vmaps ``reset`` over a batch of keys and ``step`` over (keys, state, action);
    here the env is natively batched, so this wrapper only fixes the batch size: taken from ``num_envs`` or from
    the leading axis of the ``key`` batch the caller passes (keys are otherwise unused -- the reference's reset
    ignores them too, K:472-492)."""

    def __init__(self, env, num_envs: Optional[int] = None):
        super().__init__(env)
        self.num_envs = num_envs

    def _batch(self, key):
        if self.num_envs is not None:
            return int(self.num_envs)
        if key is not None and hasattr(key, "shape") and len(key.shape) >= 1:
            return int(key.shape[0])
        raise ValueError("VecEnv needs num_envs or a batch of keys to size the batch")

    def reset(self, key=None, params=None):
        return self._env.reset(key, params, num_envs=self._batch(key))

    def step(self, key, state, action, params=None):
        return self._env.step(key, state, action, params)


# -- running normalisation ------------------------------------------------------------------------------------

def merge_moments(mean, var, count, batch_mean, batch_var, batch_count):
    """Merge (mean, population variance, count) of a running stream with those of a new batch (parallel Welford;
    the update the reference spells out inline at E:636-655, 662-681, 721-741).  Returns the merged triple."""
    delta = batch_mean - mean
    total = count + batch_count
    new_mean = mean + delta * (batch_count / total)
    m2 = var * count + batch_var * batch_count + delta * delta * (count * batch_count / total)
    return new_mean, m2 / total, total


@dataclasses.dataclass
class NormalizeVecObsEnvState:
    """E:616-621."""
    mean: Any
    var: Any
    count: Any
    env_state: Any


class NormalizeVecObservation(GymnaxWrapper):
    """E:624-692: normalise (B, ...) observations by running per-feature moments over the batch axis."""

    eps = 1e-8

    def _update(self, obs, mean, var, count, env_state):
        mean, var, count = merge_moments(mean, var, count, obs.mean(dim=0), obs.var(dim=0, unbiased=False),
                                         obs.shape[0])
        state = NormalizeVecObsEnvState(mean=mean, var=var, count=count, env_state=env_state)
        return (obs - mean) / torch.sqrt(var + self.eps), state

    def reset(self, key=None, params=None):
        obs, env_state = self._env.reset(key, params)
        return self._update(obs, torch.zeros_like(obs), torch.ones_like(obs), 1e-4, env_state)

    def step(self, key, state: NormalizeVecObsEnvState, action, params=None):
        obs, env_state, reward, done, info = self._env.step(key, state.env_state, action, params)
        nobs, new = self._update(obs, state.mean, state.var, state.count, env_state)
        return nobs, new, reward, done, info


@dataclasses.dataclass
class NormalizeVecRewEnvState:
    """E:695-701."""
    mean: Any
    var: Any
    count: Any
    return_val: Any
    env_state: Any


class NormalizeVecReward(GymnaxWrapper):
    """E:704-742: scale rewards by the running std of the discounted return."""

    eps = 1e-8

    def __init__(self, env, gamma):
        super().__init__(env)
        self.gamma = gamma

    def reset(self, key=None, params=None):
        obs, env_state = self._env.reset(key, params)
        ret = torch.zeros((obs.shape[0],), dtype=obs.dtype, device=obs.device)
        return obs, NormalizeVecRewEnvState(mean=0.0, var=1.0, count=1e-4, return_val=ret, env_state=env_state)

    def step(self, key, state: NormalizeVecRewEnvState, action, params=None):
        obs, env_state, reward, done, info = self._env.step(key, state.env_state, action, params)
        ret = state.return_val * self.gamma * (~done).to(reward.dtype) + reward
        mean, var, count = merge_moments(state.mean, state.var, state.count, ret.mean(dim=0),
                                         ret.var(dim=0, unbiased=False), obs.shape[0])
        new = NormalizeVecRewEnvState(mean=mean, var=var, count=count, return_val=ret, env_state=env_state)
        return obs, new, reward / torch.sqrt(var + self.eps), done, info
