"""Gymnasium ``VectorEnv``-shaped façade over the batched Kolmogorov step.

``gymnasium`` is not installable in this image, so the class is duck-typed
(``reset(seed, options)``, ``step(actions)``, ``single_*_space``, ``num_envs``,
``close``); it follows the 5-tuple convention of HydroGym's own Gymnasium env
(hydrogym/core.py:376-568: ``terminated`` stays False, the time limit truncates) and
the termination rule of the JAX env (hydrogym/jax/env_core.py:422-423).  Environments
that hit the time limit are auto-reset to the cached post-reset state, like gymnax's
``Environment.step`` does for the reference.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from .envs.kolmogorov import KolmogorovFlow
from .spaces import Box


class KolmogorovVectorEnv:
    def __init__(self, num_envs: int, env_config: Optional[dict] = None, flow_config: Optional[dict] = None,
                 dtype=None, device=None, keep_trajectory: bool = False):
        env_config = dict(env_config or {})
        if dtype is not None:
            env_config["dtype"] = dtype
        if device is not None:
            env_config["device"] = device
        self.env = KolmogorovFlow(env_config, flow_config)
        self.params = self.env.default_params
        self.num_envs = int(num_envs)
        self.keep_trajectory = keep_trajectory
        p = self.params
        self.single_action_space = Box(p.min_action, p.max_action, (p.action_dim,))
        self.single_observation_space = Box(p.min_obs, p.max_obs, (p.obs_dim,))
        self.action_space = Box(p.min_action, p.max_action, (self.num_envs, p.action_dim))
        self.observation_space = Box(p.min_obs, p.max_obs, (self.num_envs, p.obs_dim))
        self._stepper = self.env._stepper(self.num_envs, self.params)
        self.omega_hat = None
        self.trajectory = None
        self.time = None

    @property
    def device(self):
        return self._stepper.device

    def set_state(self, omega_hat: torch.Tensor):
        """Inject initial conditions (benchmark input; not a reference feature)."""
        self._stepper._check_state(omega_hat)
        self.omega_hat = omega_hat.clone()
        self.time = torch.zeros((self.num_envs,), dtype=torch.int64, device=self.device)
        self._cache_reset()

    def _cache_reset(self):
        """Post-reset (obs, omega_hat, trajectory) of one environment -- what time-limit truncation resets to."""
        obs, final, traj = self.env._reset_single(self.params)
        self._reset_obs, self._reset_omega, self._reset_traj = obs, final, traj

    def reset(self, seed=None, options=None):
        obs, state = self.env.reset_env(seed, self.params, num_envs=self.num_envs)
        self.omega_hat, self.time = state.omega_hat, state.time
        self.trajectory = state.trajectory if self.keep_trajectory else None
        self._cache_reset()
        return obs, {}

    def step(self, actions):
        if self.omega_hat is None:
            raise RuntimeError("call reset() (or set_state()) before step()")
        p = self.params
        actions = torch.as_tensor(actions, dtype=self.env.dtype, device=self.device)
        actions = torch.clamp(actions, p.min_action, p.max_action)
        final, traj, obs, tke, nonf = self._stepper.step(self.omega_hat, actions,
                                                         want_trajectory=self.keep_trajectory, in_place=True)
        self.trajectory = traj
        self.time = self.time + 1
        reward = -(p.reward_alpha * tke + actions.abs().sum(dim=-1))
        truncated = self.time >= p.max_episode_steps
        terminated = torch.zeros_like(truncated)
        # gymnax-style auto-reset, leaf-wise select without a host sync: truncated environments take
        # the cached post-reset state, observation and trajectory; info["final_observation"] keeps the step's own
        info = {"mean_tke": tke, "nonfinite": nonf, "final_observation": obs}
        sel = truncated[:, None, None]
        self.omega_hat = torch.where(sel, self._reset_omega.expand_as(self.omega_hat), self.omega_hat)
        if self.trajectory is not None:
            self.trajectory = torch.where(sel[..., None], self._reset_traj.expand_as(self.trajectory), self.trajectory)
        obs = torch.where(truncated[:, None], self._reset_obs.expand_as(obs), obs)
        self.time = torch.where(truncated, torch.zeros_like(self.time), self.time)
        return obs, reward, terminated, truncated, info

    def close(self):
        self._stepper.close()
