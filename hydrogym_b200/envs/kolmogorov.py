"""Kolmogorov-flow environment behind the reference's gymnax-style functional API.

Mirror of ``hydrogym/jax/envs/kolmogorov.py`` (reference lines cited per method) with
torch tensors in place of jax arrays and the solve executed by the sm_100a kernels of
``libkolmo_b200.so``.  Differences a caller can see:

* ``key`` arguments are accepted and ignored (the reference ignores them too, K:472-492).
* every state leaf may carry a leading batch axis: ``step_env`` on a state whose
  ``omega_hat`` is (B, N, Mh) is what ``jax.vmap(env.step_env)`` gives in the reference.
* build-only ``env_config`` keys: ``dtype`` ("float32"/"float64", default float32 like
  JAX), ``device``, ``dealias`` (False = reference behaviour, the 2/3 mask is a no-op
  there), and ``action_time`` / ``save_time`` (the reference pins them to 10.0 / 1).
"""
from __future__ import annotations

import dataclasses
from typing import Any, Dict, Optional

import numpy as np
import torch

from .. import constants
from ..solver import KolmogorovStepper
from ..spaces import Box


class FlowConfig:
    """Grids and constants of the flow (reference K:24-183)."""

    DEFAULT_REYNOLDS = 200
    DEFAULT_WAVENUMBER = 4
    DEFAULT_GRID_SIZE = (64, 64)
    DEFAULT_DOMAIN_X = (0, 2 * np.pi)
    DEFAULT_DOMAIN_Y = (0, 2 * np.pi)
    DEFAULT_OBS_SIZE = 8

    def __init__(self, **config):
        self.k = config.get("k", self.DEFAULT_WAVENUMBER)
        self.Re = config.get("Re", self.DEFAULT_REYNOLDS)
        self.grid_size = tuple(config.get("grid_size", self.DEFAULT_GRID_SIZE))
        self.domain_x = config.get("domain_x", self.DEFAULT_DOMAIN_X)
        self.domain_y = config.get("domain_y", self.DEFAULT_DOMAIN_Y)
        self.obs_size = config.get("obs_size", self.DEFAULT_OBS_SIZE)
        if self.grid_size[0] != self.grid_size[1]:
            raise ValueError("hydrogym_b200 supports square grids only")
        constants.check_domain(self.domain_x, self.domain_y)
        if self.grid_size[0] % self.obs_size != 0:
            raise ValueError("obs_size must divide the grid size")
        self.vorticity = None

    @property
    def nu(self):  # K:152-154
        return 1 / self.Re

    @property
    def num_inputs(self):  # K:156-159
        return 2

    def load_mesh(self, name=""):  # K:46-57
        return constants.real_mesh(self.grid_size, self.domain_x, self.domain_y)

    def load_fft_mesh(self):  # K:89-101
        return constants.fft_mesh(self.grid_size, self.domain_x, self.domain_y)

    def initialize_state(self):  # K:103-129 (numpy complex128; callers cast)
        self.vorticity = constants.initial_vorticity_hat(self.grid_size, self.domain_x, self.domain_y)
        return self.vorticity

    def forcing_function(self, k, x, y):  # K:135-146
        return (np.sin(k * y), np.zeros_like(y))

    def get_observations(self, device=None, dtype=torch.float32):
        """K:70-87: |velocity| of every snapshot in ``self.vorticity`` (T, N, Mh) at the points
        ``range(0, n, int(n / (n // obs_size)))`` in x and y -- ``(n // obs_size)**2`` points with stride
        ``obs_size`` (64 points only when n = obs_size**2; reference quirk kept).  Runs the probe kernel with
        the snapshots as the batch axis.  Returns (T, points) on the GPU."""
        if self.vorticity is None:
            raise ValueError("flow.vorticity is empty: run solve() first")
        n = self.grid_size[0]
        divisor = n // self.obs_size
        if int(n / divisor) * divisor != n:
            raise ValueError("the reference's point grid needs n to be a multiple of n // obs_size")
        traj = self.vorticity if isinstance(self.vorticity, torch.Tensor) else torch.from_numpy(np.asarray(self.vorticity))
        if traj.dim() == 2:
            traj = traj.unsqueeze(0)
        rd = torch.float64 if traj.dtype == torch.complex128 else dtype
        cd = torch.complex128 if rd == torch.float64 else torch.complex64
        st = KolmogorovStepper(grid_size=self.grid_size, Re=self.Re, k=self.k, num_envs=traj.shape[0], dtype=rd,
                               device=device, obs_size=divisor, n_inner=1, n_outer=1, domain_x=self.domain_x,
                               domain_y=self.domain_y)
        obs, _tke, _nonf = st.reward_obs(traj.to(device=st.device, dtype=cd).unsqueeze(1).contiguous())
        st.close()
        return obs


@dataclasses.dataclass(frozen=True)
class KolmogorovFlowParams:
    """Reference K:310-331 (+ gymnax ``max_steps_in_episode``)."""

    max_steps_in_episode: int = 1
    min_action: float = -0.5
    max_action: float = 0.5
    min_obs: float = -np.inf
    max_obs: float = np.inf
    dt: float = 1e-3
    action_time: float = 10.0
    save_time: float = 1
    k1: int = 4
    k2: int = 5
    k3: int = 6
    k4: int = 7
    action_dim: int = 4
    obs_dim: int = 64
    max_episode_steps: int = 1000
    reward_alpha: float = 0.0
    include_grad: bool = True

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


@dataclasses.dataclass(frozen=True)
class KolmogorovFlowState:
    """Reference K:302-307; tensors may carry a leading batch axis."""

    trajectory: torch.Tensor  # (..., n_outer, N, Mh) complex
    omega_hat: torch.Tensor   # (..., N, Mh) complex
    time: torch.Tensor        # (...,) int64
    terminal: torch.Tensor    # (...,) bool

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


def _torch_dtype(name):
    if isinstance(name, torch.dtype):
        return name
    table = {"float32": torch.float32, "fp32": torch.float32, "float64": torch.float64, "fp64": torch.float64}
    if str(name) not in table:
        raise ValueError(f"unsupported dtype {name!r}")
    return table[str(name)]


def _select(done, a, b):
    """Leaf-wise ``select(done, a, b)`` with ``done`` broadcast over trailing axes (gymnax auto-reset)."""
    d = done.reshape(done.shape + (1,) * (a.dim() - done.dim()))
    return torch.where(d, a, b)


class KolmogorovFlow:
    """Reference K:334-525 (``JAXFlowEnvBase`` E:387-438 + gymnax ``Environment``)."""

    def __init__(self, env_config: Optional[Dict] = None, flow_config: Optional[Dict] = None):
        self.env_config = env_config or {}
        self.flow = FlowConfig(**(flow_config or {}))
        self.n, self.m = self.flow.grid_size
        self.x, self.y = self.flow.load_mesh("")
        self.kx, self.ky = self.flow.load_fft_mesh()
        self._dt_override = self.env_config.get("dt", None)
        self.dtype = _torch_dtype(self.env_config.get("dtype", "float32"))
        self.cdtype = torch.complex64 if self.dtype == torch.float32 else torch.complex128
        self.device = self.env_config.get("device", None)
        self.dealias = bool(self.env_config.get("dealias", False))
        self._steppers: Dict[Any, KolmogorovStepper] = {}
        self._reset_cache = None
        default = self.default_params
        # same expressions as RungeKuttaCrankNicolson.solve (S:100-103); raises for action_time < 1 (S:93-97)
        self.n_inner, self.n_outer = constants.step_counts(
            float(default.dt), float(default.action_time), int(default.save_time))
        from ..solvers.base import PseudoSpectralNavierStokes2D, RungeKuttaCrankNicolson  # noqa: WPS433
        self.equation = PseudoSpectralNavierStokes2D(self.flow, dtype=self.dtype, device=self.device,
                                                     dealias=self.dealias)
        self.integrator = RungeKuttaCrankNicolson(
            flow=self.flow, dt=float(default.dt), save_n=int(default.save_time), equation=self.equation,
            dtype=self.dtype, device=self.device, dealias=self.dealias)

    # -- static description ----------------------------------------------------
    @property
    def name(self) -> str:  # K:359-361
        return "KolmogorovFlow"

    @property
    def default_params(self) -> KolmogorovFlowParams:  # K:363-369
        kwargs = dict(action_dim=4, obs_dim=self.flow.obs_size ** 2)
        if self._dt_override is not None:
            kwargs["dt"] = float(self._dt_override)
        for key in ("action_time", "save_time"):  # build-only keys
            if self.env_config.get(key) is not None:
                kwargs[key] = self.env_config[key]
        return KolmogorovFlowParams(**kwargs)

    @property
    def num_actions(self) -> int:
        return 4

    def action_space(self, params: Optional[KolmogorovFlowParams] = None):  # K:371-377
        params = params or self.default_params
        return Box(low=params.min_action, high=params.max_action, shape=(params.action_dim,))

    def observation_space(self, params: Optional[KolmogorovFlowParams] = None):  # K:379-384
        params = params or self.default_params
        return Box(low=params.min_obs, high=params.max_obs, shape=(params.obs_dim,))

    # -- plumbing --------------------------------------------------------------
    def _stepper(self, B: int, params: KolmogorovFlowParams) -> KolmogorovStepper:
        key = (B, params.k1, params.k2, params.k3, params.k4, params.min_action, params.max_action)
        st = self._steppers.get(key)
        if st is None:
            default = self.default_params  # dt / counts come from DEFAULT params (K:407-410)
            st = KolmogorovStepper(
                grid_size=self.flow.grid_size, Re=self.flow.Re, k=self.flow.k,
                control_ks=(params.k1, params.k2, params.k3, params.k4), dt=float(default.dt),
                n_inner=self.n_inner, n_outer=self.n_outer, num_envs=B, dtype=self.dtype, device=self.device,
                dealias=self.dealias, obs_size=self.flow.obs_size, min_action=params.min_action,
                max_action=params.max_action, domain_x=self.flow.domain_x, domain_y=self.flow.domain_y,
                chunk=int(self.env_config.get("chunk", 0)), use_graph=bool(self.env_config.get("use_graph", True)))
            self._steppers[key] = st
        return st

    def _clip_action(self, action, params):  # E:419-420
        return torch.clamp(action, params.min_action, params.max_action)

    def is_terminal(self, state: KolmogorovFlowState, params: KolmogorovFlowParams):  # E:422-423
        return torch.logical_or(state.terminal, state.time >= params.max_episode_steps)

    def discount(self, state, params):  # gymnax Environment.discount
        done = self.is_terminal(state, params)
        return torch.where(done, torch.zeros((), dtype=self.dtype, device=done.device),
                           torch.ones((), dtype=self.dtype, device=done.device))

    def _rollout(self, omega_hat0, params, action=None):
        """K:397-420.  ``action`` (already clipped) replaces the reference's ``control_field``
        tuple: the field is linear in the action, so the kernels take the 4 amplitudes."""
        batched = omega_hat0.dim() == 3
        w = omega_hat0 if batched else omega_hat0.unsqueeze(0)
        st = self._stepper(w.shape[0], params)
        a = None if action is None else (action if batched else action.unsqueeze(0))
        final, traj = st.rollout(w.contiguous(), a)
        return (final, traj) if batched else (final[0], traj[0])

    def _reward_obs(self, trajectory, params):
        batched = trajectory.dim() == 4
        tj = trajectory if batched else trajectory.unsqueeze(0)
        obs, tke, nonf = self._stepper(tj.shape[0], params).reward_obs(tj)
        return (obs, tke, nonf) if batched else (obs[0], tke[0], nonf[0])

    def _trajectory_mean_obs(self, trajectory):  # K:426-446
        return self._reward_obs(trajectory, self.default_params)[0]

    def get_obs(self, state, params=None, key=None):  # K:448-454
        return self._reward_obs(state.trajectory, params or self.default_params)[0]

    def _avg_tke(self, trajectory):  # K:456-460
        return self._reward_obs(trajectory, self.default_params)[1]

    def _reward(self, action, trajectory, params):  # K:462-470
        energy = self._reward_obs(trajectory, params)[1]
        return -(params.reward_alpha * energy + torch.sum(torch.abs(action), dim=-1))

    # -- the functional env API ------------------------------------------------
    def _reset_single(self, params):
        """Post-reset (obs, state) of ONE environment, computed once and cached: the reference's
        reset ignores the key and always rolls the same analytic field forward (K:472-492)."""
        key = (params.k1, params.k2, params.k3, params.k4, params.min_action, params.max_action)
        if self._reset_cache is None or self._reset_cache[0] != key:
            st = self._stepper(1, params)
            omega0 = torch.from_numpy(self.flow.initialize_state()).to(device=st.device, dtype=self.cdtype)
            final, traj, obs, _tke, _nonf = st.step(omega0.unsqueeze(0).contiguous(), None, want_trajectory=True)
            self._reset_cache = (key, obs[0], final[0], traj[0])
        return self._reset_cache[1:]

    def reset_env(self, key=None, params: Optional[KolmogorovFlowParams] = None, num_envs: Optional[int] = None):
        """K:472-492.  ``num_envs=B`` returns the batched (vmapped) result."""
        params = params or self.default_params
        obs, final, traj = self._reset_single(params)
        dev = final.device
        if num_envs is None:
            state = KolmogorovFlowState(trajectory=traj.clone(), omega_hat=final.clone(),
                                        time=torch.zeros((), dtype=torch.int64, device=dev),
                                        terminal=torch.zeros((), dtype=torch.bool, device=dev))
            return obs.clone(), state
        B = int(num_envs)
        state = KolmogorovFlowState(
            trajectory=traj.unsqueeze(0).expand(B, *traj.shape).contiguous(),
            omega_hat=final.unsqueeze(0).expand(B, *final.shape).contiguous(),
            time=torch.zeros((B,), dtype=torch.int64, device=dev),
            terminal=torch.zeros((B,), dtype=torch.bool, device=dev))
        return obs.unsqueeze(0).expand(B, -1).contiguous(), state

    def step_env(self, key, state: KolmogorovFlowState, action, params: Optional[KolmogorovFlowParams] = None):
        """K:494-525: (obs, next_state, reward, done, info)."""
        params = params or self.default_params
        batched = state.omega_hat.dim() == 3
        w = state.omega_hat if batched else state.omega_hat.unsqueeze(0)
        st = self._stepper(w.shape[0], params)
        action = torch.as_tensor(action, dtype=self.dtype, device=st.device)
        action = self._clip_action(action, params)                                  # K:501
        a = action if batched else action.unsqueeze(0)
        final, traj, obs, tke, nonf = st.step(w.contiguous(), a, want_trajectory=True)  # K:504-508, 517-518
        if not batched:
            final, traj, obs, tke, nonf = final[0], traj[0], obs[0], tke[0], nonf[0]
        next_state = KolmogorovFlowState(trajectory=traj, omega_hat=final, time=state.time + 1,
                                         terminal=torch.zeros_like(state.terminal))  # K:510-515
        reward = -(params.reward_alpha * tke + torch.sum(torch.abs(action), dim=-1))  # K:462-470
        done = self.is_terminal(next_state, params)                                   # K:519
        info = {"discount": self.discount(next_state, params), "mean_tke": tke, "nonfinite": nonf}  # K:521-524
        return obs, next_state, reward, done, info

    def reset(self, key=None, params: Optional[KolmogorovFlowParams] = None, num_envs: Optional[int] = None):
        """gymnax ``Environment.reset``."""
        return self.reset_env(key, params or self.default_params, num_envs=num_envs)

    def step(self, key, state: KolmogorovFlowState, action, params: Optional[KolmogorovFlowParams] = None):
        """gymnax ``Environment.step``: ``step_env`` then auto-reset where ``done``.
        The reset branch copies the cached post-reset state instead of recomputing it."""
        params = params or self.default_params
        obs_st, state_st, reward, done, info = self.step_env(key, state, action, params)
        obs_re, final_re, traj_re = self._reset_single(params)  # broadcast views, no copies
        state_out = KolmogorovFlowState(
            trajectory=_select(done, traj_re.expand_as(state_st.trajectory), state_st.trajectory),
            omega_hat=_select(done, final_re.expand_as(state_st.omega_hat), state_st.omega_hat),
            time=torch.where(done, torch.zeros_like(state_st.time), state_st.time),
            terminal=torch.where(done, torch.zeros_like(state_st.terminal), state_st.terminal))
        obs = _select(done, obs_re.expand_as(obs_st), obs_st)
        return obs, state_out, reward, done, info
