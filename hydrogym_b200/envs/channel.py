"""3-D turbulent channel-flow environment over the sm_100a kernels of ``csrc/channel.cu[h]``.

Mirrors ``hydrogym/jax/envs/channel.py`` (C) of the reference: same class / method names, argument meaning and
defaults; states and observations are torch tensors on the GPU.  A leading batch axis on the state (``num_envs``)
replaces ``jax.vmap``.  There is no CPU fallback: the step is ``kc_step`` of ``libkolmo_b200.so``.

Differences that are deliberate and documented:
* the reference also differentiates the wall shear stress w.r.t. the action inside ``step_env`` and discards the
  gradient (C:497-502); only the forward pass is observable, and only that is computed here;
* the HF-hosted initial fields (C:446-455) cannot be downloaded here: pass ``env_config["initial_field_dir"]``
  (the reference's own override, C:444-445) or ``env_config["initial_fields"] = (U, V, W)``.
"""
from __future__ import annotations

import ctypes
import dataclasses
import os
from typing import Dict, Optional

import numpy as np
import torch

from .. import _lib, channel_constants as cc
from ..spaces import Box

N = cc.N


@dataclasses.dataclass(frozen=True)
class ChannelEnvParams:
    """C:30-52."""
    max_steps_in_episode: int = 5000
    action_dim: int = 24
    obs_subsample: int = 8
    obs_include_components: int = 2
    k_det: int = 9
    Nx: int = 72
    Ny: int = 72
    Nz: int = 72
    nsteps: int = 50
    dt: float = 2e-4
    min_action: float = -1
    max_action: float = 1

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


@dataclasses.dataclass
class ChannelEnvState:
    """C:55-62; ``U, V, W`` are (72, 72, 72) or (B, 72, 72, 72) tensors, ``time`` / ``terminal`` scalars or (B,)."""
    time: torch.Tensor
    U: torch.Tensor
    V: torch.Tensor
    W: torch.Tensor
    dt: torch.Tensor
    terminal: torch.Tensor

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


class ChannelStepper:
    """Owns one ``kc_handle`` (B environments): the solver-level surface of ``include/channel_b200.h``.

    ``PseudoSpectralNavierStokes3D.rhs`` (C:302-309) -> ``rhs``; ``RungeKutta4.rk4_step`` (S:126-185) ->
    ``rk4_steps``; ``to_spectral / to_physical`` (C:156-168); ``_wss_with_aux`` + probes (C:531-567, 76-92) -> ``step``.
    """

    def __init__(self, num_envs=1, dtype=torch.float32, device=None, jets_on=True, nu=cc.NU, fx=2.0,
                 target_bulk_u=8.0, k_det=9, wss_plane=4):
        if dtype not in (torch.float32, torch.float64):
            raise ValueError("dtype must be torch.float32 or torch.float64")
        self.lib = _lib.load()
        self.device = torch.device(device if device is not None else "cuda")
        if self.device.type != "cuda":
            raise RuntimeError("hydrogym_b200 has no CPU fallback: the channel step needs a CUDA device")
        self.index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", self.index)
        self.B, self.dtype = int(num_envs), dtype
        self.cdtype = torch.complex64 if dtype == torch.float32 else torch.complex128
        cfg = _lib.KcConfig(B=self.B, dtype=_lib.KF_F32 if dtype == torch.float32 else _lib.KF_F64, device=self.index,
                            nu=float(nu), lx=cc.LX, ly=cc.LY, fx=float(fx), target_bulk_u=float(target_bulk_u),
                            k_det=int(k_det), wss_plane=int(wss_plane), jets_on=1 if jets_on else 0)
        self.h = ctypes.c_void_p()
        _lib.check(self.lib.kc_create(ctypes.byref(self.h), ctypes.byref(cfg)), "kc_create")
        self._ops = cc.operators()   # keep the host arrays alive during the upload
        _lib.check(self.lib.kc_set_operators(self.h, *[_dp(a) for a in self._ops]), "kc_set_operators")
        self.z = self._ops[0]

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.kc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- plumbing ---------------------------------------------------------------------------------------
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _field(self, a):
        a = torch.as_tensor(a).to(device=self.device, dtype=self.dtype)
        if a.dim() == 3:
            a = a.unsqueeze(0)
        if tuple(a.shape) != (self.B, N, N, N):
            raise ValueError(f"expected a ({self.B}, {N}, {N}, {N}) field, got {tuple(a.shape)}")
        return a.contiguous()

    def _actions(self, actions):
        if actions is None:
            return None
        a = torch.as_tensor(actions).to(device=self.device, dtype=self.dtype)
        if a.dim() == 1:
            a = a.unsqueeze(0)
        if tuple(a.shape) != (self.B, 24):
            raise ValueError(f"expected ({self.B}, 24) actions, got {tuple(a.shape)}")
        return a.contiguous()

    def _times(self, time):
        t = np.asarray(time.detach().cpu() if isinstance(time, torch.Tensor) else time, dtype=np.float64)
        return np.ascontiguousarray(np.broadcast_to(t.reshape(-1) if t.ndim else t, (self.B,)))

    # -- the surface ------------------------------------------------------------------------------------
    def set_state(self, U, V, W):
        U, V, W = self._field(U), self._field(V), self._field(W)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_set_state(self.h, _ptr(U), _ptr(V), _ptr(W), self._stream()), "kc_set_state")

    def get_state(self):
        out = [torch.empty((self.B, N, N, N), dtype=self.dtype, device=self.device) for _ in range(3)]
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_get_state(self.h, *[_ptr(o) for o in out], self._stream()), "kc_get_state")
        return tuple(out)

    def step(self, actions, time, nsteps=50, dt=2e-4, lo=-1.0, hi=1.0):
        """``nsteps`` solver steps from the stored physical state; returns (obs (B, 128), reward (B,))."""
        a, t = self._actions(actions), self._times(time)
        obs = torch.empty((self.B, 128), dtype=self.dtype, device=self.device)
        rew = torch.empty((self.B,), dtype=self.dtype, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_step(self.h, _ptr(a) if a is not None else None, _dp(t), int(nsteps), float(dt),
                                        float(lo), float(hi), _ptr(obs), _ptr(rew), self._stream()), "kc_step")
        return obs, rew

    def to_spectral(self):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_to_spectral(self.h, self._stream()), "kc_to_spectral")

    def to_physical(self):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_to_physical(self.h, self._stream()), "kc_to_physical")

    def rk4_steps(self, actions, time, n=1, dt=2e-4, lo=-1.0, hi=1.0):
        a, t = self._actions(actions), self._times(time)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_rk4_steps(self.h, _ptr(a) if a is not None else None, _dp(t), int(n), float(dt),
                                             float(lo), float(hi), self._stream()), "kc_rk4_steps")

    def rhs(self, actions, time, lo=-1.0, hi=1.0):
        """(3, B, 72, 72, 72) complex: ``rhs`` of the internal spectral state at time ``time``."""
        a, t = self._actions(actions), self._times(time)
        out = torch.empty((3, self.B, N, N, N), dtype=self.cdtype, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_rhs(self.h, _ptr(a) if a is not None else None, _dp(t), float(lo), float(hi),
                                       _ptr(out), self._stream()), "kc_rhs")
        return out

    def spectral(self):
        out = torch.empty((3, self.B, N, N, N), dtype=self.cdtype, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kc_get_spectral(self.h, _ptr(out), self._stream()), "kc_get_spectral")
        return out

    @property
    def launch_count(self):
        return int(self.lib.kc_launch_count(self.h))

    @property
    def workspace_bytes(self):
        return int(self.lib.kc_workspace_bytes(self.h))


def _torch_dtype(name):
    return torch.float64 if name in ("float64", torch.float64, np.float64) else torch.float32


class ChannelFlowSpectralEnv:
    """C:406-571: gymnax-style functional environment, 24 wall jets, 128 probe observations, reward = -wall shear."""

    def __init__(self, env_config: Optional[Dict] = None):
        env_config = dict(env_config or {})
        self.env_config = env_config
        self.Lx, self.Ly, self.Lz, self.nu = cc.LX, cc.LY, cc.LZ, cc.NU       # C:416-419
        self.Nx = self.Ny = self.Nz = N                                      # C:420-422
        self.dtype = _torch_dtype(env_config.get("dtype", "float32"))        # C:424-425
        self.device = torch.device(env_config.get("device", "cuda"))
        if "initial_fields" in env_config:
            fields = env_config["initial_fields"]
        elif "initial_field_dir" in env_config:                              # C:444-445
            d = env_config["initial_field_dir"]
            fields = tuple(np.load(os.path.join(d, n + ".npy")) for n in "UVW")
        else:
            raise FileNotFoundError(
                "ChannelFlowSpectralEnv: the reference downloads Channel_3D_Retau180/initial_field/{U,V,W}.npy from "
                "the Hugging Face hub (C:446-455), which is not reachable here; pass env_config['initial_field_dir'] "
                "or env_config['initial_fields'] = (U, V, W)")
        self.U0, self.V0, self.W0 = (torch.as_tensor(np.asarray(f)).to(device=self.device, dtype=self.dtype).contiguous()
                                     for f in fields)
        for f in (self.U0, self.V0, self.W0):
            if tuple(f.shape) != (N, N, N):
                raise ValueError(f"initial fields must be ({N}, {N}, {N}), got {tuple(f.shape)}")
        p = self.default_params
        idx = torch.as_tensor(cc.probe_indices(p.Nx, p.obs_subsample), device=self.device)   # C:69-73
        self.Xi_obs, self.Yi_obs = torch.meshgrid(idx, idx, indexing="ij")
        self._steppers = {}

    # -- gymnax surface ---------------------------------------------------------------------------------
    @property
    def default_params(self) -> ChannelEnvParams:  # C:459-461
        return ChannelEnvParams()

    @property
    def name(self) -> str:  # C:463-465
        return "ChannelFlowSpectralEnv"

    @property
    def num_actions(self) -> int:
        return self.default_params.action_dim

    def action_space(self, params: Optional[ChannelEnvParams] = None):  # C:521-527
        params = params or self.default_params
        return Box(low=params.min_action, high=params.max_action, shape=(params.action_dim,))

    def observation_space(self, params: Optional[ChannelEnvParams] = None):  # C:529-533
        params = params or self.default_params
        return Box(low=-np.inf, high=np.inf, shape=(params.obs_subsample ** 2 * params.obs_include_components,))

    def _stepper(self, B, params) -> ChannelStepper:
        key = (int(B), int(params.k_det))
        if key not in self._steppers:
            self._steppers[key] = ChannelStepper(num_envs=B, dtype=self.dtype, device=self.device, nu=self.nu,
                                                 k_det=params.k_det)
        return self._steppers[key]

    def close(self):
        for st in self._steppers.values():
            st.close()
        self._steppers = {}

    def get_obs(self, state: ChannelEnvState, params: Optional[ChannelEnvParams] = None, key=None):
        """C:76-85 on a stored state (a gather; ``step_env`` gets the same values from ``kc_step``)."""
        params = params or self.default_params
        u, w = state.U[..., params.k_det], state.W[..., params.k_det]
        sel = (Ellipsis, self.Xi_obs, self.Yi_obs)
        return torch.stack([u[sel], w[sel]], dim=-3).reshape(*u.shape[:-2], -1)

    def is_terminal(self, state: ChannelEnvState, params: ChannelEnvParams):  # C:483-484
        return torch.logical_or(state.terminal, state.time >= params.max_steps_in_episode)

    def discount(self, state, params):
        return torch.where(self.is_terminal(state, params), 0.0, 1.0)

    def reset_env(self, key=None, params: Optional[ChannelEnvParams] = None, num_envs: Optional[int] = None):
        """C:467-481; ``num_envs=B`` returns the batched (vmapped) state."""
        params = params or self.default_params
        shape = () if num_envs is None else (int(num_envs),)

        def rep(f):
            return f.clone() if num_envs is None else f.unsqueeze(0).expand(int(num_envs), N, N, N).contiguous()
        state = ChannelEnvState(time=torch.zeros(shape, dtype=torch.int64, device=self.device),
                                U=rep(self.U0), V=rep(self.V0), W=rep(self.W0),
                                dt=torch.full(shape, params.dt, dtype=torch.float64, device=self.device),
                                terminal=torch.zeros(shape, dtype=torch.bool, device=self.device))
        return self.get_obs(state, params, key), state

    def step_env(self, key, state: ChannelEnvState, action, params: Optional[ChannelEnvParams] = None):
        """C:486-519: (obs, next_state, reward, done, info)."""
        params = params or self.default_params
        batched = state.U.dim() == 4
        B = state.U.shape[0] if batched else 1
        st = self._stepper(B, params)
        st.set_state(state.U, state.V, state.W)
        dt = float(state.dt.reshape(-1)[0])
        nsteps = int(self.default_params.nsteps)   # C:555 reads the DEFAULT params
        obs, reward = st.step(action, state.time, nsteps=nsteps, dt=dt, lo=params.min_action, hi=params.max_action)
        U1, V1, W1 = st.get_state()
        if not batched:
            obs, reward, U1, V1, W1 = obs[0], reward[0], U1[0], V1[0], W1[0]
        time = state.time + params.nsteps                                   # C:504
        terminal = time >= params.max_steps_in_episode                      # C:505
        nxt = ChannelEnvState(time=time, U=U1, V=V1, W=W1, dt=state.dt, terminal=terminal)
        done = self.is_terminal(nxt, params)
        return obs, nxt, reward, done, {"discount": self.discount(nxt, params)}

    def reset(self, key=None, params: Optional[ChannelEnvParams] = None, num_envs: Optional[int] = None):
        return self.reset_env(key, params or self.default_params, num_envs=num_envs)

    def step(self, key, state: ChannelEnvState, action, params: Optional[ChannelEnvParams] = None):
        """gymnax ``Environment.step``: ``step_env`` then auto-reset where ``done``."""
        params = params or self.default_params
        obs_st, state_st, reward, done, info = self.step_env(key, state, action, params)
        batched = state.U.dim() == 4
        obs_re, state_re = self.reset_env(key, params, num_envs=state.U.shape[0] if batched else None)

        def sel(a, b):
            d = done.reshape(done.shape + (1,) * (a.dim() - done.dim()))
            return torch.where(d, a, b)
        state_out = ChannelEnvState(time=sel(state_re.time, state_st.time), U=sel(state_re.U, state_st.U),
                                    V=sel(state_re.V, state_st.V), W=sel(state_re.W, state_st.W),
                                    dt=state_st.dt, terminal=sel(state_re.terminal, state_st.terminal))
        return sel(obs_re, obs_st), state_out, reward, done, info


def run_channel_pseudospectral(U0, V0, W0, action, dt: float, nsteps: int = 50, return_trajectory: bool = False,
                               dtype=torch.float64, device=None, time: float = 0.0):
    """C:375-403: ``nsteps`` solver steps from physical fields with a fixed action (t = n, fx = 2, bulk target 8);
    returns the final (U, V, W), or the per-step trajectories (nsteps, 72, 72, 72) x 3."""
    st = ChannelStepper(num_envs=1, dtype=dtype, device=device)
    try:
        st.set_state(U0, V0, W0)
        st.to_spectral()
        if not return_trajectory:
            st.rk4_steps(action, time, n=nsteps, dt=dt)
            st.to_physical()
            return tuple(f[0] for f in st.get_state())
        traj = [[], [], []]
        for n in range(int(nsteps)):
            st.rk4_steps(action, time + n, n=1, dt=dt)
            st.to_physical()
            for lst, f in zip(traj, st.get_state()):
                lst.append(f[0])
        return tuple(torch.stack(l) for l in traj)
    finally:
        st.close()
