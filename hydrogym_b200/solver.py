"""Device-side stepper: owns a ``kf_handle`` and launches the sm_100a kernels on torch tensors.

PyTorch is plumbing here (device memory, streams); all arithmetic of the hot path
runs in ``libkolmo_b200.so``.  Mirrors what ``KolmogorovFlow._rollout``
(hydrogym/jax/envs/kolmogorov.py:397-420) + ``RungeKuttaCrankNicolson.solve``
(hydrogym/jax/solvers/base.py:82-116) do for a ``vmap``-ed batch.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib, constants


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class KolmogorovStepper:
    """Batched pseudo-spectral RK4-CN stepper for B environments on one GPU."""

    def __init__(self, grid_size=(64, 64), Re=200, k=4, control_ks=(4, 5, 6, 7), dt=1e-3, n_inner=999,
                 n_outer=10, num_envs=1, dtype=torch.float32, device=None, dealias=False, obs_size=8,
                 min_action=-0.5, max_action=0.5, domain_x=(0, 2 * np.pi), domain_y=(0, 2 * np.pi),
                 chunk=0, use_graph=True):
        if dtype not in (torch.float32, torch.float64):
            raise ValueError("dtype must be torch.float32 or torch.float64")
        N, M = grid_size
        if N != M:
            raise ValueError("only square grids are supported")
        constants.check_domain(domain_x, domain_y)
        if not torch.cuda.is_available():
            raise RuntimeError("hydrogym_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if self.device.type != "cuda":
            raise RuntimeError("hydrogym_b200 runs on CUDA devices only")
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", dev_index)
        self.N, self.M, self.Mh = N, M, M // 2 + 1
        self.B = int(num_envs)
        self.dtype = dtype
        self.cdtype = torch.complex64 if dtype == torch.float32 else torch.complex128
        self.n_inner, self.n_outer = int(n_inner), int(n_outer)
        self.obs_size = int(obs_size)
        self.dt, self.nu = float(dt), 1.0 / Re

        self._lib = _lib.load()
        self._large = None
        if N > 1024:
            # Grids beyond the batched kernels' plans (2048^2, 4096^2, 8192^2): the single-grid slab kernels on ONE rank
            # (csrc/slab.cu: transposed K layout, one transform per CTA), environment by environment.
            if N not in (2048, 4096, 8192):
                raise ValueError("grid size must be a power of two between 64 and 8192")
            from .slab import SlabKolmogorovSolver
            self._large = SlabKolmogorovSolver(
                grid_size=(N, M), Re=Re, k=k, control_ks=control_ks, dt=dt, dtype=dtype, device=self.device,
                dealias=dealias, obs_size=obs_size, min_action=min_action, max_action=max_action, domain_y=domain_y,
                local_ranks=1, exchange="p2p")
            self._h = ctypes.c_void_p(0)
            return
        cfg = _lib.KfConfig(N=N, M=M, B=self.B, dtype=_lib.KF_F32 if dtype == torch.float32 else _lib.KF_F64,
                            nu=self.nu, dt=self.dt, n_inner=self.n_inner, n_outer=self.n_outer,
                            dealias=int(bool(dealias)), drag=0.0, device=dev_index, obs_size=self.obs_size,
                            min_action=float(min_action), max_action=float(max_action), chunk=int(chunk),
                            use_graph=int(bool(use_graph)))
        self._h = ctypes.c_void_p(0)
        _lib.check(self._lib.kf_create(ctypes.byref(self._h), ctypes.byref(cfg)), "kf_create")

        np_c = np.complex64 if dtype == torch.float32 else np.complex128
        self.forcing_row = np.ascontiguousarray(constants.forcing_row(k, grid_size, domain_y).astype(np_c))
        self.control_rows = np.ascontiguousarray(constants.control_rows(control_ks, grid_size, domain_y).astype(np_c))
        _lib.check(self._lib.kf_set_constants(self._h, self.forcing_row.ctypes.data_as(ctypes.c_void_p),
                                              self.control_rows.ctypes.data_as(ctypes.c_void_p)), "kf_set_constants")

    def set_control_rows(self, rows):
        """Replace the 4 control basis rows (row kx=0 of control_term for unit actions)."""
        self._unsupported_large("set_control_rows")
        np_c = np.complex64 if self.dtype == torch.float32 else np.complex128
        self.control_rows = np.ascontiguousarray(np.asarray(rows).astype(np_c))
        _lib.check(self._lib.kf_set_constants(self._h, self.forcing_row.ctypes.data_as(ctypes.c_void_p),
                                              self.control_rows.ctypes.data_as(ctypes.c_void_p)), "kf_set_constants")

    def set_control_field(self, control_hat):
        """Dense control spectrum (N, Mh) -- ``control_term`` of an arbitrary ``(cfx, cfy)`` field (K:259-272) --
        added to the nonlinear term of every environment in every sub-stage; ``None`` removes it."""
        if control_hat is None and self._large:
            return
        self._unsupported_large("a dense control field")
        if control_hat is None:
            _lib.check(self._lib.kf_set_control_field(self._h, ctypes.c_void_p(0)), "kf_set_control_field")
            return
        np_c = np.complex64 if self.dtype == torch.float32 else np.complex128
        if isinstance(control_hat, torch.Tensor):
            control_hat = control_hat.detach().cpu().numpy()
        arr = np.ascontiguousarray(np.asarray(control_hat).astype(np_c))
        if arr.shape != (self.N, self.Mh):
            raise ValueError(f"control_hat must be {(self.N, self.Mh)}, got {arr.shape}")
        _lib.check(self._lib.kf_set_control_field(self._h, arr.ctypes.data_as(ctypes.c_void_p)), "kf_set_control_field")

    # -- helpers -------------------------------------------------------------
    def close(self):
        if getattr(self, "_large", None) is not None:
            self._large.close()
            self._large = None
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.kf_destroy(self._h)
            self._h = ctypes.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _check_state(self, omega_hat):
        if omega_hat.shape != (self.B, self.N, self.Mh) or omega_hat.dtype != self.cdtype:
            raise ValueError(f"omega_hat must be {(self.B, self.N, self.Mh)} {self.cdtype}, "
                             f"got {tuple(omega_hat.shape)} {omega_hat.dtype}")
        if omega_hat.device != self.device or not omega_hat.is_contiguous():
            raise ValueError("omega_hat must be a contiguous tensor on the stepper's device")

    def _check_actions(self, actions):
        if actions is None:
            return None
        actions = actions.to(device=self.device, dtype=self.dtype).contiguous()
        if actions.shape != (self.B, 4):
            raise ValueError(f"actions must be {(self.B, 4)}, got {tuple(actions.shape)}")
        return actions

    @property
    def launch_count(self):
        return self._large.launch_count if self._large else int(self._lib.kf_launch_count(self._h))

    @property
    def workspace_bytes(self):
        return self._large.workspace_bytes if self._large else int(self._lib.kf_workspace_bytes(self._h))

    @property
    def chunk(self):
        return 1 if self._large else int(self._lib.kf_chunk(self._h))

    @property
    def lanes(self):
        return 1 if self._large else int(self._lib.kf_lanes(self._h))

    # -- large grids: one environment at a time through the slab kernels -------------------------------------------
    def _large_advance(self, omega_hat, actions, n_outer, n_inner, traj, obs, tke, nonf):
        """In place on ``omega_hat``; fills whichever of traj / obs / tke / nonf is given."""
        sl = self._large
        for b in range(self.B):
            sl.set_state(omega_hat[b], None if actions is None else actions[b])
            ob = torch.zeros(self.obs_size ** 2, dtype=self.dtype, device=self.device) if obs is not None else None
            tk = torch.zeros((), dtype=torch.float64, device=self.device)
            for j in range(n_outer):
                sl.rk_steps(n_inner)
                if traj is not None:
                    traj[b, j] = sl.get_state()
                if obs is not None or tke is not None:
                    _, t = sl.reward_obs(1.0 / n_outer, obs_out=ob)
                    tk += t / n_outer
            omega_hat[b] = sl.get_state()
            if obs is not None:
                obs[b] = ob
            if tke is not None:
                tke[b] = tk.to(self.dtype)
            if nonf is not None:
                nonf[b] = (~torch.isfinite(tk)).to(torch.int32)

    def _unsupported_large(self, what):
        if self._large:
            raise NotImplementedError(f"{what} is available for grids up to 1024^2 only")

    # -- device entry points ---------------------------------------------------
    def rk_steps(self, omega_hat, actions, n_steps):
        """Advance ``omega_hat`` (B, N, Mh) IN PLACE by ``n_steps`` RK solver steps."""
        self._check_state(omega_hat)
        actions = self._check_actions(actions)
        if self._large:
            self._large_advance(omega_hat, actions, 1, int(n_steps), None, None, None, None)
            return omega_hat
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_rk_steps(self._h, _ptr(omega_hat), _ptr(actions), int(n_steps), self._stream()),
                       "kf_rk_steps")
        return omega_hat

    def rollout(self, omega_hat, actions=None, want_trajectory=True):
        """``_rollout``: returns (final_state, trajectory[B, n_outer, N, Mh]); input is not modified."""
        self._check_state(omega_hat)
        actions = self._check_actions(actions)
        final = omega_hat.clone()
        traj = (torch.empty((self.B, self.n_outer, self.N, self.Mh), dtype=self.cdtype, device=self.device)
                if want_trajectory else None)
        if self._large:
            self._large_advance(final, actions, self.n_outer, self.n_inner, traj, None, None, None)
            return final, traj
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_rollout(self._h, _ptr(final), _ptr(actions), _ptr(traj), self._stream()),
                       "kf_rollout")
        return final, traj

    def reward_obs(self, trajectory):
        """(obs[B, obs_size^2], mean_tke[B], nonfinite[B]) of a trajectory (B, T, N, Mh)."""
        if trajectory.dim() != 4 or trajectory.shape[0] != self.B or trajectory.shape[2:] != (self.N, self.Mh):
            raise ValueError("trajectory must be (B, T, N, Mh)")
        trajectory = trajectory.contiguous()
        obs = torch.empty((self.B, self.obs_size ** 2), dtype=self.dtype, device=self.device)
        tke = torch.empty((self.B,), dtype=self.dtype, device=self.device)
        nonf = torch.empty((self.B,), dtype=torch.int32, device=self.device)
        if self._large:
            T = int(trajectory.shape[1])
            obs.zero_()
            for b in range(self.B):
                tk = torch.zeros((), dtype=torch.float64, device=self.device)
                for j in range(T):
                    self._large.set_state(trajectory[b, j], None)
                    _, t = self._large.reward_obs(1.0 / T, obs_out=obs[b])
                    tk += t / T
                tke[b] = tk.to(self.dtype)
                nonf[b] = (~torch.isfinite(tk)).to(torch.int32)
            return obs, tke, nonf
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_reward_obs(self._h, _ptr(trajectory), int(trajectory.shape[1]), _ptr(obs),
                                               _ptr(tke), _ptr(nonf), self._stream()), "kf_reward_obs")
        return obs, tke, nonf

    def step(self, omega_hat, actions=None, want_trajectory=True, in_place=False):
        """Rollout + reward/obs in one call: (final, trajectory|None, obs, mean_tke, nonfinite)."""
        self._check_state(omega_hat)
        actions = self._check_actions(actions)
        final = omega_hat if in_place else omega_hat.clone()
        traj = (torch.empty((self.B, self.n_outer, self.N, self.Mh), dtype=self.cdtype, device=self.device)
                if want_trajectory else None)
        obs = torch.empty((self.B, self.obs_size ** 2), dtype=self.dtype, device=self.device)
        tke = torch.empty((self.B,), dtype=self.dtype, device=self.device)
        nonf = torch.empty((self.B,), dtype=torch.int32, device=self.device)
        if self._large:
            self._large_advance(final, actions, self.n_outer, self.n_inner, traj, obs, tke, nonf)
            return final, traj, obs, tke, nonf
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_step(self._h, _ptr(final), _ptr(actions), _ptr(traj), _ptr(obs), _ptr(tke),
                                         _ptr(nonf), self._stream()), "kf_step")
        return final, traj, obs, tke, nonf

    def physical_fields(self, omega_hat, nfields=2):
        """(nfields, B, N, N) real: u, v[, d(omega)/dx, d(omega)/dy] in physical space (irfftn of the spectral
        fields of utils.py:53-69 / kolmogorov.py:240-247)."""
        self._unsupported_large("physical_fields")
        self._check_state(omega_hat)
        out = torch.empty((int(nfields), self.B, self.N, self.M), dtype=self.dtype, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_physical_fields(self._h, _ptr(omega_hat), _ptr(out), int(nfields), self._stream()),
                       "kf_physical_fields")
        return out

    def nonlinear_terms(self, omega_hat, actions=None):
        """``PseudoSpectralNavierStokes2D.nonlinear_terms`` (K:222-257) of (B, N, Mh): advection + forcing + control
        (``actions`` (B, 4) through the control rows, plus the dense field of ``set_control_field`` if set)."""
        self._unsupported_large("nonlinear_terms")
        self._check_state(omega_hat)
        actions = self._check_actions(actions)
        out = torch.empty_like(omega_hat)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_nonlinear_terms(self._h, _ptr(omega_hat), _ptr(actions), _ptr(out), self._stream()),
                       "kf_nonlinear_terms")
        return out

    def profile_kernels(self, reps=20):
        """(ms per column-kernel launch, ms per row-kernel launch) on scratch workspace data."""
        self._unsupported_large("profile_kernels")
        a, b = ctypes.c_float(0), ctypes.c_float(0)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.kf_profile_kernels(self._h, int(reps), ctypes.byref(a), ctypes.byref(b),
                                                    self._stream()), "kf_profile_kernels")
        return float(a.value), float(b.value)

    def step_host(self, omega_hat_host, actions_host, obs_host, tke_host, nonf_host):
        """``kf_step_host``: all arguments are HOST tensors (pinned for speed); synchronous, in place."""
        self._unsupported_large("step_host")
        _lib.check(self._lib.kf_step_host(self._h, _ptr(omega_hat_host), _ptr(actions_host), _ptr(obs_host),
                                          _ptr(tke_host), _ptr(nonf_host)), "kf_step_host")


def _spectral_pointwise(name, omega_hat, nu, *extra):
    if not isinstance(omega_hat, torch.Tensor) or not omega_hat.is_cuda:
        raise ValueError("omega_hat must be a CUDA tensor (hydrogym_b200 has no CPU fallback)")
    if omega_hat.dtype not in (torch.complex64, torch.complex128):
        raise ValueError("omega_hat must be complex64 or complex128")
    N, Mh = omega_hat.shape[-2:]
    if Mh != N // 2 + 1:
        raise ValueError("omega_hat must be (..., N, N//2+1)")
    w = omega_hat.contiguous()
    out = torch.empty_like(w)
    B = w.numel() // (N * Mh)
    lib = _lib.load()
    with torch.cuda.device(w.device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(w.device).cuda_stream)
        dt = _lib.KF_F32 if w.dtype == torch.complex64 else _lib.KF_F64
        _lib.check(getattr(lib, name)(dt, int(N), int(B), float(nu), *extra, _ptr(w), _ptr(out), stream), name)
    return out


def linear_terms(omega_hat, nu):
    """K:209-211 on device: ``nu * (2 pi i)^2 (kx^2 + ky^2) * omega_hat`` for (..., N, N//2+1)."""
    return _spectral_pointwise("kf_linear_terms", omega_hat, nu)


def implicit_timestep(omega_hat, nu, time_step):
    """K:213-220 on device: ``omega_hat / (1 - time_step * nu * (2 pi i)^2 (kx^2 + ky^2))``."""
    return _spectral_pointwise("kf_implicit_timestep", omega_hat, nu, float(time_step))
