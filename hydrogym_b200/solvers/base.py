"""Direct-solver surface: the equation and integrator classes the docs / notebook call.

Reference: hydrogym/jax/solvers/base.py:24-116 (``RungeKuttaCrankNicolson``: ``RK4_CN``, ``step``, ``solve``) and
hydrogym/jax/envs/kolmogorov.py:193-292 (``PseudoSpectralNavierStokes2D``: ``linear_terms``, ``implicit_timestep``,
``nonlinear_terms``, ``control_term``, ``forcing_term``).  States are torch tensors on the GPU; every term is
evaluated by the sm_100a kernels of ``libkolmo_b200.so`` (``kf_nonlinear_terms``, ``kf_linear_terms``,
``kf_implicit_timestep``, ``kf_rk_steps`` / ``kf_rollout``) -- there is no host or torch arithmetic on the state.
The only host arithmetic is on CONSTANTS of a solve, in fp64 and once per call, like the forcing row: the spectrum
``control_term`` of a user-supplied physical control field (K:259-272, two ``rfftn`` of fixed arrays).
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Tuple

import numpy as np
import torch

from .. import constants
from .. import solver as _solver
from ..solver import KolmogorovStepper

# hydrogym/jax/solvers/base.py:13-15 (the kernels hold their own copy; exported for callers)
_alpha_RK4 = [0, 0.1496590219993, 0.3704009573644, 0.6222557631345, 0.9582821306748, 1]
_beta_RK4 = [0, -0.4178904745, -1.192151694643, -1.697784692471, -1.514183444257]
_gammas_RK4 = [0.1496590219993, 0.3792103129999, 0.8229550293869, 0.6994504559488, 0.1530572479681]


def _as_numpy(a):
    return np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a, dtype=np.float64)


def control_spectrum(control_field, flow):
    """``control_term`` (K:259-272) of a physical control field ``(cfx, cfy)`` (each (N, M) real, or ``None``) as a
    dense (N, Mh) complex128 array: ``kx * rfftn(cfy) - ky * rfftn(cfx)`` on the cycles-per-unit wavenumbers of
    ``load_fft_mesh`` -- no 2*pi*i factor, as in the reference."""
    N, M = flow.grid_size
    if control_field is None:
        return np.zeros((N, M // 2 + 1), dtype=np.complex128)
    cfx, cfy = control_field
    cfx, cfy = np.broadcast_to(_as_numpy(cfx), (N, M)), np.broadcast_to(_as_numpy(cfy), (N, M))
    kx, ky = flow.load_fft_mesh()
    return kx * np.fft.rfftn(cfy) - ky * np.fft.rfftn(cfx)


class PseudoSpectralNavierStokes2D:
    """K:193-292.  ``linear_terms`` / ``implicit_timestep`` / ``nonlinear_terms`` take and return GPU tensors of
    shape (N, Mh) or (B, N, Mh); ``forcing_term`` / ``control_term`` return the (constant) spectra as numpy
    complex128 arrays, which is what the kernels are fed with."""

    def __init__(self, flow, dtype=torch.float32, device=None, dealias=False):
        self.flow = flow
        self.grid = flow.load_fft_mesh()
        self.real_grid = flow.load_mesh("name")
        self.kx, self.ky = self.grid
        self.x, self.y = self.real_grid
        self.dtype, self.device, self.dealias = dtype, device, dealias
        self._steppers = {}

    # -- plumbing --------------------------------------------------------------
    def _state(self, omega_hat):
        cdtype = torch.complex64 if self.dtype == torch.float32 else torch.complex128
        if not isinstance(omega_hat, torch.Tensor):
            omega_hat = torch.from_numpy(np.asarray(omega_hat))
        batched = omega_hat.dim() == 3
        w = omega_hat if batched else omega_hat.unsqueeze(0)
        dev = self.device if self.device is not None else (w.device if w.is_cuda else "cuda")
        return w.to(device=dev, dtype=cdtype).contiguous(), batched

    def _stepper(self, B):
        if B not in self._steppers:
            self._steppers[B] = KolmogorovStepper(
                grid_size=self.flow.grid_size, Re=self.flow.Re, k=self.flow.k, num_envs=B, dtype=self.dtype,
                device=self.device, dealias=self.dealias, obs_size=self.flow.obs_size, n_inner=1, n_outer=1,
                min_action=-np.inf, max_action=np.inf, domain_x=self.flow.domain_x, domain_y=self.flow.domain_y)
        return self._steppers[B]

    def close(self):
        for st in self._steppers.values():
            st.close()
        self._steppers = {}

    # -- the reference's methods -------------------------------------------------
    def linear_terms(self, omega_hat):
        """K:209-211."""
        w, batched = self._state(omega_hat)
        out = _solver.linear_terms(w, self.flow.nu)
        return out if batched else out[0]

    def implicit_timestep(self, omega_hat, time_step):
        """K:213-220."""
        w, batched = self._state(omega_hat)
        out = _solver.implicit_timestep(w, self.flow.nu, time_step)
        return out if batched else out[0]

    def nonlinear_terms(self, omega_hat, control_field=None):
        """K:222-257: advection (4 ``irfftn`` + product + ``rfftn``) + forcing + control, one launch of the
        x-space-field, row and forward-column kernels."""
        w, batched = self._state(omega_hat)
        st = self._stepper(w.shape[0])
        st.set_control_field(None if control_field is None else control_spectrum(control_field, self.flow))
        out = st.nonlinear_terms(w, None)
        return out if batched else out[0]

    def forcing_term(self):
        """K:274-292 as a dense (N, Mh) complex128 array (non-zero on row kx=0 only)."""
        N, M = self.flow.grid_size
        out = np.zeros((N, M // 2 + 1), dtype=np.complex128)
        out[0] = constants.forcing_row(self.flow.k, self.flow.grid_size, self.flow.domain_y)
        return out

    def control_term(self, omega_hat=None, control_field=None):
        """K:259-272 for any ``(cfx, cfy)``; zeros for ``None``."""
        return control_spectrum(control_field, self.flow)


class RungeKuttaCrankNicolson:
    """S:24-116.  ``solve`` returns (final_state, trajectory) as torch tensors on the GPU."""

    def __init__(self, flow, dt: float, save_n: int, equation=None, dtype=torch.float32, device=None,
                 dealias=False, **kwargs):
        self.save_n, self.dt, self.flow, self.equation = save_n, dt, flow, equation
        self.dtype, self.device, self.dealias = dtype, device, dealias
        self._steppers = {}

    def _stepper(self, B, n_inner, n_outer):
        key = (B, n_inner, n_outer)
        if key not in self._steppers:
            self._steppers[key] = KolmogorovStepper(
                grid_size=self.flow.grid_size, Re=self.flow.Re, k=self.flow.k, dt=self.dt, n_inner=n_inner,
                n_outer=n_outer, num_envs=B, dtype=self.dtype, device=self.device, dealias=self.dealias,
                obs_size=self.flow.obs_size, min_action=-np.inf, max_action=np.inf,
                domain_x=self.flow.domain_x, domain_y=self.flow.domain_y)
        return self._steppers[key]

    def _prepare(self, state, n_inner, n_outer, control_field):
        cdtype = torch.complex64 if self.dtype == torch.float32 else torch.complex128
        if not isinstance(state, torch.Tensor):
            state = torch.from_numpy(np.asarray(state))
        batched = state.dim() == 3
        w = state if batched else state.unsqueeze(0)
        st = self._stepper(w.shape[0], n_inner, n_outer)
        w = w.to(device=st.device, dtype=cdtype).contiguous()
        # the control spectrum is loop-invariant (the reference recomputes it per sub-stage, XLA hoists it)
        st.set_control_field(None if control_field is None else control_spectrum(control_field, self.flow))
        return st, w, batched

    def RK4_CN(self, control_field=None):
        """S:32-59: returns ``time_step_fn(u) -> u`` advancing a state by ONE solver step (5 IMEX sub-stages) of
        size ``self.dt``; the input tensor is not modified."""
        def time_step_fn(u):
            st, w, batched = self._prepare(u, 1, 1, control_field)
            out = st.rk_steps(w.clone(), None, 1)
            return out if batched else out[0]
        return time_step_fn

    def step(self, flow=None, dt: float = None, save_n: int = 1, callbacks: Callable = None, control_field=None):
        """S:61-80: returns ``inner_scan(state) -> state`` applying ``save_n`` solver steps (the reference's
        ``lax.scan`` of ``RK4_CN`` with ``length=save_n``; ``dt`` is unused there too)."""
        n = int(save_n)

        def inner_scan(initialization):
            st, w, batched = self._prepare(initialization, 1, 1, control_field)
            out = st.rk_steps(w.clone(), None, n)
            return out if batched else out[0]
        return inner_scan

    def solve(self, dt: float, flow=None, t_span: Tuple[float, float] = (0.0, 10.0),
              callbacks: Iterable[Callable] = (), controller: Optional[Callable] = None, save_n: int = 1,
              initial_state=None, control_field=None):
        """S:82-116.  ``dt`` sets the step counts; the RK step itself uses ``self.dt`` (S:54-55).
        ``control_field`` is any physical ``(cfx(x, y), cfy(x, y))`` pair (K:259-272)."""
        flow = flow or self.flow
        n_inner, n_outer = constants.step_counts(dt, t_span[1], save_n)  # raises ValueError if end < 1
        if initial_state is None:
            initial_state = flow.initialize_state()
        st, w, batched = self._prepare(initial_state, n_inner, n_outer, control_field)
        final, traj = st.rollout(w, None)
        if not batched:
            final, traj = final[0], traj[0]
        flow.vorticity = traj  # S:110
        for cb in callbacks:   # S:114-115
            cb(flow)
        return final, traj
