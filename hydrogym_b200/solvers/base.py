"""Direct-solver surface: ``RungeKuttaCrankNicolson.solve`` as the docs/notebook call it.

Reference: hydrogym/jax/solvers/base.py:24-116 (integrator) and
hydrogym/jax/envs/kolmogorov.py:193-292 (equation).  The arithmetic runs in the same
sm_100a kernels as the environment step; this module is host glue only.
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Tuple

import numpy as np
import torch

from .. import constants
from ..solver import KolmogorovStepper

# hydrogym/jax/solvers/base.py:13-15 (the kernels hold their own copy; exported for callers)
_alpha_RK4 = [0, 0.1496590219993, 0.3704009573644, 0.6222557631345, 0.9582821306748, 1]
_beta_RK4 = [0, -0.4178904745, -1.192151694643, -1.697784692471, -1.514183444257]
_gammas_RK4 = [0.1496590219993, 0.3792103129999, 0.8229550293869, 0.6994504559488, 0.1530572479681]


class PseudoSpectralNavierStokes2D:
    """Grid/constant holder mirroring K:193-292; the terms themselves are fused into the kernels."""

    def __init__(self, flow):
        self.flow = flow
        self.grid = flow.load_fft_mesh()
        self.real_grid = flow.load_mesh("name")
        self.kx, self.ky = self.grid
        self.x, self.y = self.real_grid

    def forcing_term(self):
        """K:274-292 as a dense (N, Mh) complex128 array (non-zero on row kx=0 only)."""
        N, M = self.flow.grid_size
        out = np.zeros((N, M // 2 + 1), dtype=np.complex128)
        out[0] = constants.forcing_row(self.flow.k, self.flow.grid_size, self.flow.domain_y)
        return out

    def control_term(self, omega_hat=None, control_field=None):
        """K:259-272 for an x-independent ``(cfx, 0)`` control field."""
        N, M = self.flow.grid_size
        out = np.zeros((N, M // 2 + 1), dtype=np.complex128)
        if control_field is not None:
            out[0] = _control_row(control_field, self.flow)
        return out


def _control_row(control_field, flow):
    cfx, cfy = control_field
    cfx = np.asarray(cfx.cpu() if isinstance(cfx, torch.Tensor) else cfx, dtype=np.float64)
    cfy = np.asarray(cfy.cpu() if isinstance(cfy, torch.Tensor) else cfy, dtype=np.float64)
    if np.any(cfy != 0) or np.any(cfx != cfx[:1]):
        raise NotImplementedError(
            "only x-independent control fields (cfx(y), 0) are supported -- the form KolmogorovFlow._control_field "
            "produces (kolmogorov.py:386-395)")
    M = flow.grid_size[1]
    ky = np.fft.rfftfreq(M, flow.domain_y[1] / M)
    return -ky * np.fft.rfftn(cfx)[0]


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

    def solve(self, dt: float, flow=None, t_span: Tuple[float, float] = (0.0, 10.0),
              callbacks: Iterable[Callable] = (), controller: Optional[Callable] = None, save_n: int = 1,
              initial_state=None, control_field=None):
        """S:82-116.  ``dt`` sets the step counts; the RK step itself uses ``self.dt`` (S:54-55)."""
        flow = flow or self.flow
        n_inner, n_outer = constants.step_counts(dt, t_span[1], save_n)  # raises ValueError if end < 1
        cdtype = torch.complex64 if self.dtype == torch.float32 else torch.complex128
        if initial_state is None:
            initial_state = flow.initialize_state()
        if not isinstance(initial_state, torch.Tensor):
            initial_state = torch.from_numpy(np.asarray(initial_state))
        batched = initial_state.dim() == 3
        w = initial_state if batched else initial_state.unsqueeze(0)
        st = self._stepper(w.shape[0], n_inner, n_outer)
        w = w.to(device=st.device, dtype=cdtype).contiguous()
        actions = None
        if control_field is not None:
            # G = F + 1 * row: put the control row in slot 0 and drive it with a unit action
            rows = np.zeros((4, st.Mh), dtype=np.complex128)
            rows[0] = _control_row(control_field, flow)
            st.set_control_rows(rows)
            actions = torch.zeros((w.shape[0], 4), dtype=self.dtype, device=st.device)
            actions[:, 0] = 1.0
        final, traj = st.rollout(w, actions)
        if not batched:
            final, traj = final[0], traj[0]
        flow.vorticity = traj  # S:110
        for cb in callbacks:   # S:114-115
            cb(flow)
        return final, traj
