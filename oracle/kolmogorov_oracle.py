"""NumPy restatement of the reference's Kolmogorov-flow env step (test oracle).

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.

PINNED against the reference's own source: ``tests/golden/make_reference_golden.py``
imports the UNMODIFIED reference files from ``/root/reference`` over NumPy stand-ins
for the absent third-party modules (jax / flax / gymnax / chex / tree_math,
``tests/golden/jax_numpy_shim.py``) and records their fp64 outputs in
``tests/golden/reference_shim_64_re100.npz``; ``tests/test_reference_shim.py`` holds
every function below to those vectors at round-off (<= 1e-12 per solver step,
<= 1e-10 after 3 960 solver steps).  Not pinned: real jaxlib/XLA floating-point
ordering (JAX itself is not installable here) -- fp64 round-off level only.

Every function cites the reference lines it follows, with the tags of SURVEY.md:
  K = hydrogym/jax/envs/kolmogorov.py      S = hydrogym/jax/solvers/base.py
  U = hydrogym/jax/utils/utils.py          E = hydrogym/jax/env_core.py
The arithmetic is written the way the reference writes it (cycles-per-unit
wavenumbers, ``(2*pi*1j)**2`` factors, endpoint-inclusive grid, no-op
de-aliasing), not the simplified integer-wavenumber form the kernels use.
All functions accept leading batch axes (FFTs act on the last two axes), which
is what ``jax.vmap`` (E:609-613) does to the reference.
"""
from __future__ import annotations

import numpy as np

# S:13-15
ALPHA_RK4 = [0, 0.1496590219993, 0.3704009573644, 0.6222557631345, 0.9582821306748, 1]
BETA_RK4 = [0, -0.4178904745, -1.192151694643, -1.697784692471, -1.514183444257]
GAMMA_RK4 = [0.1496590219993, 0.3792103129999, 0.8229550293869, 0.6994504559488, 0.1530572479681]


def _ctype(dtype):
    return np.complex64 if np.dtype(dtype) == np.float32 else np.complex128


def _rfft2(a):
    return np.fft.rfftn(a, axes=(-2, -1))


def _irfft2(a):
    return np.fft.irfftn(a, axes=(-2, -1))


class FlowConfig:
    """K:24-183 (grids, initial state, forcing)."""

    def __init__(self, k=4, Re=200, grid_size=(64, 64), domain_x=(0, 2 * np.pi),
                 domain_y=(0, 2 * np.pi), obs_size=8, dtype=np.float64):
        # K:33-38
        self.k, self.Re, self.grid_size = k, Re, tuple(grid_size)
        self.domain_x, self.domain_y, self.obs_size = domain_x, domain_y, obs_size
        self.dtype = np.dtype(dtype)

    @property
    def nu(self):  # K:152-154
        return 1 / self.Re

    def load_mesh(self):
        """K:46-57 -- endpoint-INCLUSIVE linspace."""
        x0, xn, nx = self.domain_x[0], self.domain_x[1], self.grid_size[1]
        y0, yn, ny = self.domain_y[0], self.domain_y[1], self.grid_size[0]
        x = np.linspace(x0, xn, nx, dtype=self.dtype)
        y = np.linspace(y0, yn, ny, dtype=self.dtype)
        return np.meshgrid(x, y, indexing="ij")

    def load_fft_mesh(self):
        """K:89-101 -- cycles-per-unit wavenumbers k_int/(2*pi)."""
        N, M = self.grid_size
        dx = self.domain_x[1] / N
        dy = self.domain_y[1] / M
        kx = np.fft.fftfreq(N, dx).astype(self.dtype)
        ky = np.fft.rfftfreq(M, dy).astype(self.dtype)
        return np.meshgrid(kx, ky, indexing="ij")

    def initialize_state(self):
        """K:103-129: omega = d/dx cos(x) - d/dy(-sin y) = -sin X + cos Y ; rfftn."""
        X, Y = self.load_mesh()
        du_dy = -np.cos(Y)  # grad of -sin(y)
        dv_dx = -np.sin(X)  # grad of cos(x)
        vorticity = (dv_dx - du_dy).astype(self.dtype)
        return _rfft2(vorticity).astype(_ctype(self.dtype))

    def forcing_function(self, k, x, y):
        """K:135-146."""
        return (np.sin(k * y).astype(self.dtype), np.zeros_like(y))


def dealiasing(advection_term, apply_intended=False):
    """U:72-86.  As written this is the IDENTITY: the two ``.at[].set`` results
    are discarded.  ``apply_intended=True`` gives what the docstring
    meant (opt-in, not reference behaviour)."""
    if not apply_intended:
        return advection_term
    n, m = advection_term.shape[-2], advection_term.shape[-1]
    kn, km = int(n // 2 * 2 // 3), int(m * 2 // 3)
    out = advection_term.copy()
    out[..., kn:2 * kn, :] = 0.0
    out[..., :, km:] = 0.0
    return out


def compute_velocity_fft(omega_hat, kx, ky):
    """U:53-69."""
    double_derivative = (2 * np.pi * 1j) ** 2 * (abs(kx) ** 2 + abs(ky) ** 2)
    double_derivative = double_derivative.astype(omega_hat.dtype)
    double_derivative[0, 0] = 1
    psi_hat = -1 * omega_hat / double_derivative
    vxhat = (2 * np.pi * 1j) * ky * psi_hat
    vyhat = (-1 * 2 * np.pi * 1j) * kx * psi_hat
    return vxhat.astype(omega_hat.dtype), vyhat.astype(omega_hat.dtype)


def compute_tke(omega_hat, kx, ky, n):
    """U:175-194 (batched over leading axes)."""
    uhat, vhat = compute_velocity_fft(omega_hat, kx, ky)
    u = _irfft2(uhat)
    v = _irfft2(vhat)
    U = np.mean(u, axis=(-2, -1), keepdims=True)
    V = np.mean(v, axis=(-2, -1), keepdims=True)
    avg_tke = 0.5 * (np.abs(u - U) ** 2 + np.abs(v - V) ** 2) * (1 / n**2)
    return np.sum(avg_tke, axis=(-2, -1))


def compute_energy_dissipation(omega_hat, kx, ky, nu, n):
    """U:145-172 (np.gradient == jnp.gradient: unit spacing, one-sided edges)."""
    uhat, vhat = compute_velocity_fft(omega_hat, kx, ky)
    ureal, vreal = _irfft2(uhat), _irfft2(vhat)
    du_dy = np.gradient(ureal, axis=1)
    dv_dx = np.gradient(vreal, axis=0)
    du_dx = np.gradient(ureal, axis=0)
    dv_dy = np.gradient(vreal, axis=1)
    avg_epsilon = 2 * nu * ((du_dx) ** 2 + (dv_dy) ** 2 + (du_dy + dv_dx) ** 2)
    return np.sum(avg_epsilon) / (4 * np.pi ** 2)


def compute_divergence(omega_hat, kx, ky):
    """U:213-229."""
    uhat, vhat = compute_velocity_fft(omega_hat, kx, ky)
    ureal, vreal = _irfft2(uhat), _irfft2(vhat)
    return np.gradient(ureal, axis=0) + np.gradient(vreal, axis=1)


def compute_real_velocity_point(uhat, vhat, x_idx, y_idx):
    """U:129-142."""
    ureal, vreal = _irfft2(uhat), _irfft2(vhat)
    return np.sqrt(np.abs(ureal[x_idx, y_idx]) ** 2 + np.abs(vreal[x_idx, y_idx]) ** 2)


def compute_energy_mode(uhat, vhat, kx, ky, n, m):
    """U:89-101 -- energy of one mode (kx, ky integer wavenumbers)."""
    return 0.5 * (np.abs(uhat[..., kx % n, ky % m]) ** 2 + np.abs(vhat[..., kx % n, ky % m]) ** 2) / float(n * m) ** 2


def compute_velocity_mode(uhat, vhat, kx, ky, n, m):
    """U:104-121 (kx, ky integer wavenumbers)."""
    return np.sqrt((np.abs(uhat[..., kx % n, ky % m]) ** 2 + np.abs(vhat[..., kx % n, ky % m]) ** 2) / np.float64(n * m) ** 2)


def compute_reward(omega_hat, kx, ky, nu, n, actions):
    """U:197-210: TKE minus the plain sum of the four actions."""
    return compute_tke(omega_hat, kx, ky, n) - (actions[0] + actions[1] + actions[2] + actions[3])


def energy_spectrum(omega_hat, kx, ky, nbins=None):
    """Radial sum of the per-mode energies of U:89-101 over the FULL plane: bin = round(|k|) on integer
    wavenumbers, the conjugate half (not stored by rfft) counted through weights 2 on 0 < ky < m/2.
    The binning is test infrastructure of this repo (the reference ships the per-mode quantity only)."""
    n = omega_hat.shape[-2]
    mh = omega_hat.shape[-1]
    m = 2 * (mh - 1)
    uhat, vhat = compute_velocity_fft(omega_hat, kx, ky)
    e = 0.5 * (np.abs(uhat) ** 2 + np.abs(vhat) ** 2).astype(np.float64) / float(n * m) ** 2
    p = np.rint(2 * np.pi * np.asarray(kx, dtype=np.float64))
    q = np.rint(2 * np.pi * np.asarray(ky, dtype=np.float64))
    wgt = np.where((q == 0) | (q == mh - 1), 1.0, 2.0)
    bins = np.floor(np.sqrt(p * p + q * q) + 0.5).astype(np.int64)
    nbins = int(nbins) if nbins is not None else n
    out = np.zeros(omega_hat.shape[:-2] + (nbins,), dtype=np.float64)
    keep = (bins < nbins) & (bins > 0)
    flat = (e * wgt)[..., keep]
    idx = bins[keep]
    for b in range(nbins):
        sel = idx == b
        if sel.any():
            out[..., b] = flat[..., sel].sum(axis=-1)
    return out


class PseudoSpectralNavierStokes2D:
    """K:193-292."""

    def __init__(self, flow: FlowConfig, dealias=False):
        self.flow = flow
        self.kx, self.ky = flow.load_fft_mesh()
        self.x, self.y = flow.load_mesh()
        self.dealias = dealias
        self.cdtype = _ctype(flow.dtype)
        self._forcing = self.forcing_term()

    def linear_terms(self, omega_hat):
        """K:209-211."""
        lin = self.flow.nu * (2j * np.pi) ** 2 * (self.kx**2 + self.ky**2)
        return lin.astype(self.cdtype) * omega_hat

    def implicit_timestep(self, omega_hat, time_step):
        """K:213-220."""
        double_derivative = ((2j * np.pi) ** 2 * (self.kx**2 + self.ky**2)).astype(self.cdtype)
        return 1 / (1 - time_step * self.flow.nu * double_derivative) * omega_hat

    def forcing_term(self):
        """K:274-292."""
        fx, fy = self.flow.forcing_function(k=self.flow.k, x=self.x, y=self.y)
        fx_hat, fy_hat = _rfft2(fx), _rfft2(fy)
        derivative_term = 2j * np.pi
        return (derivative_term * (fy_hat * self.kx - fx_hat * self.ky)).astype(self.cdtype)

    def control_term(self, omega_hat, control_field=None):
        """K:259-272 (no 2*pi*i factor)."""
        if control_field is None:
            return np.zeros_like(omega_hat)
        cfx, cfy = control_field
        cfx_hat = _rfft2(cfx)
        cfy_hat = _rfft2(cfy)
        return (self.kx * cfy_hat - self.ky * cfx_hat).astype(self.cdtype)

    def nonlinear_terms(self, omega_hat, control_hat=None):
        """K:222-257.  ``control_hat`` is the loop-invariant value of
        ``control_term`` (hoisted; XLA does the same), or None."""
        kx, ky = self.kx, self.ky
        uhat, vhat = compute_velocity_fft(omega_hat, kx, ky)  # K:236-241
        u, v = _irfft2(uhat), _irfft2(vhat)  # K:243
        grad_x_hat = (2j * np.pi * kx * omega_hat).astype(self.cdtype)  # K:245
        grad_y_hat = (2j * np.pi * ky * omega_hat).astype(self.cdtype)  # K:246
        grad_x, grad_y = _irfft2(grad_x_hat), _irfft2(grad_y_hat)  # K:247
        advection = -(grad_x * u + grad_y * v)  # K:249
        advection_hat = _rfft2(advection).astype(self.cdtype)  # K:250
        advection_hat = dealiasing(advection_hat, self.dealias)  # K:254
        terms = advection_hat + self._forcing  # K:256
        if control_hat is not None:
            terms = terms + control_hat
        return terms


def step_counts(dt, end_time, save_n=1):
    """S:100-103 with Python float floor-division."""
    step_to_save = int(save_n // dt)
    total_steps = int(end_time // dt)
    outer_steps = int(total_steps // step_to_save)
    return step_to_save, outer_steps


class RungeKuttaCrankNicolson:
    """S:24-116."""

    def __init__(self, flow, dt, save_n, equation):
        self.flow, self.dt, self.save_n, self.equation = flow, dt, save_n, equation

    def time_step(self, u, control_hat=None):
        """S:50-57 -- one RK solver step (5 IMEX sub-stages)."""
        eq = self.equation
        h = 0
        for k in range(5):
            h = eq.nonlinear_terms(u, control_hat) + BETA_RK4[k] * h
            mu = 0.5 * self.dt * (ALPHA_RK4[k + 1] - ALPHA_RK4[k])
            yn = u + GAMMA_RK4[k] * self.dt * h + mu * eq.linear_terms(u)
            u = eq.implicit_timestep(yn, mu).astype(eq.cdtype)
        return u

    def rk_steps(self, u, n, control_hat=None):
        for _ in range(n):
            u = self.time_step(u, control_hat)
        return u

    def solve(self, dt, t_span, save_n=1, initial_state=None, control_field=None):
        """S:82-116.  Returns (final_state, trajectory[outer, ...]); snapshot j is
        the state after (j+1)*step_to_save RK steps.  Note the RK
        closure uses ``self.dt`` (fixed at construction, F6), ``dt`` only sets
        the step counts."""
        end_time = t_span[1]
        if end_time < 1:
            raise ValueError(
                "This flow configuration requires the end time to be at least 1. "
                "Please adjust t_span and run again.")
        u = self.flow.initialize_state() if initial_state is None else initial_state
        step_to_save, outer_steps = step_counts(dt, end_time, save_n)
        control_hat = None
        if control_field is not None:
            control_hat = self.equation.control_term(u, control_field)
        outputs = []
        for _ in range(outer_steps):
            u = self.rk_steps(u, step_to_save, control_hat)
            outputs.append(u)
        # stack on the axis right before (N, Mh) so batched states give (B, T, N, Mh)
        return u, np.stack(outputs, axis=-3)


class Params:
    """K:310-331 defaults."""

    def __init__(self, **kw):
        self.min_action, self.max_action = -0.5, 0.5
        self.min_obs, self.max_obs = -np.inf, np.inf
        self.dt, self.action_time, self.save_time = 1e-3, 10.0, 1
        self.k1, self.k2, self.k3, self.k4 = 4, 5, 6, 7
        self.action_dim, self.obs_dim = 4, 64
        self.max_episode_steps, self.reward_alpha = 1000, 0.0
        self.__dict__.update(kw)

    def replace(self, **kw):
        d = dict(self.__dict__)
        d.update(kw)
        return Params(**d)


class KolmogorovFlow:
    """K:334-525.  ``env_config`` understands ``dt`` (K:348) and, as an
    oracle-only convenience for short test horizons, ``action_time`` and
    ``save_time`` (the reference fixes these to 10.0 / 1 via default_params)."""

    def __init__(self, env_config=None, flow_config=None, dtype=np.float64, dealias=False):
        self.env_config = env_config or {}
        self.flow = FlowConfig(dtype=dtype, **(flow_config or {}))
        self.n, self.m = self.flow.grid_size
        self.x, self.y = self.flow.load_mesh()
        self.kx, self.ky = self.flow.load_fft_mesh()
        default = self.default_params
        self.equation = PseudoSpectralNavierStokes2D(self.flow, dealias=dealias)
        self.integrator = RungeKuttaCrankNicolson(
            self.flow, float(default.dt), int(default.save_time), self.equation)

    @property
    def default_params(self):
        """K:363-369."""
        kw = dict(action_dim=4, obs_dim=self.flow.obs_size**2)
        for key in ("dt", "action_time", "save_time"):
            if self.env_config.get(key) is not None:
                kw[key] = self.env_config[key]
        return Params(**kw)

    def _clip_action(self, action, params):
        """E:419-420."""
        return np.clip(action, params.min_action, params.max_action)

    def is_terminal(self, state, params):
        """E:422-423."""
        return np.logical_or(state["terminal"], state["time"] >= params.max_episode_steps)

    def _control_field(self, action, params):
        """K:386-395; ``action`` is (4,) or (B,4)."""
        a = np.asarray(action, dtype=self.flow.dtype)[..., None, None]
        y = self.y
        forcing_x = (a[..., 0, :, :] * np.sin(params.k1 * y) + a[..., 1, :, :] * np.sin(params.k2 * y)
                     + a[..., 2, :, :] * np.sin(params.k3 * y) + a[..., 3, :, :] * np.sin(params.k4 * y))
        forcing_x = forcing_x.astype(self.flow.dtype)
        return forcing_x, np.zeros_like(forcing_x)

    def _rollout(self, omega_hat0, params, control_field=None):
        """K:397-420 -- dt/save_time/action_time from DEFAULT params (F6)."""
        default = self.default_params
        return self.integrator.solve(
            dt=float(default.dt), t_span=(0.0, float(default.action_time)),
            save_n=int(default.save_time), initial_state=omega_hat0, control_field=control_field)

    def _trajectory_mean_obs(self, trajectory):
        """K:426-446; trajectory (..., T, N, Mh) -> (..., obs_size**2)."""
        stride_x = max(1, self.n // self.flow.obs_size)
        stride_y = max(1, self.m // self.flow.obs_size)
        uhat, vhat = compute_velocity_fft(trajectory, self.kx, self.ky)
        ureal, vreal = _irfft2(uhat), _irfft2(vhat)
        u_s = ureal[..., ::stride_x, ::stride_y]
        v_s = vreal[..., ::stride_x, ::stride_y]
        obs = np.sqrt(np.abs(u_s) ** 2 + np.abs(v_s) ** 2)
        obs = obs.reshape(obs.shape[:-2] + (-1,))
        return np.mean(obs, axis=-2)

    def get_obs(self, state, params=None, key=None):
        """K:448-454."""
        return self._trajectory_mean_obs(state["trajectory"])

    def _avg_tke(self, trajectory):
        """K:456-460."""
        return np.mean(compute_tke(trajectory, self.kx, self.ky, self.n), axis=-1)

    def _reward(self, action, trajectory, params):
        """K:462-470."""
        energy = self._avg_tke(trajectory)
        action_penalty = np.sum(np.abs(action), axis=-1)
        return -(params.reward_alpha * energy + action_penalty)

    def reset_env(self, key=None, params=None):
        """K:472-492 -- key unused; full uncontrolled rollout (F7)."""
        params = params or self.default_params
        omega0 = self.flow.initialize_state()
        final_state, trajectory = self._rollout(omega0, params, None)
        state = dict(trajectory=trajectory, omega_hat=final_state, time=np.array(0),
                     terminal=np.array(False))
        return self.get_obs(state, params, key), state

    def step_env(self, key, state, action, params=None):
        """K:494-525."""
        params = params or self.default_params
        action = self._clip_action(np.asarray(action, dtype=self.flow.dtype), params)
        control_field = self._control_field(action, params)
        final_state, trajectory = self._rollout(state["omega_hat"], params, control_field)
        next_state = dict(trajectory=trajectory, omega_hat=final_state,
                          time=state["time"] + 1, terminal=np.array(False))
        obs = self.get_obs(next_state, params, key)
        reward = self._reward(action, trajectory, params)
        done = self.is_terminal(next_state, params)
        info = {"discount": np.where(done, 0.0, 1.0), "mean_tke": self._avg_tke(trajectory)}
        return obs, next_state, reward, done, info
