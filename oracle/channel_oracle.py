"""NumPy restatement of the reference's 3-D channel-flow environment step (TEST INFRASTRUCTURE).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this file; the product
(`hydrogym_b200/`) never does.  Every function cites the reference lines it follows:

    C = hydrogym/jax/envs/channel.py      S = hydrogym/jax/solvers/base.py      U = hydrogym/jax/utils/utils.py

Pinned against the reference's OWN source (imported unmodified over the NumPy stand-ins of
`tests/golden/jax_numpy_shim.py`) by `tests/golden/make_reference_channel.py` ->
`tests/golden/reference_shim_channel.npz`, checked in `tests/test_channel_oracle.py`.

The path: Fourier x Fourier x Chebyshev pseudo-spectral DNS on 72^3 points, state = (u, v, w) FFT'd over (x, y) and
physical (collocation) in z; classical RK4 of `rhs = nonlinear + linear`, then a pressure projection with precomputed
per-mode inverses, then a constant-mass-flux correction; wall jets (24 actuators) enter through `enforce_noslip`.
One environment step = 50 such steps; reward = -(wall shear stress); observation = (U, W) at 8 x 8 probes of the
plane z-index 9.  `step_env` also differentiates the shear stress w.r.t. the action and DISCARDS the gradient
(C:497-500, 502: only `wss` and the fields are used), so the forward pass is the whole observable behaviour.
"""
from __future__ import annotations

import numpy as np

NX = NY = NZ = 72          # C:420-422 (hard-coded in the env, the params' Nx.. are only used for the probe grid)
LX, LY, LZ = 2.0 * np.pi, 1.0 * np.pi, 2.0   # C:416-418
NU = 1.9e-3                # C:419
DT = 2e-4                  # C:44
NSTEPS = 50                # C:43
ACTION_DIM = 24            # C:34
K_DET = 9                  # C:37
OBS_SUBSAMPLE = 8          # C:35
MAX_STEPS = 5000           # C:47
FX = 2.0                   # C:541 (streamwise body force)
TARGET_BULK_U = 8.0        # C:550


def dealias_mask_2_3(nx, ny):
    """U:8-19: 1 where |kx| <= nx//3 and |ky| <= ny//3 (integer wavenumbers in fftfreq order)."""
    kx = np.fft.fftfreq(nx) * nx
    ky = np.fft.fftfreq(ny) * ny
    return ((np.abs(kx) <= nx // 3)[:, None] & (np.abs(ky) <= ny // 3)[None, :]).astype(np.float32)


def cheb_D_matrices(n, lz):
    """U:22-50: Chebyshev-Gauss-Lobatto points ascending on [-1, 1], first-derivative matrix with the
    negative-sum-trick diagonal, second derivative = D @ D, both scaled by 2/lz; z = lz/2 * x."""
    k = np.arange(n)
    x = np.cos(np.pi * k / (n - 1))[::-1]
    c = np.ones(n)
    c[0] = c[-1] = 2.0
    c = c * (-1.0) ** np.arange(n)
    dx = x[:, None] - x[None, :]
    d = (c[:, None] / c[None, :]) / (dx + np.eye(n))
    d = d - np.diag(d.sum(axis=1))
    s = 2.0 / lz
    return (lz / 2.0) * x, s * d, (s * s) * (d @ d)


def synthetic_initial_fields(dtype=np.float64):
    """Deterministic stand-in for the HF-hosted `U/V/W.npy` (Channel_3D_Retau180/initial_field, unavailable here:
    no network): a flat turbulent-like mean profile with bulk mean 8 plus smooth O(0.5) three-dimensional
    perturbations that vanish at the walls.  Used for the fixture, the tests and the bench alike."""
    z = cheb_D_matrices(NZ, LZ)[0]
    x = np.arange(NX) * LX / NX
    y = np.arange(NY) * LY / NY
    X, Y, Z = np.meshgrid(x, y, z, indexing="ij")
    env = 1.0 - Z ** 2
    prof = 1.0 - Z ** 8
    U = 8.0 * prof / prof.mean() + 0.6 * env * np.sin(2 * X) * np.cos(2 * Y) * np.cos(0.5 * np.pi * Z)
    V = 0.5 * env * np.cos(X + 2 * Y) * (1.0 + Z)
    W = 0.3 * env ** 2 * np.sin(3 * X) * np.sin(2 * Y)
    return U.astype(dtype), V.astype(dtype), W.astype(dtype)


class VelocityState:
    """S:17-21 (a NamedTuple of three arrays there)."""

    def __init__(self, u, v, w):
        self.u, self.v, self.w = u, v, w

    def __iter__(self):
        return iter((self.u, self.v, self.w))


class PseudoSpectralNavierStokes3D:
    """C:103-372."""

    def __init__(self, nx=NX, ny=NY, nz=NZ, lx=LX, ly=LY, lz=LZ, nu=NU, dtype=np.float64):
        self.Nx, self.Ny, self.Nz, self.Lx, self.Ly, self.Lz, self.nu, self.dtype = nx, ny, nz, lx, ly, lz, nu, dtype
        kx = (np.fft.fftfreq(nx, d=lx / nx) * 2.0 * np.pi).astype(dtype)          # C:114-118
        ky = (np.fft.fftfreq(ny, d=ly / ny) * 2.0 * np.pi).astype(dtype)
        self.kx, self.ky = kx, ky
        self.ikx = (1j * kx)[:, None, None]                                        # C:119-120
        self.iky = (1j * ky)[None, :, None]
        self.k2 = kx[:, None] ** 2 + ky[None, :] ** 2                              # C:121
        z, dz, dzz = cheb_D_matrices(nz, lz)                                       # C:123-125
        self.z, self.Dz, self.Dzz = z, dz.astype(dtype), dzz.astype(dtype)
        self.dealias = dealias_mask_2_3(nx, ny)[:, :, None].astype(dtype)          # C:126
        # C:128-148: per-mode pressure operator Dzz - k2 I with Neumann rows (Dz[0], Dz[-1]) and, for the mean
        # mode, the middle row replaced by a Dirichlet pin; inverted once
        A = self.Dzz[None] - self.k2.reshape(-1)[:, None, None] * np.eye(nz, dtype=dtype)[None]
        A[:, 0, :] = self.Dz[0][None]
        A[:, -1, :] = self.Dz[-1][None]
        mid = nz // 2
        A[0, mid, :] = 0.0
        A[0, mid, mid] = 1.0
        self.A_inv = np.linalg.inv(A)

    # -- transforms / derivatives ------------------------------------------------------------------------
    def fft_xy(self, f):                      # C:150-151
        return np.fft.fftn(f, axes=(0, 1))

    def ifft_xy(self, f_hat):                 # C:153-154
        return np.fft.ifftn(f_hat, axes=(0, 1)).real

    def to_spectral(self, s):                 # C:156-161
        return VelocityState(self.fft_xy(s.u), self.fft_xy(s.v), self.fft_xy(s.w))

    def to_physical(self, s):                 # C:163-168
        return VelocityState(self.ifft_xy(s.u), self.ifft_xy(s.v), self.ifft_xy(s.w))

    def dz_phys(self, f):                     # C:176-177
        return np.einsum("ij,xyj->xyi", self.Dz, f)

    def dzz_phys(self, f):                    # C:179-180
        return np.einsum("ij,xyj->xyi", self.Dzz, f)

    # -- boundary conditions ------------------------------------------------------------------------------
    def jet_pattern(self, vjets, slit_length_x=3.0, slit_width_y=2.0, x_span_frac=2 / 3, nx_jets=6, ny_jets=4):
        """The zero-mean (x, y) blowing/suction pattern of C:194-217: 6 x 4 Gaussian slits with amplitudes
        `vjets` (24,), centres on a regular lattice of grid INDICES."""
        nx, ny = self.Nx, self.Ny
        x_max = int(nx * x_span_frac)
        xc = np.linspace(3.0, float(x_max - 4), nx_jets)
        yc = np.linspace(4.0, float(ny - 5), ny_jets)
        X = np.arange(nx)[:, None]
        Y = np.arange(ny)[None, :]
        dx2 = (X[None, None] - xc[:, None, None, None]) ** 2
        dy2 = (Y[None, None] - yc[None, :, None, None]) ** 2
        jets = np.exp(-0.5 * (dx2 / slit_length_x ** 2 + dy2 / slit_width_y ** 2))
        mask = np.sum(np.asarray(vjets).reshape(nx_jets, ny_jets, 1, 1) * jets, axis=(0, 1))
        return mask - mask.mean()

    def apply_jets_v(self, v, vjets, z0=1, jet_thickness=5):
        """C:182-225: walls zeroed, then the pattern written into the planes z0 .. z0+thickness-1."""
        v = v.copy()
        v[:, :, 0] = 0.0
        v[:, :, -1] = 0.0
        mask = self.jet_pattern(vjets)
        for k in range(z0, min(z0 + jet_thickness, self.Nz - 1)):
            v[:, :, k] = mask
        return v

    @staticmethod
    def jet_gain(t, action_time=50.0):
        """C:231-234: 0.3, ramped linearly over the first and last 10 time units of `action_time` (t is the
        step COUNT the env passes, C:538; beyond action_time the ramp goes negative -- kept as is)."""
        gain = 0.3
        if t < 10:
            gain = gain * (t * 0.1)
        if t >= action_time - 10:
            gain = gain * ((action_time - t) * 0.1)
        return gain

    def enforce_noslip(self, u, v, w, action=None, action_time=50.0, t=0.0):
        """C:227-241."""
        u, v = u.copy(), v.copy()
        u[:, :, 0] = 0.0
        u[:, :, -1] = 0.0
        v[:, :, 0] = 0.0
        v[:, :, -1] = 0.0
        if action is not None:
            w = self.apply_jets_v(w, -self.jet_gain(t, action_time) * np.asarray(action))
        else:
            w = w.copy()
            w[:, :, 0] = 0.0
            w[:, :, -1] = 0.0
        return u, v, w

    # -- right-hand side ----------------------------------------------------------------------------------
    def nonlinear_terms(self, s_hat, action=None, t=0.0, fx=0.0, fy=0.0, fz=0.0):
        """C:243-284: advective form u . grad(u) of the wall-corrected field, x / y derivatives spectral, z
        derivative by the collocation matrix, 2/3 mask on the transformed products, uniform body force."""
        p = self.to_physical(s_hat)
        u, v, w = self.enforce_noslip(p.u, p.v, p.w, action=action, t=t)
        uh, vh, wh = self.fft_xy(u), self.fft_xy(v), self.fft_xy(w)
        ux, uy = self.ifft_xy(self.ikx * uh), self.ifft_xy(self.iky * uh)
        vx, vy = self.ifft_xy(self.ikx * vh), self.ifft_xy(self.iky * vh)
        wx, wy = self.ifft_xy(self.ikx * wh), self.ifft_xy(self.iky * wh)
        uz, vz, wz = self.dz_phys(u), self.dz_phys(v), self.dz_phys(w)
        n_u = u * ux + v * uy + w * uz
        n_v = u * vx + v * vy + w * vz
        n_w = u * wx + v * wy + w * wz

        def forcing_hat(f):
            f_phys = np.ones((self.Nx, self.Ny, self.Nz), dtype=self.dtype) * f if np.ndim(f) == 0 else f
            return self.fft_xy(f_phys)

        return VelocityState(-self.fft_xy(n_u) * self.dealias + forcing_hat(fx),
                             -self.fft_xy(n_v) * self.dealias + forcing_hat(fy),
                             -self.fft_xy(n_w) * self.dealias + forcing_hat(fz))

    def linear_terms(self, s_hat, action=None, t=0.0):
        """C:286-300: nu * (-k2 * state + fft(Dzz wall-corrected field)) -- the k2 part acts on the UNcorrected
        spectral state."""
        p = self.to_physical(s_hat)
        u, v, w = self.enforce_noslip(p.u, p.v, p.w, action=action, t=t)
        k2 = self.k2[:, :, None]
        return VelocityState(self.nu * (-k2 * s_hat.u + self.fft_xy(self.dzz_phys(u))),
                             self.nu * (-k2 * s_hat.v + self.fft_xy(self.dzz_phys(v))),
                             self.nu * (-k2 * s_hat.w + self.fft_xy(self.dzz_phys(w))))

    def rhs(self, s_hat, action=None, t=0.0, fx=0.0, fy=0.0, fz=0.0):
        """C:302-309."""
        n = self.nonlinear_terms(s_hat, action=action, t=t, fx=fx, fy=fy, fz=fz)
        l = self.linear_terms(s_hat, action=action, t=t)
        return VelocityState(n.u + l.u, n.v + l.v, n.w + l.w)

    def project(self, s_hat, dt, action=None, t=0.0):
        """C:311-372: pressure Poisson solve per (kx, ky) mode with Neumann data w/dt at both walls, velocity
        correction, walls re-imposed in physical space."""
        u_hat, v_hat, w_hat = s_hat.u, s_hat.v, s_hat.w
        w = self.ifft_xy(w_hat)
        div = self.ifft_xy(self.ikx * u_hat) + self.ifft_xy(self.iky * v_hat) + self.dz_phys(w)
        rhs = (self.fft_xy(div) / dt).reshape(self.Nx * self.Ny, self.Nz).copy()
        rhs[:, 0] = self.fft_xy(w[:, :, 0] / dt).reshape(-1)
        rhs[:, -1] = self.fft_xy(w[:, :, -1] / dt).reshape(-1)
        rhs[0, self.Nz // 2] = 0.0
        p_hat = np.einsum("mij,mj->mi", self.A_inv, rhs).reshape(self.Nx, self.Ny, self.Nz)
        u_new = u_hat - dt * (self.ikx * p_hat)
        v_new = v_hat - dt * (self.iky * p_hat)
        w_new = w_hat - dt * self.fft_xy(self.dz_phys(self.ifft_xy(p_hat)))
        u, v, w = self.enforce_noslip(self.ifft_xy(u_new), self.ifft_xy(v_new), self.ifft_xy(w_new), action=action, t=t)
        return VelocityState(self.fft_xy(u), self.fft_xy(v), self.fft_xy(w))


class RungeKutta4:
    """S:119-185."""

    def __init__(self, equation, dt, save_n=1):
        self.equation, self.dt, self.save_n = equation, dt, int(save_n)

    def rk4_step(self, s, dt=None, action=None, t=0.0, fx=0.0, fy=0.0, fz=0.0, enforce_const_massflux=True,
                 target_bulk_u=8.0):
        eq = self.equation
        dt = self.dt if dt is None else dt

        def axpy(a, b, alpha):
            return VelocityState(a.u + alpha * b.u, a.v + alpha * b.v, a.w + alpha * b.w)

        k1 = eq.rhs(s, action=action, t=t, fx=fx, fy=fy, fz=fz)
        k2 = eq.rhs(axpy(s, k1, 0.5 * dt), action=action, t=t + 0.5 * dt, fx=fx, fy=fy, fz=fz)
        k3 = eq.rhs(axpy(s, k2, 0.5 * dt), action=action, t=t + 0.5 * dt, fx=fx, fy=fy, fz=fz)
        k4 = eq.rhs(axpy(s, k3, dt), action=action, t=t + dt, fx=fx, fy=fy, fz=fz)
        star = VelocityState(s.u + (dt / 6.0) * (k1.u + 2 * k2.u + 2 * k3.u + k4.u),
                             s.v + (dt / 6.0) * (k1.v + 2 * k2.v + 2 * k3.v + k4.v),
                             s.w + (dt / 6.0) * (k1.w + 2 * k2.w + 2 * k3.w + k4.w))
        new = eq.project(star, dt, action=action, t=t + dt)
        if enforce_const_massflux:   # S:166-183: shift u so that its grid mean is the target, walls again
            p = eq.to_physical(new)
            delta = target_bulk_u - np.mean(p.u)
            u, v, w = eq.enforce_noslip(p.u + delta, p.v, p.w, action=action, t=t + dt)
            new = eq.to_spectral(VelocityState(u, v, w))
        return new


def make_obs_grid_indices(nx, ny, n):
    """C:69-73."""
    xs = np.linspace(0, nx - 1, n).astype(np.int64)
    ys = np.linspace(0, ny - 1, n).astype(np.int64)
    return np.meshgrid(xs, ys, indexing="ij")


def wss_compute(k, U, nu=NU, z=None):
    """C:88-92: one-sided difference of U between plane k and the wall, averaged over the wall."""
    return np.mean(nu * (U[:, :, k] - U[:, :, 0]) / (z[k] - z[0]))


class ChannelFlowSpectralEnv:
    """C:406-571 with the state as a plain dict {time, U, V, W, dt, terminal}."""

    def __init__(self, initial_fields=None, dtype=np.float64, nsteps=NSTEPS):
        self.dtype, self.nsteps = dtype, int(nsteps)
        self.equation = PseudoSpectralNavierStokes3D(dtype=dtype)
        self.integrator = RungeKutta4(self.equation, DT, 1)
        U0, V0, W0 = initial_fields if initial_fields is not None else synthetic_initial_fields(dtype)
        self.U0, self.V0, self.W0 = (np.asarray(a).astype(dtype) for a in (U0, V0, W0))
        self.Xi, self.Yi = make_obs_grid_indices(NX, NY, OBS_SUBSAMPLE)

    def get_obs(self, state):
        """C:76-85: (U, W) at the 8 x 8 probe lattice of the plane k_det, U block first."""
        return np.stack([state["U"][:, :, K_DET][self.Xi, self.Yi], state["W"][:, :, K_DET][self.Xi, self.Yi]],
                        axis=0).reshape(-1)

    def reset_env(self):
        """C:471-481."""
        state = dict(time=0, U=self.U0, V=self.V0, W=self.W0, dt=DT, terminal=False)
        return self.get_obs(state), state

    def rollout(self, state, action):
        """C:531-567 (`_wss_with_aux`): nsteps RK4 steps from the physical fields, t = state.time + n."""
        eq = self.equation
        s = eq.to_spectral(VelocityState(state["U"], state["V"], state["W"]))
        for n in range(self.nsteps):
            s = self.integrator.rk4_step(s, dt=state["dt"], action=action, t=state["time"] + n, fx=FX,
                                         target_bulk_u=TARGET_BULK_U)
        p = eq.to_physical(s)
        return wss_compute(4, p.u, z=eq.z), (p.u, p.v, p.w)

    def step_env(self, state, action):
        """C:486-519: clip, roll out, reward = -wss, time += nsteps (the PARAMS' value, 50), terminal at the
        horizon."""
        action = np.clip(np.asarray(action, dtype=np.float64), -1.0, 1.0)
        wss, (U1, V1, W1) = self.rollout(state, action)
        time = state["time"] + NSTEPS
        terminal = time >= MAX_STEPS
        nxt = dict(time=time, U=U1, V=V1, W=W1, dt=state["dt"], terminal=terminal)
        done = bool(terminal)
        return self.get_obs(nxt), nxt, -wss, done, {"discount": 0.0 if done else 1.0}
