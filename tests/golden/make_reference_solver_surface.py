"""Golden vectors for the equation / integrator METHODS the reference documents, from the reference's own source
(run in the build container, from the repo root):

    python tests/golden/make_reference_solver_surface.py      # writes reference_shim_solver_surface.npz

Same route as make_reference_golden.py: `/root/reference/hydrogym/jax/envs/kolmogorov.py` and `solvers/base.py` are
imported UNMODIFIED over the NumPy stand-ins of jax_numpy_shim.py (fp64, 64^2, Re = 100).  Unlike the cfg-1 vectors
(three dominant modes), everything here starts from a BROADBAND random state, and the control field is a general
x-dependent pair (cfx(x, y), cfy(x, y)) with cfy != 0 -- the case `KolmogorovFlow._control_field` never produces:

* `PseudoSpectralNavierStokes2D.linear_terms`, `implicit_timestep`, `nonlinear_terms` (without / with the field),
  `control_term` of the field                                  (hydrogym/jax/envs/kolmogorov.py:209-272);
* `RungeKuttaCrankNicolson.RK4_CN(control_field)` x 10, `.step(..., save_n=5, control_field)` and
  `.solve(dt=0.05, t_span=(0, 2), control_field=...)`             (hydrogym/jax/solvers/base.py:32-116).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import jax_numpy_shim  # noqa: E402

FLOW = dict(Re=100, grid_size=(64, 64))


def random_state(n=64, seed=77):
    """Smooth broadband random omega_hat (n, n/2+1), not Hermitian in the ky = 0 / n/2 columns (any complex array is
    a legal argument of the reference's methods)."""
    rng = np.random.default_rng(seed)
    mh = n // 2 + 1
    p = np.abs(np.fft.fftfreq(n, 1.0 / n))[:, None]
    q = np.arange(mh)[None, :]
    env = np.exp(-(p * p + q * q) / 64.0) * n * n * 0.1
    return (rng.normal(size=(n, mh)) + 1j * rng.normal(size=(n, mh))) * env


def control_field(x, y):
    cfx = 0.3 * np.sin(2 * x) * np.cos(3 * y) + 0.1 * np.cos(x)
    cfy = -0.2 * np.cos(x + 2 * y) + 0.05 * np.sin(5 * y)
    return cfx, cfy


def main():
    jax_numpy_shim.install()
    import jax.numpy as jnp

    from hydrogym.jax.envs.kolmogorov import KolmogorovFlow

    env = KolmogorovFlow(None, FLOW)
    eq, integ = env.equation, env.integrator
    w = jnp.asarray(random_state())
    cf = tuple(jnp.asarray(a) for a in control_field(env.x, env.y))
    out = {"w_rand": w, "cfx": cf[0], "cfy": cf[1]}
    out["control_term_general"] = eq.control_term(w, cf)
    out["linear_terms_rand"] = eq.linear_terms(w)
    out["implicit_rand_mu"] = eq.implicit_timestep(w, 7.5e-5)
    out["nonlinear_terms_rand"] = eq.nonlinear_terms(w, None)
    out["nonlinear_terms_rand_ctrl"] = eq.nonlinear_terms(w, cf)
    step = integ.RK4_CN(control_field=cf)
    u = w
    for _ in range(10):
        u = step(u)
    out["rk10_general_ctrl"] = u
    out["step_save5_general_ctrl"] = integ.step(env.flow, 1e-3, 5, [], control_field=cf)(w)
    fin, traj = integ.solve(dt=0.05, flow=env.flow, t_span=(0, 2.0), save_n=1, initial_state=w, control_field=cf)
    out["solve_general_final"], out["solve_general_traj"] = fin, traj
    path = os.path.join(HERE, "reference_shim_solver_surface.npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes;", {k: np.asarray(v).shape for k, v in out.items()})
    print("finite:", all(np.isfinite(np.asarray(v)).all() for v in out.values()),
          " |nonlinear| / |w| =", float(np.linalg.norm(out["nonlinear_terms_rand"]) / np.linalg.norm(w)),
          " |ctrl| =", float(np.linalg.norm(out["control_term_general"])))


if __name__ == "__main__":
    main()
