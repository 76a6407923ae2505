"""Golden vectors from the REFERENCE'S OWN SOURCE (run in the build container, from the repo root):

    python tests/golden/make_reference_golden.py

Imports `/root/reference/hydrogym/jax/envs/kolmogorov.py` (and the files it pulls in) UNMODIFIED, with the absent
third-party modules (`jax`, `flax`, `gymnax`, `chex`, `tree_math`, ...) replaced by the NumPy stand-ins of
`jax_numpy_shim.py`, and records what the reference's code computes in fp64 at BASELINE.json's configs[0] shape
(64^2, Re = 100):

* `FlowConfig.initialize_state` / `load_mesh` / `load_fft_mesh`, `forcing_term`, `control_term` for the docs' action;
* `RungeKuttaCrankNicolson.RK4_CN` time steps: 1, 10, 100 solver steps with zero control and 100 with the action;
* `KolmogorovFlow.reset_env` + three `step_env` calls at `env_config["dt"] = 1e-2` (990 solver steps each): obs, reward,
  mean TKE, done, discount, final omega_hat and trajectory snapshots;
* gymnax `step` auto-reset with `max_episode_steps = 1`, and `VecEnv` (2 environments) `reset` / `step`;
* `utils.py` diagnostics on the reset state; `FlowConfig.get_observations`;
* the wrapper stack of `env_core.py` (LogWrapper, ClipAction, NormalizeVecObservation, NormalizeVecReward) over a
  scripted environment.

* (second file, `reference_shim_sizes.npz`) the bench's own synthetic initial vorticity and action through 50 solver steps
  at 128^2 Re = 100 (configs[1]) and 20 at 256^2 Re = 1000 (configs[2]).

* (third file, `reference_shim_64_re100_fp32.npz`) 1 / 10 / 100 solver steps in SINGLE precision (JAX's default mode).

The file `reference_shim_64_re100.npz` written here is what `tests/test_reference_shim.py` holds the oracle to.
The GPU box has no `/root/reference`: only the committed `.npz` travels.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import jax_numpy_shim  # noqa: E402

ACTION = np.array([-0.25, -0.03, 0.02, 0.01])  # examples/jax/getting_started/1_kolmogorov/test_kolmogorov_env.py:33
ACTIONS = np.array([ACTION, [0.9, -0.9, 0.1, 0.0], [0, 0, 0, 0]], dtype=np.float64)
FLOW = dict(Re=100, grid_size=(64, 64))


def scripted_env_classes(jax, jnp, struct, environment):
    """A 3-float scripted gymnax environment (obs = affine in time and action) to drive the reference wrappers."""

    @struct.dataclass
    class S(environment.EnvState):
        time: int
        x: jnp.ndarray

    @struct.dataclass
    class P(environment.EnvParams):
        max_steps_in_episode: int = 3
        min_action: float = -0.5
        max_action: float = 0.5

    class Scripted(environment.Environment):
        @property
        def default_params(self):
            return P()

        def reset_env(self, key, params):
            x = jnp.asarray(key, dtype=jnp.float64)[-1] * jnp.array([0.5, -1.0, 2.0])
            return x, S(time=0, x=x)

        def step_env(self, key, state, action, params):
            a = jnp.clip(action, params.min_action, params.max_action)
            x = 0.9 * state.x + jnp.array([1.0, 2.0, -1.0]) * jnp.sum(a) + 0.1
            st = S(time=state.time + 1, x=x)
            done = st.time >= params.max_steps_in_episode
            return x, st, jnp.sum(x) - 0.3 * st.time, done, {"discount": 1.0 - done}

        def observation_space(self, params):
            return environment_spaces.Box(-jnp.inf, jnp.inf, (3,))

        def action_space(self, params=None):
            return environment_spaces.Box(-0.5, 0.5, (4,))

    from gymnax.environments import spaces as environment_spaces
    return Scripted


def main():
    jax_numpy_shim.install()
    import jax
    import jax.numpy as jnp
    from flax import struct
    from gymnax.environments import environment

    from hydrogym.jax import env_core
    from hydrogym.jax.envs.kolmogorov import KolmogorovFlow
    from hydrogym.jax.utils import utils as U

    out = {}
    # ---- constants and solver steps at the default dt = 1e-3
    env = KolmogorovFlow(None, FLOW)
    p = env.default_params
    out["x"], out["y"], out["kx"], out["ky"] = env.x, env.y, env.kx, env.ky
    w0 = env.flow.initialize_state()
    out["omega0"] = w0
    out["forcing_term"] = env.equation.forcing_term()
    cf = env._control_field(jnp.array(ACTION), p)
    out["control_term"] = env.equation.control_term(w0, cf)
    out["control_term_none"] = env.equation.control_term(w0, None)
    out["nonlinear_terms_omega0"] = env.equation.nonlinear_terms(w0, None)
    out["linear_terms_omega0"] = env.equation.linear_terms(w0)
    out["implicit_omega0_mu"] = env.equation.implicit_timestep(w0, 7.5e-5)
    step = env.integrator.RK4_CN(control_field=None)
    w, done = w0, 0
    for n in (1, 10, 100):
        for _ in range(n - done):
            w = step(w)
        done = n
        out[f"zero_action_rk{n}"] = w
    step_a = env.integrator.RK4_CN(control_field=cf)
    w = w0
    for _ in range(100):
        w = step_a(w)
    out["action_rk100"] = w
    fin, traj = env.integrator.solve(dt=0.05, flow=env.flow, t_span=(0, 3.0), save_n=1, initial_state=w0,
                                     control_field=cf)
    out["solve_dt0p05_final"], out["solve_dt0p05_traj"] = fin, traj
    out["get_observations"] = env.flow.get_observations()  # uses flow.vorticity = traj (solvers/base.py:110)

    # ---- env-level trace at env_config dt = 1e-2 (n_inner = 99, n_outer = 10)
    env = KolmogorovFlow(dict(dt=1e-2), FLOW)
    p = env.default_params
    key = jax.random.PRNGKey(0)
    obs, state = env.reset_env(key, p)
    out["env_reset_obs"], out["env_reset_omega"], out["env_reset_traj"] = obs, state.omega_hat, state.trajectory
    wr = state.omega_hat
    uh, vh = U.compute_velocity_fft(wr, env.kx, env.ky)
    out["diag_uhat"], out["diag_vhat"] = uh, vh
    out["diag_tke"] = U.compute_tke(wr, env.kx, env.ky, env.n)
    out["diag_dissipation"] = U.compute_energy_dissipation(wr, env.kx, env.ky, env.flow.nu, env.n)
    out["diag_divergence"] = U.compute_divergence(wr, env.kx, env.ky)
    out["diag_energy_mode"] = np.array([U.compute_energy_mode(uh, vh, a, b, env.n, env.m) for a, b in ((3, 4), (-2, 5), (0, 4))])
    out["diag_velocity_mode"] = np.array([U.compute_velocity_mode(uh, vh, a, b, env.n, env.m) for a, b in ((3, 4), (-2, 5), (0, 4))])
    out["diag_compute_reward"] = np.array(U.compute_reward(wr, env.kx, env.ky, env.flow.nu, env.n, jnp.array(ACTION)))
    out["diag_point_velocity"] = np.array(U.compute_real_velocity_point(uh, vh, 8, 24))
    out["dealiasing_is_identity"] = np.array(U.dealiasing(wr) is wr)
    obs_l, rew_l, tke_l, done_l, disc_l, time_l = [], [], [], [], [], []
    for a in ACTIONS:
        obs, state, reward, done, info = env.step_env(key, state, jnp.array(a), p)
        obs_l.append(obs), rew_l.append(reward), tke_l.append(info["mean_tke"]), done_l.append(done)
        disc_l.append(info["discount"]), time_l.append(state.time)
    out["env_actions"] = ACTIONS
    out["env_obs"], out["env_reward"], out["env_mean_tke"] = np.array(obs_l), np.array(rew_l), np.array(tke_l)
    out["env_done"], out["env_discount"], out["env_time"] = np.array(done_l), np.array(disc_l), np.array(time_l)
    out["env_final_omega"], out["env_final_traj"] = state.omega_hat, state.trajectory
    # reward with a non-zero alpha (K:462-470), same trajectory
    out["env_reward_alpha2"] = env._reward(jnp.clip(jnp.array(ACTIONS[2]), -0.5, 0.5), state.trajectory,
                                           p.replace(reward_alpha=2.0))
    # gymnax auto-reset: max_episode_steps = 1 -> done after the first step, state/obs replaced by the reset ones
    p1 = p.replace(max_episode_steps=1)
    obs_r, state_r = env.reset(key, p1)
    obs_s, state_s, rew_s, done_s, info_s = env.step(key, state_r, jnp.array(ACTION), p1)
    out["auto_obs"], out["auto_omega"], out["auto_time"] = obs_s, state_s.omega_hat, np.array(state_s.time)
    out["auto_reward"], out["auto_done"], out["auto_mean_tke"] = np.array(rew_s), np.array(done_s), info_s["mean_tke"]
    out["auto_discount"] = np.array(info_s["discount"])
    # VecEnv (E:609-613): two environments, different actions
    venv = env_core.VecEnv(env)
    keys = jax.random.split(key, 2)
    vobs, vstate = venv.reset(keys, p)
    vacts = jnp.array([ACTIONS[0], ACTIONS[1]])
    vobs2, vstate2, vrew, vdone, vinfo = venv.step(keys, vstate, vacts, p)
    out["vec_reset_obs"], out["vec_actions"] = vobs, vacts
    out["vec_obs"], out["vec_omega"], out["vec_reward"], out["vec_done"] = vobs2, vstate2.omega_hat, vrew, vdone
    out["vec_mean_tke"], out["vec_time"] = vinfo["mean_tke"], vstate2.time

    # ---- wrapper stack over a scripted env (E:449-742)
    Scripted = scripted_env_classes(jax, jnp, struct, environment)
    base = Scripted()
    wenv = env_core.LogWrapper(base)
    wenv = env_core.ClipAction(wenv, low=-0.5, high=0.5) if _takes(env_core.ClipAction, "low") else env_core.ClipAction(wenv)
    wenv = env_core.VecEnv(wenv)
    wenv = env_core.NormalizeVecObservation(wenv)
    wenv = env_core.NormalizeVecReward(wenv, 0.99)
    B, T = 3, 8
    keys = jnp.array([[0, 1], [0, 2], [0, 3]], dtype=np.uint32)
    rng = np.random.default_rng(7)
    acts = rng.uniform(-1.0, 1.0, (T, B, 4))
    obs, st = wenv.reset(keys, base.default_params)
    wobs, wrew, wdone, wret, wlen, wretd = [np.array(obs)], [], [], [], [], []
    for t in range(T):
        obs, st, rew, done, info = wenv.step(keys, st, jnp.array(acts[t]), base.default_params)
        wobs.append(np.array(obs)), wrew.append(np.array(rew)), wdone.append(np.array(done))
        wret.append(np.array(info["returned_episode_returns"])), wlen.append(np.array(info["returned_episode_lengths"]))
        wretd.append(np.array(info["returned_episode"]))
    out["wrap_keys"], out["wrap_actions"] = np.array(keys), acts
    out["wrap_obs"], out["wrap_reward"], out["wrap_done"] = np.array(wobs), np.array(wrew), np.array(wdone)
    out["wrap_returns"], out["wrap_lengths"], out["wrap_returned"] = np.array(wret), np.array(wlen), np.array(wretd)

    # FlattenObservationWrapper / TransformObservation / TransformReward (E:460-495, 585-607) over the same scripted
    # dynamics with a 2-D observation (3, 2), one un-batched environment, 5 steps (one episode end)
    class Scripted2D(Scripted):
        def reset_env(self, key, params):
            x, st = super().reset_env(key, params)
            return jnp.outer(x, jnp.array([1.0, -2.0])), st

        def step_env(self, key, state, action, params):
            x, st, r, d, info = super().step_env(key, state, action, params)
            return jnp.outer(x, jnp.array([1.0, -2.0])), st, r, d, info

        def observation_space(self, params):
            from gymnax.environments import spaces
            return spaces.Box(-jnp.inf, jnp.inf, (3, 2))

    fenv = env_core.FlattenObservationWrapper(Scripted2D())
    assert fenv.observation_space(base.default_params).shape == (6,)
    fenv = env_core.TransformReward(env_core.TransformObservation(fenv, lambda o: 2.0 * o + 1.0), lambda r: r - 1.0)
    key1 = jnp.array([0, 5], dtype=np.uint32)
    obs, st = fenv.reset(key1, base.default_params)
    fobs, frew, fdone = [np.array(obs)], [], []
    for t in range(5):
        obs, st, rew, done, info = fenv.step(key1, st, jnp.array(acts[t][0]), base.default_params)
        fobs.append(np.array(obs)), frew.append(np.array(rew)), fdone.append(np.array(done))
    out["flat_key"], out["flat_obs"], out["flat_reward"], out["flat_done"] = np.array(key1), np.array(fobs), np.array(frew), np.array(fdone)

    path = os.path.join(HERE, "reference_shim_64_re100.npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")

    # ---- the bench's own synthetic inputs (bench.synthetic_omega_hat / synthetic_actions, regenerated from their seeds
    # by the tests) through the reference's solver at configs[1] (128^2, Re = 100) and configs[2] (256^2, Re = 1000)
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import bench
    sizes = {}
    for N, Re, n_rk in ((128, 100, 50), (256, 1000, 20)):
        env = KolmogorovFlow(None, dict(Re=Re, grid_size=(N, N)))
        w = jnp.asarray(bench.synthetic_omega_hat(N, 0, np.float64))
        act = bench.synthetic_actions(1, 1, np.float64)[0][0]
        cf = env._control_field(jnp.clip(jnp.array(act), -0.5, 0.5), env.default_params)
        step = env.integrator.RK4_CN(control_field=cf)
        for _ in range(n_rk):
            w = step(w)
        w = np.asarray(w)
        sizes[f"n{N}_steps"] = np.array(n_rk)
        sizes[f"n{N}_action"] = act
        sizes[f"n{N}_norm"] = np.array(np.linalg.norm(w))
        sizes[f"n{N}_sum"] = np.array(w.sum())
        if N == 128:
            sizes["n128_omega"] = w
        else:  # low-wavenumber block |kx| <= 32, ky <= 32 (the fixture stays small) + the two scalars above
            sizes["n256_omega_low"] = w[np.r_[0:33, N - 32:N]][:, :33]
    path = os.path.join(HERE, "reference_shim_sizes.npz")
    np.savez_compressed(path, **sizes)
    print("wrote", path, os.path.getsize(path), "bytes")


def main_fp32():
    """The same reference source in SINGLE precision (JAX's default mode, `jax_enable_x64` off): a separate interpreter
    (`python make_reference_golden.py --fp32`, spawned by main) because the precision switch is global to the stand-ins."""
    jax_numpy_shim.install(x64=False)
    import jax.numpy as jnp

    from hydrogym.jax.envs.kolmogorov import KolmogorovFlow
    env = KolmogorovFlow(None, FLOW)
    w0 = env.flow.initialize_state()
    assert w0.dtype == np.complex64 and env.kx.dtype == np.float32
    out = {"omega0": w0}
    cf = env._control_field(jnp.array(ACTION), env.default_params)
    for name, control in (("zero_action", None), ("action", cf)):
        step = env.integrator.RK4_CN(control_field=control)
        w, done = w0, 0
        for n in (1, 10, 100):
            for _ in range(n - done):
                w = step(w)
            done = n
            out[f"{name}_rk{n}"] = w
        assert w.dtype == np.complex64
    path = os.path.join(HERE, "reference_shim_64_re100_fp32.npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes")


def _takes(cls, name):
    import inspect
    return name in inspect.signature(cls.__init__).parameters


if __name__ == "__main__":
    if "--fp32" in sys.argv:
        main_fp32()
    else:
        main()
        import subprocess
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--fp32"])
