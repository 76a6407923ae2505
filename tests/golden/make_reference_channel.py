"""Golden vectors for the 3-D channel-flow environment step from the reference's own source
(run in the build container, from the repo root; ~6 minutes of CPU):

    python tests/golden/make_reference_channel.py      # writes tests/golden/reference_shim_channel.npz

Same route as make_reference_golden.py: `/root/reference/hydrogym/jax/envs/channel.py`, `solvers/base.py`
(`RungeKutta4`) and `utils/utils.py` (`cheb_D_matrices`, `dealias_mask_2_3`) are imported UNMODIFIED over the NumPy
stand-ins of jax_numpy_shim.py, in fp64 (`env_config["dtype"] = "float64"`).  The HF-hosted initial fields are not
available offline; the env's own `initial_field_dir` override (channel.py:444-445) points at `.npy` files written from
`oracle.channel_oracle.synthetic_initial_fields()` -- the tests and the bench rebuild the same fields from that function.
`jax.value_and_grad(..., has_aux=True)` returns (value, None): `step_env` discards the gradient (channel.py:497-502).

Stored (physical 72^3 arrays as every-6th (x, y) column, spectral ones as their |kx|, |ky| <= 5 block, both with
their 2-norms; everything else in full):
* operator constants: z, Dz, Dzz, dealias mask, A_inv of four modes, the jet pattern of the test action;
* `nonlinear_terms` / `linear_terms` of the initial state with the action at t = 3;
* `rk4_step` x 1 and x 2 (t = 0, 1), with and without action;
* two consecutive `step_env` calls (2 x 50 RK4 steps; the second runs at t >= 50 where the reference's gain ramp is
  negative): observation, reward, done, time, the fields.
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import jax_numpy_shim  # noqa: E402

SUB = (slice(None, None, 6), slice(None, None, 6), slice(None))          # physical fields: every 6th column
LOW = np.r_[0:6, 67:72]                                                    # spectral fields: the |k| <= 5 block
MODES = (0, 1, 73, 36 * 72 + 36)


def test_action():
    a = 0.9 * np.sin(1.3 * np.arange(24) + 0.4)
    a[5] = 1.7     # clipped to 1 by step_env
    a[17] = -1.2   # clipped to -1
    return a


def pack(out, name, a):
    a = np.asarray(a)
    out[name + "_sub"] = a[SUB].copy() if not np.iscomplexobj(a) else a[np.ix_(LOW, LOW, np.arange(72))].copy()
    out[name + "_norm"] = np.sqrt((np.abs(a) ** 2).sum())


def main():
    jax_numpy_shim.install()
    import jax
    jax.checkpoint = lambda f, **k: f   # rematerialisation hint only (channel.py:392, 553)
    from hydrogym.jax.envs import channel as ch

    from oracle import channel_oracle as orc

    out = {}
    with tempfile.TemporaryDirectory() as d:
        for n, a in zip("UVW", orc.synthetic_initial_fields()):
            np.save(os.path.join(d, n + ".npy"), a)
        env = ch.ChannelFlowSpectralEnv({"initial_field_dir": d, "dtype": "float64"})
    eq, integ, params = env.equation, env.integrator, env.default_params
    act = test_action()
    clipped = np.clip(act, -1, 1)
    out.update(action=act, z=eq.z, Dz=eq.Dz, Dzz=eq.Dzz, dealias=eq.dealias[:, :, 0],
               A_inv_modes=np.stack([eq.A_inv[m] for m in MODES]), modes=np.array(MODES),
               jets_plane=eq.apply_jets_v(jax_numpy_shim._wrap(np.zeros((72, 72, 72))), clipped)[:, :, 2])
    key = jax.random.PRNGKey(0)
    obs0, st = env.reset_env(key, params)
    out["obs0"] = obs0
    s0 = eq.to_spectral(ch.VelocityState(st.U, st.V, st.W))
    N = eq.nonlinear_terms(s0, action=clipped, t=3.0, fx=2.0)
    L = eq.linear_terms(s0, action=clipped, t=3.0)
    for nm, a in zip(("Nu", "Nv", "Nw", "Lu", "Lv", "Lw"), (N.u, N.v, N.w, L.u, L.v, L.w)):
        pack(out, nm, a)
    for tag, a in (("act", clipped), ("noact", None)):
        s = s0
        for n in range(2):
            s = integ.rk4_step(s, dt=st.dt, action=a, t=n, fx=2.0, target_bulk_u=8.0)
            for c, f in zip("uvw", (s.u, s.v, s.w)):
                pack(out, f"rk{n + 1}_{tag}_{c}", f)
        print("rk4 x2", tag, "done", flush=True)
    for i in (1, 2):
        obs, st, reward, done, info = env.step_env(key, st, act, params)
        out[f"step{i}_obs"], out[f"step{i}_reward"], out[f"step{i}_done"] = obs, reward, done
        out[f"step{i}_time"], out[f"step{i}_discount"] = st.time, info["discount"]
        for c, f in zip("UVW", (st.U, st.V, st.W)):
            pack(out, f"step{i}_{c}", f)
        out[f"step{i}_U_plane4"] = np.asarray(st.U)[:, :, 4]
        print("step_env", i, "reward", float(reward), "time", int(st.time), flush=True)
    path = os.path.join(HERE, "reference_shim_channel.npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes")
    print("finite:", all(np.isfinite(np.asarray(v)).all() for v in out.values()))


if __name__ == "__main__":
    main()
