"""Generate the golden vectors in this directory from the oracle (run from the repo root):

    python tests/golden/make_golden.py

The reference ships no golden data for the JAX Kolmogorov path and JAX cannot be
installed here, so these vectors are ORACLE-generated.  The oracle itself is pinned against the
reference's own source by make_reference_golden.py / reference_shim_64_re100.npz.
They freeze the oracle's behaviour (including the fp32 cases): CPU tests check the oracle still reproduces them,
GPU tests check the CUDA path against them.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import kolmogorov_oracle as orc  # noqa: E402

ACTION = np.array([-0.25, -0.03, 0.02, 0.01])  # examples/jax/getting_started/1_kolmogorov/test_kolmogorov_env.py:33
FLOW = dict(Re=100, grid_size=(64, 64))


def main():
    out = {}
    for name, dt in (("f64", np.float64), ("f32", np.float32)):
        env = orc.KolmogorovFlow(flow_config=FLOW, dtype=dt)
        w = env.flow.initialize_state()
        out[f"omega0_{name}"] = w
        done = 0
        for n in (1, 10, 100):
            w = env.integrator.rk_steps(w, n - done)
            done = n
            out[f"zero_action_rk{n}_{name}"] = w
        ch = env.equation.control_term(w, env._control_field(ACTION.astype(dt), env.default_params))
        out[f"action_rk100_{name}"] = env.integrator.rk_steps(env.flow.initialize_state(), 100, ch)
    # env-level trace: dt=1e-2, action_time=2 (n_inner=99, n_outer=2), fp64
    env = orc.KolmogorovFlow(dict(dt=1e-2, action_time=2.0), FLOW)
    obs, state = env.reset_env()
    out["env_reset_obs"] = obs
    out["env_reset_omega"] = state["omega_hat"]
    acts = np.array([ACTION, [0.9, -0.9, 0.1, 0.0], [0, 0, 0, 0]], dtype=np.float64)
    out["env_actions"] = acts
    obs_l, rew_l, tke_l = [], [], []
    for a in acts:
        obs, state, reward, done, info = env.step_env(None, state, a)
        obs_l.append(obs), rew_l.append(reward), tke_l.append(info["mean_tke"])
    out["env_obs"], out["env_reward"], out["env_mean_tke"] = np.array(obs_l), np.array(rew_l), np.array(tke_l)
    out["env_final_omega"] = state["omega_hat"]
    # sanity anchors from the long uncontrolled run: TKE after 200 / 1000 / 2000 RK steps
    env = orc.KolmogorovFlow(flow_config=FLOW)
    w = env.flow.initialize_state()
    tke = []
    for n in (200, 800, 1000):
        w = env.integrator.rk_steps(w, n)
        tke.append(float(orc.compute_tke(w, env.kx, env.ky, 64)))
    out["tke_200_1000_2000"] = np.array(tke)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kolmogorov_64_re100.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
