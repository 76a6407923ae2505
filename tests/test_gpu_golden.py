"""GPU path vs the committed golden vectors (tests/golden/kolmogorov_64_re100.npz) and size-independent properties."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "kolmogorov_64_re100.npz"))
ACTION = np.array([-0.25, -0.03, 0.02, 0.01])


def rel(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def stepper(**kw):
    from hydrogym_b200 import KolmogorovStepper
    args = dict(grid_size=(64, 64), Re=100, num_envs=1, device="cuda:0")
    args.update(kw)
    return KolmogorovStepper(**args)


@pytest.mark.parametrize("name,dtype,tol", [("f64", torch.float64, 1e-9), ("f32", torch.float32, 1e-3)])
def test_golden_rk_trajectories(name, dtype, tol):
    st = stepper(dtype=dtype)
    w = torch.from_numpy(GOLD[f"omega0_{name}"][None]).cuda()
    done = 0
    for n in (1, 10, 100):
        st.rk_steps(w, None, n - done)
        done = n
        assert rel(w.cpu().numpy()[0], GOLD[f"zero_action_rk{n}_{name}"]) < tol
    w = torch.from_numpy(GOLD[f"omega0_{name}"][None]).cuda()
    a = torch.tensor(ACTION[None], dtype=dtype).cuda()
    st.rk_steps(w, a, 100)
    assert rel(w.cpu().numpy()[0], GOLD[f"action_rk100_{name}"]) < tol
    # north-star tolerance, fp32 path against the fp64 golden trajectory
    assert rel(w.cpu().numpy()[0], GOLD["action_rk100_f64"]) < 1e-3


def test_golden_env_trace():
    from hydrogym_b200 import KolmogorovFlow
    env = KolmogorovFlow(dict(dt=1e-2, action_time=2.0, dtype="float64", device="cuda:0"), dict(Re=100, grid_size=(64, 64)))
    obs, state = env.reset_env(None, env.default_params)
    assert np.allclose(obs.cpu().numpy(), GOLD["env_reset_obs"], rtol=1e-8)
    assert rel(state.omega_hat.cpu().numpy(), GOLD["env_reset_omega"]) < 1e-9
    for i, a in enumerate(GOLD["env_actions"]):
        obs, state, reward, done, info = env.step_env(None, state, torch.from_numpy(a), env.default_params)
        assert np.allclose(obs.cpu().numpy(), GOLD["env_obs"][i], rtol=1e-8)
        assert float(reward) == pytest.approx(float(GOLD["env_reward"][i]), abs=1e-12)
        assert float(info["mean_tke"]) == pytest.approx(float(GOLD["env_mean_tke"][i]), rel=1e-9)
    assert rel(state.omega_hat.cpu().numpy(), GOLD["env_final_omega"]) < 1e-8


@pytest.mark.parametrize("N,B,dtype", [(256, 256, torch.float32), (256, 32, torch.float64), (128, 1024, torch.float32),
                                       (1024, 16, torch.float32)])
def test_full_size_properties(N, B, dtype):
    """At BASELINE.json's full sizes (no oracle run): real-field consistency, env independence (a batch of
    identical envs stays identical; permuting envs permutes results) and agreement of the fused reward
    path with the trajectory path."""
    from hydrogym_b200 import KolmogorovStepper
    from bench import synthetic_omega_hat
    npdt = np.float32 if dtype == torch.float32 else np.float64
    st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=dtype, device="cuda:0", n_inner=3, n_outer=2)
    base = [torch.from_numpy(synthetic_omega_hat(N, i, npdt)) for i in range(3)]
    idx = torch.arange(B) % 3
    w0 = torch.stack([base[i] for i in idx.tolist()]).cuda()
    acts = torch.zeros((B, 4), dtype=dtype, device="cuda")
    acts[:, 0] = (idx.to(dtype) - 1.0).cuda() * 0.3
    final, traj, obs, tke, nonf = st.step(w0, acts, want_trajectory=True)
    assert int(nonf.sum()) == 0 and torch.isfinite(torch.view_as_real(final)).all()
    for k in range(3):                                   # identical envs give bit-identical results
        sel = (idx == k).nonzero().flatten().cuda()
        assert torch.equal(final[sel], final[sel[:1]].expand(len(sel), -1, -1))
        assert torch.equal(obs[sel], obs[sel[:1]].expand(len(sel), -1))
    assert torch.equal(traj[:, -1], final)               # last snapshot == final state
    obs2, tke2, _ = st.reward_obs(traj)                   # fused-in-rollout reductions == trajectory reductions
    assert torch.allclose(obs2, obs, rtol=1e-5, atol=1e-6) and torch.allclose(tke2, tke, rtol=1e-5)
    # the field stays the transform of a REAL field: the ky=0 column keeps Hermitian symmetry in kx
    col0 = final[:, :, 0]
    mirror = torch.conj(col0[:, (-torch.arange(N)) % N])
    scale = col0.abs().max()
    assert (col0 - mirror).abs().max() < (1e-4 if dtype == torch.float32 else 1e-10) * scale
