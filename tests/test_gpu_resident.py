"""GPU: the on-chip-resident RK kernel (csrc/resident.cuh: one CTA per 64^2 environment, every sub-stage of every RK
step in shared memory) against the NumPy oracle and against the two-kernel path (KF_RESIDENT=0) on the same inputs."""
import os

import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc
from test_gpu_parity import make_stepper, random_states, rel_l2

pytestmark = pytest.mark.gpu


def _resident(on, fn):
    old = os.environ.get("KF_RESIDENT")
    os.environ["KF_RESIDENT"] = "1" if on else "0"
    try:
        return fn()
    finally:
        if old is None:
            os.environ.pop("KF_RESIDENT", None)
        else:
            os.environ["KF_RESIDENT"] = old


@pytest.mark.parametrize("dealias", [False, True])
@pytest.mark.parametrize("N,steps,dtype,tol", [(64, 20, torch.float64, 1e-11), (64, 20, torch.float32, 2e-5),
                                               (128, 6, torch.float32, 2e-5), (256, 3, torch.float32, 2e-5)])
def test_resident_rk_steps_match_oracle(N, steps, dtype, tol, dealias):
    """64^2: one CTA per environment; 128^2 / 256^2 fp32: clusters of 2 / 8 CTAs over distributed shared memory."""
    B = 5
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=npdt, dealias=dealias)
    w0 = random_states(ref, B, seed=21).astype(orc._ctype(npdt))
    actions = np.random.default_rng(5).uniform(-0.7, 0.7, size=(B, 4)).astype(npdt)
    clipped = ref._clip_action(actions, ref.default_params)
    ch = ref.equation.control_term(w0, ref._control_field(clipped, ref.default_params))
    want = ref.integrator.rk_steps(w0.copy(), steps, ch)

    def run():
        st = make_stepper(N, 100, B, dtype, dealias=dealias)
        w = torch.from_numpy(w0).cuda()
        st.rk_steps(w, torch.from_numpy(actions).cuda(), steps)
        out = w.cpu().numpy()
        st.close()
        return out

    got = _resident(True, run)
    two_kernel = _resident(False, run)
    assert np.isfinite(got).all()
    for b in range(B):
        assert rel_l2(got[b], want[b]) < tol, (b, rel_l2(got[b], want[b]))
        assert rel_l2(got[b], two_kernel[b]) < tol, (b, rel_l2(got[b], two_kernel[b]))


def test_resident_many_environments_and_step_outputs():
    """More environments than the GPU holds clusters at once, and the full env-step outputs (trajectory, obs, TKE)
    against the two-kernel path."""
    N, B = 64, 700
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float32)
    w0 = random_states(ref, B, seed=3).astype(np.complex64)
    acts = np.random.default_rng(9).uniform(-0.5, 0.5, (B, 4)).astype(np.float32)

    def run():
        st = make_stepper(N, 100, B, torch.float32, n_inner=6, n_outer=3)
        final, traj, obs, tke, nonf = st.step(torch.from_numpy(w0).cuda(), torch.from_numpy(acts).cuda())
        out = [t.cpu().numpy() for t in (final, traj, obs, tke, nonf)]
        launches = st.launch_count
        st.close()
        return out, launches

    (res, l_res), (two, l_two) = _resident(True, run), _resident(False, run)
    assert l_res < l_two  # 3 resident launches + reductions instead of 10 launches per RK step and chunk
    assert (res[4] == 0).all() and np.isfinite(res[0]).all()
    assert rel_l2(res[0], two[0]) < 2e-5 and rel_l2(res[1], two[1]) < 2e-5
    assert np.abs(res[2] - two[2]).max() < 1e-4 and np.abs(res[3] - two[3]).max() < 1e-5
    # and three of them against the oracle
    clipped = ref._clip_action(acts[:3], ref.default_params)
    ch = ref.equation.control_term(w0[:3], ref._control_field(clipped, ref.default_params))
    want = ref.integrator.rk_steps(w0[:3].copy(), 18, ch)
    for b in range(3):
        assert rel_l2(res[0][b], want[b]) < 2e-5
