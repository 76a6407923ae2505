"""The two-lane / L2 time-blocked schedule (kf_create's default for fp32 batches of >= 64 envs, i.e. the
configuration bench.py measures) against the single-lane schedule and against the oracle.  Environments are
independent, so every schedule must give BIT-identical results."""
import os

import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc
from test_gpu_parity import make_stepper, random_states, rel_l2

pytestmark = pytest.mark.gpu


def _with_lanes(value, fn):
    """Run fn with KF_LANES forced and the on-chip-resident path off (these tests are about the lanes of the
    two-kernel path; 64^2 would otherwise take the resident kernel, which has no lanes)."""
    old, old_res = os.environ.get("KF_LANES"), os.environ.get("KF_RESIDENT")
    os.environ["KF_RESIDENT"] = "0"
    if value is None:
        os.environ.pop("KF_LANES", None)
    else:
        os.environ["KF_LANES"] = value
    try:
        return fn()
    finally:
        for key, val in (("KF_LANES", old), ("KF_RESIDENT", old_res)):
            if val is None:
                os.environ.pop(key, None)
            else:
                os.environ[key] = val


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("B,chunk", [(7, 0), (7, 2), (2, 0), (5, 4)])
def test_two_lanes_bitwise_equal_one_lane(dtype, B, chunk):
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(64, 64)), dtype=npdt)
    w0 = random_states(ref, B, seed=B).astype(orc._ctype(npdt))
    acts = np.random.default_rng(2).uniform(-0.5, 0.5, (B, 4)).astype(npdt)

    def run(expect_lanes, **kw):
        st = make_stepper(64, 100, B, dtype, n_inner=7, n_outer=3, dealias=True, **kw)
        assert st.lanes == expect_lanes, (st.lanes, st.chunk)
        w = torch.from_numpy(w0).cuda()
        st.rk_steps(w, torch.from_numpy(acts).cuda(), 23)  # crosses the graph-batch boundary
        final, traj, obs, tke, nonf = st.step(w, torch.from_numpy(acts).cuda())
        out = [t.cpu().numpy() for t in (w, final, traj, obs, tke, nonf)]
        st.close()
        return out

    one = _with_lanes("1", lambda: run(1, chunk=chunk))
    two = _with_lanes("2", lambda: run(2, chunk=chunk))
    for i, (a, b) in enumerate(zip(one, two)):
        if i in (3, 4):  # obs / mean TKE are summed with fp64 atomics: order-dependent in the last bits
            assert np.allclose(a, b, rtol=1e-12 if dtype == torch.float64 else 1e-6, atol=0)
        else:
            assert np.array_equal(a, b)
    assert np.isfinite(one[1]).all() and (one[5] == 0).all()


def test_default_schedule_of_the_bench_configuration_matches_oracle():
    """256^2 fp32, >= 64 envs, no explicit chunk: two lanes of 32 envs (2 x 32 x 7 spectra fit 95 % of L2)."""
    N, B, steps = 256, 96, 3
    ref = orc.KolmogorovFlow(flow_config=dict(Re=1000, grid_size=(N, N)), dtype=np.float32, dealias=True)
    w0 = random_states(ref, B, seed=9).astype(np.complex64)
    acts = np.random.default_rng(4).uniform(-0.5, 0.5, (B, 4)).astype(np.float32)
    st = _with_lanes(None, lambda: make_stepper(N, 1000, B, torch.float32, dealias=True))
    assert st.lanes == 2 and st.chunk == 32
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, torch.from_numpy(acts).cuda(), steps)
    got = w.cpu().numpy()
    st.close()
    pick = [0, 31, 32, 63, 64, 95]  # first / last env of each of the three chunks (lane 0, lane 1, lane 0)
    ch = ref.equation.control_term(w0[pick], ref._control_field(acts[pick], ref.default_params))
    want = ref.integrator.rk_steps(w0[pick].copy(), steps, ch)
    for i, b in enumerate(pick):
        assert rel_l2(got[b], want[i]) < 2e-5, (b, rel_l2(got[b], want[i]))
    # and bit-identical to the single-lane, single-chunk schedule
    st1 = _with_lanes("1", lambda: make_stepper(N, 1000, B, torch.float32, dealias=True, chunk=B))
    assert st1.lanes == 1 and st1.chunk == B
    w1 = torch.from_numpy(w0).cuda()
    st1.rk_steps(w1, torch.from_numpy(acts).cuda(), steps)
    assert np.array_equal(w1.cpu().numpy(), got)
    st1.close()
