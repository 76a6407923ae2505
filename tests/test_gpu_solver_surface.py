"""GPU: the equation / integrator methods the reference documents (kolmogorov.py:209-272, base.py:32-116) on the
product classes, against the reference's own source run on a broadband state with a general control field
(tests/golden/reference_shim_solver_surface.npz) and against the cfg-1 vectors."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
SURF = np.load(os.path.join(HERE, "golden", "reference_shim_solver_surface.npz"))
REF = np.load(os.path.join(HERE, "golden", "reference_shim_64_re100.npz"))
FLOW = dict(Re=100, grid_size=(64, 64))


def rel(a, b):
    a = a.cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def make(dtype=torch.float64, dt=1e-3):
    from hydrogym_b200 import FlowConfig
    from hydrogym_b200.solvers import PseudoSpectralNavierStokes2D, RungeKuttaCrankNicolson
    flow = FlowConfig(**FLOW)
    eq = PseudoSpectralNavierStokes2D(flow, dtype=dtype, device="cuda:0")
    integ = RungeKuttaCrankNicolson(flow, dt=dt, save_n=1, equation=eq, dtype=dtype, device="cuda:0")
    return flow, eq, integ


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 5e-6)])
def test_equation_terms_vs_reference_source(dtype, tol):
    flow, eq, _ = make(dtype)
    cd = torch.complex128 if dtype == torch.float64 else torch.complex64
    w = torch.from_numpy(SURF["w_rand"]).to("cuda", cd)
    cf = (SURF["cfx"], SURF["cfy"])
    assert rel(eq.linear_terms(w), SURF["linear_terms_rand"]) < tol
    assert rel(eq.implicit_timestep(w, 7.5e-5), SURF["implicit_rand_mu"]) < tol
    assert rel(eq.nonlinear_terms(w, None), SURF["nonlinear_terms_rand"]) < 10 * tol
    assert rel(eq.nonlinear_terms(w, cf), SURF["nonlinear_terms_rand_ctrl"]) < 10 * tol
    # clearing the field again gives the uncontrolled term; batched input gives batched output
    wb = torch.stack([w, 2 * w])
    nb = eq.nonlinear_terms(wb, None)
    assert nb.shape == wb.shape and rel(nb[0], SURF["nonlinear_terms_rand"]) < 10 * tol
    # the cfg-1 vectors (three dominant modes) and the constants
    w0 = torch.from_numpy(REF["omega0"]).to("cuda", cd)
    assert rel(eq.nonlinear_terms(w0, None), REF["nonlinear_terms_omega0"]) < 10 * tol
    assert rel(eq.linear_terms(w0), REF["linear_terms_omega0"]) < tol
    assert rel(eq.implicit_timestep(w0, 7.5e-5), REF["implicit_omega0_mu"]) < tol
    assert rel(eq.forcing_term(), REF["forcing_term"]) < 1e-13
    assert rel(eq.control_term(None, cf), SURF["control_term_general"]) < 1e-14
    eq.close()


def test_integrator_methods_with_general_control_field():
    flow, eq, integ = make(torch.float64)
    w = torch.from_numpy(SURF["w_rand"]).cuda()
    cf = (SURF["cfx"], SURF["cfy"])
    step = integ.RK4_CN(control_field=cf)
    u = w
    for _ in range(10):
        u = step(u)
    assert rel(u, SURF["rk10_general_ctrl"]) < 1e-11
    assert torch.equal(w, torch.from_numpy(SURF["w_rand"]).cuda())  # functional: the input is not modified
    assert rel(integ.step(flow, 1e-3, 5, [], control_field=cf)(w), SURF["step_save5_general_ctrl"]) < 1e-11
    fin, traj = integ.solve(0.05, flow, (0, 2.0), save_n=1, initial_state=w, control_field=cf)
    assert tuple(traj.shape) == SURF["solve_general_traj"].shape
    assert rel(fin, SURF["solve_general_final"]) < 1e-10 and rel(traj, SURF["solve_general_traj"]) < 1e-10
    # and without a field the dense spectrum is really gone (zero-control cfg-1 vector)
    u = torch.from_numpy(REF["omega0"]).cuda()
    step0 = integ.RK4_CN(control_field=None)
    for _ in range(10):
        u = step0(u)
    assert rel(u, REF["zero_action_rk10"]) < 1e-11


def test_dense_control_field_on_two_lanes_matches_oracle():
    """The dense spectrum is shared by both lanes of a large fp32 batch (graph replay included)."""
    from oracle import kolmogorov_oracle as orc
    from hydrogym_b200 import KolmogorovStepper
    N, B, steps = 128, 128, 9
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float32)
    rng = np.random.default_rng(3)
    p = np.abs(np.fft.fftfreq(N, 1.0 / N))[:, None]
    q = np.arange(N // 2 + 1)[None, :]
    w0 = ((rng.normal(size=(B, N, N // 2 + 1)) + 1j * rng.normal(size=(B, N, N // 2 + 1)))
          * np.exp(-(p * p + q * q) / 64.0) * N * N * 0.1).astype(np.complex64)
    x, y = ref.x, ref.y
    cf = (0.3 * np.sin(2 * x) * np.cos(3 * y), -0.2 * np.cos(x + 2 * y))
    ch = ref.equation.control_term(w0[0], cf)
    want = ref.integrator.rk_steps(w0[:3].copy(), steps, ch)
    st = KolmogorovStepper(grid_size=(N, N), Re=100, num_envs=B, dtype=torch.float32, device="cuda:0")
    assert st.lanes == 2
    st.set_control_field(ch)
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, None, steps)
    got = w.cpu().numpy()
    for b in range(3):
        assert rel(got[b], want[b]) < 2e-5
    assert rel(got[B - 1], ref.integrator.rk_steps(w0[B - 1:].copy(), steps, ch)[0]) < 2e-5
    st.set_control_field(None)
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, None, steps)
    assert rel(w[0].cpu().numpy(), ref.integrator.rk_steps(w0[:1].copy(), steps, None)[0]) < 2e-5
    st.close()
