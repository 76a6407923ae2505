"""GPU parity of the slab-decomposed single-grid solver (kf_slab_*): against the NumPy oracle, against the
batched kernels, and across decompositions.  ``local_ranks=P`` plays P ranks on one device with the same
kernels and layouts as the NCCL driver (tests/test_gpu_slab_nccl.py covers the collective itself)."""
import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc
from test_gpu_parity import ACTION, random_states, rel_l2

pytestmark = pytest.mark.gpu


def make_slab(N, Re, dtype, P, **kw):
    from hydrogym_b200.slab import SlabKolmogorovSolver
    return SlabKolmogorovSolver(grid_size=(N, N), Re=Re, dtype=dtype, device="cuda:0", local_ranks=P, **kw)


@pytest.mark.parametrize("P", [1, 2, 4, 8])
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-11), (torch.float32, 2e-5)])
def test_slab_256_matches_oracle(P, dtype, tol):
    N, steps = 256, 3
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=npdt)
    w0 = random_states(ref, 1, seed=11).astype(orc._ctype(npdt))
    action = np.array([0.7, -0.3, 0.1, -0.6], dtype=npdt)  # first one outside the clip range
    clipped = ref._clip_action(action[None], ref.default_params)
    ch = ref.equation.control_term(w0, ref._control_field(clipped, ref.default_params))
    want = ref.integrator.rk_steps(w0.copy(), steps, ch)[0]
    sl = make_slab(N, 100, dtype, P)
    sl.set_state(torch.from_numpy(w0[0]).cuda(), torch.from_numpy(action).cuda())
    sl.rk_steps(steps)
    got = sl.get_state().cpu().numpy()
    sl.close()
    assert np.isfinite(got).all()
    assert rel_l2(got, want) < tol, rel_l2(got, want)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-11), (torch.float32, 2e-5)])
def test_slab_1024_matches_batched_kernels(dtype, tol):
    from hydrogym_b200 import KolmogorovStepper
    N, steps = 1024, 2
    npdt = np.float64 if dtype == torch.float64 else np.float32
    ref = orc.KolmogorovFlow(flow_config=dict(Re=1000, grid_size=(N, N)), dtype=npdt)
    w0 = random_states(ref, 1, seed=5).astype(orc._ctype(npdt))
    act = torch.tensor(ACTION, dtype=dtype, device="cuda")
    st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=1, dtype=dtype, device="cuda:0")
    w = torch.from_numpy(w0).cuda()
    st.rk_steps(w, act[None], steps)
    want = w[0].cpu().numpy()
    st.close()
    for P in (1, 4):
        sl = make_slab(N, 1000, dtype, P)
        sl.set_state(torch.from_numpy(w0[0]).cuda(), act)
        sl.rk_steps(steps)
        got = sl.get_state().cpu().numpy()
        sl.close()
        assert rel_l2(got, want) < tol, (P, rel_l2(got, want))


@pytest.mark.parametrize("P", [1, 2, 8])
def test_slab_p2p_exchange_matches_all_to_all(P):
    """Peer-to-peer mode (kernels store straight into the owning rank's receive buffer) vs the all-to-all mode."""
    N, steps = 256, 2
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float64)
    w0 = torch.from_numpy(random_states(ref, 1, seed=4).astype(np.complex128)[0]).cuda()
    act = torch.tensor(ACTION, dtype=torch.float64, device="cuda")
    out = {}
    for mode in ("nccl", "p2p"):
        sl = make_slab(N, 100, torch.float64, P, exchange=mode)
        sl.set_state(w0, act)
        sl.rk_steps(steps)
        obs, tke = sl.reward_obs()
        out[mode] = (sl.get_state().cpu().numpy(), obs.cpu().numpy(), float(tke))
        sl.close()
    assert np.array_equal(out["nccl"][0], out["p2p"][0])
    assert np.abs(out["nccl"][1] - out["p2p"][1]).max() < 1e-13
    assert abs(out["nccl"][2] - out["p2p"][2]) < 1e-13


@pytest.mark.parametrize("P", [1, 4])
def test_slab_reward_obs_matches_oracle(P):
    N = 256
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float64)
    w0 = random_states(ref, 1, seed=2).astype(np.complex128)[0]
    sl = make_slab(N, 100, torch.float64, P)
    sl.set_state(torch.from_numpy(w0).cuda(), None)
    obs, tke = sl.reward_obs()
    sl.close()
    want_obs = ref._trajectory_mean_obs(w0[None])
    want_tke = ref._avg_tke(w0[None])
    assert np.abs(obs.cpu().numpy() - want_obs).max() < 1e-10
    assert abs(float(tke) - float(want_tke)) < 1e-10 * max(1.0, abs(float(want_tke)))


def _band_limited(N, K, seed, dtype):
    """Hermitian-consistent random state with |kx|, ky <= K on an N x N grid, as omega_hat of rfft2 scaling."""
    rng = np.random.default_rng(seed)
    w = np.zeros((N, N // 2 + 1), dtype=np.complex128)
    kx = np.r_[0:K + 1, N - K:N]
    blk = rng.normal(size=(len(kx), K + 1)) + 1j * rng.normal(size=(len(kx), K + 1))
    w[np.ix_(kx, np.arange(K + 1))] = blk
    w = np.fft.rfft2(np.fft.irfft2(w, s=(N, N)))  # make it the transform of a real field
    w[0, 0] = 0
    return (w * (N * N) * 0.2).astype(dtype)


def test_slab_8192_band_limited_matches_1024():
    """No oracle runs at 8192^2 in seconds.  A state band-limited to K = 8 stays alias-free through one RK
    step (five products: 32 K = 256 < 512) on BOTH a 1024^2 and an 8192^2 grid, so the two grids must give
    the same spectrum up to the (N^2) transform scaling.  Forcing / control rows are zeroed (they are
    grid-dependent by the reference's endpoint-inclusive sampling)."""
    from hydrogym_b200 import KolmogorovStepper
    K, Ns, Nl = 8, 1024, 8192
    ws = _band_limited(Ns, K, 3, np.complex64)
    wl = np.zeros((Nl, Nl // 2 + 1), dtype=np.complex64)
    kxs = np.r_[0:K + 1, Ns - K:Ns]
    kxl = np.r_[0:K + 1, Nl - K:Nl]
    wl[np.ix_(kxl, np.arange(K + 1))] = ws[np.ix_(kxs, np.arange(K + 1))] * (Nl / Ns) ** 2
    st = KolmogorovStepper(grid_size=(Ns, Ns), Re=1000, num_envs=1, dtype=torch.float32, device="cuda:0")
    st.forcing_row[:] = 0
    st.set_control_rows(np.zeros_like(st.control_rows))
    w = torch.from_numpy(ws[None]).cuda()
    st.rk_steps(w, None, 1)
    want = w[0].cpu().numpy()
    st.close()
    for P in (1, 8):
        sl = make_slab(Nl, 1000, torch.float32, P, zero_forcing=True)
        sl.set_state(torch.from_numpy(wl).cuda(), None)
        sl.rk_steps(1)
        got = sl.get_state().cpu().numpy()
        sl.close()
        kk = 32 * K
        a = got[np.ix_(np.r_[0:kk + 1, Nl - kk:Nl], np.arange(kk + 1))] / (Nl / Ns) ** 2
        b = want[np.ix_(np.r_[0:kk + 1, Ns - kk:Ns], np.arange(kk + 1))]
        assert rel_l2(a, b) < 5e-5, (P, rel_l2(a, b))
        # nothing outside the band
        mask = np.ones(got.shape, dtype=bool)
        mask[np.ix_(np.r_[0:kk + 1, Nl - kk:Nl], np.arange(kk + 1))] = False
        assert np.abs(got[mask]).max() < 1e-4 * np.abs(a).max() * (Nl / Ns) ** 2


def test_slab_4096_fp64_band_limited_matches_1024():
    """fp64 at 4096^2 (the largest fp64 slab grid built): same alias-free band-limited property as the 8192^2 fp32 test,
    against the batched fp64 kernels at 1024^2, at fp64 tolerance; emulated ranks P = 1 and 4."""
    from hydrogym_b200 import KolmogorovStepper
    K, Ns, Nl = 8, 1024, 4096
    ws = _band_limited(Ns, K, 5, np.complex128)
    wl = np.zeros((Nl, Nl // 2 + 1), dtype=np.complex128)
    kxs = np.r_[0:K + 1, Ns - K:Ns]
    kxl = np.r_[0:K + 1, Nl - K:Nl]
    wl[np.ix_(kxl, np.arange(K + 1))] = ws[np.ix_(kxs, np.arange(K + 1))] * (Nl / Ns) ** 2
    st = KolmogorovStepper(grid_size=(Ns, Ns), Re=1000, num_envs=1, dtype=torch.float64, device="cuda:0")
    st.forcing_row[:] = 0
    st.set_control_rows(np.zeros_like(st.control_rows))
    w = torch.from_numpy(ws[None]).cuda()
    st.rk_steps(w, None, 1)
    want = w[0].cpu().numpy()
    st.close()
    kk = 32 * K
    for P in (1, 4):
        sl = make_slab(Nl, 1000, torch.float64, P, zero_forcing=True)
        sl.set_state(torch.from_numpy(wl).cuda(), None)
        sl.rk_steps(1)
        got = sl.get_state().cpu().numpy()
        sl.close()
        a = got[np.ix_(np.r_[0:kk + 1, Nl - kk:Nl], np.arange(kk + 1))] / (Nl / Ns) ** 2
        b = want[np.ix_(np.r_[0:kk + 1, Ns - kk:Ns], np.arange(kk + 1))]
        assert rel_l2(a, b) < 1e-11, (P, rel_l2(a, b))
        mask = np.ones(got.shape, dtype=bool)
        mask[np.ix_(np.r_[0:kk + 1, Nl - kk:Nl], np.arange(kk + 1))] = False
        assert np.abs(got[mask]).max() < 1e-11 * np.abs(a).max() * (Nl / Ns) ** 2


def test_slab_8192_matches_oracle_golden():
    """ONE solver step at 8192^2 against the NumPy oracle itself (tests/golden/make_slab8192_golden.py, run in the
    build container: fp64, reference forcing, control action ON, BROADBAND state with populated Nyquist row and
    column -- every twiddle of the 8192-point plans and the forcing / control rows at that size are exercised).
    The fixture holds a summary of the 537 MB result: norm, the |kx|, ky <= 32 block, 64 row-band and 64 column-band
    norms, the Nyquist-column norm and 512 scattered entries (64 on the Nyquist row, 64 on the Nyquist column)."""
    import os
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "golden"))
    import make_slab8192_golden as G
    gold = np.load(os.path.join(here, "golden", "slab8192_rk1_fp64.npz"))
    N = int(gold["N"])
    w0 = G.slab_state(N, int(gold["seed"]))
    s0 = G.summarise(w0, N)
    assert np.array_equal(s0["scattered"], gold["in_scattered"]) and s0["norm"] == float(gold["in_norm"])  # same input
    w0_64 = w0
    # fp32 on 1 and 8 (emulated) ranks; fp64 (LEAN packed-column / observation kernels, see slab.cuh) on 2 ranks
    for dtype, P in ((torch.float32, 1), (torch.float32, 8), (torch.float64, 2)):
        f32 = dtype == torch.float32
        w0 = torch.from_numpy(w0_64.astype(np.complex64) if f32 else w0_64).cuda()
        act = torch.from_numpy(gold["action"].astype(np.float32 if f32 else np.float64)).cuda()
        sl = make_slab(N, float(gold["Re"]), dtype, P, dt=float(gold["dt"]))
        sl.set_state(w0, act)
        del w0
        sl.rk_steps(1)
        obs, tke = sl.reward_obs()
        assert torch.isfinite(obs).all() and torch.isfinite(tke)
        got = sl.get_state().cpu().numpy().astype(np.complex128)
        sl.close()
        torch.cuda.empty_cache()
        s = G.summarise(got, N)
        errs = dict(
            norm=abs(s["norm"] / float(gold["out_norm"]) - 1),
            low=rel_l2(s["low"], gold["out_low"]),
            scattered=rel_l2(s["scattered"], gold["out_scattered"]),
            nyq_row=rel_l2(s["scattered"][384:448], gold["out_scattered"][384:448]),
            nyq_col=rel_l2(s["scattered"][448:], gold["out_scattered"][448:]),
            row_bands=float(np.abs(s["row_bands"] / gold["out_row_bands"] - 1).max()),
            col_bands=float(np.abs(s["col_bands"] / gold["out_col_bands"] - 1).max()),
            nyq_col_norm=abs(s["nyq_col_norm"] / float(gold["out_nyq_col_norm"]) - 1),
            # the step changed the state by 2.4 %: hold the CHANGE, not only the state, on the scattered entries
            delta=rel_l2(s["scattered"] - gold["in_scattered"], gold["out_scattered"] - gold["in_scattered"]),
            delta_low=rel_l2(s["low"] - gold["in_low"], gold["out_low"] - gold["in_low"]))
        print(f"[slab 8192 vs oracle] {'fp32' if f32 else 'fp64'} P={P}: " + " ".join(f"{k}={v:.2e}" for k, v in errs.items()))
        s_ = 1.0 if f32 else 1e-6  # fp64: the same bounds six orders of magnitude down (measured ~1e-15 .. 1e-11)
        assert errs["norm"] < 1e-6 * s_ and errs["low"] < 2e-5 * s_ and errs["scattered"] < 2e-5 * s_, errs
        assert errs["nyq_row"] < 5e-5 * s_ and errs["nyq_col"] < 5e-5 * s_, errs
        assert errs["row_bands"] < 1e-5 * s_ and errs["col_bands"] < 1e-5 * s_ and errs["nyq_col_norm"] < 1e-5 * s_, errs
        assert errs["delta"] < 1e-3 * s_ and errs["delta_low"] < 5e-3 * s_, errs


def test_slab_graph_replay_matches_host_loop():
    """kf_slab_rk_steps (one CUDA graph per RK step, single rank) vs the host-issued sub-stage loop: bit-identical."""
    N, steps = 256, 5
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float32)
    w0 = torch.from_numpy(random_states(ref, 1, seed=8).astype(np.complex64)[0]).cuda()
    act = torch.tensor(ACTION, dtype=torch.float32, device="cuda")
    out = {}
    for use_graph in (True, False):
        sl = make_slab(N, 100, torch.float32, 1, exchange="p2p", use_graph=use_graph)
        sl.set_state(w0, act)
        sl.rk_steps(steps)
        sl.rk_steps(1)  # a second call replays the instantiated graph / falls back to direct launches
        out[use_graph] = sl.get_state().cpu().numpy()
        launches = sl.launch_count
        sl.close()
        assert launches > 0
    assert np.isfinite(out[True]).all() and np.array_equal(out[True], out[False])


def test_slab_pipelined_column_pass_is_bit_identical():
    """KF_SLAB_COL_CHUNKS = n > 1 (off by default): the column pass in n pieces with each piece's field transposes on a
    side stream -- same kernels on the same data, so the result must not change by a bit (4 emulated ranks, p2p)."""
    import os
    N, steps = 256, 3
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float32)
    w0 = torch.from_numpy(random_states(ref, 1, seed=12).astype(np.complex64)[0]).cuda()
    act = torch.tensor(ACTION, dtype=torch.float32, device="cuda")
    out = {}
    old = os.environ.get("KF_SLAB_COL_CHUNKS")
    try:
        for pieces in ("1", "2", "4"):
            os.environ["KF_SLAB_COL_CHUNKS"] = pieces
            sl = make_slab(N, 100, torch.float32, 4, exchange="p2p")
            sl.set_state(w0, act)
            sl.rk_steps(steps)
            out[pieces] = sl.get_state().cpu().numpy()
            sl.close()
    finally:
        if old is None:
            os.environ.pop("KF_SLAB_COL_CHUNKS", None)
        else:
            os.environ["KF_SLAB_COL_CHUNKS"] = old
    assert np.isfinite(out["1"]).all()
    assert np.array_equal(out["1"], out["2"]) and np.array_equal(out["1"], out["4"])


@pytest.mark.parametrize("dtype,P,N", [(torch.float32, 1, 256), (torch.float32, 4, 256), (torch.float32, 8, 256),
                                       (torch.float32, 2, 1024), (torch.float64, 4, 256)])
def test_slab_direct_field_exchange_is_bit_identical(dtype, P, N):
    """KF_SLAB_DIRECT=1 (the default with more than one rank; 0 = sender-side transposes): the column kernels store their x-space field columns straight into the destination ranks'
    receive buffers and the receiver transposes them locally, instead of sender-side transposes with remote stores.
    Same arithmetic on the same values -- state, reward and observation must not change by a bit."""
    import os
    steps = 2
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float32)
    cdt = np.complex64 if dtype == torch.float32 else np.complex128
    w0 = torch.from_numpy(random_states(ref, 1, seed=21).astype(cdt)[0]).cuda()
    act = torch.tensor(ACTION, dtype=dtype, device="cuda")
    out = {}
    old = os.environ.get("KF_SLAB_DIRECT")
    try:
        for mode in ("0", "1"):
            os.environ["KF_SLAB_DIRECT"] = mode
            sl = make_slab(N, 100, dtype, P, exchange="p2p")
            sl.set_state(w0, act)
            sl.rk_steps(steps)
            obs, r = sl.reward_obs()
            out[mode] = (sl.get_state().cpu().numpy(), float(r), obs.cpu().numpy())
            sl.close()
    finally:
        if old is None:
            os.environ.pop("KF_SLAB_DIRECT", None)
        else:
            os.environ["KF_SLAB_DIRECT"] = old
    assert np.isfinite(out["0"][0]).all()
    assert np.array_equal(out["0"][0], out["1"][0])
    # the reward's partial sums are atomic adds (order not fixed): last-bit differences are allowed there
    assert abs(out["0"][1] - out["1"][1]) <= 1e-12 * abs(out["0"][1]) + (0 if dtype == torch.float64 else 1e-6)
    assert np.array_equal(out["0"][2], out["1"][2])


def test_slab_phase_times_advance_like_rk_steps():
    """``phase_times`` (per-phase events for the bench's slab record) runs the same kernels as ``rk_steps``."""
    N = 256
    ref = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(N, N)), dtype=np.float32)
    w0 = torch.from_numpy(random_states(ref, 1, seed=3).astype(np.complex64)[0]).cuda()
    act = torch.tensor(ACTION, dtype=torch.float32, device="cuda")
    out = []
    for timed in (False, True):
        sl = make_slab(N, 100, torch.float32, 2, exchange="p2p")
        sl.set_state(w0, act)
        if timed:
            ph = sl.phase_times(2)
            assert set(ph) == {"row", "wait_row", "col", "wait_col"} and all(v >= 0 for v in ph.values())
            assert ph["row"] > 0 and ph["col"] > 0
        else:
            sl.rk_steps(2)
        out.append(sl.get_state().cpu().numpy())
        sl.close()
    assert np.array_equal(out[0], out[1])
