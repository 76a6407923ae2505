"""GPU: ``kf_reward_obs_noalloc`` (allocation-free variant of ``kf_reward_obs``: its scratch sums live in a static device
array inside the library; no counterpart in the reference) against the
existing entry point, through the C ABI, on two calls in a row.  Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ENV_SNAPSHOTS = 16384  # KF_NOALLOC_MAX_ENV_SNAPSHOTS in include/kolmo_b200.h


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _trajectory(B, T, N, cdtype, seed, amp):
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randn(B, T, N, N // 2 + 1, dtype=cdtype, device="cuda", generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]
    q = torch.arange(N // 2 + 1, device="cuda")[None, :]
    return (w * torch.exp(-(p * p + q * q) / 64.0) * (N * N * amp)).contiguous()


def _call(st, name, traj):
    from hydrogym_b200 import _lib
    B, T = traj.shape[0], traj.shape[1]
    obs = torch.full((B, st.obs_size ** 2), float("nan"), dtype=st.dtype, device=st.device)
    tke = torch.full((B,), float("nan"), dtype=st.dtype, device=st.device)
    nonf = torch.full((B,), -1, dtype=torch.int32, device=st.device)
    stream = ctypes.c_void_p(torch.cuda.current_stream(st.device).cuda_stream)
    rc = getattr(st._lib, name)(st._h, _ptr(traj), int(T), _ptr(obs), _ptr(tke), _ptr(nonf), stream)
    _lib.check(rc, name)
    torch.cuda.synchronize()
    return obs.cpu().numpy(), tke.cpu().numpy(), nonf.cpu().numpy()


@pytest.mark.parametrize("dtype,rtol", [(torch.float64, 1e-12), (torch.float32, 1e-5)])
def test_noalloc_reward_obs_matches_existing_entry_point_twice(dtype, rtol):
    from hydrogym_b200 import KolmogorovStepper
    N, B, T = 64, 6, 5
    cd = torch.complex128 if dtype == torch.float64 else torch.complex64
    st = KolmogorovStepper(grid_size=(N, N), Re=100, num_envs=B, dtype=dtype, device="cuda:0", n_inner=1, n_outer=T)
    try:
        # second call: a different, much weaker trajectory -- sums left over from the first call would dominate it
        for seed, amp in ((1, 0.3), (2, 0.003)):
            traj = _trajectory(B, T, N, cd, seed, amp)
            want = _call(st, "kf_reward_obs", traj)
            got = _call(st, "kf_reward_obs_noalloc", traj)
            assert np.isfinite(got[0]).all() and np.isfinite(got[1]).all() and (got[1] > 0).all()
            # the sums are fp64 atomics: order-dependent in the last bits, nothing more
            assert np.allclose(got[0], want[0], rtol=rtol, atol=rtol * np.abs(want[0]).max())
            assert np.allclose(got[1], want[1], rtol=rtol, atol=0)
            assert np.array_equal(got[2], want[2]) and (got[2] == 0).all()
    finally:
        st.close()


def test_noalloc_reward_obs_rejects_sizes_beyond_the_static_scratch():
    from hydrogym_b200 import KolmogorovStepper
    st = KolmogorovStepper(grid_size=(64, 64), Re=100, num_envs=2, dtype=torch.float32, device="cuda:0", n_inner=1, n_outer=1)
    try:
        dummy = torch.zeros((2, 1, 64, 33), dtype=torch.complex64, device="cuda")  # never read: the size check comes first
        stream = ctypes.c_void_p(torch.cuda.current_stream(st.device).cuda_stream)
        n_snap = MAX_ENV_SNAPSHOTS // 2 + 1  # B * n_snap = MAX + 2
        rc = st._lib.kf_reward_obs_noalloc(st._h, _ptr(dummy), n_snap, None, None, None, stream)
        assert rc != 0
        msg = st._lib.kf_last_error().decode()
        assert "exceeds the static scratch" in msg and str(MAX_ENV_SNAPSHOTS) in msg
        # the limit itself is accepted by the size check (a null trajectory is what stops this call)
        rc = st._lib.kf_reward_obs_noalloc(st._h, None, MAX_ENV_SNAPSHOTS // 2, None, None, None, stream)
        assert rc != 0 and "null argument" in st._lib.kf_last_error().decode()
    finally:
        st.close()
