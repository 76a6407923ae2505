"""CPU tests of the product's host side: init-time constants vs the oracle, step counts, the C-ABI
library (loads, exports every symbol include/kolmo_b200.h declares), loud failure without a GPU."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest
import torch

from oracle import kolmogorov_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_constants_match_oracle():
    from hydrogym_b200 import constants
    for N in (64, 128):
        env = orc.KolmogorovFlow(flow_config=dict(grid_size=(N, N)))
        assert np.allclose(constants.forcing_row(4, (N, N)), env.equation._forcing[0], rtol=1e-12, atol=1e-9)
        unit = np.eye(4)
        rows = constants.control_rows((4, 5, 6, 7), (N, N))
        for i in range(4):
            ref = env.equation.control_term(env.equation._forcing, env._control_field(unit[i], env.default_params))
            assert np.allclose(rows[i], ref[0], rtol=1e-12, atol=1e-9)
        assert np.allclose(constants.initial_vorticity_hat((N, N)), env.flow.initialize_state(), atol=1e-9)
        kx, ky = constants.fft_mesh((N, N))
        assert np.allclose(kx, env.kx) and np.allclose(ky, env.ky)
        assert np.allclose(kx[:, 0] * 2 * np.pi, np.fft.fftfreq(N, 1 / N))          # 2*pi*kx is the integer p


def test_step_counts_and_end_time_guard():
    from hydrogym_b200 import constants
    assert constants.step_counts(1e-3, 10.0, 1) == (999, 10)
    assert constants.step_counts(0.1, 10.0, 1) == (9, 11)
    with pytest.raises(ValueError):
        constants.step_counts(1e-3, 0.5, 1)


def test_library_exports_every_declared_symbol():
    from hydrogym_b200 import _lib
    header = open(os.path.join(ROOT, "include", "kolmo_b200.h")).read()
    declared = set(re.findall(r"\b(kf_[a-z0-9_]+)\s*\(", header))
    channel = open(os.path.join(ROOT, "include", "channel_b200.h")).read()
    declared |= set(re.findall(r"\b(kc_[a-z0-9_]+)\s*\(", channel))
    assert declared == set(_lib.SYMBOLS)
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert all(re.search(rf" T {n}\b", nm) for n in declared)


def test_config_struct_matches_header_field_order():
    from hydrogym_b200 import _lib
    header = open(os.path.join(ROOT, "include", "kolmo_b200.h")).read()
    body = header[header.index("typedef struct kf_config {"):header.index("} kf_config;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in re.findall(r"(?:int|double)\s+([^;]+);", body):
        names += [n.strip() for n in decl.split(",")]
    assert names == [f[0] for f in _lib.KfConfig._fields_]


def test_channel_config_struct_matches_header_field_order():
    from hydrogym_b200 import _lib
    header = open(os.path.join(ROOT, "include", "channel_b200.h")).read()
    body = header[header.index("typedef struct {"):header.index("} kc_config;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in re.findall(r"(?:int|double)\s+([^;]+);", body):
        names += [n.strip() for n in decl.split(",")]
    assert names == [f[0] for f in _lib.KcConfig._fields_]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_channel_env_fails_loudly_without_gpu():
    """No CPU fallback: without a CUDA device the channel stepper raises instead of computing elsewhere."""
    import numpy as np
    from hydrogym_b200 import ChannelStepper
    with pytest.raises(RuntimeError):
        ChannelStepper(num_envs=1, device="cuda:0")
    with pytest.raises(RuntimeError):
        ChannelStepper(num_envs=1, device="cpu")
    del np


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu():
    from hydrogym_b200 import KolmogorovFlow, KolmogorovStepper, _lib
    with pytest.raises(RuntimeError):
        KolmogorovStepper(num_envs=1)
    env = KolmogorovFlow(dict(dt=1e-2), dict(grid_size=(64, 64)))      # host-only construction works
    assert (env.n_inner, env.n_outer) == (99, 10) and env.default_params.obs_dim == 64
    with pytest.raises(RuntimeError):
        env.reset_env(None, env.default_params)
    # the C ABI itself refuses too (no device), with a message
    lib = _lib.load()
    h = ctypes.c_void_p(0)
    cfg = _lib.KfConfig(N=64, M=64, B=1, dtype=0, nu=0.01, dt=1e-3, n_inner=9, n_outer=1, dealias=0, drag=0.0,
                        device=0, obs_size=8, min_action=-0.5, max_action=0.5, chunk=0, use_graph=0)
    assert lib.kf_create(ctypes.byref(h), ctypes.byref(cfg)) != 0
    assert len(lib.kf_last_error()) > 0


def test_env_api_surface_without_gpu():
    from hydrogym_b200 import KolmogorovFlow
    env = KolmogorovFlow(None, None)
    p = env.default_params
    assert env.name == "KolmogorovFlow" and (env.n, env.m) == (64, 64)
    assert (p.dt, p.action_time, p.save_time, p.max_episode_steps, p.reward_alpha) == (1e-3, 10.0, 1, 1000, 0.0)
    assert (p.k1, p.k2, p.k3, p.k4, p.action_dim, p.obs_dim) == (4, 5, 6, 7, 4, 64)
    assert p.replace(reward_alpha=2.0).reward_alpha == 2.0
    a, o = env.action_space(p), env.observation_space(p)
    assert a.shape == (4,) and (a.low, a.high) == (-0.5, 0.5) and o.shape == (64,) and o.low == -np.inf
    with pytest.raises(ValueError):
        KolmogorovFlow(dict(action_time=0.5), None)                     # solve()'s end_time < 1 guard
    with pytest.raises(ValueError):
        KolmogorovFlow(None, dict(grid_size=(64, 128)))


def test_radix_butterflies_on_host():
    """The in-register DFT butterflies are __host__ __device__: compile them for the CPU and check vs a naive DFT."""
    exe = "/tmp/kf_radix_host_test"
    src = os.path.join(ROOT, "tests", "cpu", "radix_host_test.cu")
    subprocess.check_call(["nvcc", "-std=c++17", "-O2", "-Wno-deprecated-gpu-targets", src, "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout


def test_non_default_domain_is_refused_not_silently_mixed():
    """The kernels work on integer wavenumbers (period 2*pi): any other domain must raise instead of combining
    host-side constants of one domain with device-side operators of another (kolmogorov.py:33-34 defaults)."""
    import numpy as np
    import pytest
    from hydrogym_b200 import constants
    from hydrogym_b200.envs.kolmogorov import FlowConfig
    constants.check_domain((0, 2 * np.pi), (0, 2 * np.pi))
    FlowConfig(grid_size=(64, 64))
    for bad in (dict(domain_x=(0, 4 * np.pi)), dict(domain_y=(0, np.pi)), dict(domain_x=(1.0, 1.0 + 2 * np.pi))):
        with pytest.raises(ValueError, match="default domain"):
            FlowConfig(grid_size=(64, 64), **bad)
