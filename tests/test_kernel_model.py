"""CPU check of the kernels' ALGORITHM (layout packing, two-rows-per-FFT, Stockham indexing) against the oracle."""
import numpy as np
import pytest

import kernel_model as km
from oracle import kolmogorov_oracle as orc


@pytest.mark.parametrize("L,radices", [(64, (8, 8)), (128, (16, 8)), (128, (8, 2, 8)), (256, (16, 16)), (256, (8, 4, 8)),
                                       (512, (8, 8, 8)), (1024, (16, 4, 16)), (1024, (8, 2, 8, 8))])
def test_stockham_index_formulas(L, radices):
    rng = np.random.default_rng(L)
    x = rng.normal(size=(2, L)) + 1j * rng.normal(size=(2, L))
    assert np.abs(km.stockham_fft(x, radices, -1) - np.fft.fft(x)).max() < 1e-11
    assert np.abs(km.stockham_fft(x, radices, +1) - np.fft.ifft(x) * L).max() < 1e-11


def test_packed_layout_rk_step_equals_oracle_even_for_non_hermitian_states():
    env = orc.KolmogorovFlow(flow_config=dict(Re=100, grid_size=(64, 64)))
    rng = np.random.default_rng(0)
    w0 = env.flow.initialize_state() + (rng.normal(size=(64, 33)) + 1j * rng.normal(size=(64, 33))) * 5
    a = np.array([-0.25, -0.03, 0.02, 0.01])
    ch = env.equation.control_term(w0, env._control_field(a, env.default_params))
    ref = env.integrator.time_step(w0, ch)
    G = (env.equation._forcing + ch)[0]
    m, n = km.rk_step(*km.import_state(w0), G, env.flow.nu, 1e-3)
    out = km.export_state(m, n)
    assert np.linalg.norm(out - ref) / np.linalg.norm(ref) < 1e-13
