"""Development timing of the on-chip-resident kernel against the two-kernel path (KF_RESIDENT=1 / 0):
ms per RK step, RK-steps*env/s and the 70C algorithmic fraction for small grids."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydrogym_b200 import KolmogorovStepper

CASES = [(64, 1, "f64", 2000), (64, 1, "f32", 2000), (64, 64, "f32", 1000), (64, 296, "f32", 1000), (64, 1024, "f32", 400),
         (64, 4096, "f32", 200), (64, 1024, "f64", 200), (128, 1, "f32", 1000), (128, 74, "f32", 400), (128, 1024, "f32", 200),
         (256, 1, "f32", 400), (256, 18, "f32", 200), (256, 256, "f32", 100)]


def run(N, B, dt, steps):
    dtype = torch.float32 if dt == "f32" else torch.float64
    cd = torch.complex64 if dt == "f32" else torch.complex128
    out = {}
    for res in ("1", "0"):
        os.environ["KF_RESIDENT"] = res
        st = KolmogorovStepper(grid_size=(N, N), Re=100, num_envs=B, dtype=dtype, device="cuda:0")
        g = torch.Generator(device="cuda").manual_seed(0)
        w = torch.randn(B, N, N // 2 + 1, dtype=cd, device="cuda", generator=g)
        p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]
        q = torch.arange(N // 2 + 1, device="cuda")[None, :]
        w = (w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.1).contiguous()
        st.rk_steps(w, None, max(8, steps // 10))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st.rk_steps(w, None, steps)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        C = N * (N // 2 + 1) * (8 if dt == "f32" else 16)
        out[res] = (ms, 70 * C * B / (ms * 1e-3) / 1e9 / 6552.6, bool(torch.isfinite(torch.view_as_real(w)).all()))
        st.close()
    r, k = out["1"], out["0"]
    print(f"N={N:4d} B={B:5d} {dt}: resident {r[0]*1e3:9.2f} us/RK step (frac {r[1]:.3f}, {B/r[0]*1e3:11.0f} RK*env/s)   "
          f"two-kernel {k[0]*1e3:9.2f} us (frac {k[1]:.3f})   speed-up {k[0]/r[0]:.2f}x  finite={r[2] and k[2]}", flush=True)


if __name__ == "__main__":
    for case in CASES:
        if case[0] == 128 and os.environ.get("KF_SKIP_128"):
            continue
        run(*case)
