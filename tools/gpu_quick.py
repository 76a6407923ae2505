"""Quick GPU timing of named configuration presets (development aid, not the bench): ms per RK step, the 70C
algorithmic GB/s and the isolated column / row kernel times."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hydrogym_b200 import KolmogorovStepper

def run(N, B, dtype, steps=20, **kw):
    st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=dtype, device="cuda:0", **kw)
    cd = torch.complex64 if dtype == torch.float32 else torch.complex128
    g = torch.Generator(device="cuda").manual_seed(0)
    w = torch.randn(B, N, N // 2 + 1, dtype=cd, device="cuda", generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]; q = torch.arange(N // 2 + 1, device="cuda")[None, :]
    w = w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.5
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        st.rk_steps(w, None, 8)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); st.rk_steps(w, None, steps); e1.record(s)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    w_bytes = 8 if dtype == torch.float32 else 16
    C = N * (N // 2 + 1) * w_bytes
    gbs = 70 * C * B / (ms * 1e-3) / 1e9
    print(f"N={N} B={B} {dtype} {kw}: {ms:.3f} ms/RK-step  {B/ms*1e3:.0f} RKsteps*env/s  alg {gbs:.0f} GB/s ({gbs/6545*100:.1f}%) finite={bool(torch.isfinite(torch.view_as_real(w)).all())}", flush=True)
    mc, mr = st.profile_kernels(20)
    print(f"      col {mc*1e3:.1f} us  row {mr*1e3:.1f} us per launch (chunk {st.chunk})", flush=True)
    st.close()

PRESETS = {
    # name: list of (N, B, dtype, steps, kwargs)
    "flagship": [(256, 256, "f32", 40, {})],
    "configs": [(256, 256, "f32", 40, {}), (256, 256, "f64", 20, {}), (128, 1024, "f32", 20, {}),
                (1024, 16, "f32", 10, {}), (64, 1, "f64", 200, {})],
    "chunks": [(256, 256, "f32", 40, dict(chunk=c)) for c in (0, 16, 32, 37, 64, 74, 128)],
    "other": [(1024, 16, "f32", 10, {}), (512, 64, "f32", 10, {}), (64, 1024, "f32", 20, {}), (128, 1024, "f64", 10, {})],
    "nograph": [(256, 256, "f32", 40, dict(use_graph=False))],
    "c1024": [(1024, 16, "f32", 10, dict(chunk=c)) for c in (0, 8, 4, 2)],
    "d256": [(256, 256, "f64", 20, dict(chunk=c)) for c in (0, 128, 32, 16)],
    "dealias": [(256, 256, "f32", 40, dict(dealias=True))],
}

if __name__ == "__main__":
    # usage: tools/gpu_quick.py [preset ...]   (default: flagship); presets: see PRESETS
    for name in (sys.argv[1:] or ["flagship"]):
        for N, B, dt, steps, kw in PRESETS[name]:
            run(N, B, torch.float32 if dt == "f32" else torch.float64, steps=steps, **kw)
