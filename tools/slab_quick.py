"""Scratch GPU timing of the slab solver on one device (P emulated ranks)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydrogym_b200.slab import SlabKolmogorovSolver

def run(N, P, dtype=torch.float32, steps=3):
    sl = SlabKolmogorovSolver(grid_size=(N, N), Re=1000, dtype=dtype, device="cuda:0", local_ranks=P, dealias=True)
    cd = torch.complex64 if dtype == torch.float32 else torch.complex128
    g = torch.Generator(device="cuda").manual_seed(0)
    w = torch.randn(N, N // 2 + 1, dtype=cd, device="cuda", generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]; q = torch.arange(N // 2 + 1, device="cuda")[None, :]
    w = w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.5
    sl.set_state(w, None)
    sl.rk_steps(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(); sl.rk_steps(steps); e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    C = N * (N // 2 + 1) * (8 if dtype == torch.float32 else 16)
    out = sl.get_state()
    print(f"slab N={N} P={P} {dtype}: {ms:.2f} ms/RK-step  alg(70C) {70*C/ms/1e6:.0f} GB/s  finite={bool(torch.isfinite(torch.view_as_real(out)).all())}", flush=True)
    sl.close()

if __name__ == "__main__":
    run(1024, 1); run(4096, 1); run(8192, 1); run(8192, 2, steps=2)
