"""Scratch: two half-batches on two streams vs one full batch (do kernel tails / phases overlap?)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydrogym_b200 import KolmogorovStepper

def state(B, N, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randn(B, N, N // 2 + 1, dtype=torch.complex64, device="cuda", generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]; q = torch.arange(N // 2 + 1, device="cuda")[None, :]
    return w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.5

def run(parts, N=256, B=256, steps=40):
    Bp = B // parts
    sts = [KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=Bp, dtype=torch.float32, device="cuda:0") for _ in range(parts)]
    ws = [state(Bp, N, i) for i in range(parts)]
    streams = [torch.cuda.Stream() for _ in range(parts)]
    for st, w, s in zip(sts, ws, streams):
        with torch.cuda.stream(s):
            st.rk_steps(w, None, 8)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for st, w, s in zip(sts, ws, streams):
        with torch.cuda.stream(s):
            st.rk_steps(w, None, steps)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    C = N * (N // 2 + 1) * 8
    print(f"{parts} stream(s) x {Bp} envs: {ms:.3f} ms per RK step of all {B} envs  ({70*C*B/ms/1e6/6534*100:.1f}% of roofline)", flush=True)
    for st in sts: st.close()

if __name__ == "__main__":
    run(1); run(2); run(4); run(1, steps=80); run(2, steps=80)
