"""Scratch: SUSTAINED throughput (several seconds per point, power-capped clocks) of the flagship config."""
import os, subprocess, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hydrogym_b200 import KolmogorovStepper

def clocks(stop, out):
    while not stop.is_set():
        try:
            r = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"],
                               capture_output=True, text=True, timeout=5).stdout.strip().split(",")
            out.append((float(r[0]), float(r[1])))
        except Exception:
            pass
        time.sleep(0.2)

def run(N, B, dtype, warm=1500, steps=3000, **kw):
    st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=dtype, device="cuda:0", dealias=True, **kw)
    cd = torch.complex64 if dtype == torch.float32 else torch.complex128
    g = torch.Generator(device="cuda").manual_seed(0)
    w = torch.randn(B, N, N // 2 + 1, dtype=cd, device="cuda", generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]; q = torch.arange(N // 2 + 1, device="cuda")[None, :]
    w = w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.05
    st.rk_steps(w, None, warm)
    torch.cuda.synchronize()
    stop, samples = threading.Event(), []
    th = threading.Thread(target=clocks, args=(stop, samples)); th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); st.rk_steps(w, None, steps); e1.record()
    torch.cuda.synchronize()
    stop.set(); th.join()
    ms = e0.elapsed_time(e1) / steps
    C = N * (N // 2 + 1) * (8 if dtype == torch.float32 else 16)
    mhz = sorted(s[0] for s in samples)[len(samples) // 2] if samples else 0
    pw = max((s[1] for s in samples), default=0)
    print(f"N={N} B={B} {kw} lanes={os.environ.get('KF_LANES','auto')}: {ms:.4f} ms/RK-step  {70*C*B/ms/1e6/6534*100:.1f}% roofline  "
          f"sm {mhz:.0f} MHz  {pw:.0f} W  finite={bool(torch.isfinite(torch.view_as_real(w)).all())}", flush=True)
    st.close()

if __name__ == "__main__":
    N, B = int(os.environ.get("KF_N", 256)), int(os.environ.get("KF_B", 256))
    steps = int(os.environ.get("KF_STEPS", 3000))
    for ch in [int(a) for a in sys.argv[1:]] or [0]:
        run(N, B, torch.float64 if os.environ.get("KF_DTYPE") == "fp64" else torch.float32, chunk=ch, warm=steps // 2, steps=steps)
