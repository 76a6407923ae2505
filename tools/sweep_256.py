"""Development sweep of the flagship configuration (256^2 x 256 fp32, literal arithmetic): ms per RK step for a list of
(chunk, KF_L2_HINTS) settings, ~600 RK steps per point after warm-up.  `python tools/sweep_256.py` runs every
variants/lib_*.so (KF_LIB) in its own interpreter; `--here` runs the settings against the library already selected."""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
SETTINGS = [(0, 0, 0), (0, 1, 0), (37, 1, 0), (0, 1, 1)]  # (chunk, KF_L2_HINTS, KF_RESIDENT)


def here():
    import torch
    from hydrogym_b200 import KolmogorovStepper
    N, B = 256, 256
    for chunk, hints, colsm in SETTINGS:
        os.environ["KF_L2_HINTS"] = str(hints)
        os.environ["KF_RESIDENT"] = str(colsm)
        st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=torch.float32, device="cuda:0", chunk=chunk)
        g = torch.Generator(device="cuda").manual_seed(0)
        w = torch.randn(B, N, N // 2 + 1, dtype=torch.complex64, device="cuda", generator=g)
        p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]
        q = torch.arange(N // 2 + 1, device="cuda")[None, :]
        w = w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.25
        st.rk_steps(w, None, 200)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st.rk_steps(w, None, 600)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 600
        C = N * (N // 2 + 1) * 8
        mc, mr = st.profile_kernels(20)
        print(f"  chunk {st.chunk:3d} (asked {chunk:2d}) lanes {st.lanes} hints {hints} resident(cluster of 8) {colsm}: {ms:.4f} ms/RK step  "
              f"frac {70 * C * B / (ms * 1e-3) / 1e9 / 6552.6:.3f}  finite={bool(torch.isfinite(torch.view_as_real(w)).all())}  col {mc*1e3:.1f} us row {mr*1e3:.1f} us",
              flush=True)
        st.close()


if __name__ == "__main__":
    if "--here" in sys.argv:
        here()
    else:
        for lib in sorted(glob.glob(os.path.join(root, "variants", "lib_*.so"))):
            print("==", os.path.basename(lib), flush=True)
            subprocess.run([sys.executable, os.path.abspath(__file__), "--here"], env=dict(os.environ, KF_LIB=lib), cwd=root)
