"""Scratch: does the 256^2 Re=1000 run stay finite? (dealias on/off, fp32/fp64, synthetic vs reference IC)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import synthetic_omega_hat, synthetic_actions
from hydrogym_b200 import KolmogorovStepper, constants

N, B, STEPS = 256, 16, 6
for dtype in (torch.float32, torch.float64):
    npdt = np.float32 if dtype == torch.float32 else np.float64
    for dealias in (False, True):
        for ic in ("synthetic", "reference"):
            st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=dtype, device="cuda:0", dealias=dealias)
            if ic == "synthetic":
                w = torch.from_numpy(np.stack([synthetic_omega_hat(N, i, npdt) for i in range(B)])).cuda()
            else:
                w0 = constants.initial_vorticity_hat((N, N)).astype(np.complex64 if npdt == np.float32 else np.complex128)
                w = torch.from_numpy(np.stack([w0] * B)).cuda()
            acts = torch.from_numpy(synthetic_actions(STEPS, B, npdt)).cuda()
            out = []
            for s in range(STEPS):
                _, _, obs, tke, nonf = st.step(w, acts[s], want_trajectory=False, in_place=True)
                vort = torch.fft.irfft2(w, s=(N, N))
                out.append((int(nonf.sum()), float(tke[torch.isfinite(tke)].mean()) if torch.isfinite(tke).any() else float("nan"),
                            float(vort[torch.isfinite(vort)].abs().max()) if torch.isfinite(vort).any() else float("nan")))
            print(dtype, "dealias", dealias, ic, " (nonfinite, mean tke, max|w|) per env-step:", out, flush=True)
            st.close()
