"""Short run of the flagship configuration for ncu (256^2, Re=1000, fp32 by default)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from hydrogym_b200 import KolmogorovStepper  # noqa: E402

N = int(os.environ.get("KF_N", 256))
B = int(os.environ.get("KF_B", 256))
dtype = torch.float64 if os.environ.get("KF_DTYPE", "fp32") == "fp64" else torch.float32
steps = int(os.environ.get("KF_STEPS", 4))
chunk = int(os.environ.get("KF_CHUNK", 0))
st = KolmogorovStepper(grid_size=(N, N), Re=1000, num_envs=B, dtype=dtype, device="cuda:0", use_graph=False, chunk=chunk)
cd = torch.complex64 if dtype == torch.float32 else torch.complex128
g = torch.Generator(device="cuda").manual_seed(0)
w = torch.randn(B, N, N // 2 + 1, dtype=cd, device="cuda", generator=g)
p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]
q = torch.arange(N // 2 + 1, device="cuda")[None, :]
w = w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.5
st.rk_steps(w, None, steps)
torch.cuda.synchronize()
print("finite:", bool(torch.isfinite(torch.view_as_real(w)).all()))
