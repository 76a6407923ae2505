"""Scratch A/B of programmatic dependent launch (KF_PDL) on the flagship configuration: one process per setting."""
import os, subprocess, sys
code = "import sys; sys.path.insert(0, 'tools'); import torch; from gpu_sustained import run; run(256, 256, torch.float32, warm=1000, steps=3000); run(128, 1024, torch.float32, warm=1000, steps=3000)"
for v in ("0", "1", "0", "1"):
    env = dict(os.environ, KF_PDL=v)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    print("KF_PDL=" + v, out.stdout.strip().replace("\n", " | "), out.stderr.strip()[-300:], flush=True)
