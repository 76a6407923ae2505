"""Scratch: time .so build variants (KF_LIB) on the flagship config."""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = ("import sys; sys.path.insert(0, %r); import torch; from tools.gpu_quick import run; "
        "%s") % (root, os.environ.get("KF_VARIANT_RUN", "run(256, 256, torch.float32, steps=40)"))
for lib in sorted(glob.glob(os.path.join(root, "variants", "lib_*.so"))):
    print("==", os.path.basename(lib), flush=True)
    subprocess.run([sys.executable, "-c", code], env=dict(os.environ, KF_LIB=lib), cwd=root)
