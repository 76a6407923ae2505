"""Scratch: sustained timing of .so build variants (KF_LIB)."""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for lib in sorted(glob.glob(os.path.join(root, "variants", "lib_*.so"))):
    print("==", os.path.basename(lib), flush=True)
    subprocess.run([sys.executable, os.path.join(root, "tools", "gpu_sustained.py")] + sys.argv[1:], env=dict(os.environ, KF_LIB=lib), cwd=root)
