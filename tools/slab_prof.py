import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from slab_quick import run
if __name__ == "__main__":
    run(int(sys.argv[1]) if len(sys.argv) > 1 else 8192, 1, steps=1)
