import sys, torch
sys.path.insert(0, __file__.rsplit("/", 2)[0] + "/tools")
from gpu_quick import run
if __name__ == "__main__":
    run(1024, 16, torch.float32, steps=10)
    run(512, 64, torch.float32, steps=10)
    run(64, 1024, torch.float32, steps=20)
    run(128, 1024, torch.float64, steps=10)
