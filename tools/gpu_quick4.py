import sys, torch
sys.path.insert(0, __file__.rsplit("/", 2)[0] + "/tools")
from gpu_quick import run
if __name__ == "__main__":
    for ch in (0, 37, 74, 111, 148, 18):
        run(256, 256, torch.float32, steps=40, chunk=ch)
    run(256, 259, torch.float32, steps=40, chunk=37)
    run(256, 256, torch.float32, steps=40, chunk=37, dealias=True)
    run(256, 256, torch.float32, steps=40, chunk=0, dealias=True)
