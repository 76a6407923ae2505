"""Slab-decomposed single-grid bench (BASELINE.json configs[4]): RK steps/s of ONE N x N flow over the ranks of
`torchrun` (or one GPU without it).  Prints one JSON line on rank 0.  Not the driver's bench (that is bench.py)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from hydrogym_b200.slab import SlabKolmogorovSolver

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--exchange", default="p2p", choices=["nccl", "p2p"])
    ap.add_argument("--dt", type=float, default=0.0, help="0 = CFL-safe default 0.8/N (1e-4 at 8192)")
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--phases", type=int, default=0, help="RK steps of per-phase event timing after the run (p2p)")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N = a.grid
    dt = a.dt if a.dt > 0 else 0.8 / N  # u ~ 1: CFL = u dt N / (2 pi) ~ 0.13
    rd = torch.float32 if a.dtype == "fp32" else torch.float64
    cd = torch.complex64 if a.dtype == "fp32" else torch.complex128
    sl = SlabKolmogorovSolver(grid_size=(N, N), Re=1000, dtype=rd, device=f"cuda:{local}", dealias=True, dt=dt, exchange=a.exchange)
    g = torch.Generator(device="cuda").manual_seed(0)
    w = torch.randn(N, N // 2 + 1, dtype=cd, device="cuda", generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device="cuda")[:, None]; q = torch.arange(N // 2 + 1, device="cuda")[None, :]
    w = w * torch.exp(-(p * p + q * q) / 64.0) * N * N * 0.5
    sl.set_state(w, None)
    del w
    sl.rk_steps(a.warmup)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); sl.rk_steps(a.steps); e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    phases = None
    if a.phases > 0 and a.exchange == "p2p":
        ph = sl.phase_times(a.phases)
        t = torch.tensor(list(ph.values()), device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        phases = {"rank0": ph, "max_over_ranks": dict(zip(ph, t.tolist()))}
    obs, tke = sl.reward_obs()
    C = N * (N // 2 + 1) * (8 if a.dtype == "fp32" else 16)
    if rank == 0:
        print(json.dumps({"metric": "RK steps/s (single grid, slab-decomposed)", "grid": N, "dtype": a.dtype, "n_gpus": world, "exchange": a.exchange, "dt": dt, "rk_steps_run": a.steps + a.warmup,
                          "ms_per_rk_step": float(ms), "value": 1e3 / float(ms), "unit": "RK steps/s",
                          "algorithmic_GBps_per_gpu": 70 * C / world / float(ms) / 1e6,
                          "a2a_bytes_per_gpu_per_step": 5 * 5 * C // world * (world - 1) // world,
                          "phases_us_per_substage": phases, "tke": float(tke), "finite": bool(torch.isfinite(obs).all()),
                          "workspace_bytes_per_gpu": sl.workspace_bytes}), flush=True)
    sl.close()
    if world > 1:
        dist.destroy_process_group()

if __name__ == "__main__":
    main()
