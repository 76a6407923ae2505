#!/usr/bin/env python
"""Benchmark of the Kolmogorov-flow env step (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One "step" = one env-step (n_outer * n_inner RK solver steps, 9 990 at the default
dt = 1e-3) of a batch of 256 environments at 256^2, Re = 1000 per GPU (weak scaling:
independent environments shard by batch, no collectives on the data path).
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for the definitions.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "env-steps/s (batched, 256^2 Re=1000)"
UNIT = "env-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--grid", type=int, default=256)
    ap.add_argument("--batch", type=int, default=256, help="environments per GPU")
    ap.add_argument("--re", type=float, default=1000.0)
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--dt", type=float, default=1e-3)
    ap.add_argument("--action-time", type=float, default=10.0)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--no-dealias", dest="dealias", action="store_false",
                    help="reference-literal arithmetic (the reference's 2/3 mask is a no-op); at 256^2 Re=1000 that "
                         "diverges within one env-step in fp32 AND fp64 (profiles/r1_stability_probe.txt)")
    ap.set_defaults(dealias=True)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# synthetic inputs: band-limited Gaussian random vorticity, rms(omega) = 4
# ---------------------------------------------------------------------------------------------
def synthetic_omega_hat(N, env_index, dtype=np.float32):
    rng = np.random.default_rng(1234 + env_index)
    Mh = N // 2 + 1
    p = np.abs(np.fft.fftfreq(N, 1.0 / N))[:, None]
    q = np.arange(Mh)[None, :]
    k = np.sqrt(p * p + q * q)
    amp = np.exp(-(k / 8.0) ** 2) * ((k >= 1) & (k <= N / 3))
    w = (rng.normal(size=(N, Mh)) + 1j * rng.normal(size=(N, Mh))) * amp
    omega = np.fft.irfft2(w, s=(N, N))
    omega *= 4.0 / np.sqrt(np.mean(omega ** 2))
    what = np.fft.rfft2(omega)
    what[0, 0] = 0.0
    return what.astype(np.complex64 if dtype == np.float32 else np.complex128)


def synthetic_actions(steps, B, dtype=np.float32):
    return np.random.default_rng(4321).uniform(-0.5, 0.5, (steps, B, 4)).astype(dtype)


# ---------------------------------------------------------------------------------------------
# CPU baseline: the oracle (a NumPy restatement of the reference, "port") on the host cores,
# environments spread over worker processes (they are independent), on a bounded sample.
# ---------------------------------------------------------------------------------------------
def _cpu_worker(job):
    N, Re, dt, npdt, env_ids, n_rk = job
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import kolmogorov_oracle as orc
    dtype = np.float32 if npdt == "fp32" else np.float64
    env = orc.KolmogorovFlow(dict(dt=dt), dict(Re=Re, grid_size=(N, N)), dtype=dtype)
    w = np.stack([synthetic_omega_hat(N, i, dtype) for i in env_ids])
    acts = synthetic_actions(1, max(env_ids) + 1, dtype)[0][list(env_ids)]
    ch = env.equation.control_term(w, env._control_field(acts, env.default_params))
    t0 = time.perf_counter()
    w = env.integrator.rk_steps(w, n_rk, ch)
    return time.perf_counter() - t0, bool(np.isfinite(w).all())


def cpu_baseline(args, rk_per_env_step, seconds, nproc=None):
    """Returns dict(value=env-steps/s on the host, cores, kind, sample)."""
    import multiprocessing as mp
    nproc = nproc or max(1, min(os.cpu_count() or 1, 64))
    N = args.grid
    # calibrate: one env, 2 RK steps, in this process
    t_cal, _ = _cpu_worker((N, args.re, args.dt, args.dtype, (0,), 2))
    per_rk = max(t_cal / 2, 1e-4)
    n_rk = int(max(4, min(5000, seconds / per_rk)))
    jobs = [(N, args.re, args.dt, args.dtype, (i,), n_rk) for i in range(nproc)]
    t0 = time.perf_counter()
    if nproc > 1:
        with mp.get_context("fork").Pool(nproc) as pool:
            res = pool.map(_cpu_worker, jobs)
    else:
        res = [_cpu_worker(jobs[0])]
    wall = time.perf_counter() - t0
    env_rk_per_s = nproc * n_rk / wall
    return dict(value=env_rk_per_s / rk_per_env_step, unit=UNIT, cores=nproc, kind="port",
                rk_steps_env_per_s=env_rk_per_s,
                sample=f"{nproc} envs (one per process) x {n_rk} RK steps at {N}^2 {args.dtype}, NumPy oracle, "
                       f"wall {wall:.1f}s; converted with {rk_per_env_step} RK steps per env-step",
                finite=all(r[1] for r in res))


# ---------------------------------------------------------------------------------------------
# clocks during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])), mx.append(float(f[2])), power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        return dict(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                    power_w_max=float(max(power)), samples=len(sm))


def _clean(x):
    """JSON has no NaN/Inf: map non-finite floats to None."""
    if isinstance(x, float):
        return x if np.isfinite(x) else None
    if isinstance(x, dict):
        return {k: _clean(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_clean(v) for v in x]
    return x


def dist_env():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    return rank, world, local


def rk_counts(args):
    from hydrogym_b200 import constants
    n_inner, n_outer = constants.step_counts(args.dt, args.action_time, 1)
    return n_inner, n_outer


def base_config(args, n_inner, n_outer, world):
    return {
        "workload": f"Kolmogorov flow {args.grid}^2 Re={args.re:g}, batch of {args.batch} envs per GPU, "
                    f"reward and probe observations every step (BASELINE.json configs[2])",
        "grid": args.grid, "Re": args.re, "envs_per_gpu": args.batch, "global_envs": args.batch * world,
        "dt": args.dt, "action_time": args.action_time, "n_inner": n_inner, "n_outer": n_outer,
        "rk_steps_per_env_step": n_inner * n_outer, "dealias": bool(args.dealias),
        "parallelism": f"env-batch sharding x{world}, no collectives",
    }


# ---------------------------------------------------------------------------------------------
# reference arm: the reference's CPU implementation of the path.  JAX is not installable in
# this image, so the arm times the oracle port on all host cores.
# ---------------------------------------------------------------------------------------------
def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    n_inner, n_outer = rk_counts(args)
    rk = n_inner * n_outer
    per_step_seconds = max(3.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    vals = []
    cb = None
    for i in range(args.warmup + args.steps):
        cb = cpu_baseline(args, rk, per_step_seconds)
        if i >= args.warmup:
            vals.append(cb["value"])
    value = float(np.mean(vals)) if vals else cb["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / value if value > 0 else None,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32" if args.dtype == "fp32" else "f64",
        "data": "synthetic", "config": base_config(args, n_inner, n_outer, 1),
        "cpu_baseline": dict(cb, value=value),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference arm = NumPy oracle port of the reference on host cores (JAX not installable here); "
                "each step is a bounded sample, converted to env-steps/s",
    }
    print(json.dumps(_clean(line)), flush=True)


# ---------------------------------------------------------------------------------------------
# native arm
# ---------------------------------------------------------------------------------------------
def run_native(args):
    import torch
    import torch.distributed as dist

    from hydrogym_b200 import KolmogorovStepper

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; hydrogym_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n_inner, n_outer = rk_counts(args)
    rk = n_inner * n_outer
    N, B = args.grid, args.batch
    tdt = torch.float32 if args.dtype == "fp32" else torch.float64
    npdt = np.float32 if args.dtype == "fp32" else np.float64
    wbytes = 4 if args.dtype == "fp32" else 8
    st = KolmogorovStepper(grid_size=(N, N), Re=args.re, dt=args.dt, n_inner=n_inner, n_outer=n_outer, num_envs=B,
                           dtype=tdt, device=dev, dealias=args.dealias, chunk=args.chunk)

    # inputs: each rank owns envs [rank*B, (rank+1)*B) of the global batch
    w_host = torch.from_numpy(np.stack([synthetic_omega_hat(N, rank * B + i, npdt) for i in range(B)])).pin_memory()
    total = args.warmup + args.steps
    acts_np = np.stack([synthetic_actions(total + 1, world * B, npdt)[:, rank * B:(rank + 1) * B]])[0]
    acts_host = torch.from_numpy(np.ascontiguousarray(acts_np)).pin_memory()
    w = w_host.to(dev)
    acts = acts_host.to(dev)

    # ---- device-resident timing: `value` ----
    for i in range(args.warmup):
        st.step(w, acts[i], want_trajectory=False, in_place=True)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = st.launch_count
    stream = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    last = None
    for i in range(args.warmup, total):
        last = st.step(w, acts[i], want_trajectory=False, in_place=True)
    e1.record(stream)
    barrier()
    t_dev = max_over_ranks(e0.elapsed_time(e1) * 1e-3)
    launches = st.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    nonfinite = int(last[4].sum().item()) if last is not None else 0
    mean_tke = float(last[3].double().mean().item()) if last is not None else float("nan")
    value = world * B * args.steps / t_dev

    # ---- end-to-end through the C ABI with host buffers: `e2e` ----
    obs_h = torch.empty((B, st.obs_size ** 2), dtype=tdt).pin_memory()
    tke_h = torch.empty((B,), dtype=tdt).pin_memory()
    nonf_h = torch.empty((B,), dtype=torch.int32).pin_memory()
    w_e2e = w.cpu().pin_memory()
    st.step_host(w_e2e, acts_host[total], obs_h, tke_h, nonf_h)  # warm-up (allocates staging)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        st.step_host(w_e2e, acts_host[args.warmup + i], obs_h, tke_h, nonf_h)
    barrier()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    e2e_value = world * B * args.steps / t_e2e
    h2d = w_e2e.numel() * w_e2e.element_size() + B * 4 * wbytes
    d2h = w_e2e.numel() * w_e2e.element_size() + obs_h.numel() * wbytes + B * wbytes + B * 4

    # ---- roofline of the dominant kernel (column kernel), CUDA events on the launch stream ----
    ms_col, ms_row = st.profile_kernels(reps=20)
    C = N * (N // 2 + 1) * 2 * wbytes            # one spectral field
    chunk = st.chunk
    lanes = st.lanes
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    col_gbs = 9 * C * chunk / (ms_col * 1e-3) / 1e9
    row_gbs = 5 * C * chunk / (ms_row * 1e-3) / 1e9
    step_gbs = 70.0 * C * B * rk * args.steps / t_dev / 1e9  # per GPU (t_dev is the max over ranks)
    traffic = None
    prof_json = os.path.join(ROOT, "profiles", "ncu_summary.json")
    if os.path.exists(prof_json):
        try:
            pj = json.load(open(prof_json))
            key = f"{N}_{args.dtype}_B{B}_chunk{chunk}"
            traffic = pj.get("col_kernel_dram_bytes_per_launch", {}).get(key)
        except Exception:
            traffic = None

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if args.dtype == "fp32" else "f64",
            "data": "synthetic",
            "config": dict(base_config(args, n_inner, n_outer, world), chunk=chunk, lanes=lanes,
                           l2_policy="inputs larger than L2: per-GPU working set "
                                     f"{7 * C * chunk * lanes / 1e6:.0f} MB ({lanes} lane(s) x {chunk} envs) vs 126 MB L2"
                                     if 7 * C * chunk * lanes > 126e6 else
                                     f"working set {7 * C * chunk * lanes / 1e6:.0f} MB is L2-resident BY DESIGN (time blocking)"),
            "rk_steps_env_per_s": value * rk,
            "roofline": {"bound": "hbm", "kernel": "col_tma_kernel (sub-stage epilogue+prologue, 9C per env), timed ALONE "
                                                   "on one lane; in the step the two lanes overlap it with the other "
                                                   "lane's row kernel, which is why whole_step.frac is higher",
                         "achieved": col_gbs, "peak": peak, "unit": "GB/s", "frac": col_gbs / peak,
                         "traffic": traffic, "peak_source": peak_src, "ms_per_launch": ms_col,
                         "row_kernel": {"achieved": row_gbs, "frac": row_gbs / peak, "ms_per_launch": ms_row},
                         "whole_step": {"achieved": step_gbs, "frac": step_gbs / peak,
                                        "definition": "70*C*B*n_rk / t (BASELINE.md section 2), per GPU"}},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": t_e2e / args.steps * 1e3, "api": "kf_step_host (C ABI, pinned host buffers)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "check": {"nonfinite_envs": nonfinite, "mean_tke": mean_tke},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, rk, args.cpu_seconds)
        print(json.dumps(_clean(line)), flush=True)
    st.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
