#!/usr/bin/env python
"""Benchmark of the Kolmogorov-flow env step (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One "step" = one env-step (n_outer * n_inner RK solver steps, 9 990 at the default dt = 1e-3) of a batch of 256
environments at 256^2, Re = 1000 per GPU (weak scaling: independent environments shard by batch, no collectives on
the data path).  The headline runs the REFERENCE'S LITERAL ARITHMETIC (its 2/3 de-aliasing mask is a no-op,
hydrogym/jax/utils/utils.py:72-86): every step starts from the reference's own initial state (kolmogorov.py:103-129)
under fresh random actions -- the first env-step of an episode -- which stays finite.  The line also carries
  masked        : the same workload on synthetic random initial fields with the intended 2/3 mask ON (random fields
                  blow up within one env-step without it, profiles/r1_stability_probe.txt) -- NOT the reference's
                  arithmetic, labelled as such;
  extra.configs : BASELINE.json configs[1..3] at reduced horizons (256^2 fp64, 128^2 x 1024 fp32, 1024^2 x 16 fp32);
  extra.slab    : configs[4], ONE 8192^2 flow slab-decomposed over the N ranks (NCCL all-to-all / peer stores), with a
                  parity check of both exchanges against the batched kernels;
  extra.channel : SURVEY 8(f)-4, the 3-D channel-flow env step (72^3, 50 RK4 + projection steps), fp32 / fp64, 1 and 16
                  envs per GPU, one fp64 env step checked against the reference-source fixture, NumPy oracle timed beside.
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
    ap.add_argument("--dealias", action="store_true",
                    help="headline with the intended 2/3 mask ON and synthetic random initial fields (round-1 headline; "
                         "not the reference's arithmetic).  Default: reference-literal arithmetic, mask off")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of each cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip masked / extra.configs / extra.slab records")
    ap.add_argument("--masked-steps", type=int, default=3)
    ap.add_argument("--e2e-steps", type=int, default=10, help="upper bound on the end-to-end steps (<= --steps)")
    ap.add_argument("--slab-grid", type=int, default=8192)
    ap.add_argument("--slab-steps", type=int, default=20)
    ap.add_argument("--cpu-variant-job", default=None, help=argparse.SUPPRESS)  # internal: see cpu_variant
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


def reference_initial_state(N, dtype=np.float32):
    """kolmogorov.py:103-129: rfftn(-sin X + cos Y) on the endpoint-inclusive mesh."""
    from hydrogym_b200 import constants
    return constants.initial_vorticity_hat((N, N)).astype(np.complex64 if dtype == np.float32 else np.complex128)


def initial_states(N, env_ids, dtype, literal):
    if literal:
        w0 = reference_initial_state(N, dtype)
        return np.broadcast_to(w0, (len(env_ids),) + w0.shape).copy()
    return np.stack([synthetic_omega_hat(N, i, dtype) for i in env_ids])


# ---------------------------------------------------------------------------------------------
# CPU baseline: the oracle (a NumPy restatement of the reference, "port") on the host cores, on a
# bounded sample of the same workload.  Two variants, the faster one is the reported value:
#   numpy-procs : one environment per worker process, np.fft (pocketfft, one thread each)
#   scipy-fft   : ONE process, a batch of environments, scipy.fft.rfftn/irfftn(workers=cores)
#   torch-mkl   : ONE process, a batch of environments as torch CPU tensors (oracle/torch_cpu_port.py:
#                 torch.fft + threaded pointwise kernels, cores intra-op threads)
# ---------------------------------------------------------------------------------------------
def _cpu_worker(job):
    N, Re, dt, npdt, env_ids, n_rk, dealias, fft_workers = job
    if fft_workers < 0:
        return _cpu_worker_torch(job)
    if fft_workers:
        import scipy.fft as sfft
    else:
        os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import kolmogorov_oracle as orc
    if fft_workers:  # same restatement, multithreaded FFT back-end
        orc._rfft2 = lambda a: sfft.rfftn(a, axes=(-2, -1), workers=fft_workers)
        orc._irfft2 = lambda a: sfft.irfftn(a, axes=(-2, -1), workers=fft_workers)
    dtype = np.float32 if npdt == "fp32" else np.float64
    env = orc.KolmogorovFlow(dict(dt=dt), dict(Re=Re, grid_size=(N, N)), dtype=dtype, dealias=dealias)
    w = initial_states(N, env_ids, dtype, literal=not dealias)
    acts = synthetic_actions(1, max(env_ids) + 1, dtype)[0][list(env_ids)]
    ch = env.equation.control_term(w, env._control_field(acts, env.default_params))
    t0 = time.perf_counter()
    w = env.integrator.rk_steps(w, n_rk, ch)
    return time.perf_counter() - t0, bool(np.isfinite(w).all())


def _cpu_worker_torch(job):
    N, Re, dt, npdt, env_ids, n_rk, dealias, threads = job
    from oracle import kolmogorov_oracle as orc
    from oracle.torch_cpu_port import TorchCpuSolver
    dtype = np.float32 if npdt == "fp32" else np.float64
    env = orc.KolmogorovFlow(dict(dt=dt), dict(Re=Re, grid_size=(N, N)), dtype=dtype, dealias=dealias)
    w = initial_states(N, env_ids, dtype, literal=not dealias)
    acts = synthetic_actions(1, max(env_ids) + 1, dtype)[0][list(env_ids)]
    ch = env.equation.control_term(w, env._control_field(acts, env.default_params))
    solver = TorchCpuSolver(env, threads=-threads)
    t0 = time.perf_counter()
    w = solver.rk_steps(w, n_rk, ch)
    return time.perf_counter() - t0, bool(np.isfinite(w).all())


def _run_jobs(jobs, nproc):
    import multiprocessing as mp
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(nproc) as pool:
        res = pool.map(_cpu_worker, jobs)
    return time.perf_counter() - t0, res


def _cpu_variant_here(a):
    """Runs in a FRESH interpreter (see cpu_variant): no CUDA context, no thread pools inherited through fork."""
    N, variant, ncores, seconds, dealias = a["grid"], a["variant"], a["ncores"], a["seconds"], a["dealias"]
    if variant == "numpy-procs":
        res = [_cpu_worker((N, a["re"], a["dt"], a["dtype"], (0,), 2, dealias, 0))]
        per_rk = max(res[0][0] / 2, 1e-4)
        n_rk = int(max(4, min(5000, seconds / per_rk)))
        jobs = [(N, a["re"], a["dt"], a["dtype"], (i,), n_rk, dealias, 0) for i in range(ncores)]
        _, res = _run_jobs(jobs, ncores)
        envs, what = ncores, f"{ncores} envs (one per process, np.fft single-threaded each)"
    else:
        torch_variant = variant == "torch-mkl"
        envs = 4 * ncores if torch_variant else ncores
        ids = tuple(range(envs))
        thr = -ncores if torch_variant else ncores
        res = [_cpu_worker((N, a["re"], a["dt"], a["dtype"], ids, 2, dealias, thr))]
        per_rk = max(res[0][0] / 2, 1e-4)
        n_rk = int(max(4, min(5000, seconds / per_rk)))
        res = [_cpu_worker((N, a["re"], a["dt"], a["dtype"], ids, n_rk, dealias, thr))]
        what = (f"{envs} envs batched in one process, torch CPU tensors, {ncores} intra-op threads" if torch_variant
                else f"{envs} envs batched in one process, scipy.fft workers={ncores}")
    compute = max(r[0] for r in res)  # the timed rk_steps loop (process start-up and state set-up excluded)
    env_rk_per_s = envs * n_rk / compute
    return dict(variant=variant, rk_steps_env_per_s=env_rk_per_s, threads=ncores,
                sample=f"{what} x {n_rk} RK steps at {N}^2 {a['dtype']}, {compute:.1f}s", finite=all(r[1] for r in res))


def cpu_variant(args, rk_per_env_step, seconds, variant, ncores):
    """One CPU variant, timed in a fresh interpreter: the bench process holds a CUDA context and torch / OpenMP thread
    pools, and a fork()ed child that touches a threaded library again can dead-lock."""
    job = dict(grid=args.grid, re=args.re, dt=args.dt, dtype=args.dtype, dealias=bool(args.dealias), variant=variant,
               ncores=ncores, seconds=seconds)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    limit = 60.0 + 6.0 * seconds
    try:
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-variant-job", json.dumps(job)],
                             capture_output=True, text=True, timeout=limit, env=env, cwd=ROOT)
        r = json.loads(out.stdout.strip().splitlines()[-1])
    except subprocess.TimeoutExpired:
        return dict(variant=variant, value=0.0, rk_steps_env_per_s=0.0, threads=ncores, finite=False,
                    sample=f"timed out after {limit:.0f}s")
    except Exception as e:
        return dict(variant=variant, value=0.0, rk_steps_env_per_s=0.0, threads=ncores, finite=False,
                    sample=f"failed: {type(e).__name__}: {e}")
    r["value"] = r["rk_steps_env_per_s"] / rk_per_env_step
    return r


def cpu_baseline(args, rk_per_env_step, seconds, variants=("numpy-procs", "scipy-fft", "torch-mkl")):
    """dict(value=env-steps/s on the host, cores, kind, sample, variants=[...]); value = the fastest variant."""
    ncores = max(1, min(os.cpu_count() or 1, 64))
    runs = [cpu_variant(args, rk_per_env_step, seconds, v, ncores) for v in variants]
    best = max(runs, key=lambda r: r["value"])
    return dict(value=best["value"], unit=UNIT, cores=ncores, kind="port", variant=best["variant"],
                rk_steps_env_per_s=best["rk_steps_env_per_s"],
                sample=best["sample"] + f"; NumPy oracle (port of the reference), dealias={bool(args.dealias)}; "
                                        f"converted with {rk_per_env_step} RK steps per env-step",
                finite=best["finite"], variants=runs)


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
    if isinstance(x, (np.floating, np.integer)):
        return _clean(x.item())
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
    """Identical on both arms (native / --impl reference): the workload, nothing about how it is run."""
    literal = not args.dealias
    return {
        "workload": f"Kolmogorov flow {args.grid}^2 Re={args.re:g}, batch of {args.batch} envs per GPU, "
                    f"reward and probe observations every step (BASELINE.json configs[2])",
        "grid": args.grid, "Re": args.re, "envs_per_gpu": args.batch, "global_envs": args.batch * world,
        "dt": args.dt, "action_time": args.action_time, "n_inner": n_inner, "n_outer": n_outer,
        "rk_steps_per_env_step": n_inner * n_outer, "dealias": bool(args.dealias),
        "arithmetic": "reference-literal (2/3 mask is a no-op, utils.py:72-86)" if literal
                      else "intended 2/3 mask ON (not what the reference executes)",
        "initial_state": "reference initial state rfftn(-sin X + cos Y) (kolmogorov.py:103-129) in every env, re-seeded "
                         "every step; random actions per env and step (seed 4321)" if literal
                         else "synthetic band-limited random fields, rms vorticity 4 (seeds 1234+env), carried from step "
                              "to step; random actions (seed 4321)",
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
    # pick the faster CPU variant once (short calibration), then time it per step on bounded samples
    cal = cpu_baseline(args, rk, 4.0)
    per_step_seconds = max(3.0, min(20.0, 140.0 / max(1, args.steps + args.warmup)))
    vals, cb = [], cal
    for i in range(args.warmup + args.steps):
        cb = cpu_baseline(args, rk, per_step_seconds, variants=(cal["variant"],))
        if i >= args.warmup:
            vals.append(cb["value"])
    value = float(np.mean(vals)) if vals else cb["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / value if value > 0 else None,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32" if args.dtype == "fp32" else "f64",
        "data": "synthetic", "config": base_config(args, n_inner, n_outer, 1),
        "cpu_baseline": dict(cb, value=value, calibration=cal["variants"]),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference arm = NumPy oracle port of the reference on host cores (JAX not installable here), the "
                "fastest of {one env per process, batched scipy.fft workers, batched torch-CPU}; each step is a bounded "
                "sample, "
                "converted to env-steps/s",
    }
    print(json.dumps(_clean(line)), flush=True)


# ---------------------------------------------------------------------------------------------
# native arm
# ---------------------------------------------------------------------------------------------
def _peak():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _device_random_states(torch, B, N, cdtype, dev, seed, amp=0.25):
    """Smooth broadband random omega_hat on the device (extra records: any finite field times the same)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    w = torch.randn(B, N, N // 2 + 1, dtype=cdtype, device=dev, generator=g)
    p = torch.fft.fftfreq(N, 1.0 / N, device=dev)[:, None]
    q = torch.arange(N // 2 + 1, device=dev)[None, :]
    return (w * torch.exp(-(p * p + q * q) / 64.0) * (N * N * amp)).contiguous()


def time_rk_steps(torch, N, B, Re, tdt, dev, n_rk, dealias=False, warm=24):
    """ms per RK step of B envs at N^2 (device-resident, CUDA events on the launch stream), literal arithmetic."""
    from hydrogym_b200 import KolmogorovStepper
    cd = torch.complex64 if tdt == torch.float32 else torch.complex128
    st = KolmogorovStepper(grid_size=(N, N), Re=Re, num_envs=B, dtype=tdt, device=dev, dealias=dealias)
    w = _device_random_states(torch, B, N, cd, dev, seed=N + B)
    st.rk_steps(w, None, warm)
    stream = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    st.rk_steps(w, None, n_rk)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / n_rk
    finite = bool(torch.isfinite(torch.view_as_real(w)).all())
    chunk, lanes = st.chunk, st.lanes
    st.close()
    del w
    return ms, finite, chunk, lanes


def extra_configs(torch, dev, world, max_over_ranks, peak):
    """BASELINE.json configs[1..3] beside the headline: ms per RK step and the 70C whole-step roofline fraction."""
    out = []
    for name, N, B, Re, dt_name, n_rk in (("configs[2] fp64", 256, 256, 1000.0, "fp64", 200),
                                          ("configs[1]", 128, 1024, 100.0, "fp32", 400),
                                          ("configs[3]", 1024, 16, 1000.0, "fp32", 200),
                                          ("configs[0] grid (the reference's default 64^2), batched fp32", 64, 1024, 100.0, "fp32", 1000),
                                          ("configs[0] grid, batched fp64", 64, 1024, 100.0, "fp64", 400),
                                          ("configs[0]: 64^2 Re=100 single env fp64", 64, 1, 100.0, "fp64", 2000)):
        tdt = torch.float32 if dt_name == "fp32" else torch.float64
        C = N * (N // 2 + 1) * 2 * (4 if dt_name == "fp32" else 8)
        rec = {"config": name, "grid": N, "envs_per_gpu": B, "Re": Re, "dtype": dt_name, "dealias": False,
               "rk_steps_timed": n_rk}
        if N == 64:
            # default path: the on-chip-resident kernel (one CTA per environment, csrc/resident.cuh); the two-kernel
            # path (KF_RESIDENT=0) is timed beside it.  The resident kernel moves almost nothing through HBM, so its
            # fraction of the 70C ALGORITHMIC roofline may exceed 1.
            old = os.environ.get("KF_RESIDENT")
            os.environ["KF_RESIDENT"] = "0"
            try:
                ms2 = max_over_ranks(time_rk_steps(torch, N, B, Re, tdt, dev, n_rk)[0])
            finally:
                if old is None:
                    os.environ.pop("KF_RESIDENT", None)
                else:
                    os.environ["KF_RESIDENT"] = old
            rec["two_kernel_path"] = {"ms_per_rk_step": ms2, "frac": 70.0 * C * B / (ms2 * 1e-3) / 1e9 / peak}
            rec["path"] = "resident_kernel (on-chip, one CTA per environment)"
        ms, finite, chunk, lanes = time_rk_steps(torch, N, B, Re, tdt, dev, n_rk)
        ms = max_over_ranks(ms)
        gbs = 70.0 * C * B / (ms * 1e-3) / 1e9
        rec.update({"ms_per_rk_step": ms, "rk_steps_env_per_s": world * B / (ms * 1e-3),
                    "whole_step": {"achieved": gbs, "frac": gbs / peak}, "finite": finite, "chunk": chunk, "lanes": lanes})
        out.append(rec)
    return out


def slab_record(torch, dist, args, dev, rank, world, max_over_ranks, peak):
    """configs[4]: ONE N^2 flow slab-decomposed over the `world` ranks, both exchanges, with a parity check of the
    decomposition (1024^2, 2 RK steps, against the batched kernels on the same state)."""
    from hydrogym_b200 import KolmogorovStepper
    from hydrogym_b200.slab import SlabKolmogorovSolver
    N = args.slab_grid
    rec = {"grid": N, "dtype": "fp32", "n_ranks": world, "Re": 1000.0, "dealias": False}
    # ---- parity of the decomposition through the real collective --------------------------------------------
    Nc, steps = 1024, 2
    wc = _device_random_states(torch, 1, Nc, torch.complex64, dev, seed=5, amp=0.1)
    act = torch.tensor([-0.25, -0.03, 0.02, 0.01], dtype=torch.float32, device=dev)
    st = KolmogorovStepper(grid_size=(Nc, Nc), Re=1000, num_envs=1, dtype=torch.float32, device=dev)
    want = wc.clone()
    st.rk_steps(want, act[None], steps)
    st.close()
    check = {"grid": Nc, "rk_steps": steps, "against": "batched kernels (KolmogorovStepper) on the same state and action"}
    for mode in ("nccl", "p2p"):
        sl = SlabKolmogorovSolver(grid_size=(Nc, Nc), Re=1000, dtype=torch.float32, device=dev, exchange=mode)
        sl.set_state(wc[0], act)
        sl.rk_steps(steps)
        got = sl.get_state()
        sl.close()
        err = float((torch.linalg.norm(got - want[0]) / torch.linalg.norm(want[0])).item())
        check[f"rel_l2_{mode}"] = max_over_ranks(err)
    check["pass"] = bool(check["rel_l2_nccl"] < 2e-5 and check["rel_l2_p2p"] < 2e-5)
    rec["check"] = check
    # ---- timing --------------------------------------------------------------------------------------------
    dt = 0.8 / N  # u ~ 1: CFL = u dt N / (2 pi) ~ 0.13
    C = N * (N // 2 + 1) * 8
    a2a = 5 * 5 * C // world * (world - 1) // world  # bytes leaving each GPU per RK step (4 fields + advection, 5 sub-stages)
    w = _device_random_states(torch, 1, N, torch.complex64, dev, seed=0, amp=0.5)[0]
    rec["exchanges"] = {}
    for mode in (("p2p", "nccl") if world > 1 else ("p2p",)):
        sl = SlabKolmogorovSolver(grid_size=(N, N), Re=1000, dtype=torch.float32, device=dev, dt=dt, exchange=mode)
        sl.set_state(w, None)
        sl.rk_steps(3)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        stream = torch.cuda.current_stream(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = sl.launch_count
        e0.record(stream)
        sl.rk_steps(args.slab_steps)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ms = max_over_ranks(e0.elapsed_time(e1) / args.slab_steps)
        _, tke = sl.reward_obs()
        finite = bool(torch.isfinite(tke).all())
        r = {"ms_per_rk_step": ms, "rk_steps_run": args.slab_steps + 3, "rk_steps_per_s": 1e3 / ms,
             "frac_of_70C_roofline": 70.0 * C / world / (ms * 1e-3) / 1e9 / peak,
             "a2a_bytes_per_gpu_per_step": a2a, "nvlink_gbs_per_gpu": a2a / (ms * 1e-3) / 1e9,
             "gpu_launches_per_rank": int(sl.launch_count - l0), "finite": finite,
             "workspace_bytes_per_gpu": int(sl.workspace_bytes)}
        if mode == "p2p":
            # where a sub-stage goes (per-phase entry points with events between them, no graph): mean and max over ranks
            ph = sl.phase_times(2)
            r["phases_us_per_substage"] = {k: {"rank0": v, "max_over_ranks": max_over_ranks(v)} for k, v in ph.items()}
            r["phases_largest"] = max(ph, key=lambda k: r["phases_us_per_substage"][k]["max_over_ranks"])
        rec["exchanges"][mode] = r
        sl.close()
    best = min(rec["exchanges"], key=lambda m: rec["exchanges"][m]["ms_per_rk_step"])
    rec.update({"exchange": best, "dt": dt, **rec["exchanges"][best]})
    del w
    # ---- the same grid in fp64 (peer-store exchange, fewer steps) ---------------------------------------------
    try:
        w64 = _device_random_states(torch, 1, N, torch.complex128, dev, seed=0, amp=0.5)[0]
        sl = SlabKolmogorovSolver(grid_size=(N, N), Re=1000, dtype=torch.float64, device=dev, dt=dt, exchange="p2p")
        sl.set_state(w64, None)
        del w64
        sl.rk_steps(2)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n64 = max(2, args.slab_steps // 4)
        e0.record(torch.cuda.current_stream(dev))
        sl.rk_steps(n64)
        e1.record(torch.cuda.current_stream(dev))
        torch.cuda.synchronize(dev)
        ms = max_over_ranks(e0.elapsed_time(e1) / n64)
        _, tke = sl.reward_obs()
        rec["fp64"] = {"ms_per_rk_step": ms, "rk_steps_run": n64 + 2, "exchange": "p2p",
                       "frac_of_70C_roofline": 70.0 * 2 * C / world / (ms * 1e-3) / 1e9 / peak,
                       "finite": bool(torch.isfinite(tke).all()), "workspace_bytes_per_gpu": int(sl.workspace_bytes)}
        sl.close()
    except Exception as e:
        rec["fp64"] = {"error": f"{type(e).__name__}: {e}"}
    return rec


def channel_record(torch, args, dev, rank, world, max_over_ranks, peak):
    """the 3-D channel-flow env step (72^3, 50 RK4 + projection steps per env step), env batch sharded
    over the ranks like the headline.  One fp64 env step is checked against the reference-source fixture first."""
    import numpy as np

    from hydrogym_b200 import ChannelStepper, channel_constants as cc
    fields = cc.synthetic_initial_fields(np.float64)
    rec = {"grid": [72, 72, 72], "rk4_steps_per_env_step": 50, "dt": 2e-4,
           "initial_state": "synthetic (channel_constants.synthetic_initial_fields): the HF-hosted fields are not "
                            "reachable offline"}
    gold_path = os.path.join(ROOT, "tests", "golden", "reference_shim_channel.npz")
    if os.path.exists(gold_path):
        gold = np.load(gold_path)
        st = ChannelStepper(num_envs=1, dtype=torch.float64, device=dev)
        st.set_state(*[torch.from_numpy(f) for f in fields])
        obs, rew = st.step(torch.from_numpy(gold["action"]), 0.0, nsteps=50, dt=2e-4)
        err_o = float(np.linalg.norm(obs[0].cpu().numpy() - gold["step1_obs"]) / np.linalg.norm(gold["step1_obs"]))
        err_r = abs(float(rew[0]) - float(gold["step1_reward"])) / abs(float(gold["step1_reward"]))
        st.close()
        rec["check"] = {"against": "reference source over NumPy stand-ins, fp64, one env step (fixture)",
                        "rel_err_obs": max_over_ranks(err_o), "rel_err_reward": max_over_ranks(err_r)}
        rec["check"]["pass"] = bool(rec["check"]["rel_err_obs"] < 1e-9 and rec["check"]["rel_err_reward"] < 1e-9)
    # algorithmic traffic of the pass structure: complex 72^3 arrays moved per solver step and env
    arrays = 198.0
    rec["configs"] = []
    for dn, B, steps in (("fp32", 1, 3), ("fp32", 16, 2), ("fp64", 1, 2), ("fp64", 16, 1)):
        dt_ = torch.float32 if dn == "fp32" else torch.float64
        st = ChannelStepper(num_envs=B, dtype=dt_, device=dev)
        U, V, W = (torch.from_numpy(f).to(dev, dt_).unsqueeze(0).expand(B, 72, 72, 72).contiguous() for f in fields)
        st.set_state(U, V, W)
        act = torch.from_numpy(np.random.default_rng(rank).uniform(-1, 1, (B, 24))).to(dev, dt_)
        st.step(act, 0.0)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = st.launch_count
        e0.record(torch.cuda.current_stream(dev))
        for i in range(steps):
            obs, rew = st.step(act, 0.0)   # every step at the start of an episode's clock (jet gain ramp 0 -> 0.3)
        e1.record(torch.cuda.current_stream(dev))
        torch.cuda.synchronize(dev)
        ms = max_over_ranks(e0.elapsed_time(e1) / steps)
        nbytes = arrays * 72 ** 3 * (8 if dn == "fp32" else 16) * 50 * B
        rec["configs"].append({"dtype": dn, "envs_per_gpu": B, "ms_per_env_step": ms,
                               "env_steps_per_s": world * B * 1e3 / ms, "us_per_rk4_step_per_env": 1e3 * ms / 50 / B,
                               "gpu_launches_per_env_step": int((st.launch_count - l0) // steps),
                               "frac_of_hbm_roofline": nbytes / (ms * 1e-3) / 1e9 / peak,
                               "finite": bool(torch.isfinite(obs).all() and torch.isfinite(rew).all())})
        st.close()
    if world == 1 and not args.no_cpu_baseline:
        # CPU leg: the NumPy oracle's RungeKutta4.rk4_step on this box's host cores -- as written (np.fft, one thread)
        # and with its two transform methods on scipy.fft with all cores (the rest of the arithmetic unchanged)
        import time

        from oracle import channel_oracle as orc
        ncores = os.cpu_count() or 1

        def time_oracle(env, n=3):
            s = env.equation.to_spectral(orc.VelocityState(*fields))
            a = np.zeros(24)
            s = env.integrator.rk4_step(s, dt=orc.DT, action=a, t=0, fx=2.0)
            t0 = time.perf_counter()
            for i in range(n):
                s = env.integrator.rk4_step(s, dt=orc.DT, action=a, t=1 + i, fx=2.0)
            return (time.perf_counter() - t0) / n

        variants = {"numpy (1 thread)": (time_oracle(orc.ChannelFlowSpectralEnv(initial_fields=fields)), 1)}
        try:
            import scipy.fft as sfft
            env = orc.ChannelFlowSpectralEnv(initial_fields=fields)
            env.equation.fft_xy = lambda f: sfft.fftn(f, axes=(0, 1), workers=ncores)
            env.equation.ifft_xy = lambda f: sfft.ifftn(f, axes=(0, 1), workers=ncores).real
            variants[f"scipy.fft workers={ncores}"] = (time_oracle(env), ncores)
        except Exception as e:  # scipy missing: keep the NumPy number
            variants["scipy.fft"] = (float("inf"), 0)
            rec["cpu_baseline_note"] = f"{type(e).__name__}: {e}"
        best = min(variants, key=lambda k: variants[k][0])
        sec, cores = variants[best]
        rec["cpu_baseline"] = {"kind": "port", "variant": best, "cores": cores,
                               "sample": "3 RK4 solver steps of one env (oracle/channel_oracle.py, fp64)",
                               "s_per_rk4_step": sec, "value": 1.0 / (50 * sec), "unit": "env-steps/s",
                               "variants_s_per_rk4_step": {k: v[0] for k, v in variants.items()}}
    return rec


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
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n_inner, n_outer = rk_counts(args)
    rk = n_inner * n_outer
    N, B = args.grid, args.batch
    tdt = torch.float32 if args.dtype == "fp32" else torch.float64
    npdt = np.float32 if args.dtype == "fp32" else np.float64
    wbytes = 4 if args.dtype == "fp32" else 8
    C = N * (N // 2 + 1) * 2 * wbytes            # one spectral field
    peak, peak_src = _peak()
    total = args.warmup + args.steps
    acts_np = synthetic_actions(total + 1, world * B, npdt)[:, rank * B:(rank + 1) * B]
    acts_host = torch.from_numpy(np.ascontiguousarray(acts_np)).pin_memory()
    acts = acts_host.to(dev)

    def timed_steps(st, w, w_seed, first, count, want_trajectory):
        """`count` env-steps on the launch stream between two CUDA events; returns (seconds, launches, non-finite envs
        summed over the steps, mean TKE of the last step).  w_seed != None: every step restarts from it."""
        stream = torch.cuda.current_stream(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nonf = torch.zeros((), dtype=torch.int64, device=dev)
        l0 = st.launch_count
        last = None
        e0.record(stream)
        for i in range(first, first + count):
            if w_seed is not None:
                w.copy_(w_seed)
            last = st.step(w, acts[i], want_trajectory=want_trajectory, in_place=True)
            nonf += last[4].sum()
        e1.record(stream)
        barrier()
        t = max_over_ranks(e0.elapsed_time(e1) * 1e-3)
        return t, st.launch_count - l0, int(nonf.item()), float(last[3].double().mean().item())

    # ---- headline: device-resident timing `value` ------------------------------------------------------------
    literal = not args.dealias
    st = KolmogorovStepper(grid_size=(N, N), Re=args.re, dt=args.dt, n_inner=n_inner, n_outer=n_outer, num_envs=B,
                           dtype=tdt, device=dev, dealias=args.dealias, chunk=args.chunk)
    w_host = torch.from_numpy(initial_states(N, range(rank * B, (rank + 1) * B), npdt, literal)).pin_memory()
    w = w_host.to(dev)
    w_seed = w.clone() if literal else None
    log(f"rank {rank}: headline warm-up ({args.warmup} env-steps)")
    timed_steps(st, w, w_seed, 0, args.warmup, True)
    barrier()
    log(f"headline timed region ({args.steps} env-steps)")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t_dev, launches, nonfinite, mean_tke = timed_steps(st, w, w_seed, args.warmup, args.steps, True)
    clocks = sampler.stop() if rank == 0 else None
    value = world * B * args.steps / t_dev
    step_gbs = 70.0 * C * B * rk * args.steps / t_dev / 1e9  # per GPU (t_dev is the max over ranks)

    log(f"value = {value:.3f} env-steps/s, whole-step {step_gbs / peak:.3f} of peak, nonfinite {nonfinite}; e2e next")
    # ---- end-to-end through the C ABI with host buffers: `e2e` ------------------------------------------------
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    obs_h = torch.empty((B, st.obs_size ** 2), dtype=tdt).pin_memory()
    tke_h = torch.empty((B,), dtype=tdt).pin_memory()
    nonf_h = torch.empty((B,), dtype=torch.int32).pin_memory()
    w_e2e = (w_host.clone() if literal else w.cpu()).pin_memory()
    st.step_host(w_e2e, acts_host[total], obs_h, tke_h, nonf_h)  # warm-up (allocates staging)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        if literal:
            w_e2e.copy_(w_host)  # host-side re-seed, inside the timed region
        st.step_host(w_e2e, acts_host[args.warmup + i], obs_h, tke_h, nonf_h)
    barrier()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    e2e_value = world * B * e2e_steps / t_e2e
    e2e_nonfinite = int(nonf_h.sum().item())
    h2d = w_e2e.numel() * w_e2e.element_size() + B * 4 * wbytes
    d2h = w_e2e.numel() * w_e2e.element_size() + obs_h.numel() * wbytes + B * wbytes + B * 4

    log(f"e2e = {e2e_value:.3f} env-steps/s; kernel profile next")
    # ---- roofline of the dominant kernel (column kernel), CUDA events inside the real launch sequence ---------
    ms_col, ms_row = st.profile_kernels(reps=40)
    chunk, lanes = st.chunk, st.lanes
    col_gbs = 9 * C * chunk / (ms_col * 1e-3) / 1e9
    row_gbs = 5 * C * chunk / (ms_row * 1e-3) / 1e9
    traffic = None
    prof_json = os.path.join(ROOT, "profiles", "ncu_summary.json")
    if os.path.exists(prof_json):
        try:
            pj = json.load(open(prof_json))
            traffic = pj.get("col_kernel_dram_bytes_per_launch", {}).get(f"{N}_{args.dtype}_B{B}_chunk{chunk}")
        except Exception:
            traffic = None
    st.close()
    del w, w_seed

    # ---- the same workload with the intended 2/3 mask ON, synthetic random fields (NOT the reference's arithmetic) --
    masked = None
    if literal and not args.no_extras and args.masked_steps > 0:
        log("masked record")
        stm = KolmogorovStepper(grid_size=(N, N), Re=args.re, dt=args.dt, n_inner=n_inner, n_outer=n_outer, num_envs=B,
                                dtype=tdt, device=dev, dealias=True, chunk=args.chunk)
        wm = torch.from_numpy(initial_states(N, range(rank * B, (rank + 1) * B), npdt, False)).to(dev)
        timed_steps(stm, wm, None, 0, 1, False)
        k = min(args.masked_steps, total - 1)
        tm, _, nfm, tkem = timed_steps(stm, wm, None, 1, k, False)
        mcol, mrow = stm.profile_kernels(reps=40)
        # Bytes that do NOT move although the 70 C figure counts them: the RK register in sub-stages 0 (not read) and
        # 4 (not written), and with the mask on the advection columns ky >= km -- never written by the row kernel
        # (MH - km of MH columns) nor read by the column kernel (whole tiles of 8 columns at or above km).
        MH, TCOLS = N // 2, 8
        km = ((MH + 1) * 2) // 3
        row_skip = (MH - km) / MH
        col_skip = (MH - -(-km // TCOLS) * TCOLS) / MH
        moved_step = 70.0 - 2.0 - 5.0 * (row_skip + col_skip)
        moved_col = 9.0 - col_skip
        gb = lambda coef, secs: coef * C * B * rk * k / secs / 1e9
        masked = {"dealias": True, "arithmetic": "intended 2/3 mask ON -- not what the reference executes",
                  "initial_state": "synthetic band-limited random fields (seeds 1234+env), carried from step to step",
                  "steps": k, "value": world * B * k / tm, "unit": UNIT, "ms_per_step": tm / k * 1e3,
                  "whole_step": {"frac": gb(70.0, tm) / peak, "frac_moved": gb(moved_step, tm) / peak,
                                 "bytes_per_rk_step_over_C": {"algorithmic": 70.0, "moved": moved_step}},
                  "col_kernel": {"ms_per_launch": mcol, "frac": 9 * C * stm.chunk / (mcol * 1e-3) / 1e9 / peak,
                                 "frac_moved": moved_col * C * stm.chunk / (mcol * 1e-3) / 1e9 / peak},
                  "row_kernel": {"ms_per_launch": mrow},
                  "check": {"nonfinite_envs": nfm, "mean_tke": tkem}}
        stm.close()
        del wm

    # ---- extra records: the other BASELINE.json configurations and the slab-decomposed single grid -------------
    extra = {}
    if not args.no_extras:
        log("extra.configs")
        try:
            extra["configs"] = extra_configs(torch, dev, world, max_over_ranks, peak)
        except Exception as e:  # never lose the headline line to an extra record
            extra["configs_error"] = f"{type(e).__name__}: {e}"
        log("extra.slab")
        try:
            extra["slab"] = slab_record(torch, dist, args, dev, rank, world, max_over_ranks, peak)
        except Exception as e:
            extra["slab_error"] = f"{type(e).__name__}: {e}"
        log("extra.channel")
        try:
            extra["channel"] = channel_record(torch, args, dev, rank, world, max_over_ranks, peak)
        except Exception as e:
            extra["channel_error"] = f"{type(e).__name__}: {e}"

    if rank == 0:
        ws = 7 * C * chunk * lanes
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if args.dtype == "fp32" else "f64",
            "data": "synthetic",
            "config": base_config(args, n_inner, n_outer, world),
            "run": {"chunk": chunk, "lanes": lanes, "trajectory": "returned every step (B x n_outer x N x Mh) in the "
                    "device-timed region; kf_step_host (e2e) returns state, obs, mean TKE, flags but no trajectory",
                    "l2_policy": f"inputs larger than L2: per-GPU working set {ws / 1e6:.0f} MB ({lanes} lane(s) x "
                                 f"{chunk} envs) vs 126 MB L2" if ws > 126e6 else
                                 f"working set {ws / 1e6:.0f} MB is L2-resident BY DESIGN (time blocking)"},
            "rk_steps_env_per_s": value * rk,
            "roofline": {"bound": "hbm", "kernel": "col_tma_kernel (sub-stage epilogue+prologue, 9C per env), timed with "
                                                   "CUDA events inside the real row/column launch sequence of one lane",
                         "achieved": col_gbs, "peak": peak, "unit": "GB/s", "frac": col_gbs / peak,
                         "traffic": traffic, "peak_source": peak_src, "ms_per_launch": ms_col,
                         "row_kernel": {"achieved": row_gbs, "frac": row_gbs / peak, "ms_per_launch": ms_row},
                         "whole_step": {"achieved": step_gbs, "frac": step_gbs / peak,
                                        "frac_moved": step_gbs / peak * 68.0 / 70.0,
                                        "definition": "70*C*B*n_rk / t (BASELINE.md section 2), per GPU; frac_moved "
                                                      "counts the 68*C the kernels move with the mask off (the RK "
                                                      "register is not read in sub-stage 0, not written in 4)"}},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": t_e2e / e2e_steps * 1e3,
                    "api": "kf_step_host (C ABI, pinned host buffers)", "nonfinite_envs": e2e_nonfinite},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "check": {"nonfinite_envs": nonfinite, "mean_tke": mean_tke, "valid": nonfinite == 0},
        }
        if masked is not None:
            line["masked"] = masked
        if extra:
            line["extra"] = extra
        if world == 1 and not args.no_cpu_baseline:
            log("cpu_baseline")
            line["cpu_baseline"] = cpu_baseline(args, rk, args.cpu_seconds)
        print(json.dumps(_clean(line)), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def log(msg):
    """Progress on stderr (stdout carries the ONE JSON line)."""
    print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


def main():
    args = parse_args()
    if args.cpu_variant_job:
        print(json.dumps(_clean(_cpu_variant_here(json.loads(args.cpu_variant_job)))), flush=True)
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
