"""Channel-flow env step timing: env-steps/s and time per RK4 solver step for a few batch sizes.
Prints one JSON line per configuration.  Not the driver's bench (bench.py carries the `extra.channel` record)."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import channel_oracle as orc   # synthetic initial fields only (tools/ is not the product)
from hydrogym_b200 import ChannelStepper

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="1,4,16,64")
    ap.add_argument("--dtypes", default="fp32,fp64")
    ap.add_argument("--steps", type=int, default=3, help="timed env steps (50 RK4 steps each)")
    ap.add_argument("--nsteps", type=int, default=50)
    a = ap.parse_args()
    fields = orc.synthetic_initial_fields(np.float64)
    for dn in a.dtypes.split(","):
        dt = torch.float32 if dn == "fp32" else torch.float64
        for B in [int(b) for b in a.batches.split(",")]:
            st = ChannelStepper(num_envs=B, dtype=dt, device="cuda:0")
            U, V, W = (torch.from_numpy(f).to("cuda", dt).unsqueeze(0).expand(B, 72, 72, 72).contiguous() for f in fields)
            st.set_state(U, V, W)
            act = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (B, 24))).to("cuda", dt)
            st.step(act, 0.0, nsteps=a.nsteps)   # warm-up
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = st.launch_count
            e0.record()
            for i in range(a.steps):
                obs, rew = st.step(act, 50.0 * (i + 1) % 40, nsteps=a.nsteps)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.steps
            print(json.dumps({"workload": f"channel 72^3 B={B} {dn}", "ms_per_env_step": ms, "env_steps_per_s": B * 1e3 / ms,
                              "us_per_rk4_step": 1e3 * ms / a.nsteps, "us_per_rk4_step_per_env": 1e3 * ms / a.nsteps / B,
                              "launches_per_env_step": (st.launch_count - l0) // a.steps, "reward0": float(rew[0]),
                              "finite": bool(torch.isfinite(obs).all()), "workspace_MB": st.workspace_bytes / 1e6}), flush=True)
            st.close()

if __name__ == "__main__":
    main()
