"""world_size-2 gloo test (CPU) of the multi-process env sharding logic."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hydrogym_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, global_envs, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(global_envs, rank, world)
    local_obs = torch.arange(lo, hi, dtype=torch.float32)[:, None].repeat(1, 4) + 0.5   # "obs" tagged by env id
    full = sharding.gather_env_outputs(local_obs, global_envs)
    tmax = sharding.max_over_ranks(1.0 + rank)
    if rank == 0:
        out.put((full[:, 0].tolist(), tmax))
    dist.barrier()
    dist.destroy_process_group()


def test_shards_partition_the_batch():
    for n, w in ((256, 8), (1024, 8), (7, 2), (5, 8), (16, 4)):
        ranges = [sharding.shard_range(n, r, w) for r in range(w)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
        sizes = [hi - lo for lo, hi in ranges]
        assert max(sizes) - min(sizes) <= 1


def test_two_rank_gather_and_timing_reduce():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 7, out)) for r in range(2)]
    [p.start() for p in procs]
    ids, tmax = out.get(timeout=120)
    [p.join(timeout=120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert ids == [i + 0.5 for i in range(7)]
    assert tmax == 2.0
