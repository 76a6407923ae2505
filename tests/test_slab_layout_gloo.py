"""World-size-2 gloo test of the slab exchange pattern (hydrogym_b200/slab.py layout algebra): K layout ->
send blocks -> all_to_all_single -> X rows -> send blocks -> all_to_all_single -> K layout reproduces the
global transposes a slab-decomposed 2-D FFT needs.  CPU only."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hydrogym_b200 import slab


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, N, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        MH, Q, X = N // 2, N // 2 // world, N // world
        rng = np.random.default_rng(7)
        A = rng.standard_normal((N, MH)).astype(np.float32)       # global x-space array [x][k]
        mine_k = np.ascontiguousarray(A[:, rank * Q:(rank + 1) * Q].T)  # K layout [Q][N]
        send = torch.from_numpy(np.ascontiguousarray(slab.k_to_send_blocks(mine_k, world)))
        recv = torch.empty_like(send)
        dist.all_to_all_single(recv.view(-1), send.view(-1))
        rows = slab.recv_blocks_to_rows(recv).numpy()
        ok1 = np.array_equal(rows, A[rank * X:(rank + 1) * X, :])
        # and back: X rows -> K layout
        send2 = torch.from_numpy(np.ascontiguousarray(slab.rows_to_send_blocks(rows, world)))
        recv2 = torch.empty_like(send2)
        dist.all_to_all_single(recv2.view(-1), send2.view(-1))
        back = slab.recv_blocks_to_k(recv2.numpy())
        ok2 = np.array_equal(back, mine_k)
        out.put((rank, bool(ok1), bool(ok2)))
    finally:
        dist.destroy_process_group()


def test_slab_exchange_pattern_world2():
    world, N = 2, 64
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, N, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True, True), (1, True, True)]


def test_layout_algebra_single_process():
    """The same identities for P = 4 without a process group (blocks moved by indexing)."""
    P, N = 4, 64
    MH, Q, X = N // 2, N // 2 // P, N // P
    A = np.random.default_rng(3).standard_normal((N, MH))
    sends = [slab.k_to_send_blocks(np.ascontiguousarray(A[:, r * Q:(r + 1) * Q].T), P) for r in range(P)]
    for r in range(P):
        recv = np.stack([sends[s][r] for s in range(P)])
        assert np.array_equal(slab.recv_blocks_to_rows(recv), A[r * X:(r + 1) * X])
    sends2 = [slab.rows_to_send_blocks(A[r * X:(r + 1) * X], P) for r in range(P)]
    for r in range(P):
        recv = np.stack([sends2[s][r] for s in range(P)])
        assert np.array_equal(slab.recv_blocks_to_k(recv), A[:, r * Q:(r + 1) * Q].T)
