"""Slab-decomposed single-grid Kolmogorov solver: one N x N flow spread over P GPUs (BASELINE.json configs[4]).

The reference has no multi-device path; the arithmetic per RK sub-stage is the reference's
(``RK4_CN.time_step_fn`` hydrogym/jax/solvers/base.py:50-57 over ``nonlinear_terms`` /
``implicit_timestep`` hydrogym/jax/envs/kolmogorov.py:209-257), cut where a 2-D FFT changes direction:

* K layout -- rank r owns Q = M/2/P packed ky columns (all kx): x-direction FFTs, RK update (``kf_slab_col``);
* X layout -- rank r owns X = N/P x rows (all ky): y-direction FFTs and the advection product (``kf_slab_row``);

and the two all-to-alls per sub-stage that move the 4 x-space fields K -> X and the advection rows X -> K
run through ``torch.distributed.all_to_all_single`` (NCCL over NVLink).  All arithmetic is in
``libkolmo_b200.so``; torch is device memory, streams and the collective.

Two drivers share the code: one process per GPU (``group`` given or the default process group), and
``local_ranks=P`` which plays P ranks on ONE device with block copies instead of NCCL -- the layouts and
kernels are identical, so the decomposition is testable on a single GPU.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, constants


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class _SlabRank:
    """One rank's ``kf_slab`` handle and its four exchange buffers."""

    def __init__(self, lib, cfg, rank, nranks, dtype, device, forcing_row, control_rows, p2p=False):
        self.lib, self.rank, self.nranks = lib, rank, nranks
        self.h = ctypes.c_void_p(0)
        _lib.check(lib.kf_slab_create(ctypes.byref(self.h), ctypes.byref(cfg), rank, nranks), "kf_slab_create")
        _lib.check(lib.kf_slab_set_constants(self.h, forcing_row.ctypes.data_as(ctypes.c_void_p),
                                             control_rows.ctypes.data_as(ctypes.c_void_p)), "kf_slab_set_constants")
        self.acc = torch.zeros(4, dtype=torch.float64, device=device)
        self.ipc = None
        if p2p:  # receive buffers live in the library (cudaMalloc: exportable through CUDA IPC); no send buffers
            self.ipc = (ctypes.c_ubyte * 128)()
            _lib.check(lib.kf_slab_p2p_init(self.h, ctypes.byref(self.ipc)), "kf_slab_p2p_init")
            self.fld_send = self.fld_recv = self.adv_send = self.adv_recv = None
            return
        real = 4 if dtype == torch.float32 else 8
        nf = lib.kf_slab_exchange_bytes(self.h, 0) // real
        na = lib.kf_slab_exchange_bytes(self.h, 1) // real
        self.fld_send = torch.zeros(nf, dtype=dtype, device=device)
        self.adv_send = torch.zeros(na, dtype=dtype, device=device)
        if nranks == 1:  # the all-to-all is the identity
            self.fld_recv, self.adv_recv = self.fld_send, self.adv_send
        else:
            self.fld_recv = torch.zeros(nf, dtype=dtype, device=device)
            self.adv_recv = torch.zeros(na, dtype=dtype, device=device)

    def close(self):
        if self.h.value:
            self.lib.kf_slab_destroy(self.h)
            self.h = ctypes.c_void_p(0)


class SlabKolmogorovSolver:
    """RK4-CN pseudo-spectral solve of ONE Kolmogorov flow, slab-decomposed over ``world`` ranks."""

    def __init__(self, grid_size=(8192, 8192), Re=1000, k=4, control_ks=(4, 5, 6, 7), dt=1e-3,
                 dtype=torch.float32, device=None, dealias=False, obs_size=8, min_action=-0.5, max_action=0.5,
                 domain_y=(0, 2 * np.pi), group=None, local_ranks=None, zero_forcing=False, exchange="nccl",
                 use_graph=True, device_barrier=True):
        if exchange not in ("nccl", "p2p"):
            raise ValueError("exchange must be 'nccl' (all_to_all_single) or 'p2p' (stores into peer-mapped buffers)")
        if dtype not in (torch.float32, torch.float64):
            raise ValueError("dtype must be torch.float32 or torch.float64")
        N, M = grid_size
        if N != M:
            raise ValueError("only square grids are supported")
        constants.check_domain((0, 2 * np.pi), domain_y)
        if not torch.cuda.is_available():
            raise RuntimeError("hydrogym_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        dev = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if dev.type != "cuda":
            raise RuntimeError("hydrogym_b200 runs on CUDA devices only")
        self.device = torch.device("cuda", dev.index if dev.index is not None else torch.cuda.current_device())
        self.N, self.M, self.Mh = N, M, M // 2 + 1
        self.dtype = dtype
        self.cdtype = torch.complex64 if dtype == torch.float32 else torch.complex128
        self.dt, self.nu, self.obs_size = float(dt), 1.0 / Re, int(obs_size)
        self.group = group
        if local_ranks is not None:
            self.local, self.world, my_ranks = True, int(local_ranks), list(range(int(local_ranks)))
        else:
            self.local = False
            if dist.is_initialized():
                self.world, my_ranks = dist.get_world_size(group), [dist.get_rank(group)]
            else:
                self.world, my_ranks = 1, [0]
        self.p2p = exchange == "p2p"
        self.use_graph, self.device_barrier = bool(use_graph), bool(device_barrier)
        self._flag = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._lib = _lib.load()
        cfg = _lib.KfConfig(N=N, M=M, B=1, dtype=_lib.KF_F32 if dtype == torch.float32 else _lib.KF_F64,
                            nu=self.nu, dt=self.dt, n_inner=1, n_outer=1, dealias=int(bool(dealias)), drag=0.0,
                            device=self.device.index, obs_size=self.obs_size, min_action=float(min_action),
                            max_action=float(max_action), chunk=0, use_graph=0)
        np_c = np.complex64 if dtype == torch.float32 else np.complex128
        forcing_row = np.ascontiguousarray(constants.forcing_row(k, grid_size, domain_y).astype(np_c))
        control_rows = np.ascontiguousarray(constants.control_rows(control_ks, grid_size, domain_y).astype(np_c))
        if zero_forcing:  # decaying flow (tests: grid-independent dynamics)
            forcing_row[:] = 0
            control_rows[:] = 0
        self._ranks = []
        with torch.cuda.device(self.device):
            for r in my_ranks:
                self._ranks.append(_SlabRank(self._lib, cfg, r, self.world, dtype, self.device, forcing_row,
                                             control_rows, p2p=self.p2p))
            if self.p2p:
                self._attach_peers()
        self.rk_steps_done = 0

    # -- plumbing --------------------------------------------------------------
    def close(self):
        for r in getattr(self, "_ranks", []):
            r.close()
        self._ranks = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _attach_peers(self):
        """Map every rank's receive buffers into every other rank (CUDA IPC over NVLink, or raw pointers when the
        ranks are emulated in this process)."""
        if self.local:
            arr = (ctypes.c_void_p * self.world)(*[r.h for r in self._ranks])
            for r in self._ranks:
                _lib.check(self._lib.kf_slab_p2p_attach_local(r.h, arr), "kf_slab_p2p_attach_local")
            return
        r = self._ranks[0]
        mine = torch.frombuffer(bytearray(bytes(r.ipc)), dtype=torch.uint8).to(self.device)
        if self.world > 1:
            allh = torch.empty(self.world * 128, dtype=torch.uint8, device=self.device)
            dist.all_gather_into_tensor(allh, mine, group=self.group)
        else:
            allh = mine
        blob = allh.cpu().numpy().tobytes()
        buf = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
        _lib.check(self._lib.kf_slab_p2p_attach(r.h, ctypes.byref(buf)), "kf_slab_p2p_attach")
        if self.world > 1:
            dist.barrier(group=self.group)

    def _rank_barrier(self):
        """Stream-ordered cross-rank barrier: every rank's earlier kernels (and their peer stores) are complete
        before any rank's later kernels start.  Peer-to-peer mode: one small kernel per rank on epoch flags in the
        peer-mapped buffers (``kf_slab_barrier``); otherwise a 1-element NCCL all-reduce."""
        if self.local or self.world == 1:
            return
        if self.p2p and self.device_barrier:
            _lib.check(self._lib.kf_slab_barrier(self._ranks[0].h, self._stream()), "kf_slab_barrier")
        else:
            dist.all_reduce(self._flag, group=self.group)

    def _all_to_all(self, which):
        """Block d of every rank's send buffer goes to rank d; block s of a recv buffer comes from rank s.
        In peer-to-peer mode the producing kernels have already stored the blocks remotely: only a barrier."""
        if self.p2p:
            self._rank_barrier()
            return
        if self.world == 1:
            return
        if self.local:
            P = self.world
            sends = [getattr(r, which + "_send").view(P, -1) for r in self._ranks]
            for r in self._ranks:
                recv = getattr(r, which + "_recv").view(P, -1)
                for s in range(P):
                    recv[s].copy_(sends[s][r.rank])
        else:
            r = self._ranks[0]
            dist.all_to_all_single(getattr(r, which + "_recv"), getattr(r, which + "_send"), group=self.group)

    def _all_reduce(self, name):
        """Sum a small per-rank tensor attribute over the ranks; returns the total."""
        if self.local:
            return torch.stack([getattr(r, name) for r in self._ranks]).sum(0)
        t = getattr(self._ranks[0], name).clone()
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        return t

    @property
    def launch_count(self):
        return sum(int(self._lib.kf_slab_launch_count(r.h)) for r in self._ranks)

    @property
    def workspace_bytes(self):
        per = [int(self._lib.kf_slab_workspace_bytes(r.h)) +
               sum(t.numel() * t.element_size() for t in {id(t): t for t in
                                                          (r.fld_send, r.fld_recv, r.adv_send, r.adv_recv)
                                                          if t is not None}.values())
               for r in self._ranks]
        return max(per)

    # -- solver ----------------------------------------------------------------
    def set_state(self, omega_hat, action=None):
        """Load the FULL state (N, Mh) complex (every rank passes the same array) and the action (4,) or None."""
        if omega_hat.shape != (self.N, self.Mh) or omega_hat.dtype != self.cdtype:
            raise ValueError(f"omega_hat must be {(self.N, self.Mh)} {self.cdtype}")
        omega_hat = omega_hat.to(self.device).contiguous()
        if action is not None:
            action = torch.as_tensor(action).to(device=self.device, dtype=self.dtype).contiguous()
            if action.shape != (4,):
                raise ValueError("action must have shape (4,)")
        with torch.cuda.device(self.device):
            s = self._stream()
            if self.p2p:
                self._rank_barrier()  # nobody is still reading the receive buffers the transposes will overwrite
            for r in self._ranks:
                _lib.check(self._lib.kf_slab_import(r.h, _ptr(omega_hat), s), "kf_slab_import")
                _lib.check(self._lib.kf_slab_set_action(r.h, _ptr(action), s), "kf_slab_set_action")
                _lib.check(self._lib.kf_slab_col(r.h, -1, None, _ptr(r.fld_send), s), "kf_slab_col")
            self._all_to_all("fld")

    def set_action(self, action):
        action = None if action is None else torch.as_tensor(action).to(device=self.device, dtype=self.dtype).contiguous()
        with torch.cuda.device(self.device):
            for r in self._ranks:
                _lib.check(self._lib.kf_slab_set_action(r.h, _ptr(action), self._stream()), "kf_slab_set_action")

    def rk_steps(self, n_steps):
        """Advance the loaded state by ``n_steps`` RK solver steps (5 sub-stages, 10 all-to-alls each).  In
        peer-to-peer mode with one rank per process the whole step is ONE CUDA graph per rank replayed by
        ``kf_slab_rk_steps`` (device-side barriers, no host call per sub-stage)."""
        with torch.cuda.device(self.device):
            s = self._stream()
            if self.p2p and self.use_graph and not (self.local and self.world > 1):
                _lib.check(self._lib.kf_slab_rk_steps(self._ranks[0].h, int(n_steps), s), "kf_slab_rk_steps")
                self.rk_steps_done += int(n_steps)
                return
            for _ in range(int(n_steps)):
                for stage in range(5):
                    for r in self._ranks:
                        _lib.check(self._lib.kf_slab_row(r.h, _ptr(r.fld_recv), _ptr(r.adv_send), s), "kf_slab_row")
                    self._all_to_all("adv")
                    for r in self._ranks:
                        _lib.check(self._lib.kf_slab_col(r.h, stage, _ptr(r.adv_recv), _ptr(r.fld_send), s),
                                   "kf_slab_col")
                    self._all_to_all("fld")
        self.rk_steps_done += int(n_steps)

    def phase_times(self, n_steps=4):
        """Where one sub-stage's time goes on THIS rank: advances the state by ``n_steps`` RK steps through the
        per-phase entry points (no graph) with CUDA events between them and returns the mean microseconds per
        sub-stage of ``row`` (received-field transposes + row kernel, advection stored to the peers), ``wait_row``
        (cross-rank barrier after it: the slowest rank's row phase + the flag round trip), ``col`` (advection
        transpose + column kernels, fields stored to the peers) and ``wait_col``.  Peer-to-peer mode only."""
        if not self.p2p:
            raise ValueError("phase_times needs exchange='p2p'")
        names = ("row", "wait_row", "col", "wait_col")
        n = int(n_steps) * 5
        with torch.cuda.device(self.device):
            s = self._stream()
            stream = torch.cuda.current_stream(self.device)
            ev = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(n)]
            for i in range(n):
                stage = i % 5
                ev[i][0].record(stream)
                for r in self._ranks:
                    _lib.check(self._lib.kf_slab_row(r.h, _ptr(r.fld_recv), _ptr(r.adv_send), s), "kf_slab_row")
                ev[i][1].record(stream)
                self._all_to_all("adv")
                ev[i][2].record(stream)
                for r in self._ranks:
                    _lib.check(self._lib.kf_slab_col(r.h, stage, _ptr(r.adv_recv), _ptr(r.fld_send), s), "kf_slab_col")
                ev[i][3].record(stream)
                self._all_to_all("fld")
                ev[i][4].record(stream)
            torch.cuda.synchronize(self.device)
        self.rk_steps_done += int(n_steps)
        return {nm: 1e3 * sum(e[k].elapsed_time(e[k + 1]) for e in ev) / n for k, nm in enumerate(names)}

    def get_state(self):
        """The full state (N, Mh) complex, assembled on every rank (all-reduce of the per-rank column slabs)."""
        out = torch.zeros((self.N, self.Mh), dtype=self.cdtype, device=self.device)
        with torch.cuda.device(self.device):
            for r in self._ranks:
                _lib.check(self._lib.kf_slab_export(r.h, _ptr(out), self._stream()), "kf_slab_export")
        if not self.local and self.world > 1:
            flat = torch.view_as_real(out)
            dist.all_reduce(flat, group=self.group)
        return out

    def reward_obs(self, obs_weight=1.0, obs_out=None):
        """(obs[obs_size^2], tke) of the CURRENT state: partial sums per rank, summed over the ranks
        (compute_tke hydrogym/jax/utils/utils.py:175-194; probes kolmogorov.py:426-446)."""
        n2 = self.obs_size ** 2
        with torch.cuda.device(self.device):
            for r in self._ranks:
                r.acc.zero_()
                r.obs = torch.zeros(n2, dtype=self.dtype, device=self.device)
                _lib.check(self._lib.kf_slab_obs(r.h, _ptr(r.fld_recv), _ptr(r.acc), _ptr(r.obs), float(obs_weight),
                                                 self._stream()), "kf_slab_obs")
        acc = self._all_reduce("acc")
        obs = self._all_reduce("obs")
        cnt = float(self.N) * float(self.M)
        tke = 0.5 * ((acc[2] - acc[0] * acc[0] / cnt) + (acc[3] - acc[1] * acc[1] / cnt)) / (float(self.N) ** 2)
        if obs_out is not None:
            obs_out += obs
        return obs, tke

    def rollout(self, omega_hat, action=None, n_outer=10, n_inner=999):
        """``_rollout`` + reward/obs of one environment (kolmogorov.py:397-446): returns
        (final omega_hat, mean obs[obs_size^2], mean_tke) over the n_outer snapshots."""
        self.set_state(omega_hat, action)
        obs = torch.zeros(self.obs_size ** 2, dtype=self.dtype, device=self.device)
        tke = torch.zeros((), dtype=torch.float64, device=self.device)
        for _ in range(int(n_outer)):
            self.rk_steps(n_inner)
            _, t = self.reward_obs(1.0 / n_outer, obs_out=obs)
            tke += t / n_outer
        return self.get_state(), obs, tke.to(self.dtype)


# -- layout algebra (host restatement of what transpose_kernel / the blocked index maps do) ------------------
# Used by the CPU tests (gloo, world size 2) to pin the exchange pattern; the GPU path does the same moves in
# transpose_kernel (K rows -> [P][X][Q] send blocks) and in the row kernels' blocked addressing.

def k_to_send_blocks(fk, world):
    """K-layout array [Q][N] (this rank's columns, x contiguous) -> send blocks [P][X][Q]; block d holds the
    rows x in [d*X, (d+1)*X) that rank d owns in the X layout."""
    Q, N = fk.shape
    return fk.T.reshape(world, N // world, Q)


def recv_blocks_to_rows(recv):
    """Received blocks [P][X][Q] (block s = columns owned by rank s) -> this rank's rows [X][P*Q]."""
    P, X, Q = recv.shape
    return recv.transpose(1, 0, 2).reshape(X, P * Q) if isinstance(recv, np.ndarray) else \
        recv.permute(1, 0, 2).reshape(X, P * Q)


def rows_to_send_blocks(rows, world):
    """This rank's rows [X][MH] -> send blocks [P][X][Q]; block d holds the columns rank d owns."""
    X, MH = rows.shape
    Q = MH // world
    r = rows.reshape(X, world, Q)
    return r.transpose(1, 0, 2) if isinstance(rows, np.ndarray) else r.permute(1, 0, 2)


def recv_blocks_to_k(recv):
    """Received blocks [P][X][Q] (block s = rows x in [s*X, (s+1)*X)) -> K-layout array [Q][N]."""
    P, X, Q = recv.shape
    return recv.reshape(P * X, Q).T
