"""Multi-GPU plumbing of the sweep path: one process per GPU over torch.distributed (NCCL on the
GPU box, gloo in the CPU tests).

The sweep itself never communicates: walkers (Markov chains) are independent
(reference source/main.cpp:130-133,156-157), so they shard across ranks with no data-path
collective (`shard_walkers`).  Communication exists in exactly two places, as in SURVEY.md 8e:

* replica exchange (reference source/update.cpp:34-117, one MPI rank per beta): `ReplicaExchange`
  keeps the reference's protocol -- fields both ways, own and cross actions, the lower rank decides,
  verdict to the partner, rebuild on reject -- but batched over the W ladders a rank holds and with
  int8 fields (W*nt*N bytes per direction instead of the reference's int64 nt*N per chain), as grouped
  point-to-point operations (`batch_isend_irecv`: ncclSend/ncclRecv pairs over NVLink);
* measurement: `reduce_bin` all-reduces per-bin sums of the equal-time observables, `jackknife`
  restates scripts/analysis.py:232-298 for the pooled bins (the reference pools files offline).

Engines are duck-typed so the host logic can be exercised on CPU: an engine provides
`fields_buffer()`, `install_fields(t)`, `rebuild()`, `global_action()` (see `ContextEngine`).
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.distributed as dist


def shard_walkers(total: int, rank: int, world: int):
    """Contiguous block partition of `total` independent walkers: (offset, count) of `rank`."""
    base, rem = divmod(total, world)
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def partner_rank(rank: int, world_size: int, exchange_attempt: int) -> int:
    """update::partner_rank (source/update.cpp:34-45): periodic even/odd pairing.  Pure integer
    arithmetic; the C ABI exports the same function as dqmcb_partner_rank."""
    even_attempt = exchange_attempt % 2 == 0
    even_rank = rank % 2 == 0
    off = 1 if even_attempt == even_rank else -1
    return (rank + off + world_size) % world_size


class _DevPtr:
    """Zero-copy torch view of a raw device pointer (dqmcb_device_ptr) via __cuda_array_interface__."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3}


class ContextEngine:
    """Adapter of a capi.Context (W walkers of one replica on this rank's GPU) to the exchange protocol."""

    def __init__(self, ctx, device: int):
        from . import capi

        self.ctx, self.capi, self.device = ctx, capi, device
        dev = torch.device("cuda", device)
        ptr = ctx.device_ptr(capi.PTR_FIELDS)
        self._fields = torch.as_tensor(_DevPtr(ptr, (ctx.W, ctx.nt, ctx.N), "|i1"), device=dev)
        # the rest of a walker's state (LDR stack, G, log det M): zero-copy views for snapshot / restore around an exchange
        W, N, ns = ctx.W, ctx.N, ctx.n_stack
        view = lambda which, shape: torch.as_tensor(_DevPtr(ctx.device_ptr(which), shape, "<f8"), device=dev)
        self._state = [view(capi.PTR_STACK_L, (W, ns * N * N)), view(capi.PTR_STACK_D, (W, ns * N)), view(capi.PTR_STACK_R, (W, ns * N * N)),
                       view(capi.PTR_G, (W, N * N)), view(capi.PTR_LOGDET, (W,))]
        self._snap = None

    @property
    def n_walkers(self):
        return self.ctx.W

    def fields_buffer(self) -> torch.Tensor:
        """int8 [W, nt, N] view of the device-resident fields (no copy)."""
        self.ctx.synchronize()
        return self._fields

    def install_fields(self, t: torch.Tensor) -> None:
        if t.data_ptr() != self._fields.data_ptr():
            self._fields.copy_(t)
        torch.cuda.synchronize(self.device)

    def snapshot(self) -> None:
        """Device-side copy of every walker's derived state (stacks, G, log det M) -- 0.5 GB at C1 / 1024 walkers, a
        fraction of a millisecond of HBM time against the ~50 ms of a rebuild."""
        self.ctx.synchronize()
        if self._snap is None:
            self._snap = [torch.empty_like(t) for t in self._state]
        for d, s_ in zip(self._snap, self._state):
            d.copy_(s_)

    def restore(self, walkers: torch.Tensor) -> None:
        """Put the snapshot back for the walkers selected by the bool mask `walkers` (their exchange was rejected)."""
        idx = torch.nonzero(walkers.to(self._fields.device), as_tuple=False).flatten()
        if idx.numel():
            for d, s_ in zip(self._state, self._snap):
                d.index_copy_(0, idx, s_.index_select(0, idx))
        torch.cuda.synchronize(self.device)

    def rebuild(self) -> None:
        """DQMC::init_stacks + init_greenfunctions (source/update.cpp:76-80,109-115)."""
        self.ctx.init_stacks()
        self.ctx.init_greenfunctions()

    def global_action(self) -> torch.Tensor:
        return torch.from_numpy(self.ctx.global_action())


class ReplicaExchange:
    """update::replica_exchange (source/update.cpp:47-117) for the W ladders held by this rank."""

    def __init__(self, engine, rank: int, world_size: int, group=None, comm_device=None, seed: int = 0):
        self.e, self.rank, self.ws, self.group = engine, rank, world_size, group
        self.attempt = 0          # exchange_attempt (first call uses 1, source/update.cpp:52)
        self.accepted = 0         # per-pair accepts seen by this rank (the reference counts on rank 0 only)
        self.proposed = 0
        self.comm_device = comm_device  # where scalars travel (cuda for nccl, cpu for gloo)
        self._gen = np.random.default_rng(seed * 7919 + rank)
        self.profile = None       # set to a dict to collect per-phase wall times of step() (development aid; adds syncs)

    def _p2p(self, send: torch.Tensor, recv: torch.Tensor, partner: int):
        ops = [dist.P2POp(dist.isend, send, partner, self.group), dist.P2POp(dist.irecv, recv, partner, self.group)]
        if self.rank > partner:  # same global order on both sides
            ops.reverse()
        for r in dist.batch_isend_irecv(ops):
            r.wait()

    def step(self, u_decider: np.ndarray | None = None) -> np.ndarray:
        """One exchange attempt.  `u_decider[W]` (optional) = the canonical doubles the deciding rank's
        bernoulli would draw (parity runs); otherwise a rank-local generator is used.
        Returns accept[W] (bool) as seen by this rank."""
        self.attempt += 1
        partner = partner_rank(self.rank, self.ws, self.attempt)
        W = self.e.n_walkers
        import time as _time

        _t = [_time.perf_counter()]

        def _mark(name):
            if self.profile is not None:
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                now = _time.perf_counter()
                self.profile[name] = self.profile.get(name, 0.0) + now - _t[0]
                _t[0] = now

        if partner == self.rank or partner < 0 or partner >= self.ws:
            return np.zeros(W, dtype=bool)
        own = self.e.fields_buffer()
        mine = own.clone()
        theirs = torch.empty_like(own)
        self._p2p(mine, theirs, partner)                      # MPI_Sendrecv of the fields (update.cpp:64-66)
        _mark("swap_fields")
        S_own = self.e.global_action().to(torch.float64)      # S_r({s}_r)          (update.cpp:73)
        can_snapshot = hasattr(self.e, "snapshot")
        if can_snapshot:
            self.e.snapshot()                                 # own stacks / G / log det M, in case the swap is rejected
        _mark("own_action_snapshot")
        self.e.install_fields(theirs)
        self.e.rebuild()                                      # stacks + G from scratch (update.cpp:76-80)
        S_cross = self.e.global_action().to(torch.float64)    # S_r({s}_partner)    (update.cpp:81)
        _mark("rebuild_cross_action")
        dev = self.comm_device if self.comm_device is not None else own.device
        pack = torch.stack([S_cross, S_own]).to(dev)          # both scalars in one message (update.cpp:84-90)
        other = torch.empty_like(pack)
        self._p2p(pack, other, partner)
        verdict = torch.empty(W, dtype=torch.uint8, device=dev)
        if self.rank < partner:                               # the lower rank decides (update.cpp:93-102)
            dS = (pack[0] + other[0]) - (pack[1] + other[1])
            prob = torch.clamp(torch.exp(-dS), max=1.0).cpu().numpy()
            u = np.asarray(u_decider, dtype=np.float64) if u_decider is not None else self._gen.random(W)
            acc = u < prob
            verdict.copy_(torch.from_numpy(acc.astype(np.uint8)))
            for r in dist.batch_isend_irecv([dist.P2POp(dist.isend, verdict, partner, self.group)]):
                r.wait()
        else:
            for r in dist.batch_isend_irecv([dist.P2POp(dist.irecv, verdict, partner, self.group)]):
                r.wait()
            acc = verdict.cpu().numpy().astype(bool)
        _mark("actions_verdict")
        self.proposed += W
        self.accepted += int(acc.sum())
        if not acc.all():                                     # restore rejected walkers (update.cpp:109-115)
            keep = torch.from_numpy(acc).to(own.device).view(W, 1, 1)
            self.e.install_fields(torch.where(keep, theirs, mine))
            if can_snapshot:
                # the reference rebuilds stacks and G from the restored fields (init_stacks + init_greenfunctions, a pure
                # function of the fields); the state saved before the swap IS that result, so the rejected walkers get it
                # back by a masked device copy and nothing is recomputed
                self.e.restore(torch.from_numpy(~acc))
            else:
                self.e.rebuild()
        _mark("restore_rejected")
        return acc


def reduce_bin(local_sum: torch.Tensor, local_count: int, group=None) -> torch.Tensor:
    """Mean over all walkers of all ranks of per-walker observable sums for one bin:
    all_reduce(SUM) of [count, values...] (a few KB)."""
    buf = torch.cat([torch.tensor([float(local_count)], dtype=torch.float64, device=local_sum.device), local_sum.to(torch.float64).flatten()])
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return (buf[1:] / buf[0]).view(local_sum.shape)


def jackknife(bins: np.ndarray):
    """Jackknife mean and error over bins (axis 0), scripts/analysis.py:232-298."""
    bins = np.asarray(bins, dtype=np.float64)
    n = bins.shape[0]
    if n < 2:
        raise ValueError("Need at least 2 bins for jackknife analysis")
    full = bins.mean(axis=0)
    jk = (n * full - bins) / (n - 1)
    var = ((jk - jk.mean(axis=0)) ** 2).sum(axis=0) * (n - 1) / n
    return full, np.sqrt(var)
