"""world_size-2 (gloo, CPU) tests of the multi-rank host logic in dqmc_b200/parallel.py: walker
sharding, the batched replica-exchange protocol, and the measurement reduction.  The exchange is run
over an oracle-backed engine (the CUDA engine needs a GPU; the protocol code is the same) and compared
with the single-process restatement of update::replica_exchange, which tests/test_oracle.py pins
against the reference's own run."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from dqmc_b200 import parallel as P  # noqa: E402
from oracle import dqmc_oracle as O  # noqa: E402

BETAS = [2.0, 1.6]
CFG = dict(L1=4, L2=4, U=4.0, t=1.0, mu=0.0, nt=16, n_stab=4)
W = 2  # ladders per rank


class OracleEngine:
    """W independent chains of one replica, backed by the CPU oracle (test infrastructure)."""

    def __init__(self, beta, seeds):
        self.sims = []
        for s in seeds:
            m = O.Model(CFG["L1"], CFG["L2"], CFG["U"], CFG["t"], CFG["mu"], beta, CFG["nt"], fields=O.random_fields(s, CFG["nt"], 16))
            d = O.DQMC(m, CFG["n_stab"])
            d.init_stacks()
            d.init_greenfunctions()
            self.sims.append(d)

    @property
    def n_walkers(self):
        return len(self.sims)

    def fields_buffer(self):
        return torch.from_numpy(np.stack([d.m.fields for d in self.sims]).astype(np.int8))

    def install_fields(self, t):
        for d, f in zip(self.sims, t.numpy()):
            d.m.fields = f.astype(np.int64).copy()

    def rebuild(self):
        for d in self.sims:
            d.init_stacks()
            d.init_greenfunctions()

    def global_action(self):
        return torch.tensor([d.m.global_action(d.log_det_M) for d in self.sims], dtype=torch.float64)


def _seeds(rank):
    return [500 + 10 * rank + w for w in range(W)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, u_table, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        eng = OracleEngine(BETAS[rank], _seeds(rank))
        ex = P.ReplicaExchange(eng, rank, world, comm_device=torch.device("cpu"))
        accs = []
        for a in range(2):
            accs.append(ex.step(u_decider=u_table[a]))
        f = eng.fields_buffer()
        gathered = [torch.empty_like(f) for _ in range(world)]
        dist.all_gather(gathered, f)
        # measurement reduction: mean over all walkers of all ranks
        local = torch.tensor([float(rank + 1), 2.0 * (rank + 1)], dtype=torch.float64) * W
        red = P.reduce_bin(local, W)
        if rank == 0:
            out.put((np.stack([g.numpy() for g in gathered]), np.stack(accs), red.numpy(), ex.attempt, ex.accepted))
    finally:
        dist.destroy_process_group()


def test_shard_walkers_partition():
    for total in (1024, 1000, 7):
        for world in (1, 2, 4, 8):
            parts = [P.shard_walkers(total, r, world) for r in range(world)]
            assert sum(c for _, c in parts) == total
            off = 0
            for o, c in parts:
                assert o == off
                off += c
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1


def test_partner_rank_matches_reference_table():
    g = np.load(os.path.join(ROOT, "tests", "golden", "misc.npz"))
    tab = g["partner_rank_ws248_r8_a1to4"]
    for wi, ws in enumerate((2, 4, 8)):
        for r in range(ws):
            for ai, a in enumerate(range(1, 5)):
                assert P.partner_rank(r, ws, a) == tab[wi, r, ai]


def test_jackknife_matches_reference_formula():
    rng = np.random.default_rng(3)
    x = rng.standard_normal((12, 5))
    mean, err = P.jackknife(x)
    n = 12
    for j in range(5):  # scripts/analysis.py:232-262, written out
        full = x[:, j].mean()
        jk = np.array([(n * full - x[i, j]) / (n - 1) for i in range(n)])
        e = np.sqrt(np.sum((jk - jk.mean()) ** 2) * (n - 1) / n)
        assert abs(mean[j] - full) < 1e-15 and abs(err[j] - e) < 1e-14
    with pytest.raises(ValueError):
        P.jackknife(x[:1])


def test_replica_exchange_world2_gloo():
    world = 2
    rng = np.random.default_rng(11)
    u_table = rng.random((2, W))
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, u_table, out)) for r in range(world)]
    for p in procs:
        p.start()
    fields, accs, red, attempt, accepted = out.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # expected: the single-process restatement of update::replica_exchange, ladder by ladder
    for w in range(W):
        sims = []
        for r in range(world):
            m = O.Model(CFG["L1"], CFG["L2"], CFG["U"], CFG["t"], CFG["mu"], BETAS[r], CFG["nt"],
                        fields=O.random_fields(_seeds(r)[w], CFG["nt"], 16))
            d = O.DQMC(m, CFG["n_stab"])
            d.init_stacks()
            d.init_greenfunctions()
            sims.append(d)
        for a in range(2):
            res = O.replica_exchange(sims, a + 1, {0: float(u_table[a, w]), 1: float(u_table[a, w])})
            assert bool(accs[a][w]) == res[0]
        for r in range(world):
            assert np.array_equal(fields[r, w], sims[r].m.fields), f"ladder {w} rank {r}"
    assert attempt == 2 and accepted == int(accs.sum())
    assert np.allclose(red, [1.5, 3.0])  # mean over 4 walkers of (1,2),(1,2),(2,4),(2,4)


# ---- pooled measurement bins over ranks (dqmc_b200/measure.py) -----------------------------------------
class _FakeBinCtx:
    """Stands in for capi.Context's bin_* calls (which need a GPU): rank-dependent sums and counts."""

    def __init__(self, rank):
        self.N, self.rank = 16, rank
        self.calls = 0

    def bin_configure(self, L1, L2):
        pass

    def bin_measure(self):
        self.calls += 1

    def bin_fetch(self, reset=True):
        n = 10 * (self.rank + 1) * self.calls
        sc = np.array([1.0, 2.0, 3.0]) * n * (self.rank + 1)
        chi = np.arange(16, dtype=np.float64).reshape((4, 4), order="F") * n * (self.rank + 2)
        self.calls = 0
        return sc, chi, n


def _bin_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dqmc_b200 import capi, measure

        # the Fourier transform of a bin is a device call; no GPU here, so the test substitutes the oracle's
        capi.chi_r_to_chi_k = lambda r, L1, L2, device=0: O.chi_r_to_chi_k(r, L1, L2)
        mm = measure.MeasurementManager(_FakeBinCtx(rank), 4, 4, rank=rank, pool_ranks=True)
        mm.measure()
        mm.measure()
        new = mm.accumulate()
        if rank == 0:
            out.put({k: v for k, v in new.items()})
    finally:
        dist.destroy_process_group()


def test_pooled_measurement_bin_world2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bin_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    new = out.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # rank 0: n=20, sc = (1,2,3)*20 ; rank 1: n=40, sc = (1,2,3)*80 -> pooled mean (1,2,3)*100/60
    assert np.allclose([new[f"/bin_0/scalar/{k}"][0] for k in ("density", "doubleOcc", "swave")], np.array([1, 2, 3]) * 100 / 60)
    base = np.arange(16, dtype=np.float64).reshape((4, 4), order="F")
    want = base * (20 * 2 + 40 * 3) / 60
    assert np.allclose(new["/bin_0/equaltime/densityCorr"][:, :, 0], want)
    k = O.chi_r_to_chi_k(want[:, :, None], 4, 4)
    assert np.allclose(new["/binK_0/equaltime/densityCorr"][..., 0], k.real) and new["/binK_0/equaltime/densityCorr"].shape == (4, 4, 1, 2)
