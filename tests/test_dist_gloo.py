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
