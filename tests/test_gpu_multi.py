"""Multi-GPU parity of the two places where the sweep path communicates, on real NCCL: replica exchange
(update::replica_exchange / partner_rank, source/update.cpp:34-117) and the pooled measurement bin.  Skipped below two
devices (the single-GPU box of the round-end run); on an N-GPU box it launches one process per GPU with torchrun and
reads the checks benchmarks/replica_exchange_run.py makes after every attempt: periodic pairing is an involution that
matches dqmcb_partner_rank, accepted ladders hold exactly the partner's previous fields and rejected ones their own
(field conservation), the engine state equals a rebuild from the installed fields, and the all-reduce of the REAL device
bins (dqmcb_bin_*) equals the mean of the gathered per-rank sums."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch

    return torch.cuda.device_count()


@pytest.mark.parametrize("ladder", ["beta", "U"])
def test_replica_exchange_and_pooled_bins_over_nccl(ladder):
    n = min(_ngpu(), 8)
    if n < 2:
        pytest.skip("needs at least two GPUs")
    n -= n % 2  # the reference requires an even number of replicas (source/main.cpp:57)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29631" if ladder == "beta" else "29632", os.path.join(ROOT, "benchmarks", "replica_exchange_run.py"),
           "--ladder", ladder, "--walkers", "64", "--sweeps", "4", "--every", "1", "--nt", "20", "--beta0", "2.0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    d = json.loads(line)
    assert d["checks_ok"] and d["pooled_bin_allreduce_ok"] and d["exchanges"] == 4
    for p in d["pairings"]:
        pr = p["partner_of_rank"]
        assert all(pr[pr[r]] == r for r in range(n))
    if n == 8:  # rank 0 meets rank 7 on one parity of the attempt counter and rank 1 on the other
        assert {p["partner_of_rank"][0] for p in d["pairings"]} == {1, 7}
