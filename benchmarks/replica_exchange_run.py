"""Config C4 (BASELINE.json): replica exchange across the GPUs of one box, one replica per rank,
W ladders batched per rank, NCCL point-to-point field swaps (dqmc_b200/parallel.py).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/replica_exchange_run.py [--ladder beta|U] [--walkers W] [--sweeps S] [--every E]

beta ladder = what the reference supports (source/main.cpp:47-67); U ladder = extension.
Checks after every exchange: the multiset of field configurations of each ladder is conserved across
ranks, accepted ladders hold exactly the partner's previous fields, and the rebuilt G is the one the
engine recomputes from the installed fields.  Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dqmc_b200 import capi, host, parallel  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ladder", default="beta", choices=["beta", "U"])
    ap.add_argument("--walkers", type=int, default=256)
    ap.add_argument("--sweeps", type=int, default=8)
    ap.add_argument("--every", type=int, default=2)
    ap.add_argument("--L", type=int, default=8)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L, W = args.L, args.walkers
    if args.ladder == "beta":
        betas = np.linspace(8.0, 2.0, world) if world > 1 else np.array([8.0])
        beta, U = float(betas[rank]), 4.0
        nt = 80  # nt is shared along the ladder: each replica has its own dtau (source/model.cpp:23)
    else:
        Us = np.linspace(2.0, 8.0, world) if world > 1 else np.array([4.0])
        beta, U, nt = 8.0, float(Us[rank]), 80
    model = host.AttractiveHubbard(L, L, U, 1.0, 0.0, beta, nt, 10, W, device=local)
    ctx = model.ctx
    rng = np.random.default_rng(17)  # same initial fields on every rank would be degenerate: offset by rank
    f0 = np.random.default_rng(1000 + rank).integers(0, 4, size=(W, nt, L * L), dtype=np.int8)
    ctx.set_fields(f0)
    ctx.rng_mode(capi.RNG_PHILOX, 4242 + rank)
    ctx.init_stacks()
    ctx.init_greenfunctions()
    eng = parallel.ContextEngine(ctx, local)
    ex = parallel.ReplicaExchange(eng, rank, world, comm_device=torch.device("cuda", local), seed=5)
    ok = True
    t_sweep = t_ex = 0.0
    n_ex = 0
    for s in range(args.sweeps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.sweep_0_to_beta()
        ctx.sweep_beta_to_0()
        ctx.synchronize()
        t_sweep += time.perf_counter() - t0
        if world > 1 and (s + 1) % args.every == 0:
            before = eng.fields_buffer().clone()
            gathered_b = [torch.empty_like(before) for _ in range(world)]
            dist.all_gather(gathered_b, before)
            dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            acc = ex.step()
            ctx.synchronize()
            torch.cuda.synchronize()
            t_ex += time.perf_counter() - t0
            n_ex += 1
            after = eng.fields_buffer().clone()
            partner = parallel.partner_rank(rank, world, ex.attempt)
            accd = torch.from_numpy(acc).to(after.device)
            # accepted ladders hold the partner's previous fields, rejected ones their own
            want = torch.where(accd.view(W, 1, 1), gathered_b[partner], before)
            ok &= bool(torch.equal(after, want))
            # the engine state is the one rebuilt from the installed fields
            G1 = ctx.get_G()
            ctx.init_stacks()
            ctx.init_greenfunctions()
            ok &= bool(np.max(np.abs(ctx.get_G() - G1)) < 1e-12)
    stats = torch.tensor([t_sweep, t_ex, float(ex.accepted), float(ex.proposed), float(ok)], dtype=torch.float64, device="cuda")
    mx = stats.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = stats.clone()
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    if rank == 0:
        N = L * L
        print(json.dumps({
            "config": f"C4 replica exchange, {args.ladder} ladder over {world} GPUs, {L}x{L}, nt={nt}, {W} ladders batched per rank",
            "sweeps": args.sweeps, "exchanges": n_ex,
            "sweep_s_max": float(mx[0]), "exchange_s_per_attempt_max": float(mx[1]) / max(n_ex, 1),
            "site_flips_per_sec": world * W * args.sweeps * 2 * nt * N / float(mx[0]),
            "exchange_accept_rate": float(sm[2]) / max(float(sm[3]), 1.0),
            "swap_bytes_per_attempt_per_rank": W * nt * N + 2 * W * 8 + W,
            "checks_ok": bool(sm[4] == world),
        }))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
