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
    ap.add_argument("--beta0", type=float, default=4.0, help="beta of rank 0 (beta ladder) / of every rank (U ladder)")
    ap.add_argument("--dbeta", type=float, default=0.1, help="ladder spacing in beta: close enough for a non-zero exchange rate")
    ap.add_argument("--U0", type=float, default=4.0)
    ap.add_argument("--dU", type=float, default=0.05)
    ap.add_argument("--nt", type=int, default=40)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L, W = args.L, args.walkers
    nt = args.nt  # nt is shared along the ladder: each replica has its own dtau (source/model.cpp:23)
    if args.ladder == "beta":
        beta, U = args.beta0 - args.dbeta * rank, args.U0
    else:
        beta, U = args.beta0, args.U0 + args.dU * rank
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
    pairings = []
    t_first = 0.0
    worst_dG = worst_dld = 0.0
    ok_state = True
    n_untimed = 2  # the first attempt of each pairing parity sets up NCCL's point-to-point channels: not in the average
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
            if n_ex == n_untimed and os.environ.get("DQMCB_EX_PROF"):
                ex.profile = {}
            t0 = time.perf_counter()
            acc = ex.step()
            ctx.synchronize()
            torch.cuda.synchronize()
            if n_ex < n_untimed:
                t_first += time.perf_counter() - t0  # includes NCCL's lazy set-up of the point-to-point channels
            else:
                t_ex += time.perf_counter() - t0
            n_ex += 1
            after = eng.fields_buffer().clone()
            partner = parallel.partner_rank(rank, world, ex.attempt)
            # periodic even/odd pairing (source/update.cpp:34-45): an involution over all ranks; with 8 ranks rank 0 meets
            # rank 7 on one parity of the attempt counter and rank 1 on the other
            pt = torch.tensor([partner], device=after.device)
            allp = [torch.empty_like(pt) for _ in range(world)]
            dist.all_gather(allp, pt)
            allp = [int(x.item()) for x in allp]
            ok &= all(allp[allp[r]] == r and allp[r] != r for r in range(world))
            ok &= allp[rank] == capi.lib().dqmcb_partner_rank(rank, world, ex.attempt)
            if rank == 0:
                pairings.append({"attempt": ex.attempt, "partner_of_rank": allp})
            accd = torch.from_numpy(acc).to(after.device)
            # accepted ladders hold the partner's previous fields, rejected ones their own
            want = torch.where(accd.view(W, 1, 1), gathered_b[partner], before)
            ok &= bool(torch.equal(after, want))
            # the engine state is the one a rebuild from the installed fields gives (rejected ladders were restored from the
            # snapshot taken before the swap: the same function of the fields up to rounding)
            G1, ld1 = ctx.get_G(), ctx.get_logdet()
            ctx.init_stacks()
            ctx.init_greenfunctions()
            dG = float(np.max(np.abs(ctx.get_G() - G1)))
            dld = float(np.max(np.abs(ctx.get_logdet() - ld1) / np.abs(ld1)))
            worst_dG, worst_dld = max(worst_dG, dG), max(worst_dld, dld)
            ok_state &= bool(dG < 1e-8 and dld < 1e-10)
    # pooled measurement bin over the ranks with the REAL device bins (dqmcb_bin_*; include/measurementh5.h:189-274 pooled as
    # in SURVEY 8e): all-reduce of [count, sums] must equal the mean of the gathered per-rank sums
    ctx.bin_configure(L, L)
    ctx.bin_measure()
    sc, chi, nsamp = ctx.bin_fetch()
    local = torch.from_numpy(np.concatenate([sc, chi.ravel()])).to("cuda")
    pooled = parallel.reduce_bin(local, nsamp)
    every = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(every, local)
    want = torch.stack(every).sum(0) / (nsamp * world)
    pool_ok = bool(torch.allclose(pooled, want, rtol=1e-13, atol=1e-13)) and nsamp == W
    stats = torch.tensor([t_sweep, t_ex, float(ex.accepted), float(ex.proposed), float(ok and ok_state), float(pool_ok), worst_dG, worst_dld],
                         dtype=torch.float64, device="cuda")
    mx = stats.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = stats.clone()
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    if rank == 0:
        N = L * L
        print(json.dumps({
            "config": f"C4 replica exchange, {args.ladder} ladder over {world} GPUs, {L}x{L}, nt={nt}, {W} ladders batched per rank",
            "sweeps": args.sweeps, "exchanges": n_ex,
            "sweep_s_max": float(mx[0]), "exchange_s_per_attempt_max": float(mx[1]) / max(n_ex - n_untimed, 1), "first_two_attempts_s": t_first,
            "exchange_phases_s_rank0": ex.profile,
            "site_flips_per_sec": world * W * args.sweeps * 2 * nt * N / float(mx[0]),
            "exchange_accept_rate": float(sm[2]) / max(float(sm[3]), 1.0),
            "swap_bytes_per_attempt_per_rank": W * nt * N + 2 * W * 8 + W,
            "checks_ok": bool(sm[4] == world), "pooled_bin_allreduce_ok": bool(sm[5] == world),
            "state_vs_rebuild": {"max_abs_dG": float(mx[6]), "max_rel_dlogdet": float(mx[7]), "bar": "1e-8 / 1e-10"},
            "sweep_s_per_sweep_max": float(mx[0]) / max(args.sweeps, 1),
            "exchange_cost_in_sweeps": (float(mx[1]) / max(n_ex - n_untimed, 1)) / (float(mx[0]) / max(args.sweeps, 1)),
            "ladder": ({"beta": [args.beta0 - args.dbeta * r for r in range(world)], "U": args.U0} if args.ladder == "beta"
                       else {"U": [args.U0 + args.dU * r for r in range(world)], "beta": args.beta0}),
            "pairings": pairings[:4],
        }))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
