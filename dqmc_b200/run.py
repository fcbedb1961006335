"""Driver: the run loop of the reference's source/main.cpp:36-214 over the batched device engine.

    python -m dqmc_b200.run parameters.in [--walkers W] [--seed S] [--out results] [--checkpoint-every B] [--resume]
    python -m torch.distributed.run --nproc-per-node N ... -m dqmc_b200.run parameters.in ...     (one rank per GPU)

Reads the reference's parameter file (examples/parameters.in, same keys and parser rules), thermalises
`n_therms` full sweeps, then runs `n_bins` bins of `n_sweeps` sweeps: replica exchange every
`sweep_steps` sweeps when [ParallelTempering] enabled (one beta per rank, source/main.cpp:47-67,130-140),
sweep_0_to_beta, sweep_beta_to_0, sweep_unequalTime (if isMeasureUnequalTime), measure; accumulate per bin
(source/main.cpp:156-171).  Every rank writes `<out>/data_<rank>.npz` with the dataset tree of the reference's
results/data_<rank>.h5 (HDF5 is not available here) and rank 0 writes `<out>/info` (include/lattice.h:110-136).

Differences from the reference, all additive: W walkers per rank instead of one chain (every measure() adds W
samples to the bin); Philox counter-based random numbers by default -- `--rng mt19937` regenerates the
reference's libstdc++ stream on the device instead, walker w of rank r = utility::random(seed + r*W + w) with the
reference's random start, so that `--walkers 1 --rng mt19937 --seed S` writes the bins the reference program
writes for rng(S) (tests/golden/program.npz); checkpoint/resume of fields, bins and counters (absent in the reference; a resumed run
continues with a fresh random stream: statistically equivalent, not bit-identical).  `symmetric` is read and
ignored exactly as the reference's measurements ignore the half-warped copy (source/main.cpp:161-165 measures
`greens`, not `greens_symm`).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

from . import capi, host, measure, parallel
from .params import Parameters


def save_info(path: str, L1: int, L2: int) -> None:
    """Lattice::save_info (include/lattice.h:110-136) for the square lattice of source/main.cpp:84-88."""
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        f.write(f"L1 {L1}\nL2 {L2}\nn_orb 1\na1_x 1\na1_y 0\na2_x 0\na2_y 1\n")


def main(argv=None) -> dict:
    ap = argparse.ArgumentParser(prog="dqmc_b200.run")
    ap.add_argument("parameters")
    ap.add_argument("--walkers", type=int, default=1024, help="independent chains per rank (the reference runs 1)")
    ap.add_argument("--seed", type=int, default=None, help="default: time-based like source/main.cpp:37")
    ap.add_argument("--out", default="results")
    ap.add_argument("--checkpoint-every", type=int, default=0, help="bins between checkpoints (0 = off)")
    ap.add_argument("--resume", action="store_true")
    ap.add_argument("--rng", default="philox", choices=["philox", "mt19937"],
                    help="mt19937: walker w of rank r runs the reference's libstdc++ stream of utility::random(seed + r*W + w), "
                         "random start included -- a --walkers 1 run then reproduces the reference program's bins")
    ap.add_argument("--delay-k", type=int, default=-1, help="delayed rank-k updates (default: 32 for N >= 128, else off)")
    args = ap.parse_args(argv)

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    p = Parameters(args.parameters)
    L1, L2 = p.getInt("Lattice", "L1"), p.getInt("Lattice", "L2")
    U, t, mu = p.getDouble("hubbard", "U"), p.getDouble("hubbard", "t"), p.getDouble("hubbard", "mu")
    nt, n_stab = p.getInt("simulation", "nt"), p.getInt("simulation", "n_stab")
    n_sweeps, n_therms, n_bins = p.getInt("simulation", "n_sweeps"), p.getInt("simulation", "n_therms"), p.getInt("simulation", "n_bins")
    unequal = p.getBool("simulation", "isMeasureUnequalTime")
    p.getBool("simulation", "symmetric", False)
    pt = p.getBool("ParallelTempering", "enabled", False)
    if pt:  # source/main.cpp:47-67
        betas = p.getDoubleVector("ParallelTempering", "betas")
        if len(betas) != world:
            sys.stderr.write(f"ERROR: The number of betas ({len(betas)}) in parameters.in must match the number of processes ({world}).\n")
            raise SystemExit(1)
        if world % 2 != 0:
            sys.stderr.write(f"ERROR: currently number of processor ( nprocs = {world}) need to be even for replica exchange\n")
            raise SystemExit(1)
        beta, exchange_step = betas[rank], p.getInt("ParallelTempering", "sweep_steps")
    else:
        beta, exchange_step = p.getDouble("simulation", "beta"), 0
    # time-based default seed (source/main.cpp:37): taken ONCE, on rank 0, and broadcast -- ranks that start a second apart
    # would otherwise draw the same seed + rank and run duplicate chains
    seed0 = int(time.time()) if args.seed is None else args.seed
    if dist is not None and args.seed is None:
        import torch

        tsd = torch.tensor([seed0], dtype=torch.int64, device="cuda")
        dist.broadcast(tsd, src=0)
        seed0 = int(tsd.item())
    seed = seed0 + rank
    N = L1 * L2
    delay_k = args.delay_k if args.delay_k >= 0 else (0 if N < 128 else (16 if N <= 160 else 32))
    if unequal:
        # the unequal-time sweep keeps Gtt / Gt0 / G0t of every slice of every walker on the device: fail BEFORE thermalisation,
        # naming the walker count that fits, instead of with a bare cudaMalloc error in the first bin
        import torch

        need = args.walkers * 3 * (nt + 1) * N * N * 8
        free = torch.cuda.mem_get_info(local)[0] if torch.cuda.is_available() else 0
        if free and need > 0.9 * free:
            fit = int(0.9 * free // (3 * (nt + 1) * N * N * 8))
            sys.stderr.write(f"ERROR: isMeasureUnequalTime needs {need / 1e9:.1f} GB of device memory for {args.walkers} walkers "
                             f"({free / 1e9:.1f} GB free): use --walkers {fit} or fewer\n")
            raise SystemExit(1)
    model = host.AttractiveHubbard(L1, L2, U, t, mu, beta, nt, n_stab, args.walkers, device=local, delay_k=delay_k)
    ctx = model.ctx
    ckpt = os.path.join(args.out, f"checkpoint_{rank}.npz")
    mm = measure.MeasurementManager(ctx, L1, L2, rank=rank, out_dir=args.out, device=local, unequal_time=unequal)
    start_bin, therm_done = 0, False
    have = int(args.resume and os.path.exists(ckpt))
    next_bin = int(np.load(ckpt, allow_pickle=False)["next_bin"]) if have else 0
    if dist is not None and args.resume:
        # every rank must resume from the same bin (or none): a missing or older checkpoint on one rank would desynchronise
        # the sweep counts and deadlock the exchange barriers
        import torch

        tv = torch.tensor([have, next_bin, -next_bin], dtype=torch.int64, device="cuda")
        dist.all_reduce(tv, op=dist.ReduceOp.MIN)
        if int(tv[0]) != have or int(tv[1]) != next_bin or -int(tv[2]) != next_bin:
            sys.stderr.write(f"ERROR: rank {rank}: checkpoints disagree across ranks (have={have}, next_bin={next_bin}); not resuming\n")
            raise SystemExit(1)
    ex_state = None
    if have:
        z = np.load(ckpt, allow_pickle=False)
        if "ex_state" in z.files:
            ex_state = z["ex_state"]
        ctx.set_fields(z["fields"])
        start_bin, therm_done = int(z["next_bin"]), True
        mm.current_bin = start_bin
        mm.datasets = {k[3:]: z[k] for k in z.files if k.startswith("ds:")}
        seed += 7919 * (start_bin + 1)  # fresh stream after a restart
    elif args.rng == "mt19937":
        base = seed0 + rank * args.walkers
        ctx.set_fields(np.stack([host.reference_fields(base + w, nt, N) for w in range(args.walkers)]))
    else:
        model.random_fields(seed)       # GHQField's random start (include/field.h:52-57), one draw per walker
    if args.rng == "mt19937":
        base = seed0 + rank * args.walkers + (7919 * start_bin if therm_done else 0)
        ctx.rng_seed_mt19937(np.arange(base, base + args.walkers, dtype=np.uint64).astype(np.uint32))
    else:
        ctx.rng_mode(capi.RNG_PHILOX, seed)
    ctx.init_stacks()
    ctx.init_greenfunctions()
    ex = None
    if pt:
        import torch

        eng = parallel.ContextEngine(ctx, local)
        ex = parallel.ReplicaExchange(eng, rank, world, comm_device=torch.device("cuda", local), seed=seed)
        if ex_state is not None:  # attempt parity and counters continue across a restart
            ex.attempt, ex.accepted, ex.proposed = (int(x) for x in ex_state)
    if rank == 0:
        save_info(os.path.join(args.out, "info"), L1, L2)
        print(f"{'Parallel Tempering enabled.' if pt else 'Standard DQMC run (Parallel Tempering disabled).'}", flush=True)

    def one_sweep(isweep: int, exchange: bool = True) -> None:
        if exchange and ex is not None and isweep % exchange_step == 0:
            dist.barrier()
            ex.step()
        ctx.sweep_0_to_beta()
        ctx.sweep_beta_to_0()

    t0 = time.perf_counter()
    if not therm_done:
        for isweep in range(1, n_therms + 1):  # source/main.cpp:130-133: no exchange while thermalising
            one_sweep(isweep, exchange=False)
        ctx.synchronize()
    t_therm = time.perf_counter() - t0
    t0 = time.perf_counter()
    for b in range(start_bin, n_bins):          # source/main.cpp:145-171
        for s in range(1, n_sweeps + 1):
            one_sweep(b * n_sweeps + s)
            if unequal:
                ctx.sweep_unequal_time()
            mm.measure()
        mm.accumulate()
        if args.checkpoint_every and (b + 1) % args.checkpoint_every == 0 and b + 1 < n_bins:
            os.makedirs(args.out, exist_ok=True)
            tmp = ckpt + ".tmp.npz"  # written beside and renamed: a kill during the write never leaves a truncated checkpoint
            np.savez(tmp, fields=ctx.get_fields(), next_bin=b + 1,
                     ex_state=np.array([ex.attempt, ex.accepted, ex.proposed] if ex is not None else [0, 0, 0], dtype=np.int64),
                     **{"ds:" + k: v for k, v in mm.datasets.items()})
            os.replace(tmp, ckpt)
    ctx.synchronize()
    t_meas = time.perf_counter() - t0
    path = mm.save()
    acc, emax, emean = ctx.get_stats()
    if unequal:  # the reference folds the unequal-time check_error values into max_err() too (source/dqmc.cpp:264-280)
        emax = np.maximum(emax, ctx.get_unequal_err())
    n_half = 2.0 * ((n_bins - start_bin) * n_sweeps + (0 if therm_done else n_therms))
    summary = {
        "rank": rank, "beta": beta, "walkers": args.walkers, "bins": n_bins, "file": path,
        "thermalisation_s": t_therm, "measurement_s": t_meas,
        "acceptance_rate": float(acc.mean() / max(n_half, 1.0)),             # source/main.cpp:183
        "max_precision_error": float(emax.max()), "mean_precision_error": float(np.nanmean(emean)),
        "site_flips_per_sec": args.walkers * (n_bins - start_bin) * n_sweeps * 2 * nt * N / max(t_meas, 1e-9),
    }
    if ex is not None:
        summary["exchange_rate"] = ex.accepted / max(ex.proposed, 1)        # source/main.cpp:203-208
    print(json.dumps(summary), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return summary


if __name__ == "__main__":
    main()
