#!/usr/bin/env python
"""bench.py -- headline benchmark of the DQMC sweep hot path (BASELINE.json metric:
site-flips/sec and sweeps/sec per box).

A "step" is one full sweep (sweep_0_to_beta + sweep_beta_to_0, source/main.cpp:156-157) of every
walker on every GPU.  Workload at N=1 GPU: config C1 of BASELINE.json ("Hubbard 8x8 U=4 beta=8
half-filling, 1024 batched walkers per GPU"), nt = 80 (dtau = 0.1), n_stab = 10.  Walkers shard
across ranks with no data-path collective (weak scaling: 1024 walkers per GPU).

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, through the C ABI)
    python bench.py --impl reference --gpus N ...            # reference CPU arm (oracle/_ref)

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (L1, L2, U, t, mu, beta, nt, n_stab, walkers_per_gpu)
    "C1": (8, 8, 4.0, 1.0, 0.0, 8.0, 80, 10, 1024),
    "C0": (6, 6, 4.0, 1.0, -0.1, 4.0, 40, 10, 2048),  # 2048 walkers: the LDR stacks (172 MB) exceed the L2
    "C2": (12, 12, 4.0, 1.0, 0.0, 10.0, 100, 10, 296),
    "C3": (16, 16, 6.0, 1.0, 0.0, 16.0, 160, 10, 148),
}
METRIC = "site_flips_per_sec"
UNIT = "site-flips/s"


def workload_desc(name, wl, n_gpus):
    L1, L2, U, t, mu, beta, nt, n_stab, W = wl
    return {
        "workload": f"{name}: attractive Hubbard {L1}x{L2} U={U:g} mu={mu:g} beta={beta:g} nt={nt} n_stab={n_stab}, "
                    f"{W} walkers/GPU, 1 step = 1 full sweep (fwd+bwd) of every walker",
        "lattice": f"{L1}x{L2}", "U": U, "beta": beta, "nt": nt, "n_stab": n_stab,
        "walkers_per_gpu": W, "walkers_total": W * n_gpus,
        "proposals_per_step": 2 * nt * L1 * L2 * W * n_gpus,
        "l2_policy": (lambda mb: ("working set (LDR stacks %.0f MB/GPU) larger than the 126 MB L2, no flush" % mb) if mb > 126 else
                      ("WARNING: LDR stacks %.0f MB/GPU fit in the 126 MB L2 and no flush is done -- raise --walkers" % mb))(
                          W * ((nt + n_stab - 1) // n_stab) * (2 * (L1 * L2) ** 2 + L1 * L2) * 8 / 1e6),
        "parallelism": f"walkers sharded over {n_gpus} GPU(s), no data-path collective",
    }


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock / throttle reasons with nvidia-smi while the timed region runs."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 3 + i and s[3 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.samples[0][1]),
                "reasons": reasons, "samples": len(self.samples)}


def host_cores():
    """(hardware threads, physical cores) of this host; physical = distinct (physical id, core id) pairs of /proc/cpuinfo."""
    threads = os.cpu_count() or 1
    phys = set()
    try:
        pid = cid = None
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("physical id"):
                pid = ln.split(":")[1].strip()
            elif ln.startswith("core id"):
                cid = ln.split(":")[1].strip()
            elif not ln.strip():
                if cid is not None:
                    phys.add((pid, cid))
                pid = cid = None
        if cid is not None:
            phys.add((pid, cid))
    except OSError:
        pass
    return threads, (len(phys) or threads)


def measured_peaks():
    p = {}
    for fn in ("MEASURED_PEAKS.json", os.path.join("profiles", "fp64_smem_peaks.json")):
        try:
            with open(os.path.join(ROOT, fn)) as f:
                p.update(json.load(f))
        except Exception:
            pass
    return p


# ------------------------------------------------------------------------------------------------
def cpu_reference_run(wl, n_chains, n_sweeps, warm_sweeps, seed0=1000):
    """Times the reference's own CPU sweep loop (oracle/_ref = its unmodified sources) on
    n_chains host threads, one independent chain each.  Returns (flips/s, sweeps/s, seconds)."""
    from oracle import ref

    if not ref.available():
        raise RuntimeError("oracle/_ref/libdqmc_ref.so missing")
    L1, L2, U, t, mu, beta, nt, n_stab, _ = wl
    import io
    import contextlib

    sims = [ref.RefSim(L1, L2, U, t, mu, beta, nt, n_stab, seed0 + i) for i in range(n_chains)]
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(2)
    os.dup2(devnull, 2)  # the reference prints a precision warning per stabilisation from random starts
    try:
        if warm_sweeps:
            ref.run_sweeps(sims, warm_sweeps)
        secs = ref.run_sweeps(sims, n_sweeps)
    finally:
        os.dup2(saved, 2)
        os.close(devnull)
        os.close(saved)
    sweeps = n_chains * n_sweeps
    return sweeps * 2 * nt * L1 * L2 / secs, sweeps / secs, secs


def run_reference_arm(args, wl, name):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores, phys = host_cores()
    chains = cores
    per_step = {36: 32, 64: 8, 144: 1, 256: 1}.get(wl[0] * wl[1], 1)  # bounded sample: the run ends within minutes
    L1, L2, U, t, mu, beta, nt, n_stab, W = wl
    from oracle import ref

    if not ref.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/libdqmc_ref.so was not built (needs /root/reference at build time)"})
        return
    sims = [ref.RefSim(L1, L2, U, t, mu, beta, nt, n_stab, 1000 + i) for i in range(chains)]
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(2)
    os.dup2(devnull, 2)
    try:
        for _ in range(args.warmup):
            ref.run_sweeps(sims, per_step)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ref.run_sweeps(sims, per_step)
        secs = time.perf_counter() - t0
    finally:
        os.dup2(saved, 2)
        os.close(devnull)
        os.close(saved)
    sweeps = chains * per_step * args.steps
    flips = sweeps * 2 * nt * L1 * L2
    value = flips / secs
    from oracle import ref as _r

    sample = (f"{chains} independent chains (one per hardware thread; {phys} physical cores) x {per_step} full sweep(s) per step -- a bounded sample of "
              f"the workload, NOT {W} walkers -- reference sources compiled unmodified, {'-O3 -ffast-math' if _r.VARIANT == 'fast' else '-O3'}, "
              f"OpenBLAS (not MKL: absent here) pinned to 1 thread per chain")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_desc(name, wl, args.gpus),
        "sweeps_per_sec": sweeps / secs,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": chains, "physical_cores": phys, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
_REAL_STDOUT = None


def _quiet_stdout():
    """Libraries (NCCL prints its version banner on the first communicator) write to fd 1; the contract is ONE
    JSON line on stdout, so everything else is sent to stderr and the line goes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=120)  # >= 2 s of timed region at C1
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C1", choices=list(WORKLOADS))
    ap.add_argument("--walkers", type=int, default=0, help="override walkers per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--therm", type=int, default=100, help="untimed thermalisation sweeps before the timed region (SURVEY 8d: >= 100)")
    ap.add_argument("--n-stab", type=int, default=0, help="override n_stab (C3 is reported at 10 and at 5, SURVEY H2)")
    ap.add_argument("--delay-k", type=int, default=-1, help="delayed rank-k Metropolis updates (default: 32 for N >= 128, else rank-1)")
    args = ap.parse_args()
    wl = list(WORKLOADS[args.workload])
    if args.walkers:
        wl[8] = args.walkers
    if args.n_stab:
        wl[7] = args.n_stab
    wl = tuple(wl)
    if args.impl == "reference":
        run_reference_arm(args, wl, args.workload)
        return

    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the dqmc_b200 product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from dqmc_b200 import capi, host

    L1, L2, U, t, mu, beta, nt, n_stab, W = wl
    N = L1 * L2
    # delayed rank-k updates where G is not register resident: k = 16 keeps the N = 144 tile in shared memory next to the
    # pending factors (k = 32 pushes it out to L2: measured 0.56 vs 0.30 ms per slice), k = 32 at N = 256 (G lives in L2 anyway)
    delay_k = args.delay_k if args.delay_k >= 0 else (0 if N < 128 else (16 if N <= 160 else 32))
    model = host.AttractiveHubbard(L1, L2, U, t, mu, beta, nt, n_stab, W, device=local_rank, delay_k=delay_k)
    ctx = model.ctx
    stream = torch.cuda.Stream(device=local_rank)
    ctx.set_stream(stream.cuda_stream)
    model.random_fields(1234 + rank)
    ctx.rng_mode(capi.RNG_PHILOX, 99 + rank)
    ctx.init_stacks()
    ctx.init_greenfunctions()
    ctx.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        ctx.sweep_0_to_beta()
        ctx.sweep_beta_to_0()

    # ---- thermalisation (untimed): acceptance and the conditioning of the stacks change over the first sweeps
    t_th = time.perf_counter()
    for _ in range(args.therm):
        step()
    ctx.synchronize()
    t_th = time.perf_counter() - t_th
    ctx.reset_stats()
    # ---- kernel-resident throughput (`value`) ----
    # product configuration: walkers split into groups on separate streams (no per-launch events)
    for _ in range(args.warmup):
        step()
    barrier()
    n_launch0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(args.steps):
                step()
            e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    gpu_launches = ctx.launch_count() - n_launch0
    tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    total_walkers = W * world
    flips_per_step = 2 * nt * N * total_walkers
    value = flips_per_step * args.steps / (ms_max * 1e-3)
    sweeps_per_sec = total_walkers * args.steps / (ms_max * 1e-3)
    acc, emax, emean = ctx.get_stats()

    # ---- per-kernel-class launch times for the roofline: CUDA events around every launch on the
    # launching stream (this serialises the walker groups, so it is a separate, untimed-for-`value` pass)
    ctx.timing_enable(True)
    n_cls = max(1, min(args.steps, 3))
    for _ in range(n_cls):
        step()
    ctx.synchronize()
    kms, kn = ctx.timing_get()
    ctx.timing_enable(False)

    # ---- end to end through the C ABI with HOST buffers (`e2e`) ----
    # per step: pinned host RNG tape for the sweep (the reference's rng argument of local_update)
    # H2D, fwd+bwd sweep, measurement, then fields + observables D2H.
    gen = np.random.default_rng(7 + rank)
    visits = 2 * nt
    pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
    t_perm, t_k, t_u = pin((W, visits, N), torch.uint8), pin((W, visits, N), torch.uint8), pin((W, visits, N), torch.float64)
    base = np.tile(np.arange(N, dtype=np.uint8), (W, visits, 1))
    t_perm.numpy()[:] = gen.permuted(base, axis=-1)
    t_k.numpy()[:] = gen.integers(0, 3, size=(W, visits, N), dtype=np.uint8)
    t_u.numpy()[:] = gen.random((W, visits, N))
    h2d = t_perm.numel() + t_k.numel() + t_u.numel() * 8
    d2h = W * nt * N + W * 3 * 8

    # pipelined: the tape of step k+1 goes H2D on a copy stream while step k computes (dqmcb_set_tape_async /
    # dqmcb_tape_commit); every timed step still uploads one full tape and downloads its own results
    a_perm, a_k, a_u = t_perm.numpy(), t_k.numpy(), t_u.numpy()
    ctx.set_tape_async(a_perm, a_k, a_u)
    ctx.tape_commit()

    def e2e_step():
        ctx.set_tape_async(a_perm, a_k, a_u)
        out = ctx.run_sweeps_host(None, 1)
        ctx.tape_commit()
        return out

    for _ in range(max(1, min(args.warmup, 2))):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = flips_per_step * args.steps / float(te.item())
    ctx.rng_mode(capi.RNG_PHILOX, 99 + rank)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel class ----
    peaks = measured_peaks()
    dom = int(np.argmax(kms[:2]))
    dom_name = ["propagate (wrap + Metropolis)", "stabilize (Bbar + LDR stack + G)"][dom]
    acc_rate = float(acc.mean()) / (2.0 * (args.steps + args.warmup)) if acc.size else 0.5  # stats since the end of thermalisation
    a = min(max(acc_rate, 0.0), 1.0)
    n3 = float(N) ** 3
    # algorithmic FP64 flops per launch: propagate launch = n_stab slices x
    # (wrap 4N^3 + accepted flips a*N*2N^2) per walker; stabilize launch = Bbar 2N^3(n_stab-1) +
    # ~18N^3 (2 GEMM + QRP + Q + 2 LU solves + 3 GEMM) per walker
    flops_prop = W * n_stab * (4.0 * n3 + a * 2.0 * n3)
    flops_stab = W * (2.0 * n3 * (n_stab - 1) + 18.0 * n3)
    fl = [flops_prop, flops_stab][dom]
    avg_ms = kms[dom] / max(int(kn[dom]), 1)
    achieved = fl / (avg_ms * 1e-3) / 1e12
    peak = peaks.get("fp64_dmma_m8n8k4_tflops")
    ncu = {}
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_summary.json")) as f:
            ncu = json.load(f).get(args.workload, {})
    except Exception:
        pass
    other = 1 - dom
    avg_other = kms[other] / max(int(kn[other]), 1)
    ach_other = [flops_prop, flops_stab][other] / (avg_other * 1e-3) / 1e12
    step_flops = (flops_prop + flops_stab) * 2 * ((nt + n_stab - 1) // n_stab) * world
    roofline = {
        "bound": "tensor", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
        "frac": (achieved / peak) if peak else None,
        "traffic": ncu.get(["propagate", "stabilize"][dom], {}).get("dram_bytes_per_launch"),
        "peak_source": "FP64 DMMA m8n8k4 / DFMA peak measured on this pool's B200 by benchmarks/peaks.cu (profiles/fp64_smem_peaks.json); "
                       "MEASURED_PEAKS.json carries no FP64 figure",
        "avg_launch_ms": avg_ms, "share_of_step": float(kms[dom] / max(kms[:2].sum(), 1e-9)),
        "algorithmic_flops_per_launch": fl,
        "class_ms": {"propagate": float(kms[0]), "stabilize": float(kms[1])},
        "other_kernel": {"kernel": ["propagate (wrap + Metropolis)", "stabilize (Bbar + LDR stack + G)"][other], "achieved": ach_other,
                         "frac": (ach_other / peak) if peak else None, "avg_launch_ms": avg_other},
        "whole_step": {"algorithmic_tflops": step_flops / (ms_max / args.steps * 1e-3) / 1e12,
                       "frac": (step_flops / (ms_max / args.steps * 1e-3) / 1e12 / (peak * world)) if peak else None},
        "hbm_note": "stabilize moves 3 LDR stack entries + G per walker per launch (algorithmic %.0f MB); see traffic" % (W * (3 * (2 * N * N + N) + 2 * N * N) * 8 / 1e6),
    }

    cpu_baseline = None
    if not args.no_cpu_baseline:
        try:
            cores, phys = host_cores()
            ns = {36: 160, 64: 40, 144: 4, 256: 1}.get(N, 4)  # ~15-30 CPU-seconds over all host threads
            v, sps, secs = cpu_reference_run(wl, cores, ns, 1)
            cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "physical_cores": phys, "kind": "reference", "sweeps_per_sec": sps,
                            "sample": f"{cores} independent chains (one per hardware thread; {phys} physical cores) x {ns} full sweeps of the same model "
                                      f"from a random start (a bounded sample, not {W} walkers), reference sources compiled unmodified (oracle/_ref), -O3, "
                                      f"OpenBLAS (MKL absent) 1 thread/chain, {secs:.1f} s wall"}
            # the same sample with the reference's own Release flags (-O3 -ffast-math, CMakeLists.txt:7): separate process,
            # because the variant is chosen when oracle/ref.py is imported
            try:
                env = dict(os.environ, DQMCB_REF_VARIANT="fast")
                out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", args.workload, "--steps", "1",
                                      "--warmup", "1"] + (["--n-stab", str(args.n_stab)] if args.n_stab else []),
                                     capture_output=True, text=True, timeout=900, env=env).stdout.strip().splitlines()
                fl_ = json.loads(out[-1]) if out else {}
                cpu_baseline["value_fast_math"] = fl_.get("value")
            except Exception as ex:
                cpu_baseline["value_fast_math"] = None
                cpu_baseline["fast_math_note"] = f"unavailable: {ex}"
        except Exception as ex:  # the oracle always exists in-tree; report loudly if the binary is missing
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {ex}"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_desc(args.workload, wl, world),
        "sweeps_per_sec": sweeps_per_sec, "acceptance": a, "max_wrap_err": float(emax.max()), "mean_wrap_err": float(emean.mean()),
        "thermalisation": {"sweeps": args.therm, "seconds": t_th, "note": "untimed; acceptance and wrap errors are those of the timed sweeps after it"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * world, "d2h_bytes_per_step": int(d2h) * world,
                "path": "dqmcb_set_tape_async/dqmcb_tape_commit (pinned host RNG tape of the next step uploaded on a copy stream during the current step) + dqmcb_run_sweeps_host (fwd+bwd sweep, measure, fields+observables D2H)"},
        "gpu_launches": int(gpu_launches), "walker_groups": ctx.n_groups(), "delay_k": delay_k,
        "clocks": clocks.summary(),
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
    }
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
