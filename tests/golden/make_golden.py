"""tests/golden/make_golden.py -- generates the golden fixtures tests/golden/*.npz.

The reference (kfkq/DQMC) ships no tests, golden vectors or fixtures, so the
vectors are OUTPUTS OF THE REFERENCE ITSELF: its own source/{model,dqmc,update,stablelinalg}.cpp
compiled unmodified into oracle/_ref/libdqmc_ref.so (oracle/Makefile) and driven through
oracle/ref.py with an injected seed.  Run in the build container (needs /root/reference):

    python tests/golden/make_golden.py

The fixtures are small and committed; /root/reference is not needed to USE them.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref  # noqa: E402

CASES = {
    # name: (L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps)
    "c0": (6, 6, 4.0, 1.0, -0.1, 4.0, 40, 10, 4242, 2),   # examples/parameters.in
    "r7": (7, 5, 3.0, 1.0, 0.2, 3.0, 23, 6, 777, 1),       # odd N, ragged last block
    "c1s": (8, 8, 4.0, 1.0, 0.0, 2.0, 20, 5, 99, 1),       # 8x8 (N = 64) short
}


def silence_stderr():
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(2)
    os.dup2(devnull, 2)
    return saved, devnull


def restore_stderr(tok):
    saved, devnull = tok
    os.dup2(saved, 2)
    os.close(devnull)
    os.close(saved)


def trajectory_case(name, cfg):
    L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps = cfg
    s = ref.RefSim(L1, L2, U, t, mu, beta, nt, n_stab, seed)
    n = s.ns
    out = {
        "cfg": np.array([L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps], dtype=np.float64),
        "g": np.array(s.g),
        "fields0": s.fields().astype(np.int8),
        "expK": s.expK(), "invexpK": s.invexpK(), "expK_half": s.expK_half(), "invexpK_half": s.invexpK_half(),
        "G0": s.G(0), "logdet0": np.array(s.logdet()),
        "stack0_prod": np.stack([(lambda F: (F[0] * F[1]) @ F[2])(s.stack(i)) for i in range(s.n_stack)]),
        "stack0_d": np.stack([s.stack(i)[1] for i in range(s.n_stack)]),
        "Bbar0": np.stack([s.Bbar(i) for i in range(s.n_stack)]),
        "action0": np.array(s.global_action()),
        "half_wrap0": s.half_wrap(),
    }
    o = s.observables()
    out["obs0"] = np.array(o[:3])
    out["corr0"] = o[3]
    # one teacher-forced forward wrap and one slice of Metropolis, step by step, on a scratch sim
    s2 = ref.RefSim(L1, L2, U, t, mu, beta, nt, n_stab, seed)
    s2.wrap_fwd(0)
    out["G_wrap_fwd0"] = s2.G(1)
    perm, k, u = s2.rng_tape(1, advance=False)
    f = s2.fields()
    prop = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]])
    Rs, Ds, acc = [], [], []
    for idx in range(n):
        i = int(perm[0, idx])
        new = int(prop[f[0, i], k[0, idx]])
        R, dl = s2.local_update_ratio(0, i, new)
        a = bool(u[0, idx] < min(1.0, abs(R)))
        Rs.append(R); Ds.append(dl); acc.append(a)
        if a:
            s2.update_greens_local(dl, 0, i)
            f[0, i] = new
            s2.set_fields(f, reinit=False)
    out["slice0_R"] = np.array(Rs); out["slice0_delta"] = np.array(Ds); out["slice0_accept"] = np.array(acc)
    out["slice0_G"] = s2.G(1)
    out["slice0_fields"] = f[0].astype(np.int8)
    # the reference's own local_update on a third sim must reproduce the manual walk
    s3 = ref.RefSim(L1, L2, U, t, mu, beta, nt, n_stab, seed)
    s3.wrap_fwd(0)
    s3.local_update(0)
    assert np.array_equal(s3.fields()[0], f[0]), "manual walk != update::local_update"
    assert np.max(np.abs(s3.G(1) - out["slice0_G"])) < 1e-13
    # backward wrap of the same G
    s3.wrap_bwd(0)
    out["slice0_G_wrap_bwd"] = s3.G(0)
    # free-running sweeps with the reference's own RNG stream
    tp, tk, tu = s.rng_tape(2 * nt * n_sweeps, advance=False)
    out["tape_perm"] = tp.astype(np.uint8); out["tape_k"] = tk.astype(np.uint8); out["tape_u"] = tu
    Gs, fs, lds = [], [], []
    tok = silence_stderr()
    try:
        for _ in range(n_sweeps):
            s.sweep_fwd()
            Gs.append(s.G(nt)); fs.append(s.fields().astype(np.int8)); lds.append(s.logdet())
            s.sweep_bwd()
            Gs.append(s.G(0)); fs.append(s.fields().astype(np.int8)); lds.append(s.logdet())
    finally:
        restore_stderr(tok)
    out["traj_G"] = np.stack(Gs); out["traj_fields"] = np.stack(fs); out["traj_logdet"] = np.array(lds)
    out["traj_acc_rate"] = np.array(s.acc_rate())
    out["traj_max_err"] = np.array(s.max_err()); out["traj_mean_err"] = np.array(s.mean_err())
    out["traj_action"] = np.array(s.global_action())
    o = s.observables()
    out["traj_obs"] = np.array(o[:3]); out["traj_corr"] = o[3]
    out["traj_stack_prod"] = np.stack([(lambda F: (F[0] * F[1]) @ F[2])(s.stack(i)) for i in range(s.n_stack)])
    np.savez_compressed(os.path.join(HERE, f"traj_{name}.npz"), **out)
    print(name, "acc", float(out["traj_acc_rate"]), "max_err", float(out["traj_max_err"]))


def graded(rng, n, span):
    Q1, _ = np.linalg.qr(rng.standard_normal((n, n)))
    Q2, _ = np.linalg.qr(rng.standard_normal((n, n)))
    return (Q1 * np.exp(np.linspace(span, -span, n))) @ Q2


def stablelinalg_case():
    rng = np.random.default_rng(20261104)
    out = {}
    for n, span in ((12, 5.0), (36, 8.0)):
        M1, M2 = graded(rng, n, span), graded(rng, n, span)
        F1, F2 = ref.to_LDR(M1), ref.to_LDR(M2)
        pre = f"n{n}_"
        out[pre + "M1"], out[pre + "M2"] = M1, M2
        for nm, F in (("F1", F1), ("F2", F2)):
            out[pre + nm + "_L"], out[pre + nm + "_d"], out[pre + nm + "_R"] = F
        for nm, F in (("mat_mul_ldr", ref.mat_mul_ldr(M2, F1)), ("ldr_mul_mat", ref.ldr_mul_mat(F1, M2)),
                      ("ldr_mul_ldr", ref.ldr_mul_ldr(F1, F2))):
            out[pre + nm + "_L"], out[pre + nm + "_d"], out[pre + nm + "_R"] = F
        G, ld = ref.inv_I_plus_ldr(F1)
        out[pre + "inv1_G"], out[pre + "inv1_logdet"] = G, np.array(ld)
        out[pre + "inv2_G"] = ref.inv_I_plus_ldr_mul_ldr(F1, F2)
        out[pre + "inv3_G"] = ref.inv_invldr_plus_ldr(F1, F2)
    np.savez_compressed(os.path.join(HERE, "stablelinalg.npz"), **out)


def misc_case():
    out = {"mt19937_seed5489": ref.mt19937_raw(5489, 16), "mt19937_seed1": ref.mt19937_raw(1, 16)}
    tab = np.array([[[ref.partner_rank(r, ws, a) for a in range(1, 5)] for r in range(8)] for ws in (2, 4, 8)])
    out["partner_rank_ws248_r8_a1to4"] = tab
    np.savez_compressed(os.path.join(HERE, "misc.npz"), **out)


def exchange_case():
    """update::replica_exchange (source/update.cpp:47-117) on a 4-replica beta ladder, two attempts."""
    betas = [2.0, 1.6, 1.2, 0.8]
    L1, L2, U, t, mu, nt, n_stab = 4, 4, 4.0, 1.0, 0.0, 16, 4
    sims = [ref.RefSim(L1, L2, U, t, mu, b, nt, n_stab, 300 + r) for r, b in enumerate(betas)]
    out = {"betas": np.array(betas), "cfg": np.array([L1, L2, U, t, mu, nt, n_stab], dtype=np.float64),
           "seeds": np.array([300 + r for r in range(4)])}
    tok = silence_stderr()
    try:
        for s in sims:
            s.sweep_fwd(); s.sweep_bwd()
        out["fields_before"] = np.stack([s.fields().astype(np.int8) for s in sims])
        out["action_before"] = np.array([s.global_action() for s in sims])
        # canonical double each rank's generator would produce next (decider's bernoulli draw)
        for a in range(2):
            out[f"u_next_{a}"] = np.array([s.peek_canonical() for s in sims])
            ref.replica_exchange(sims)
            out[f"fields_after_{a}"] = np.stack([s.fields().astype(np.int8) for s in sims])
            out[f"logdet_after_{a}"] = np.array([s.logdet() for s in sims])
            out[f"G_after_{a}"] = np.stack([s.G(0) for s in sims])
    finally:
        restore_stderr(tok)
    np.savez_compressed(os.path.join(HERE, "exchange.npz"), **out)


def measure_case():
    """The reference's measurement pipeline (include/measurementh5.h) on G's of a real trajectory:
    MeasurementManager wired as source/main.cpp:114-119, 2 bins x 4 measurements on the 6x6 lattice,
    datasets as the HDF5 recorder shim saw them; stand-alone transforms on random cubes (square 6x6,
    odd 5x5 and, for the fold only, 4x6 -- chi_r_to_chi_k of the reference indexes out of bounds there)."""
    out = {}
    L1 = L2 = 6
    sim = ref.RefSim(L1, L2, 4.0, 1.0, -0.1, 2.0, 20, 5, 77)
    mm = ref.RefMeasure(L1, L2, rank=2)
    Gs = []
    for b in range(2):
        for _ in range(4):
            sim.sweep_fwd()
            sim.sweep_bwd()
            G = sim.G(0)
            Gs.append(G)
            mm.measure(G)
        mm.accumulate()
    out["G"] = np.stack(Gs)
    out["file_name"] = np.array(mm.file_name())
    ds = mm.datasets()
    out["paths"] = np.array(list(ds.keys()))
    for k, v in ds.items():
        out["ds:" + k] = v
    rng = np.random.default_rng(5)
    for (a, b) in ((6, 6), (5, 5), (4, 6)):
        m2 = ref.RefMeasure(a, b)
        c = rng.standard_normal((3, a * b, a * b))
        r = m2.chi_site_to_chi_r(c)
        out[f"tf_{a}x{b}_site"] = c
        out[f"tf_{a}x{b}_r"] = r
        if a == b:
            k = m2.chi_r_to_chi_k(r)
            out[f"tf_{a}x{b}_k"] = k
        m2.close()
    out["pbc_shortest_L4567"] = np.array([[mm.pbc_shortest(d, L) if abs(d) < L else 0 for d in range(-6, 7)] for L in (4, 5, 6, 7)])
    mm.close()
    np.savez_compressed(os.path.join(HERE, "measure.npz"), **out)


def unequal_case():
    """DQMC::sweep_unequalTime of the reference (isMeasureUnequalTime = true) after one full sweep on 4x4 and a
    ragged 5x3 / nt=11 / n_stab=4 case: fields, the tape that produced them, every Gtt/Gt0/G0t, the three
    tau observables, the half-warped arrays, and the MeasurementManager datasets with unequal-time on."""
    out = {}
    for name, cfg in (("a", (4, 4, 4.0, 1.0, -0.1, 2.0, 20, 5, 7)), ("b", (5, 3, 6.0, 1.0, 0.3, 2.0, 11, 4, 32))):
        L1, L2, U, t, mu, beta, nt, n_stab, seed = cfg
        s = ref.RefSim(*cfg, unequal=True)
        out[f"{name}_cfg"] = np.array(cfg, dtype=np.float64)
        out[f"{name}_fields0"] = s.fields().astype(np.int8)
        perm, k, u = s.rng_tape(2 * nt)
        out[f"{name}_perm"], out[f"{name}_k"], out[f"{name}_u"] = perm.astype(np.uint8), k.astype(np.uint8), u
        tok = silence_stderr()
        try:
            s.sweep_fwd()
            s.sweep_bwd()
            s.sweep_unequal()
        finally:
            restore_stderr(tok)
        out[f"{name}_fields"] = s.fields().astype(np.int8)
        for wh, nm in enumerate(("Gtt", "Gt0", "G0t")):
            out[f"{name}_{nm}"] = np.stack([s.G_unequal(wh, l) for l in range(nt + 1)])
        g, d, c = s.observables_tau()
        out[f"{name}_greenTau"], out[f"{name}_doublonTau"], out[f"{name}_currxxTau"] = g, d, c
        out[f"{name}_half"] = s.half_wrap_unequal()
        out[f"{name}_logdet"] = np.array(s.logdet())
        if L1 == L2:  # MeasurementManager with unequal-time on (k-space of the reference needs a square lattice)
            mm = ref.RefMeasure(L1, L2, rank=0, unequal=True)
            mm.measure_unequal(out[f"{name}_Gtt"], out[f"{name}_Gt0"], out[f"{name}_G0t"])
            mm.accumulate()
            ds = mm.datasets()
            out[f"{name}_paths"] = np.array(list(ds.keys()))
            for k2, v in ds.items():
                out[f"{name}_ds:" + k2] = v
            mm.close()
    np.savez_compressed(os.path.join(HERE, "unequal.npz"), **out)


PROGRAM_CFG = dict(L1=4, L2=4, U=4.0, t=1.0, mu=-0.1, beta=1.0, nt=10, n_stab=5, seed=11, n_therms=3, n_sweeps=2, n_bins=3)


def program_case():
    """The whole program: the loop of source/main.cpp:126-171 (thermalise, then per sweep sweep_0_to_beta,
    sweep_beta_to_0, sweep_unequalTime, measure; accumulate per bin) run by the reference's own classes for
    utility::random rng(seed), isMeasureUnequalTime = true.  The datasets of results/data_0.h5 are the fixture
    `python -m dqmc_b200.run --rng mt19937 --walkers 1 --seed 11` must reproduce (tests/test_gpu_driver.py)."""
    c = PROGRAM_CFG
    nt = c["nt"]
    s = ref.RefSim(c["L1"], c["L2"], c["U"], c["t"], c["mu"], c["beta"], nt, c["n_stab"], c["seed"], unequal=True)
    mm = ref.RefMeasure(c["L1"], c["L2"], rank=0, unequal=True)
    out = {"cfg_keys": np.array(list(c.keys())), "cfg": np.array(list(c.values()), dtype=np.float64),
           "fields0": s.fields().astype(np.int8)}
    tok = silence_stderr()
    try:
        for _ in range(c["n_therms"]):
            s.sweep_fwd()
            s.sweep_bwd()
        for isweep in range(1, c["n_bins"] * c["n_sweeps"] + 1):
            s.sweep_fwd()
            s.sweep_bwd()
            s.sweep_unequal()
            mm.measure_unequal(*[np.stack([s.G_unequal(wh, l) for l in range(nt + 1)]) for wh in range(3)])
            if isweep % c["n_sweeps"] == 0:
                mm.accumulate()
    finally:
        restore_stderr(tok)
    out["fields"] = s.fields().astype(np.int8)
    out["acc_rate"] = np.array(s.acc_rate())
    for k, v in mm.datasets().items():
        out["ds:" + k] = v
    mm.close()
    np.savez_compressed(os.path.join(HERE, "program.npz"), **out)


if __name__ == "__main__" and not (len(sys.argv) > 1 and sys.argv[1] == "full"):
    assert ref.build(), "oracle/_ref could not be built (needs /root/reference)"
    for nm, cfg in CASES.items():
        trajectory_case(nm, cfg)
    stablelinalg_case()
    misc_case()
    exchange_case()
    measure_case()
    unequal_case()
    program_case()
    print("golden fixtures written to", HERE)


# ---------------------------------------------------------------------------------------------------
# Full-size BASELINE configurations: free-running trajectories of the reference at the
# stated beta / nt.  The RNG stream is NOT stored: it is std::mt19937(seed), which the device regenerates
# (DQMCB_RNG_MT19937) and oracle/dqmc_oracle.py restates.  Written to traj_full_<name>.npz by
#     python tests/golden/make_golden.py full
# ---------------------------------------------------------------------------------------------------
FULL_CASES = {
    # name: (L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps, keep_G = number of trailing half sweeps whose G is stored)
    "c1": (8, 8, 4.0, 1.0, 0.0, 8.0, 80, 10, 8101, 10, 20),
    "c2": (12, 12, 4.0, 1.0, 0.0, 10.0, 100, 10, 12101, 2, 4),
    "c2n5": (12, 12, 4.0, 1.0, 0.0, 10.0, 100, 5, 12105, 2, 2),
    "c3n5": (16, 16, 6.0, 1.0, 0.0, 16.0, 160, 5, 16105, 2, 2),
    # n_stab = 10 at U = 6, beta = 16: the reference's own check_error reaches O(1) within one sweep, so a
    # trajectory is meaningless there; only the start-up state (a pure function of the fields) is pinned
    "c3": (16, 16, 6.0, 1.0, 0.0, 16.0, 160, 10, 16110, 0, 1),
}


def full_case(name, cfg):
    L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps, keep = cfg
    s = ref.RefSim(L1, L2, U, t, mu, beta, nt, n_stab, seed)
    out = {
        "cfg": np.array([L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps], dtype=np.float64),
        "fields0": s.fields().astype(np.int8),
        "expK": s.expK(), "invexpK": s.invexpK(),
        "G0_diag": np.diag(s.G(0)).copy(), "logdet0": np.array(s.logdet()),
        "stack0_d": np.stack([np.sort(s.stack(i)[1]) for i in range(s.n_stack)]),
    }
    if keep >= 2 * n_sweeps:
        out["G0"] = s.G(0)
    if n_sweeps == 0:
        np.savez_compressed(os.path.join(HERE, f"traj_full_{name}.npz"), **out)
        print(name, "init only, bytes", os.path.getsize(os.path.join(HERE, f"traj_full_{name}.npz")))
        return
    fs, lds, Gs, Gd, errs = [], [], [], [], []
    tok = silence_stderr()
    try:
        for _ in range(n_sweeps):
            for half in range(2):
                (s.sweep_fwd if half == 0 else s.sweep_bwd)()
                G = s.G(nt if half == 0 else 0)
                fs.append(s.fields().astype(np.int8)); lds.append(s.logdet()); Gs.append(G); Gd.append(np.diag(G).copy())
                errs.append((s.max_err(), s.mean_err()))
    finally:
        restore_stderr(tok)
    out["traj_fields"] = np.stack(fs); out["traj_logdet"] = np.array(lds); out["traj_G_diag"] = np.stack(Gd)
    out["traj_G_tail"] = np.stack(Gs[-keep:])  # G at the end of the last `keep` half sweeps
    out["traj_err"] = np.array(errs)             # (running max, running mean) of check_error after each half sweep
    out["traj_acc_rate"] = np.array(s.acc_rate())
    out["stack_d"] = np.stack([np.sort(s.stack(i)[1]) for i in range(s.n_stack)])
    out["traj_action"] = np.array(s.global_action())
    o = s.observables()
    out["traj_obs"] = np.array(o[:3])
    np.savez_compressed(os.path.join(HERE, f"traj_full_{name}.npz"), **out)
    print(name, "acc", float(out["traj_acc_rate"]), "check_error max/mean", errs[-1], "bytes",
          os.path.getsize(os.path.join(HERE, f"traj_full_{name}.npz")))


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "full":
    for nm, cfg in FULL_CASES.items():
        if len(sys.argv) > 2 and nm not in sys.argv[2:]:
            continue
        full_case(nm, cfg)
