"""Development aid: second-generation stabilisation kernel (stab2.cuh) against the first (DQMCB_STAB2=0) block by
block on identical inputs, then the per-launch time of both.  usage: stab2_check.py [W] [beta] [n_stab]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqmc_b200 import host, capi

W = int(sys.argv[1]) if len(sys.argv) > 1 else 4
beta = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
n_stab = int(sys.argv[3]) if len(sys.argv) > 3 else 5
nt = int(round(beta / 0.1))


def make(stab2):
    if stab2:
        os.environ.pop("DQMCB_STAB2", None)
    else:
        os.environ["DQMCB_STAB2"] = "0"
    os.environ["DQMCB_STREAMS"] = "1"
    m = host.AttractiveHubbard(8, 8, 4.0, 1.0, 0.0, beta, nt, n_stab, W)
    m.random_fields(3)
    c = m.ctx
    c.rng_mode(capi.RNG_PHILOX, 42)
    c.init_stacks(); c.init_greenfunctions()
    return m, c


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def compare(tag, ca, cb):
    Ga, Gb = ca.get_G(), cb.get_G()
    out = [f"G {rel(Ga, Gb):.1e}"]
    worst_p = worst_d = 0.0
    for i in range(ca.n_stack):
        La, da, Ra = ca.get_ldr(i); Lb, db, Rb = cb.get_ldr(i)
        for w in range(W):
            pa = (La[w] * da[w]) @ Ra[w]; pb = (Lb[w] * db[w]) @ Rb[w]
            worst_p = max(worst_p, rel(pa, pb))
            worst_d = max(worst_d, float(np.max(np.abs(np.sort(da[w]) / np.sort(db[w]) - 1))))
    out.append(f"stack prod {worst_p:.1e} d {worst_d:.1e}")
    out.append(f"logdet {np.max(np.abs(ca.get_logdet() - cb.get_logdet())):.1e}")
    out.append(f"fields {'same' if np.array_equal(ca.get_fields(), cb.get_fields()) else 'DIFF'}")
    print(tag, "; ".join(out), flush=True)


ma, ca = make(True)
mb, cb = make(False)
compare("init", ca, cb)
n_stack = ca.n_stack
for i in range(n_stack):
    l0, l1 = i * n_stab, min((i + 1) * n_stab, nt)
    for c in (ca, cb):
        for l in range(l0, l1):
            c.wrap_forward(l); c.local_update(l)
        c.stabilize_forward(l1 - 1)
        c.synchronize()
    compare(f"fwd block {i}", ca, cb)
for i in range(n_stack - 1, -1, -1):
    l0, l1 = i * n_stab, min((i + 1) * n_stab, nt)
    for c in (ca, cb):
        for l in range(l1 - 1, l0 - 1, -1):
            c.local_update(l); c.wrap_backward(l)
        c.stabilize_backward(l0)
        c.synchronize()
    compare(f"bwd block {i}", ca, cb)
for c in (ca, cb):
    acc, emax, emean = c.get_stats()
    print("check_error max", float(emax.max()), "mean", float(emean.mean()))
# timing of whole sweeps, per class
for name, c in (("stab2", ca), ("stab1", cb)):
    c.timing_enable(True)
    for _ in range(2):
        c.sweep_0_to_beta(); c.sweep_beta_to_0()
    c.synchronize()
    ms, n = c.timing_get()
    print(f"{name}: propagate {ms[0]/max(n[0],1):.4f} ms x {n[0]}, stabilize {ms[1]/max(n[1],1):.4f} ms x {n[1]}", flush=True)
    c.timing_enable(False)
