"""Quick device timing of the sweep at a named config (development aid; bench.py is the contract).
usage: quick_time.py L beta W U n_sweeps [delay_k] [n_stab]"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqmc_b200 import host, capi

def main():
    a = sys.argv[1:]
    L = int(a[0]) if len(a) > 0 else 8
    beta = float(a[1]) if len(a) > 1 else 8.0
    W = int(a[2]) if len(a) > 2 else 1024
    U = float(a[3]) if len(a) > 3 else 4.0
    nsw = int(a[4]) if len(a) > 4 else 2
    dk = int(a[5]) if len(a) > 5 else 0
    n_stab = int(a[6]) if len(a) > 6 else 10
    nt = int(round(beta / 0.1))
    m = host.AttractiveHubbard(L, L, U, 1.0, 0.0, beta, nt, n_stab, W, delay_k=dk)
    m.random_fields(1)
    c = m.ctx
    c.rng_mode(capi.RNG_PHILOX, 42)
    t0 = time.time(); c.init_stacks(); c.synchronize(); t1 = time.time()
    c.init_greenfunctions(); c.synchronize(); t2 = time.time()
    print(f"init_stacks {t1-t0:.4f}s init_green {t2-t1:.4f}s")
    c.timing_enable(True)
    for s in range(nsw):
        t0 = time.time(); c.sweep_0_to_beta(); c.synchronize(); t1 = time.time()
        c.sweep_beta_to_0(); c.synchronize(); t2 = time.time()
        acc, emax, emean = c.get_stats()
        print(f"sweep {s}: fwd {t1-t0:.4f}s bwd {t2-t1:.4f}s  acc {acc.mean()/(2*(s+1)):.3f} maxerr {emax.max():.2e} meanerr {emean.mean():.2e}")
    ms, n = c.timing_get()
    N = L * L
    print(f"class ms: propagate {ms[0]:.1f} ({n[0]} launches) stabilize {ms[1]:.1f} ({n[1]})")
    print(f"L={L} W={W} delay_k={dk}: sweeps/s {W/(t2-t0):.1f} flips/s {W*2*nt*N/(t2-t0):.3e}")
main()
