"""Quick device timing of the sweep at a named config (development aid; bench.py is the contract)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqmc_b200 import host, capi
import ctypes

def main():
    L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    beta = float(sys.argv[2]) if len(sys.argv) > 2 else 8.0
    W = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    U = float(sys.argv[4]) if len(sys.argv) > 4 else 4.0
    nsw = int(sys.argv[5]) if len(sys.argv) > 5 else 2
    nt = int(round(beta / 0.1))
    m = host.AttractiveHubbard(L, L, U, 1.0, 0.0, beta, nt, 10, W)
    m.random_fields(1)
    c = m.ctx
    c.rng_mode(capi.RNG_PHILOX, 42)
    t0 = time.time(); c.init_stacks(); c.synchronize(); t1 = time.time()
    c.init_greenfunctions(); c.synchronize(); t2 = time.time()
    print(f"init_stacks {t1-t0:.4f}s init_green {t2-t1:.4f}s")
    for s in range(nsw):
        t0 = time.time(); c.sweep_0_to_beta(); c.synchronize(); t1 = time.time()
        c.sweep_beta_to_0(); c.synchronize(); t2 = time.time()
        acc, emax, emean = c.get_stats()
        print(f"sweep {s}: fwd {t1-t0:.4f}s bwd {t2-t1:.4f}s  acc {acc.mean()/(2*(s+1)):.3f} maxerr {emax.max():.2e} meanerr {emean.mean():.2e}")
    N = L * L
    print(f"sweeps/s {W/(t2-t0):.1f} flips/s {W*2*nt*N/(t2-t0):.3e}")
main()
