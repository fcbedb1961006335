"""Development aid: every configuration is run several times (optionally under host contention and with
different walker-group counts); fields and G must be bit-identical between runs.  A silent intra-CTA race
shows up here as a run-to-run difference long before it produces an error."""
import os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqmc_b200 import host, capi

CASES = [  # name, L1, L2, U, beta, nt, n_stab, W, delay_k, sweeps
    ("C0 generic N=36", 6, 6, 4.0, 4.0, 40, 10, 2048, 0, 6),
    ("F32 fast N=32", 8, 4, 5.0, 2.3, 23, 5, 2048, 0, 8),
    ("C1 fast N=64", 8, 8, 4.0, 8.0, 80, 10, 1024, 0, 6),
    ("R7 generic N=35", 7, 5, 3.0, 3.0, 23, 6, 1500, 0, 6),
    ("C2 generic N=144 k=32", 12, 12, 4.0, 2.0, 20, 10, 296, 32, 3),
    ("C2 generic N=144 k=8", 12, 12, 4.0, 2.0, 20, 10, 296, 8, 3),
    ("C2 generic N=144 k=16", 12, 12, 4.0, 2.0, 20, 10, 296, 16, 3),
    ("C3 generic N=256 k=32", 16, 16, 4.0, 1.0, 10, 5, 148, 32, 2),
    ("C3 generic N=256 k=32, 1.5 waves", 16, 16, 6.0, 1.0, 10, 5, 222, 32, 2),
    ("10x6 generic N=60 (run-time Kronecker)", 10, 6, 4.0, 1.5, 15, 5, 600, 0, 4),
]

def run(case, streams):
    name, L1, L2, U, beta, nt, n_stab, W, dk, nsw = case
    os.environ["DQMCB_STREAMS"] = str(streams)
    m = host.AttractiveHubbard(L1, L2, U, 1.0, 0.0, beta, nt, n_stab, W, delay_k=dk)
    m.random_fields(3)
    c = m.ctx
    c.rng_mode(capi.RNG_PHILOX, 17)
    c.init_stacks(); c.init_greenfunctions()
    for _ in range(nsw):
        c.sweep_0_to_beta(); c.sweep_beta_to_0()
    c.sweep_unequal_time() if W <= 300 and L1 * L2 <= 64 else None
    out = (c.get_fields(), c.get_G(), c.get_stats()[1].max())
    c.close()
    return out

burn = int(sys.argv[1]) if len(sys.argv) > 1 else 0
burners = [subprocess.Popen([sys.executable, "-c", "import time\nt=time.time()\nwhile time.time()-t<150: pass"]) for _ in range(burn)]
bad = 0
try:
    for case in CASES:
        ref = run(case, 1)
        for streams in (1, 2, 3, 2):
            got = run(case, streams)
            same = np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])
            bad += not same
            print(f"{case[0]:26s} streams={streams} {'identical' if same else 'DIFFERENT'} maxerr {got[2]:.2e}", flush=True)
finally:
    for b in burners:
        b.kill()
print("FAILED" if bad else "ALL IDENTICAL")
