"""Stress: tape-mode end-to-end steps (dqmcb_set_tape + dqmcb_run_sweeps_host) with walker groups on streams,
compared bit-for-bit with the single-stream run (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dqmc_b200 import host, capi

def run(ng, steps, W=1024):
    os.environ["DQMCB_STREAMS"] = str(ng)
    m = host.AttractiveHubbard(8, 8, 4.0, 1.0, 0.0, 8.0, 80, 10, W)
    m.random_fields(1)
    c = m.ctx
    c.rng_mode(capi.RNG_PHILOX, 42)
    c.init_stacks(); c.init_greenfunctions()
    for _ in range(3):
        c.sweep_0_to_beta(); c.sweep_beta_to_0()
    gen = np.random.default_rng(7)
    N, visits = 64, 160
    out = []
    for s in range(steps):
        perm = gen.permuted(np.tile(np.arange(N, dtype=np.uint8), (W, visits, 1)), axis=-1)
        k = gen.integers(0, 3, size=(W, visits, N), dtype=np.uint8)
        u = gen.random((W, visits, N))
        c.set_tape(perm, k, u)
        out.append(c.run_sweeps_host(None, 1))
    c.close()
    return out

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
ref = run(1, steps)
for ng in (2, 3, 4):
    got = run(ng, steps)
    ok = all(all(np.array_equal(a, b) for a, b in zip(x, y) if a is not None) for x, y in zip(ref, got))
    print("groups", ng, "identical" if ok else "DIFFERENT")
