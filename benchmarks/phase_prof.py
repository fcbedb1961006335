"""Phase-level clock64 profile of one stabilisation launch (development aid).
usage: DQMCB_PHASE_PROF=1 python benchmarks/phase_prof.py [W] [L]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DQMCB_PHASE_PROF", "1")
from dqmc_b200 import host, capi
W = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
m = host.AttractiveHubbard(L, L, 4.0, 1.0, 0.0, 4.0, 40, 10, W)
m.random_fields(1)
c = m.ctx
c.rng_mode(capi.RNG_PHILOX, 42)
c.init_stacks(); c.init_greenfunctions()
c.sweep_0_to_beta(); c.sweep_beta_to_0()
c.stabilize_forward(19)   # mid-sweep stabilisation: mat_mul_ldr + inv_I_plus_ldr_mul_ldr
c.synchronize()
c.close()
