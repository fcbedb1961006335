"""Phase-level clock64 profile of one generic-engine stabilisation + propagate launch at N = 144 / 256 (development aid).
usage: DQMCB_PHASE_PROF=1 python benchmarks/phase_prof_big.py L W [delay_k]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DQMCB_PHASE_PROF", "1")
from dqmc_b200 import host, capi
L = int(sys.argv[1]); W = int(sys.argv[2]); dk = int(sys.argv[3]) if len(sys.argv) > 3 else 32
U = 6.0 if L == 16 else 4.0
m = host.AttractiveHubbard(L, L, U, 1.0, 0.0, 4.0, 40, 10, W, delay_k=dk)
m.random_fields(1)
c = m.ctx
c.rng_mode(capi.RNG_PHILOX, 42)
c.init_stacks(); c.init_greenfunctions()
c.timing_enable(True)
c.sweep_0_to_beta(); c.sweep_beta_to_0()
c.synchronize()
ms, n = c.timing_get()
print(f"L={L} W={W} dk={dk}: propagate {ms[0]/n[0]:.3f} ms/launch, stabilize {ms[1]/n[1]:.3f} ms/launch")
c.timing_enable(False)
c.stabilize_forward(19)
c.synchronize()
c.close()
