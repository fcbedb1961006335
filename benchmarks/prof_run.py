"""Short run used under ncu: a few block steps of the sweep at C1 with W walkers."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dqmc_b200 import host, capi
W = int(sys.argv[1]) if len(sys.argv) > 1 else 296
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
beta = float(sys.argv[3]) if len(sys.argv) > 3 else 8.0
nsw = int(sys.argv[4]) if len(sys.argv) > 4 else 1
m = host.AttractiveHubbard(L, L, 4.0, 1.0, 0.0, beta, int(round(beta / 0.1)), 10, W)
m.random_fields(1)
c = m.ctx
c.rng_mode(capi.RNG_PHILOX, 42)
c.init_stacks(); c.init_greenfunctions()
for _ in range(nsw):
    c.sweep_0_to_beta(); c.sweep_beta_to_0()
c.synchronize()
print("ok", c.get_stats()[0].mean())
