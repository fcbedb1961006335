"""Short run used under ncu for the large lattices: one forward half sweep at dtau = 0.1, nt = 20 (2 blocks of 10 slices) unless given."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dqmc_b200 import host, capi
L = int(sys.argv[1]); W = int(sys.argv[2]); dk = int(sys.argv[3]); nt = int(sys.argv[4]) if len(sys.argv) > 4 else 20
U = 6.0 if L == 16 else 4.0
m = host.AttractiveHubbard(L, L, U, 1.0, 0.0, 0.1 * nt, nt, 10, W, delay_k=dk)
m.random_fields(1)
c = m.ctx
c.rng_mode(capi.RNG_PHILOX, 42)
c.init_stacks(); c.init_greenfunctions()
c.sweep_0_to_beta()
c.synchronize()
print("ok", c.get_stats()[0].mean())
