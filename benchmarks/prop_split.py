"""Development aid: wrap vs Metropolis time of the propagate kernel, one slice per launch.
usage: prop_split.py L W [delay_k] [beta]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dqmc_b200 import host, capi
L = int(sys.argv[1]); W = int(sys.argv[2]); dk = int(sys.argv[3]) if len(sys.argv) > 3 else 0
beta = float(sys.argv[4]) if len(sys.argv) > 4 else 4.0
U = 6.0 if L == 16 else 4.0
nt = int(round(beta / 0.1))
m = host.AttractiveHubbard(L, L, U, 1.0, 0.0, beta, nt, 10, W, delay_k=dk)
m.random_fields(1)
c = m.ctx
c.rng_mode(capi.RNG_PHILOX, 42)
c.init_stacks(); c.init_greenfunctions()
c.sweep_0_to_beta(); c.sweep_beta_to_0(); c.synchronize()
def timed(fn, n=8):
    c.timing_enable(True)
    for l in range(n): fn(l)
    c.synchronize()
    ms, k = c.timing_get()
    c.timing_enable(False)
    return ms[0] / max(k[0], 1)
w = timed(c.wrap_forward); wb = timed(c.wrap_backward); mu = timed(c.local_update)
acc0 = c.get_stats()[0].mean()
print(f"L={L} W={W} dk={dk}: wrap_fwd {w:.3f} ms, wrap_bwd {wb:.3f} ms, local_update {mu:.3f} ms per slice-launch")
