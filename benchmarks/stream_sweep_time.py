import sys, time, os
sys.path.insert(0, "/root/repo")
from dqmc_b200 import host, capi
m = host.AttractiveHubbard(8, 8, 4.0, 1.0, 0.0, 8.0, 80, 10, 1024)
m.random_fields(1); c = m.ctx; c.rng_mode(capi.RNG_PHILOX, 42)
c.init_stacks(); c.init_greenfunctions(); c.synchronize()
for s in range(2):
    c.sweep_0_to_beta(); c.sweep_beta_to_0()
c.synchronize()
t0 = time.time()
for s in range(5):
    c.sweep_0_to_beta(); c.sweep_beta_to_0()
c.synchronize(); t = (time.time() - t0) / 5
print(f"streams={os.environ.get('DQMCB_STREAMS','2')}: {t*1e3:.2f} ms/sweep  flips/s {1024*2*80*64/t:.3e} acc {c.get_stats()[0].mean()/14:.3f}")
