"""Development aid: time / profile the stand-alone block-product kernel (k_bbar) of the generic engine. usage: prof_bbar.py L W"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dqmc_b200 import host, capi
L = int(sys.argv[1]); W = int(sys.argv[2])
m = host.AttractiveHubbard(L, L, 4.0, 1.0, 0.0, 2.0, 20, 10, W)
m.random_fields(1)
c = m.ctx
c.block_product(0); c.synchronize()
t0 = time.perf_counter()
for _ in range(5): c.block_product(1)
c.synchronize()
print(f"L={L} W={W}: k_bbar {(time.perf_counter()-t0)/5*1e3:.3f} ms per launch (9 slice products)")
