import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np
from dqmc_b200 import host, capi
m = host.AttractiveHubbard(8, 8, 4.0, 1.0, 0.0, 2.0, 20, 10, 2)
m.random_fields(1)
c = m.ctx
c.init_stacks(); c.synchronize()
B = np.asarray(c.block_product(1))[0]
L, d, R = [np.asarray(v)[0] for v in c.get_ldr(1)]
print("LDR vs Bbar", np.abs(L @ np.diag(d) @ R - B).max() / np.abs(B).max())
G = L.T @ L
print("orth", np.abs(G - np.eye(64)).max(), "diag", np.diag(G)[:8], "offmax idx", np.unravel_index(np.abs(G-np.eye(64)).argmax(), G.shape))
Q, Rr = np.linalg.qr(B)
print("col norms of L", np.linalg.norm(L, axis=0)[:8], "row norms", np.linalg.norm(L, axis=1)[:8])
print("L^T L vs L L^T:", np.abs(L@L.T-np.eye(64)).max())
import scipy.linalg as sl
Qs, Rs, Ps = sl.qr(B, pivoting=True)
ds = np.abs(np.diag(Rs))
print("d vs scipy", np.abs(np.sort(d) - np.sort(ds)).max() / ds.max(), d[:4], ds[:4])
# R check: D R should equal R_scipy P^T up to row signs
DR = np.diag(d) @ R
RsPt = np.zeros_like(Rs); RsPt[:, Ps] = Rs
print("|D R| vs |R P^T|", np.abs(np.abs(DR) - np.abs(RsPt)).max() / ds.max())
# which Q columns are right?  B = Q (D R) -> Q = B (D R)^-1
Qtrue = B @ np.linalg.inv(DR)
err = np.abs(L - Qtrue).max(axis=0)
print("Q col err", err[:16])
errr = np.abs(L - Qtrue).max(axis=1)
print("Q row err", errr[:16])
print("R^T R vs B^T B", np.abs(DR.T @ DR - B.T @ B).max() / np.abs(B.T @ B).max())
