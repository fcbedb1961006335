"""Shared helpers for the parity tests: build matched (oracle, GPU) pairs on the same inputs."""
import numpy as np

from oracle import dqmc_oracle as O

CONFIGS = {
    # name: L1, L2, U, t, mu, beta, nt, n_stab   
    "C0": (6, 6, 4.0, 1.0, -0.1, 4.0, 40, 10),
    "C1": (8, 8, 4.0, 1.0, 0.0, 8.0, 80, 10),
    "C1s": (8, 8, 4.0, 1.0, 0.0, 2.0, 20, 5),      # short 8x8 case for quick checks
    "C2s": (12, 12, 4.0, 1.0, 0.0, 2.0, 20, 5),    # 12x12, global-workspace path
    "R7": (7, 5, 3.0, 1.0, 0.2, 3.0, 23, 6),       # odd N, ragged last block (nt % n_stab != 0)
    "F32": (8, 4, 5.0, 1.0, 0.1, 2.3, 23, 5),      # N = 32 fast path, ragged last block
    "C3s": (16, 16, 6.0, 1.0, 0.0, 1.0, 10, 5),    # 16x16 (N = 256): L2-resident tiles, DMMA generic GEMM
    # atomic limit t = 0 on the fast path: exp(-dtau K) = I, every B and G diagonal -- the pivoted QR meets columns that
    # are already zero below the diagonal (reflected with tau = 2 by the fast path, H = I in LAPACK), LU pivots sit on
    # the diagonal, and every triangular solve is trivial
    "T0": (8, 8, 4.0, 0.0, 0.0, 2.0, 20, 5),
    "T0_32": (8, 4, 3.0, 0.0, 0.0, 1.5, 15, 4),
}


def make_oracle(cfg, seed):
    L1, L2, U, t, mu, beta, nt, n_stab = cfg
    f = O.random_fields(seed, nt, L1 * L2)
    m = O.Model(L1, L2, U, t, mu, beta, nt, fields=f)
    d = O.DQMC(m, n_stab)
    return m, d


def make_ctx(cfg, n_walkers, expK, invexpK, device=0, delay_k=0):
    from dqmc_b200 import capi, host

    L1, L2, U, t, mu, beta, nt, n_stab = cfg
    import math

    g = math.sqrt(0.5 * abs(U) * (beta / nt))
    return capi.Context(L1 * L2, nt, n_stab, n_walkers, g, host.ETA, host.GAMMA, expK, invexpK, -1.0, device, delay_k)


def rel_err(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
