"""Python host layer above the C ABI, mirroring the reference's object surface for the sweep path
(AttractiveHubbard / GHQField / DQMC; include/model.h, include/field.h, include/dqmc.h) for a batch
of W walkers.  The C++ mirror with the reference's exact class names lives in include/dqmcb/.

Host-side work here is set-up only (K matrix, exp(-+dtau K) by symmetric eigendecomposition, GHQ
tables); every sweep-path computation goes through libdqmc_b200.so.
"""
from __future__ import annotations

import math

import numpy as np

from . import capi

# GHQField tables, include/field.h:33-48
_S6 = math.sqrt(6.0)
GAMMA = np.array([1.0 - _S6 / 3.0, 1.0 + _S6 / 3.0, 1.0 + _S6 / 3.0, 1.0 - _S6 / 3.0])
ETA = np.array(
    [-math.sqrt(2.0 * (3.0 + _S6)), -math.sqrt(2.0 * (3.0 - _S6)), math.sqrt(2.0 * (3.0 - _S6)), math.sqrt(2.0 * (3.0 + _S6))]
)


def build_K(L1: int, L2: int, t: float, mu: float) -> np.ndarray:
    """AttractiveHubbard::build_K_matrix (source/model.cpp:39-60) with Lattice::site_neighbors
    (include/lattice.h:100-107): nearest-neighbour hopping on the periodic L1 x L2 square lattice."""
    n = L1 * L2
    K = np.zeros((n, n))
    for i in range(n):
        ux, uy = i % L1, i // L1
        K[i, i] = -mu
        for nb in (uy * L1 + (ux + 1) % L1, ((uy + 1) % L2) * L1 + ux):
            K[i, nb] = -t
            K[nb, i] = -t
    return K


def expm_sym(A: np.ndarray) -> np.ndarray:
    w, V = np.linalg.eigh(A)
    return (V * np.exp(w)) @ V.T


class AttractiveHubbard:
    """Batched counterpart of `class AttractiveHubbard` (include/model.h:11): one model, W walkers."""

    def __init__(self, L1, L2, U, t, mu, beta, nt, n_stab, n_walkers, device=0, expK=None, invexpK=None, delay_k=0):
        self.L1, self.L2, self.U, self.t, self.mu, self.beta = L1, L2, U, t, mu, beta
        self.ns_, self.nt_ = L1 * L2, nt
        self.dtau = beta / nt
        self.g = math.sqrt(0.5 * abs(U) * self.dtau)  # source/model.cpp:27
        self.alpha = -1.0  # source/model.cpp:28
        K = build_K(L1, L2, t, mu)
        self.expK = expm_sym(-self.dtau * K) if expK is None else np.asarray(expK)
        self.invexpK = expm_sym(self.dtau * K) if invexpK is None else np.asarray(invexpK)
        self.expK_half = expm_sym(-0.5 * self.dtau * K)
        self.invexpK_half = expm_sym(0.5 * self.dtau * K)
        self.ctx = capi.Context(self.ns_, nt, n_stab, n_walkers, self.g, ETA, GAMMA, self.expK, self.invexpK, self.alpha, device, delay_k)
        self.ctx.set_half_expK(self.expK_half, self.invexpK_half)

    def ns(self):
        return self.ns_

    def nt(self):
        return self.nt_

    def random_fields(self, seed: int) -> np.ndarray:
        """i.i.d. uniform fields for every walker (throughput runs; parity runs inject the reference's)."""
        rng = np.random.default_rng(seed)
        f = rng.integers(0, 4, size=(self.ctx.W, self.nt_, self.ns_), dtype=np.int8)
        self.ctx.set_fields(f)
        return f


def reference_fields(seed: int, nt: int, n_sites: int) -> np.ndarray:
    """The random start of the reference for `utility::random rng(seed)`: GHQField's constructor
    (include/field.h:52-57) fills the column-major nt x N matrix with uniform_int_distribution<int>(0, 3) drawn
    from a COPY of the generator (the stream of the sweeps then starts again at the seeded state).  For a 32-bit
    engine and a range of 4 libstdc++'s multiply-shift never rejects, so each value is the top two bits of one
    mt19937 word.  Returns int8 fields[l, site]."""
    bg = np.random.MT19937()
    bg._legacy_seeding(int(seed) & 0xFFFFFFFF)  # init_genrand(seed) = std::mt19937(seed)
    words = bg.random_raw(nt * n_sites).astype(np.uint64)
    flat = (words >> np.uint64(30)).astype(np.int8)
    return np.ascontiguousarray(flat.reshape(n_sites, nt).T)


class DQMC:
    """Batched counterpart of `class DQMC` (include/dqmc.h:21): sweeps over all walkers of a model."""

    def __init__(self, model: AttractiveHubbard):
        self.model = model
        self.ctx = model.ctx

    def init_stacks(self):
        self.ctx.init_stacks()

    def init_greenfunctions(self):
        self.ctx.init_greenfunctions()

    def sweep_0_to_beta(self):
        self.ctx.sweep_0_to_beta()

    def sweep_beta_to_0(self):
        self.ctx.sweep_beta_to_0()

    def acc_rate(self):
        return self.ctx.get_stats()[0]

    def max_err(self):
        return self.ctx.get_stats()[1]

    def mean_err(self):
        return self.ctx.get_stats()[2]
