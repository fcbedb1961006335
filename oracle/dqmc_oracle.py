"""oracle/dqmc_oracle.py -- TEST INFRASTRUCTURE, not product code.

CPU restatement (numpy + scipy's LAPACK) of the reference's sweep hot path, one function per
reference function, each citing the reference file:line it follows (paths relative to
/root/reference).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import it; the product path (dqmc_b200/) never does.

Pinning: the reference ships no tests, golden vectors or fixtures, so this
restatement is pinned against outputs of the reference itself: the reference's own
source/{model,dqmc,update,stablelinalg}.cpp compiled unmodified into oracle/_ref/ (oracle/Makefile)
and driven through oracle/ref.py.  tests/golden/*.npz holds vectors generated from that build by
tests/golden/make_golden.py; tests/test_oracle.py checks this file against them (and against the
live oracle/_ref library when it is present).

All matrices are FP64, column-major semantics are irrelevant here (numpy arrays are indexed
[row, col]); buffers crossing the C ABI are converted explicitly by the callers.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import scipy.linalg as sla

# ----------------------------------------------------------------------------------------------
# HS field tables -- include/field.h:33-48
# ----------------------------------------------------------------------------------------------
_S6 = math.sqrt(6.0)
GAMMA = np.array([1.0 - _S6 / 3.0, 1.0 + _S6 / 3.0, 1.0 + _S6 / 3.0, 1.0 - _S6 / 3.0])
ETA = np.array(
    [
        -math.sqrt(2.0 * (3.0 + _S6)),
        -math.sqrt(2.0 * (3.0 - _S6)),
        math.sqrt(2.0 * (3.0 - _S6)),
        math.sqrt(2.0 * (3.0 + _S6)),
    ]
)
PROPOSAL = np.array([[1, 2, 3], [0, 2, 3], [0, 1, 3], [0, 1, 2]], dtype=np.int64)
ALPHA = -1.0  # source/model.cpp:28


# ----------------------------------------------------------------------------------------------
# libstdc++ RNG semantics (H6/Q1 of SURVEY.md): mt19937 + std::shuffle + uniform_int + bernoulli
# ----------------------------------------------------------------------------------------------
class MT19937:
    """std::mt19937(seed): raw 32-bit stream (numpy's legacy seeding is init_genrand(seed))."""

    def __init__(self, seed: int):
        self._bg = np.random.MT19937()
        self._bg._legacy_seeding(int(seed) & 0xFFFFFFFF)
        self._buf = np.empty(0, dtype=np.uint64)
        self._pos = 0

    def next(self) -> int:
        if self._pos >= len(self._buf):
            self._buf = self._bg.random_raw(4096)
            self._pos = 0
        v = int(self._buf[self._pos])
        self._pos += 1
        return v


def _lemire(gen: MT19937, rng_range: int) -> int:
    """uniform_int_distribution over [0, rng_range) for a 32-bit URNG, libstdc++ >= 11
    (bits/uniform_int_dist.h, _S_nd): multiply-shift with rejection of the biased low part."""
    product = gen.next() * rng_range
    low = product & 0xFFFFFFFF
    if low < rng_range:
        threshold = ((1 << 32) - rng_range) % rng_range
        while low < threshold:
            product = gen.next() * rng_range
            low = product & 0xFFFFFFFF
    return product >> 32


def std_shuffle(n: int, gen: MT19937) -> np.ndarray:
    """std::shuffle(0..n-1) as in libstdc++ bits/stl_algo.h (two swap positions per draw);
    call site source/update.cpp:10-14."""
    a = list(range(n))
    if n < 2:
        return np.array(a, dtype=np.int64)
    urngrange = 0xFFFFFFFF
    if urngrange // n >= n:
        i = 1
        if n % 2 == 0:
            j = _lemire(gen, 2)
            a[i], a[j] = a[j], a[i]
            i += 1
        while i < n:
            swap_range = i + 1
            x = _lemire(gen, swap_range * (swap_range + 1))
            p0, p1 = x // (swap_range + 1), x % (swap_range + 1)
            a[i], a[p0] = a[p0], a[i]
            i += 1
            a[i], a[p1] = a[p1], a[i]
            i += 1
    else:
        for i in range(1, n):
            j = _lemire(gen, i + 1)
            a[i], a[j] = a[j], a[i]
    return np.array(a, dtype=np.int64)


def rng_tape(gen: MT19937, n_sites: int, n_calls: int):
    """Draws made by n_calls consecutive update::local_update calls (source/update.cpp:10-25):
    perm[c, idx] = visiting order, k[c, idx] = proposal index (drawn from a by-value copy of the
    generator, include/field.h:76, so it re-uses the first word of the following bernoulli draw),
    u[c, idx] = generate_canonical<double,53> used by bernoulli_distribution (include/utility.h:34-37)."""
    perm = np.empty((n_calls, n_sites), dtype=np.int64)
    k = np.empty((n_calls, n_sites), dtype=np.int64)
    u = np.empty((n_calls, n_sites), dtype=np.float64)
    for c in range(n_calls):
        perm[c] = std_shuffle(n_sites, gen)
        for idx in range(n_sites):
            x = gen.next()
            y = gen.next()
            # the by-value copy sees x first: Lemire with range 3 (rejection iff low < 1)
            prod = x * 3
            if (prod & 0xFFFFFFFF) < 1:
                prod = y * 3  # copy would redraw from ITS next word, which is y
                # (a further rejection has probability 2^-32 and is ignored: it cannot be
                #  represented without the generator state; asserted absent)
                assert (prod & 0xFFFFFFFF) >= 1
            k[c, idx] = prod >> 32
            val = (x + y * 4294967296.0) / 18446744073709551616.0
            if val >= 1.0:
                val = math.nextafter(1.0, 0.0)
            u[c, idx] = val
    return perm, k, u


def random_fields(seed: int, nt: int, n_sites: int) -> np.ndarray:
    """GHQField ctor, include/field.h:52-57: uniform_int_distribution<int>(0,3) over the
    column-major nt x N matrix, drawn from a COPY of the generator.  Returns fields[l, site]."""
    gen = MT19937(seed)
    flat = np.array([_lemire(gen, 4) for _ in range(nt * n_sites)], dtype=np.int64)
    return flat.reshape(n_sites, nt).T.copy()  # column-major linear index i -> (l=i%nt, site=i//nt)


# ----------------------------------------------------------------------------------------------
# Model -- source/model.cpp:3-159, include/lattice.h:100-107
# ----------------------------------------------------------------------------------------------
def site_neighbor(i: int, dx: int, dy: int, L1: int, L2: int) -> int:
    """include/lattice.h:100-107 (single orbital)."""
    ux, uy = i % L1, i // L1
    return ((uy + dy) % L2) * L1 + (ux + dx) % L1


def build_K(L1: int, L2: int, t: float, mu: float) -> np.ndarray:
    """source/model.cpp:39-60 (assignment, not accumulation)."""
    n = L1 * L2
    K = np.zeros((n, n))
    for i in range(n):
        K[i, i] = -mu
        nx = site_neighbor(i, 1, 0, L1, L2)
        K[i, nx] = -t
        K[nx, i] = -t
        ny = site_neighbor(i, 0, 1, L1, L2)
        K[i, ny] = -t
        K[ny, i] = -t
    return K


def expm_sym(A: np.ndarray) -> np.ndarray:
    w, V = np.linalg.eigh(A)
    return (V * np.exp(w)) @ V.T


@dataclass
class Model:
    """AttractiveHubbard, source/model.cpp:3-36."""

    L1: int
    L2: int
    U: float
    t: float
    mu: float
    beta: float
    nt: int
    fields: np.ndarray = None  # [nt, N] values in {0,1,2,3}
    expK: np.ndarray = field(default=None, repr=False)
    invexpK: np.ndarray = field(default=None, repr=False)
    expK_half: np.ndarray = field(default=None, repr=False)
    invexpK_half: np.ndarray = field(default=None, repr=False)

    def __post_init__(self):
        self.ns = self.L1 * self.L2
        self.dtau = self.beta / self.nt
        self.g = math.sqrt(0.5 * abs(self.U) * self.dtau)  # source/model.cpp:27
        K = build_K(self.L1, self.L2, self.t, self.mu)
        if self.expK is None:
            self.expK = expm_sym(-self.dtau * K)  # source/model.cpp:32-35
            self.invexpK = expm_sym(self.dtau * K)
            self.expK_half = expm_sym(-0.5 * self.dtau * K)
            self.invexpK_half = expm_sym(0.5 * self.dtau * K)

    def expV(self, l: int) -> np.ndarray:  # source/model.cpp:62-72
        return np.exp(self.g * ETA[self.fields[l]])

    def invexpV(self, l: int) -> np.ndarray:  # source/model.cpp:74-84
        return np.exp(-self.g * ETA[self.fields[l]])

    def bosonic_ratio(self, new: int, old: int):  # source/model.cpp:99-107
        d_eta = ETA[new] - ETA[old]
        boson = math.exp(ALPHA * self.g * d_eta)
        return boson, 1.0 / boson - 1.0

    def local_update_ratio(self, G: np.ndarray, l: int, i: int, new: int):  # source/model.cpp:109-122
        old = int(self.fields[l, i])
        gammaR = GAMMA[new] / GAMMA[old]
        boson, delta = self.bosonic_ratio(new, old)
        det_flv = 1.0 + (1.0 - G[i, i]) * delta  # source/model.cpp:95
        return gammaR * boson * det_flv**2, delta

    @staticmethod
    def update_greens_local(G: np.ndarray, delta: float, i: int) -> None:  # source/model.cpp:124-138
        pref = delta / (1.0 + (1.0 - G[i, i]) * delta)
        Ucol = G[:, i].copy()
        Vrow = G[i, :].copy()
        Vrow[i] -= 1.0
        G += pref * np.outer(Ucol, Vrow)

    def global_action(self, log_det_M: float) -> float:  # source/model.cpp:140-159
        S = -2.0 * log_det_M
        f = self.fields.T.reshape(-1)  # column-major traversal order of the nt x N imat
        logboson = 0.0
        loggamma = 0.0
        for v in f:
            logboson += ALPHA * self.g * ETA[v]
            loggamma += math.log(GAMMA[v])
        return S - (logboson + loggamma)


# ----------------------------------------------------------------------------------------------
# stablelinalg -- source/stablelinalg.cpp
# ----------------------------------------------------------------------------------------------
@dataclass
class LDR:  # include/stablelinalg.h:16-33
    L: np.ndarray
    d: np.ndarray
    R: np.ndarray

    def to_matrix(self) -> np.ndarray:
        return (self.L * self.d) @ self.R


def to_LDR(M: np.ndarray) -> LDR:  # source/stablelinalg.cpp:35-55
    Q, R, P = sla.qr(M, pivoting=True)  # dgeqp3 + dorgqr, as arma::qr(...,"vector")
    d = np.abs(np.diag(R))
    Rn = R / d[:, None]
    Pinv = np.argsort(P, kind="stable")
    return LDR(Q, d, Rn[:, Pinv])


def ldr_mul_mat(F: LDR, M: np.ndarray) -> LDR:  # source/stablelinalg.cpp:57-67
    Mp = F.d[:, None] * (F.R @ M)
    q = to_LDR(Mp)
    return LDR(F.L @ q.L, q.d, q.R)


def mat_mul_ldr(M: np.ndarray, F: LDR) -> LDR:  # source/stablelinalg.cpp:69-79
    Mp = (M @ F.L) * F.d[None, :]
    q = to_LDR(Mp)
    return LDR(q.L, q.d, q.R @ F.R)


def ldr_mul_ldr(F1: LDR, F2: LDR) -> LDR:  # source/stablelinalg.cpp:81-92
    Mp = F1.R @ F2.L
    Mp = F1.d[:, None] * Mp
    Mp = Mp * F2.d[None, :]
    q = to_LDR(Mp)
    return LDR(F1.L @ q.L, q.d, q.R @ F2.R)


def _split(d: np.ndarray):
    large = np.where(d >= 1.0, d, 1.0)
    small = np.where(d >= 1.0, 1.0, d)
    return large, small


def inv_I_plus_ldr(F: LDR):  # source/stablelinalg.cpp:94-126
    Dl, Ds = _split(F.d)
    n = len(F.d)
    X = np.linalg.solve(F.R, np.diag(1.0 / Dl))
    M = X + F.L * Ds[None, :]
    _, logabs = np.linalg.slogdet(M)
    logdet = float(np.sum(np.log(Dl)) + logabs)
    Gt = np.linalg.solve(M.T, X.T)
    return Gt.T, logdet


def inv_I_plus_ldr_mul_ldr(F1: LDR, F2: LDR) -> np.ndarray:  # source/stablelinalg.cpp:128-158
    D1l, D1s = _split(F1.d)
    D2l, D2s = _split(F2.d)
    X = np.linalg.solve(F2.R, np.diag(1.0 / D2l))
    TermA = (1.0 / D1l)[:, None] * (F1.L.T @ X)
    TermB = D1s[:, None] * (F1.R @ (F2.L * D2s[None, :]))
    M = TermA + TermB
    RHS = (1.0 / D1l)[:, None] * F1.L.T
    Y = np.linalg.solve(M, RHS)
    return X @ Y


def inv_invldr_plus_ldr(F1: LDR, F2: LDR) -> np.ndarray:  # source/stablelinalg.cpp:160-190
    D1l, D1s = _split(F1.d)
    D2l, D2s = _split(F2.d)
    X = np.linalg.solve(F2.R, np.diag(1.0 / D2l))
    TermA = (1.0 / D1l)[:, None] * (F1.L.T @ X)
    TermB = D1s[:, None] * (F1.R @ (F2.L * D2s[None, :]))
    M = TermA + TermB
    RHS = D1s[:, None] * F1.R
    Y = np.linalg.solve(M, RHS)
    return X @ Y


# ----------------------------------------------------------------------------------------------
# Sweep engine -- source/dqmc.cpp, source/update.cpp
# ----------------------------------------------------------------------------------------------
def partner_rank(rank: int, world_size: int, attempt: int) -> int:  # source/update.cpp:34-45
    if attempt % 2 == 0:
        off = 1 if rank % 2 == 0 else -1
    else:
        off = -1 if rank % 2 == 0 else 1
    return (rank + off + world_size) % world_size


class DQMC:
    """class DQMC, source/dqmc.cpp:5-456 (equal-time parts).  Only the two adjacent Gtt entries
    the equal-time sweeps touch are kept: self.G is 'the current' Gtt."""

    def __init__(self, model: Model, n_stab: int):
        self.m = model
        self.nt = model.nt
        self.n_stab = n_stab
        self.n_stack = -(-self.nt // n_stab)  # source/dqmc.cpp:10
        self.loc_l_end = [n_stab - 1] * self.n_stack  # source/dqmc.cpp:14-18
        if self.nt % n_stab != 0:
            self.loc_l_end[-1] = self.nt % n_stab - 1
        self.acc_rate = 0.0
        self.max_err = 0.0
        self.total_err = 0.0
        self.n_err = 0
        self.stack: list[LDR] = [None] * self.n_stack
        self.G = None  # current equal-time G
        self.log_det_M = 0.0
        self.trace = None  # optional list collecting (l, site, new, R, delta, u, accept)

    # --- B matrices ---
    def B(self, l: int) -> np.ndarray:  # source/dqmc.cpp:78-80
        return self.m.expV(l)[:, None] * self.m.expK

    def invB(self, l: int) -> np.ndarray:  # source/dqmc.cpp:84-86
        return self.m.invexpK * self.m.invexpV(l)[None, :]

    def calculate_Bbar(self, i_stack: int) -> np.ndarray:  # source/dqmc.cpp:88-105
        Bbar = np.eye(self.m.ns)
        for loc_l in range(self.loc_l_end[i_stack] + 1):
            l = i_stack * self.n_stab + loc_l
            Bbar = self.B(l) @ Bbar
        return Bbar

    # --- init ---
    def init_stacks(self) -> None:  # source/dqmc.cpp:43-59
        for i in range(self.n_stack - 1, -1, -1):
            F = to_LDR(self.calculate_Bbar(i))
            self.stack[i] = F if i == self.n_stack - 1 else ldr_mul_ldr(self.stack[i + 1], F)

    def init_greenfunctions(self) -> None:  # source/dqmc.cpp:61-72
        self.G, self.log_det_M = inv_I_plus_ldr(self.stack[0])

    # --- propagation ---
    def wrap_forward(self, G: np.ndarray, l: int) -> np.ndarray:  # source/dqmc.cpp:113-132
        return self.B(l) @ G @ self.invB(l)

    def wrap_backward(self, G: np.ndarray, l: int) -> np.ndarray:  # source/dqmc.cpp:169-187
        return self.invB(l) @ G @ self.B(l)

    def local_update(self, G: np.ndarray, l: int, perm, k, u) -> float:  # source/update.cpp:5-32
        accepted = 0
        f = self.m.fields
        for idx in range(self.m.ns):
            i = int(perm[idx])
            old = int(f[l, i])
            new = int(PROPOSAL[old, int(k[idx])])
            R, delta = self.m.local_update_ratio(G, l, i, new)
            p = min(1.0, abs(R))
            acc = bool(u[idx] < p)
            if self.trace is not None:
                self.trace.append((l, i, new, R, delta, float(u[idx]), acc))
            if acc:
                accepted += 1
                Model.update_greens_local(G, delta, i)
                f[l, i] = new
        return accepted / self.m.ns

    def check_error(self, G_wrapped: np.ndarray, G_stab: np.ndarray) -> float:  # source/dqmc.cpp:317-329
        err = float(np.max(np.abs(G_wrapped - G_stab)))
        self.max_err = max(self.max_err, err)
        self.total_err += err
        self.n_err += 1
        return err

    # --- sweeps ---
    def sweep_0_to_beta(self, tape) -> None:  # source/dqmc.cpp:337-396
        """tape = (perm[nt,N], k[nt,N], u[nt,N]) for slices in visiting order l = 0..nt-1."""
        perm, k, u = tape
        for l in range(self.nt):
            i_stack = l // self.n_stab
            self.G = self.wrap_forward(self.G, l)
            acc = self.local_update(self.G, l, perm[l], k[l], u[l])
            self.acc_rate += acc / self.nt
            if l % self.n_stab == self.loc_l_end[i_stack]:
                G_tmp = self.G
                Bbar = self.calculate_Bbar(i_stack)
                # update_stack_forward, source/dqmc.cpp:134-146
                self.stack[i_stack] = to_LDR(Bbar) if i_stack == 0 else mat_mul_ldr(Bbar, self.stack[i_stack - 1])
                # stabilize_GF_forward, source/dqmc.cpp:148-161
                if l == self.nt - 1:
                    self.G, self.log_det_M = inv_I_plus_ldr(self.stack[i_stack])
                else:
                    self.G = inv_I_plus_ldr_mul_ldr(self.stack[i_stack], self.stack[i_stack + 1])
                self.check_error(G_tmp, self.G)

    def sweep_beta_to_0(self, tape) -> None:  # source/dqmc.cpp:398-456
        """tape rows are in visiting order: row j is used for slice l = nt-1-j."""
        perm, k, u = tape
        for j, l in enumerate(range(self.nt - 1, -1, -1)):
            i_stack = l // self.n_stab
            acc = self.local_update(self.G, l, perm[j], k[j], u[j])
            self.acc_rate += acc / self.nt
            self.G = self.wrap_backward(self.G, l)
            if l % self.n_stab == 0:
                G_tmp = self.G
                Bbar = self.calculate_Bbar(i_stack)
                # update_stack_backward, source/dqmc.cpp:189-201
                self.stack[i_stack] = (
                    to_LDR(Bbar) if i_stack == self.n_stack - 1 else ldr_mul_mat(self.stack[i_stack + 1], Bbar)
                )
                # stabilize_GF_backward, source/dqmc.cpp:203-215
                if l == 0:
                    self.G, self.log_det_M = inv_I_plus_ldr(self.stack[i_stack])
                else:
                    self.G = inv_I_plus_ldr_mul_ldr(self.stack[i_stack - 1], self.stack[i_stack])
                self.check_error(G_tmp, self.G)

    def half_wrap(self) -> np.ndarray:  # source/dqmc.cpp:288-315 (equal-time entry only)
        return self.m.invexpK_half @ self.G @ self.m.expK_half

    def mean_err(self) -> float:
        return self.total_err / self.n_err if self.n_err else float("nan")

    # --- unequal time (source/dqmc.cpp:223-280, 458-514) ---
    def sweep_unequalTime(self) -> float:
        """After sweep_beta_to_0 (self.G = Gtt[0], stack = right products B(beta, tau)): fills
        self.Gtt_all / Gt0_all / G0t_all [nt+1, N, N]; returns the largest stabilisation error seen."""
        n, nt = self.m.ns, self.nt
        Gtt = np.zeros((nt + 1, n, n))
        Gt0 = np.zeros((nt + 1, n, n))
        G0t = np.zeros((nt + 1, n, n))
        Gtt[0] = self.G
        Bt0 = None
        worst = 0.0
        for l in range(nt):
            i_stack, loc_l = l // self.n_stab, l % self.n_stab
            if l == 0:  # :235-240
                Gt0[0] = Gtt[0]
                G0t[0] = Gtt[0] - np.eye(n)
            B, iB = self.B(l), self.invB(l)
            Gtt[l + 1] = B @ Gtt[l] @ iB  # :241-243
            Gt0[l + 1] = B @ Gt0[l]
            G0t[l + 1] = G0t[l] @ iB
            if loc_l == self.loc_l_end[i_stack]:
                tmp = (Gtt[l + 1].copy(), Gt0[l + 1].copy(), G0t[l + 1].copy())
                Bbar = self.calculate_Bbar(i_stack)
                Bt0 = to_LDR(Bbar) if i_stack == 0 else mat_mul_ldr(Bbar, Bt0)  # propagate_Bt0_Bbt :248-262
                if l == nt - 1:  # stabilize_unequalTime :264-280
                    Gtt[l + 1], self.log_det_M = inv_I_plus_ldr(Bt0)
                    Gt0[l + 1] = np.eye(n) - Gtt[l + 1]
                    G0t[l + 1] = -Gtt[l + 1]
                else:
                    Bbt = self.stack[i_stack + 1]
                    Gtt[l + 1] = inv_I_plus_ldr_mul_ldr(Bt0, Bbt)
                    Gt0[l + 1] = inv_invldr_plus_ldr(Bt0, Bbt)
                    G0t[l + 1] = -inv_invldr_plus_ldr(Bbt, Bt0)
                for a, b in zip(tmp, (Gtt[l + 1], Gt0[l + 1], G0t[l + 1])):
                    worst = max(worst, self.check_error(a, b))
        self.Gtt_all, self.Gt0_all, self.G0t_all = Gtt, Gt0, G0t
        return worst

    def half_wrap_unequal(self) -> np.ndarray:  # source/dqmc.cpp:288-315 with isUnequalTime_: [3, nt, N, N]
        a, b = self.m.invexpK_half, self.m.expK_half
        return np.stack([np.stack([a @ X[t] @ b for t in range(self.nt)]) for X in (self.Gtt_all, self.Gt0_all, self.G0t_all)])


# ----------------------------------------------------------------------------------------------
# Equal-time observables -- source/model.cpp:167-288
# ----------------------------------------------------------------------------------------------
def observables(G: np.ndarray):
    """Returns (density, doubleOcc, swave, densityCorr[N,N]) from G = Gtt[0]."""
    n = G.shape[0]
    Gc = np.eye(n) - G
    dg = np.diag(Gc)
    density = float(np.sum(dg + dg) / n)  # source/model.cpp:186-190
    docc = float(np.sum(dg * dg) / n)  # source/model.cpp:214-218
    swave = float(np.sum(Gc * Gc) / n)  # source/model.cpp:247-253
    ni = 2.0 * (1.0 - np.diag(G))
    navg = float(np.sum(ni) / n)
    # C(i,j) = n_i n_j + 2 (1 - G(j,i)) G(i,j) - navg^2   (note 1.0, not delta_ij: Q5; source/model.cpp:281)
    corr = np.outer(ni, ni) + 2.0 * (1.0 - G.T) * G - navg * navg
    return density, docc, swave, corr


def kinetic_spin(G: np.ndarray, L1: int, L2: int, t: float = 1.0):
    """Kinetic energy per site and displacement-folded Sz-Sz correlation from G.  PARITY UNPINNED against the reference:
    it has no code for either (source/model.cpp:167-288); restated from Wick's theorem with G_up = G_down = G
    (<c+_i c_j> = delta_ij - G_ji) and pinned on the analytic U = 0 result in tests/test_oracle.py.
    Returns (ekin, szz_r[dx, dy])."""
    n = L1 * L2
    e = 0.0
    szz = np.zeros((L1, L2))
    Gc = np.eye(n) - G.T  # <c+_i c_j>
    for i in range(n):
        x, y = i % L1, i // L1
        for j in ((x + 1) % L1 + L1 * y, x + L1 * ((y + 1) % L2)):
            e += G[j, i] + G[i, j]
        for dy in range(L2):
            for dx in range(L1):
                j = (x + dx) % L1 + L1 * ((y + dy) % L2)
                szz[dx, dy] += 0.5 * Gc[i, j] * G[i, j]
    return 2.0 * t * e / n, szz / n


# ----------------------------------------------------------------------------------------------
# Unequal-time observables -- source/model.cpp:290-392 (both spin species share G: factor 2 / square)
# ----------------------------------------------------------------------------------------------
def green_tau(Gt0: np.ndarray) -> np.ndarray:  # :290-313
    return Gt0 + Gt0


def doublon_tau(Gt0: np.ndarray) -> np.ndarray:  # :315-343
    return Gt0 * Gt0


def currxx_tau(Gtt: np.ndarray, Gt0: np.ndarray, G0t: np.ndarray, L1: int, L2: int) -> np.ndarray:  # :345-392
    """[n_tau, N, N]; ix = +x neighbour of i (include/lattice.h:100-107)."""
    n = L1 * L2
    nx = np.array([site_neighbor(i, 1, 0, L1, L2) for i in range(n)])
    idx = np.arange(n)
    G00 = Gtt[0]
    out = np.zeros_like(Gtt)
    d1j = 2.0 * G00[nx, idx]  # dc_term1_j = G00(jx, j) up + dn
    d2j = 2.0 * G00[idx, nx]  # dc_term2_j = G00(j, jx)
    for t in range(Gtt.shape[0]):
        d1i = 2.0 * Gtt[t][nx, idx]
        d2i = 2.0 * Gtt[t][idx, nx]
        A, Bm = G0t[t], Gt0[t]
        # c_termK(i, j), up + dn = 2 x
        c1 = 2.0 * A[nx][:, idx].T * Bm[nx][:, idx]          # G0t(jx, i) Gt0(ix, j)
        c2 = 2.0 * A[idx][:, idx].T * Bm[nx][:, nx]          # G0t(j, i)  Gt0(ix, jx)
        c3 = 2.0 * A[nx][:, nx].T * Bm[idx][:, idx]          # G0t(jx, ix) Gt0(i, j)
        c4 = 2.0 * A[idx][:, nx].T * Bm[idx][:, nx]          # G0t(j, ix)  Gt0(i, jx)
        t1 = np.outer(d1i, d1j) - c1
        t2 = np.outer(d1i, d2j) - c2
        t3 = np.outer(d2i, d1j) - c3
        t4 = np.outer(d2i, d2j) - c4
        out[t] = -(t1 - t2 - t3 + t4)
    return out


# ----------------------------------------------------------------------------------------------
# Measurement pipeline -- include/measurementh5.h (transform:: :12-117, MeasurementManager :120-361)
# and the k-points of include/lattice.h:36-49 (square lattice a1=(1,0), a2=(0,1), one orbital, as
# source/main.cpp:84-88 sets it up)
# ----------------------------------------------------------------------------------------------
def pbc_shortest(d: int, L: int) -> int:  # include/measurementh5.h:13-17 (C++ int division)
    if d > L // 2:
        d -= L
    if d <= -(L // 2):
        d += L
    return d


def chi_site_to_chi_r(chi_site: np.ndarray, L1: int, L2: int) -> np.ndarray:
    """include/measurementh5.h:20-66 for n_orb = 1: chi_site [n_tau, N, N] (or [N, N]) indexed
    (tau, i, j) -> chi_r [L1, L2, n_tau], bin (dx + L1/2 - 1, dy + L2/2 - 1) of the shortest
    displacement cell_j - cell_i, each entry divided by n_cells."""
    c = np.asarray(chi_site, dtype=np.float64)
    if c.ndim == 2:
        c = c[None]
    nt, n = c.shape[0], L1 * L2
    out = np.zeros((L1, L2, nt))
    for i in range(n):
        cxi, cyi = i % L1, i // L1
        for j in range(n):
            cxj, cyj = j % L1, j // L1
            dx = pbc_shortest(cxj - cxi, L1) + L1 // 2 - 1
            dy = pbc_shortest(cyj - cyi, L2) + L2 // 2 - 1
            out[dx, dy, :] += c[:, i, j] / n
    return out


def k_points(L1: int, L2: int) -> np.ndarray:  # include/lattice.h:36-49, a1=(1,0), a2=(0,1)
    b1 = np.array([2 * np.pi / L1, 0.0])
    b2 = np.array([0.0, 2 * np.pi / L2])
    pts = []
    for n in range(L1):
        qx = n - L1 // 2 + 1
        for m in range(L2):
            qy = m - L2 // 2 + 1
            pts.append(qx * b1 + qy * b2)
    return np.array(pts)


def chi_r_to_chi_k(chi_r: np.ndarray, L1: int, L2: int) -> np.ndarray:
    """include/measurementh5.h:78-116: chi_k[kx, ky, t] = sum_r chi_r[x, y, t] exp(-i k.r).
    The reference stores k-point kidx = n*L2 + m at (kidx / L1, kidx % L2), which is (n, m) only for
    L1 == L2 and indexes out of bounds otherwise (undefined behaviour there); this restatement uses
    (n, m) and is therefore pinned on square lattices."""
    c = np.asarray(chi_r, dtype=np.float64)
    if c.ndim == 2:
        c = c[:, :, None]
    nx, ny, nt = c.shape
    kp = k_points(L1, L2)
    x = np.arange(nx) - (L1 // 2 - 1)
    y = np.arange(ny) - (L2 // 2 - 1)
    out = np.zeros((L1, L2, nt), dtype=np.complex128)
    for kidx in range(L1 * L2):
        phase = kp[kidx, 0] * x[:, None] + kp[kidx, 1] * y[None, :]
        w = np.cos(phase) - 1j * np.sin(phase)
        out[kidx // L2, kidx % L2, :] = np.tensordot(w, c, axes=([0, 1], [0, 1]))
    return out


class MeasurementManager:
    """include/measurementh5.h:120-361 wired as source/main.cpp:114-119 (density, doubleOcc, swave,
    densityCorr).  `datasets` collects what the reference writes per bin: /bin_<n>/scalar/<name> [1],
    /bin_<n>/equaltime/densityCorr [L1, L2, 1], /binK_<n>/equaltime/densityCorr [L1, L2, 1, 2]."""

    SCALARS = ("density", "doubleOcc", "swave")
    UNEQUAL = ("greenTau", "doublonTau", "currxxTau")  # source/main.cpp:120-122, only with isMeasureUnequalTime

    def __init__(self, L1: int, L2: int, unequal: bool = False):
        self.L1, self.L2, self.unequal = L1, L2, unequal
        self.bin = 0
        self.datasets: dict = {}
        self._reset()

    def _reset(self):
        self.scal = np.zeros(3)
        self.corr = None
        self.count = 0
        self.tau = None

    def measure_unequal(self, Gtt, Gt0, G0t) -> None:  # :189-229 with the unequal-time calculators
        self.measure(Gtt[0])
        cubes = [green_tau(Gt0), doublon_tau(Gt0), currxx_tau(Gtt, Gt0, G0t, self.L1, self.L2)]
        self.tau = cubes if self.tau is None else [a + b for a, b in zip(self.tau, cubes)]

    def measure(self, G: np.ndarray) -> None:  # :189-229
        d, o, s, corr = observables(G)
        self.scal += (d, o, s)
        self.corr = corr.copy() if self.corr is None else self.corr + corr
        self.count += 1

    def accumulate(self) -> None:  # :231-274 + saveToHDF5 :276-361
        scal = self.scal / self.count if self.count else self.scal
        b = self.bin
        for name, v in zip(self.SCALARS, scal):
            self.datasets[f"/bin_{b}/scalar/{name}"] = np.array([v])
        if self.corr is not None:
            chi_r = chi_site_to_chi_r(self.corr / self.count, self.L1, self.L2)
            chi_k = chi_r_to_chi_k(chi_r, self.L1, self.L2)
            self.datasets[f"/bin_{b}/equaltime/densityCorr"] = chi_r
            self.datasets[f"/binK_{b}/equaltime/densityCorr"] = np.stack([chi_k.real, chi_k.imag], axis=-1)
        if self.tau is not None:
            for name, cube in zip(self.UNEQUAL, self.tau):
                chi_r = chi_site_to_chi_r(cube / self.count, self.L1, self.L2)
                chi_k = chi_r_to_chi_k(chi_r, self.L1, self.L2)
                self.datasets[f"/bin_{b}/unequaltime/{name}"] = chi_r
                self.datasets[f"/binK_{b}/unequaltime/{name}"] = np.stack([chi_k.real, chi_k.imag], axis=-1)
        self._reset()
        self.bin += 1


# ----------------------------------------------------------------------------------------------
# Replica exchange -- source/update.cpp:47-117, all replicas in one process
# ----------------------------------------------------------------------------------------------
def replica_exchange(sims: list[DQMC], attempt: int, u_decider: dict[int, float]):
    """One exchange attempt over all replicas (attempt = the value AFTER the increment at
    source/update.cpp:52).  u_decider[r] = canonical double drawn by the lower rank r of each
    pair for its bernoulli (source/update.cpp:96).  Returns {rank: accepted}."""
    ws = len(sims)
    partners = [partner_rank(r, ws, attempt) for r in range(ws)]
    own_fields = [s.m.fields.copy() for s in sims]
    SC = [s.m.global_action(s.log_det_M) for s in sims]
    SCp = []
    for r, s in enumerate(sims):
        s.m.fields = own_fields[partners[r]].copy()
        s.init_stacks()
        s.init_greenfunctions()
        SCp.append(s.m.global_action(s.log_det_M))
    result = {}
    for r in range(ws):
        p = partners[r]
        lo = min(r, p)
        dS = (SCp[lo] + SCp[partners[lo]]) - (SC[lo] + SC[partners[lo]])
        prob = min(1.0, math.exp(-dS))
        result[r] = bool(u_decider[lo] < prob)
    for r, s in enumerate(sims):
        if not result[r]:
            s.m.fields = own_fields[r].copy()
            s.init_stacks()
            s.init_greenfunctions()
    return result
