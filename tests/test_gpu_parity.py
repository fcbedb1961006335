"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on identical inputs.

Tolerance: 1e-10 relative (max-norm), the bar BASELINE.json's north_star states; accept/reject
decisions and fields must be identical.  Chaotic wrap drift is handled as the survey
prescribes: per-step checks are teacher-forced, trajectories compare decisions/fields exactly and
G at stabilisation points.
"""
import numpy as np
import pytest

from oracle import dqmc_oracle as O
from tests.common import CONFIGS, make_ctx, make_oracle, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _setup(name, seeds, delay_k=0):
    cfg = CONFIGS[name]
    pairs = [make_oracle(cfg, s) for s in seeds]
    m0 = pairs[0][0]
    ctx = make_ctx(cfg, len(seeds), m0.expK, m0.invexpK, delay_k=delay_k)
    for m, _ in pairs:  # all walkers share the model matrices bit-for-bit
        m.expK, m.invexpK = m0.expK, m0.invexpK
    ctx.set_fields(np.stack([m.fields for m, _ in pairs]).astype(np.int8))
    return cfg, pairs, ctx


@pytest.mark.parametrize("name", ["C0", "C1s", "R7", "C2s", "F32", "T0", "T0_32"])
def test_init_stacks_and_green(name):
    cfg, pairs, ctx = _setup(name, [11, 12, 13])
    ctx.init_stacks()
    ctx.init_greenfunctions()
    G = ctx.get_G()
    logdet = ctx.get_logdet()
    for w, (m, d) in enumerate(pairs):
        d.init_stacks()
        d.init_greenfunctions()
        assert rel_err(G[w], d.G) < TOL
        assert abs(logdet[w] - d.log_det_M) < TOL * max(1.0, abs(d.log_det_M))
    for i in range(ctx.n_stack):
        Lm, dd, Rm = ctx.get_ldr(i)
        for w, (m, d) in enumerate(pairs):
            F = d.stack[i]
            prod = (Lm[w] * dd[w]) @ Rm[w]
            assert rel_err(prod, F.to_matrix()) < 1e-9
            assert np.allclose(np.sort(dd[w]), np.sort(F.d), rtol=1e-9)
            assert np.allclose(Lm[w].T @ Lm[w], np.eye(m.ns), atol=1e-12)


@pytest.mark.parametrize("name", ["C0", "C1s", "R7", "C2s", "F32", "C3s"])
def test_wrap_steps_teacher_forced(name):
    cfg, pairs, ctx = _setup(name, [21, 22])
    rng = np.random.default_rng(5)
    n = pairs[0][0].ns
    G0 = np.stack([np.eye(n) * 0.5 + 0.1 * rng.standard_normal((n, n)) for _ in pairs])
    for l in (0, 3, cfg[6] - 1):
        ctx.set_G(G0)
        ctx.wrap_forward(l)
        Gf = ctx.get_G()
        ctx.set_G(G0)
        ctx.wrap_backward(l)
        Gb = ctx.get_G()
        for w, (m, d) in enumerate(pairs):
            assert rel_err(Gf[w], d.wrap_forward(G0[w], l)) < TOL
            assert rel_err(Gb[w], d.wrap_backward(G0[w], l)) < TOL


@pytest.mark.parametrize("name", ["C0", "C1s", "R7", "C2s", "F32", "C3s"])
def test_block_product(name):
    cfg, pairs, ctx = _setup(name, [31, 32])
    for i in range(ctx.n_stack):
        B = ctx.block_product(i)
        for w, (m, d) in enumerate(pairs):
            assert rel_err(B[w], d.calculate_Bbar(i)) < TOL


@pytest.mark.parametrize("name,delay_k", [("C0", 0), ("C1s", 0), ("R7", 0), ("F32", 0), ("C0", 8), ("C2s", 16), ("C3s", 0), ("C3s", 32)])
def test_local_update_teacher_forced(name, delay_k):
    """One slice of Metropolis from the oracle's G: identical decisions and fields, G to 1e-10."""
    cfg, pairs, ctx = _setup(name, [41, 42, 43], delay_k)
    n, nt = pairs[0][0].ns, cfg[6]
    gens = [O.MT19937(100 + w) for w in range(len(pairs))]
    tapes = [O.rng_tape(g, n, 2) for g in gens]
    ctx.set_tape(np.stack([t[0] for t in tapes]), np.stack([t[1] for t in tapes]), np.stack([t[2] for t in tapes]))
    Gin = []
    for m, d in pairs:
        d.init_stacks()
        d.init_greenfunctions()
        Gin.append(d.wrap_forward(d.G, 0))
    ctx.set_G(np.stack(Gin))
    for visit, l in enumerate((0, nt - 1)):
        ctx.local_update(l)
        G = ctx.get_G()
        f = ctx.get_fields()
        for w, (m, d) in enumerate(pairs):
            Gw = Gin[w] if visit == 0 else G_prev[w]
            Gw = Gw.copy()
            d.local_update(Gw, l, tapes[w][0][visit], tapes[w][1][visit], tapes[w][2][visit])
            assert np.array_equal(f[w], m.fields)
            assert rel_err(G[w], Gw) < TOL
        G_prev = G


@pytest.mark.parametrize("name,n_sweeps,delay_k", [("C0", 3, 0), ("C1s", 3, 0), ("R7", 3, 0), ("C2s", 1, 0), ("F32", 3, 0), ("C1", 1, 0),
                                                   ("C3s", 1, 0), ("R7", 2, 4), ("C2s", 1, 8), ("C0", 2, 16), ("C3s", 1, 32), ("T0", 2, 0), ("T0_32", 2, 0)])
def test_trajectory(name, n_sweeps, delay_k):
    """Free-running sweeps on the same RNG tape: identical accept decisions / fields; G compared at
    the end of each half sweep (a stabilisation point, where G is a function of the fields only).
    delay_k > 1 runs the delayed rank-k Metropolis update (new relative to the reference, whose
    rank-1 sequence is its oracle: same decisions, same fields, same G)."""
    cfg, pairs, ctx = _setup(name, [51, 52], delay_k)
    n, nt = pairs[0][0].ns, cfg[6]
    gens = [O.MT19937(200 + w) for w in range(len(pairs))]
    ctx.init_stacks()
    ctx.init_greenfunctions()
    for m, d in pairs:
        d.init_stacks()
        d.init_greenfunctions()
    for s in range(n_sweeps):
        tapes = [O.rng_tape(g, n, 2 * nt) for g in gens]
        ctx.set_tape(np.stack([t[0] for t in tapes]), np.stack([t[1] for t in tapes]), np.stack([t[2] for t in tapes]))
        ctx.sweep_0_to_beta()
        Gf = ctx.get_G()
        ff = ctx.get_fields()
        ctx.sweep_beta_to_0()
        Gb = ctx.get_G()
        fb = ctx.get_fields()
        for w, (m, d) in enumerate(pairs):
            t = tapes[w]
            d.sweep_0_to_beta((t[0][:nt], t[1][:nt], t[2][:nt]))
            assert np.array_equal(ff[w], m.fields), f"fields diverged in forward sweep {s}"
            assert rel_err(Gf[w], d.G) < 1e-9
            d.sweep_beta_to_0((t[0][nt:], t[1][nt:], t[2][nt:]))
            assert np.array_equal(fb[w], m.fields), f"fields diverged in backward sweep {s}"
            assert rel_err(Gb[w], d.G) < 1e-9
    acc, emax, emean = ctx.get_stats()
    for w, (m, d) in enumerate(pairs):
        assert abs(acc[w] - d.acc_rate) < 1e-12
        assert emax[w] < 1e-4
    logdet = ctx.get_logdet()
    S = ctx.global_action()
    sc, corr = ctx.measure_equal_time()
    for w, (m, d) in enumerate(pairs):
        assert abs(logdet[w] - d.log_det_M) < 1e-9 * abs(d.log_det_M)
        assert abs(S[w] - m.global_action(d.log_det_M)) < 1e-9 * abs(S[w])
        den, docc, sw, cc = O.observables(d.G)
        assert abs(sc[w, 0] - den) < 1e-9 and abs(sc[w, 1] - docc) < 1e-9 and abs(sc[w, 2] - sw) < 1e-9
        assert rel_err(corr[w], cc) < 1e-9


@pytest.mark.parametrize("n", [36, 64, 35, 144, 256])  # 144: staging tile; 256: TMA-staged products and solves (gemm_tma.cuh)
def test_stablelinalg_batched(n):
    """source/stablelinalg.cpp functions one by one on graded random matrices."""
    from dqmc_b200 import capi

    rng = np.random.default_rng(n)
    b = 3

    def graded():
        Q1, _ = np.linalg.qr(rng.standard_normal((n, n)))
        Q2, _ = np.linalg.qr(rng.standard_normal((n, n)))
        s = np.exp(np.linspace(6, -6, n))
        return (Q1 * s) @ Q2

    Ms = np.stack([graded() for _ in range(b)])
    Lg, dg, Rg = capi.to_LDR(Ms)
    F = []
    for i in range(b):
        Fo = O.to_LDR(Ms[i])
        F.append(Fo)
        assert rel_err((Lg[i] * dg[i]) @ Rg[i], Ms[i]) < 1e-12
        assert np.allclose(dg[i], Fo.d, rtol=1e-10)
        assert rel_err(np.abs(Lg[i]), np.abs(Fo.L)) < 1e-8
    M2 = np.stack([graded() for _ in range(b)])
    F2 = [O.to_LDR(M2[i]) for i in range(b)]
    pack = lambda Fs: (np.stack([f.L for f in Fs]), np.stack([f.d for f in Fs]), np.stack([f.R for f in Fs]))
    out = capi.mat_mul_ldr(M2, pack(F))
    for i in range(b):
        assert rel_err((out[0][i] * out[1][i]) @ out[2][i], O.mat_mul_ldr(M2[i], F[i]).to_matrix()) < 1e-10
    out = capi.ldr_mul_mat(pack(F), M2)
    for i in range(b):
        assert rel_err((out[0][i] * out[1][i]) @ out[2][i], O.ldr_mul_mat(F[i], M2[i]).to_matrix()) < 1e-10
    out = capi.ldr_mul_ldr(pack(F), pack(F2))
    for i in range(b):
        assert rel_err((out[0][i] * out[1][i]) @ out[2][i], O.ldr_mul_ldr(F[i], F2[i]).to_matrix()) < 1e-10
    G, ld = capi.inv_I_plus_ldr(pack(F))
    for i in range(b):
        Go, ldo = O.inv_I_plus_ldr(F[i])
        assert rel_err(G[i], Go) < TOL
        assert abs(ld[i] - ldo) < 1e-10 * max(1.0, abs(ldo))
    G = capi.inv_I_plus_ldr_mul_ldr(pack(F), pack(F2))
    for i in range(b):
        assert rel_err(G[i], O.inv_I_plus_ldr_mul_ldr(F[i], F2[i])) < TOL


@pytest.mark.parametrize("name", ["c0", "r7", "c1s"])
def test_per_proposal_against_reference_golden(name):
    """Rows a7/a9 one proposal at a time through the C ABI, against the reference's own numbers
    (tests/golden, generated from oracle/_ref): R, Delta, the accept decision, then G."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"traj_{name}.npz"))
    L1, L2, U, t, mu, beta, nt, n_stab = g["cfg"][:8]
    cfg = (int(L1), int(L2), U, t, mu, beta, int(nt), int(n_stab))
    ctx = make_ctx(cfg, 1, g["expK"], g["invexpK"])
    f0 = g["fields0"].astype(np.int8)
    ctx.set_fields(f0[None])
    ctx.set_G(g["G_wrap_fwd0"][None])
    perm, k, u = g["tape_perm"][0], g["tape_k"][0], g["tape_u"][0]
    f = f0[0].copy()
    for idx in range(cfg[0] * cfg[1]):
        i = int(perm[idx])
        new = int(O.PROPOSAL[f[i], k[idx]])
        R, dl = ctx.local_update_ratio(0, i, new)
        assert abs(R[0] - g["slice0_R"][idx]) <= 1e-11 * abs(g["slice0_R"][idx])
        assert abs(dl[0] - g["slice0_delta"][idx]) <= 1e-12 * abs(g["slice0_delta"][idx])
        acc = bool(u[idx] < min(1.0, abs(R[0])))
        assert acc == bool(g["slice0_accept"][idx])
        if acc:
            ctx.update_greens_local(0, dl[0], i)
            ctx.set_single_field(0, 0, i, new)
            f[i] = new
    assert np.array_equal(ctx.get_fields()[0, 0], g["slice0_fields"])
    assert rel_err(ctx.get_G()[0], g["slice0_G"]) < TOL
    # and the same slice in one launch of the Metropolis kernel
    ctx.set_fields(f0[None])
    ctx.set_G(g["G_wrap_fwd0"][None])
    ctx.set_tape(perm[None, None], k[None, None], u[None, None])
    ctx.local_update(0)
    assert np.array_equal(ctx.get_fields()[0, 0], g["slice0_fields"])
    assert rel_err(ctx.get_G()[0], g["slice0_G"]) < TOL
    ctx.wrap_backward(0)
    assert rel_err(ctx.get_G()[0], g["slice0_G_wrap_bwd"]) < TOL


def test_inv_invldr_plus_ldr():
    from dqmc_b200 import capi
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "stablelinalg.npz"))
    for n in (12, 36):
        p = f"n{n}_"
        F1 = (g[p + "F1_L"], g[p + "F1_d"], g[p + "F1_R"])
        F2 = (g[p + "F2_L"], g[p + "F2_d"], g[p + "F2_R"])
        assert rel_err(capi.inv_invldr_plus_ldr(F1, F2)[0], g[p + "inv3_G"]) < TOL
        assert rel_err(capi.inv_I_plus_ldr_mul_ldr(F1, F2)[0], g[p + "inv2_G"]) < TOL
        G, ld = capi.inv_I_plus_ldr(F1)
        assert rel_err(G[0], g[p + "inv1_G"]) < TOL and abs(ld[0] - float(g[p + "inv1_logdet"])) < 1e-10 * max(1, abs(ld[0]))


def test_walker_groups_on_streams_are_equivalent(monkeypatch):
    """Splitting the independent chains into groups on separate CUDA streams (DQMCB_STREAMS) must not
    change any walker's trajectory: same Philox keys, same kernels -> bit-identical fields and G."""
    from dqmc_b200 import capi

    cfg = CONFIGS["C1s"]
    W = 192
    m, _ = make_oracle(cfg, 5)
    rng = np.random.default_rng(9)
    f0 = rng.integers(0, 4, size=(W, cfg[6], cfg[0] * cfg[1]), dtype=np.int8)
    out = []
    for ng in ("1", "3"):
        monkeypatch.setenv("DQMCB_STREAMS", ng)
        ctx = make_ctx(cfg, W, m.expK, m.invexpK)
        assert ctx.n_groups() == int(ng)
        ctx.set_fields(f0)
        ctx.rng_mode(capi.RNG_PHILOX, 77)
        ctx.init_stacks()
        ctx.init_greenfunctions()
        n0 = ctx.launch_count()
        for _ in range(2):
            ctx.sweep_0_to_beta()
            ctx.sweep_beta_to_0()
        assert ctx.launch_count() - n0 == 2 * 2 * ctx.n_stack * 2 * int(ng)
        out.append((ctx.get_fields(), ctx.get_G(), ctx.get_stats()[0]))
        ctx.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert np.array_equal(out[0][1], out[1][1])
    assert np.array_equal(out[0][2], out[1][2])


def test_kronecker_factor_path_equals_dense_path(monkeypatch):
    """On the 8x8 lattice exp(-+dtau K) = Ey (x) Ex and the fast path applies the 8x8 factors (Bbar chain, wraps);
    DQMCB_NO_KRON=1 forces the dense products.  Same Philox keys: teacher-forced pieces (block product, one wrap)
    agree to 1e-12 / 1e-10, and a free-running sweep takes identical decisions with G equal at the stabilisation point.
    A matrix that is NOT of product form (one entry perturbed by 1e-9) must fall back to the dense path by itself."""
    from dqmc_b200 import capi

    cfg = CONFIGS["C1s"]
    W = 8
    m, _ = make_oracle(cfg, 5)
    rng = np.random.default_rng(19)
    f0 = rng.integers(0, 4, size=(W, cfg[6], cfg[0] * cfg[1]), dtype=np.int8)
    bad = m.expK.copy()
    bad[3, 17] *= 1.0 + 1e-9
    out = []
    for env, expK in ((None, m.expK), ("1", m.expK), (None, bad)):
        if env is None:
            monkeypatch.delenv("DQMCB_NO_KRON", raising=False)
        else:
            monkeypatch.setenv("DQMCB_NO_KRON", env)
        ctx = make_ctx(cfg, W, expK, m.invexpK)
        ctx.set_fields(f0)
        ctx.rng_mode(capi.RNG_PHILOX, 77)
        ctx.init_stacks()
        ctx.init_greenfunctions()
        B = ctx.block_product(1)
        ctx.wrap_forward(0)
        Gw = ctx.get_G()
        ctx.init_greenfunctions()
        ctx.sweep_0_to_beta()
        ctx.sweep_beta_to_0()
        out.append((B, Gw, ctx.get_fields(), ctx.get_G(), ctx.get_stats()[0]))
        ctx.close()
    kron, dense, fallback = out
    assert rel_err(kron[0], dense[0]) < 1e-12 and rel_err(kron[1], dense[1]) < TOL  # G carries the stack's conditioning
    assert not np.array_equal(kron[0], dense[0])  # the factor path really ran (different rounding order)
    assert np.array_equal(kron[2], dense[2]) and np.array_equal(kron[4], dense[4])
    assert rel_err(kron[3], dense[3]) < 1e-9
    # perturbed matrix: dense path on its own; compare with the forced-dense run of the exact matrix only loosely
    assert rel_err(fallback[0], dense[0]) < 1e-7


def test_runtime_masked_kronecker_path_on_a_10x6_lattice(monkeypatch):
    """Lattices that are neither 8x8, 12x12 nor 16x16 take the RUN-TIME masked Kronecker routine of the generic engine when
    N > 48 (below that the dense products are faster and are what runs: C0, R7).  10 x 6 (N = 60, extents that are not
    multiples of 4 or 8 in either direction): block product and one wrap against the dense path to 1e-12 / 1e-10 and
    against the oracle, identical decisions over a free-running sweep."""
    from dqmc_b200 import capi

    cfg = (10, 6, 4.0, 1.0, 0.1, 1.5, 15, 5)
    W = 4
    m, d = make_oracle(cfg, 3)
    rng = np.random.default_rng(23)
    f0 = rng.integers(0, 4, size=(W, cfg[6], cfg[0] * cfg[1]), dtype=np.int8)
    out = []
    for env in (None, "1"):
        if env is None:
            monkeypatch.delenv("DQMCB_NO_KRON", raising=False)
        else:
            monkeypatch.setenv("DQMCB_NO_KRON", env)
        ctx = make_ctx(cfg, W, m.expK, m.invexpK)
        ctx.set_fields(f0)
        ctx.rng_mode(capi.RNG_PHILOX, 78)
        ctx.init_stacks()
        ctx.init_greenfunctions()
        B = ctx.block_product(1)
        G0 = ctx.get_G()
        ctx.wrap_forward(0)
        Gw = ctx.get_G()
        ctx.init_greenfunctions()
        ctx.sweep_0_to_beta()
        ctx.sweep_beta_to_0()
        out.append((B, Gw, ctx.get_fields(), ctx.get_G(), ctx.get_stats()[0], G0))
        ctx.close()
    kron, dense = out
    assert rel_err(kron[0], dense[0]) < 1e-12 and rel_err(kron[1], dense[1]) < TOL
    assert not np.array_equal(kron[0], dense[0])  # the factor path really ran (different rounding order)
    assert np.array_equal(kron[2], dense[2]) and np.array_equal(kron[4], dense[4])
    assert rel_err(kron[3], dense[3]) < 1e-9
    # against the oracle: block product of block 1 and the start-up G of walker 0
    m.fields = f0[0].astype(np.int64).copy()
    assert rel_err(kron[0][0], d.calculate_Bbar(1)) < 1e-12
    d.init_stacks()
    d.init_greenfunctions()
    assert rel_err(kron[5][0], d.G) < TOL


@pytest.mark.parametrize("engine", ["fast", "generic"])
def test_product_form_matrix_that_is_not_symmetric_takes_the_dense_path(engine, monkeypatch):
    """The row-wise Kronecker passes feed F(gr, tg) where F(tg, gr) is meant: they need SYMMETRIC factors.  A caller-supplied
    exp(-dtau K) = Fy (x) Fx with non-symmetric Fx (a valid product form!) must therefore fall back to the dense products:
    wraps and block products then agree with numpy on the same matrices to 1e-10.  (With the symmetry check missing, the
    right wrap would silently use Fx^T.)"""
    from dqmc_b200 import capi, host
    import math

    if engine == "generic":
        monkeypatch.setenv("DQMCB_FORCE_GENERIC", "1")
    else:
        monkeypatch.delenv("DQMCB_FORCE_GENERIC", raising=False)
    monkeypatch.delenv("DQMCB_NO_KRON", raising=False)
    rng = np.random.default_rng(3)
    Fx = np.eye(8) + 0.1 * rng.standard_normal((8, 8))  # not symmetric
    Fy = np.eye(8) + 0.05 * rng.standard_normal((8, 8))
    Fy = 0.5 * (Fy + Fy.T)
    E = np.kron(Fy, Fx)  # site index x + 8 y
    Ei = np.linalg.inv(E)
    nt, n_stab, W, N = 6, 3, 2, 64
    g = math.sqrt(0.5 * 4.0 * 0.1)
    ctx = capi.Context(N, nt, n_stab, W, g, host.ETA, host.GAMMA, E, Ei, -1.0, 0, 0)
    f = rng.integers(0, 4, size=(W, nt, N), dtype=np.int8)
    ctx.set_fields(f)
    G0 = np.stack([np.eye(N) * 0.5 + 0.1 * rng.standard_normal((N, N)) for _ in range(W)])
    ctx.set_G(G0)
    ctx.wrap_forward(1)
    Gf = ctx.get_G()
    ctx.set_G(G0)
    ctx.wrap_backward(1)
    Gb = ctx.get_G()
    B = ctx.block_product(0)
    eta = np.asarray(host.ETA)
    for w in range(W):
        ev = np.exp(g * eta[f[w, 1]])
        Bl, Bli = ev[:, None] * E, Ei * (1.0 / ev)[None, :]
        assert rel_err(Gf[w], Bl @ G0[w] @ Bli) < TOL
        assert rel_err(Gb[w], Bli @ G0[w] @ Bl) < TOL
        want = np.eye(N)
        for l in range(n_stab):
            want = (np.exp(g * eta[f[w, l]])[:, None] * E) @ want
        assert rel_err(B[w], want) < TOL
    ctx.close()


def test_delayed_updates_with_tiles_in_global_memory_and_shared_sms():
    """Delayed rank-k updates at a size where G no longer fits beside the factors in shared memory
    (N=144, k=32: G lives in the global workspace) and with more walkers than SMs (two CTAs per SM):
    decisions must equal the rank-1 path's, G must agree to rounding.  Regression test for a
    write-after-read race on the running diagonal that only showed with warps far out of step."""
    from dqmc_b200 import capi

    cfg = (12, 12, 4.0, 1.0, 0.0, 1.0, 10, 5)
    W = 296
    m, _ = make_oracle(cfg, 5)
    f0 = np.random.default_rng(21).integers(0, 4, size=(W, cfg[6], 144), dtype=np.int8)
    out = []
    for dk in (0, 32):
        ctx = make_ctx(cfg, W, m.expK, m.invexpK, delay_k=dk)
        ctx.set_fields(f0)
        ctx.rng_mode(capi.RNG_PHILOX, 123)
        ctx.init_stacks()
        ctx.init_greenfunctions()
        for _ in range(2):
            ctx.sweep_0_to_beta()
            ctx.sweep_beta_to_0()
        out.append((ctx.get_fields(), ctx.get_G(), ctx.get_stats()))
        ctx.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert rel_err(out[1][1], out[0][1]) < 1e-8
    assert np.array_equal(out[0][2][0], out[1][2][0])
    assert out[1][2][1].max() < 1e-6


def test_walker_groups_with_offset_phases_under_host_contention(monkeypatch):
    """Regression test for a race in the fast-path QR (every warp derives the step's pivot from the shared norm
    array on its own; the owner used to retire the column before the step's barrier).  It only fired when a
    warp was held up for a few hundred cycles -- in practice when stabilisation kernels of two walker groups
    overlapped out of phase, which a busy host makes likely (7 of 8 runs failed with an illegal address).
    Busy host processes + grouped sweeps must complete and reproduce the single-stream trajectory bit for bit."""
    import subprocess
    import sys

    from dqmc_b200 import capi

    cfg = CONFIGS["C1"]
    W = 1024
    m, _ = make_oracle(cfg, 5)
    f0 = np.random.default_rng(12).integers(0, 4, size=(W, cfg[6], 64), dtype=np.int8)
    burners = [subprocess.Popen([sys.executable, "-c", "import time\nt=time.time()\nwhile time.time()-t<40: pass"]) for _ in range(24)]
    try:
        out = []
        for ng in ("2", "1"):
            monkeypatch.setenv("DQMCB_STREAMS", ng)
            ctx = make_ctx(cfg, W, m.expK, m.invexpK)
            ctx.set_fields(f0)
            ctx.rng_mode(capi.RNG_PHILOX, 31)
            ctx.init_stacks()
            ctx.init_greenfunctions()
            for s in range(12 if ng == "2" else 12):
                ctx.sweep_0_to_beta()
                ctx.sweep_beta_to_0()
            out.append((ctx.get_fields(), ctx.get_G()))
            ctx.close()
    finally:
        for b in burners:
            b.kill()
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])


def test_async_tape_pipeline_equals_synchronous_tape():
    """dqmcb_set_tape_async + dqmcb_tape_commit (next step's tape uploaded while the current step runs) must give
    exactly the trajectory of the synchronous dqmcb_set_tape, step after step."""
    cfg = CONFIGS["C1s"]
    W, nt, n = 96, cfg[6], 64
    m, _ = make_oracle(cfg, 5)
    f0 = np.random.default_rng(2).integers(0, 4, size=(W, nt, n), dtype=np.int8)
    gen = np.random.default_rng(3)
    tapes = []
    for s in range(4):
        perm = gen.permuted(np.tile(np.arange(n, dtype=np.uint8), (W, 2 * nt, 1)), axis=-1)
        tapes.append((np.ascontiguousarray(perm), gen.integers(0, 3, size=(W, 2 * nt, n), dtype=np.uint8), gen.random((W, 2 * nt, n))))
    outs = []
    for mode in ("sync", "async"):
        ctx = make_ctx(cfg, W, m.expK, m.invexpK)
        ctx.set_fields(f0)
        ctx.init_stacks()
        ctx.init_greenfunctions()
        res = []
        if mode == "async":
            ctx.set_tape_async(*tapes[0])
            ctx.tape_commit()
        for s in range(4):
            if mode == "sync":
                ctx.set_tape(*tapes[s])
            elif s + 1 < 4:
                ctx.set_tape_async(*tapes[s + 1])
            f, sc, G = ctx.run_sweeps_host(None, 1, want_G=True)
            if mode == "async" and s + 1 < 4:
                ctx.tape_commit()
            res.append((f, sc, G))
        outs.append(res)
        ctx.close()
    for a, b in zip(*outs):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


@pytest.mark.parametrize("name", ["c0", "r7", "c1s"])
def test_device_mt19937_reproduces_reference_trajectory(name):
    """DQMCB_RNG_MT19937: std::mt19937 + std::shuffle + by-value uniform_int + bernoulli regenerated on the device.
    Seeded like the reference run behind tests/golden/traj_*.npz (utility::random rng(seed)), the device must
    produce the reference's fields after every half sweep, with no host-side random numbers at all; walker 1 runs
    another seed and is checked against the oracle's restatement of the same stream."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"traj_{name}.npz"))
    L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps = g["cfg"]
    L1, L2, nt, n_stab, seed, n_sweeps = int(L1), int(L2), int(nt), int(n_stab), int(seed), int(n_sweeps)
    cfg = (L1, L2, float(U), float(t), float(mu), float(beta), nt, n_stab)
    n = L1 * L2
    m0 = O.Model(*cfg[:7], fields=g["fields0"].astype(np.int64))
    seed1 = seed + 12345
    m1 = O.Model(*cfg[:7], fields=O.random_fields(seed1, nt, n))
    d1 = O.DQMC(m1, n_stab)
    d1.init_stacks()
    d1.init_greenfunctions()
    gen1 = O.MT19937(seed1)
    ctx = make_ctx(cfg, 2, m0.expK, m0.invexpK)
    ctx.set_fields(np.stack([g["fields0"], m1.fields]).astype(np.int8))
    ctx.rng_seed_mt19937(np.array([seed, seed1], dtype=np.uint32))
    ctx.init_stacks()
    ctx.init_greenfunctions()
    for s in range(n_sweeps):
        tape = O.rng_tape(gen1, n, 2 * nt)
        ctx.sweep_0_to_beta()
        f = ctx.get_fields()
        assert np.array_equal(f[0], g["traj_fields"][2 * s]), f"walker 0 differs from the reference after forward sweep {s}"
        d1.sweep_0_to_beta((tape[0][:nt], tape[1][:nt], tape[2][:nt]))
        assert np.array_equal(f[1], m1.fields)
        ctx.sweep_beta_to_0()
        f = ctx.get_fields()
        assert np.array_equal(f[0], g["traj_fields"][2 * s + 1]), f"walker 0 differs from the reference after backward sweep {s}"
        d1.sweep_beta_to_0((tape[0][nt:], tape[1][nt:], tape[2][nt:]))
        assert np.array_equal(f[1], m1.fields)
        assert rel_err(ctx.get_G()[0], g["traj_G"][2 * s + 1]) < 1e-9
    assert abs(ctx.get_stats()[0][0] - float(g["traj_acc_rate"])) < 1e-12
    ctx.close()
