"""Equal-time measurement pipeline on the device against the reference's
MeasurementManager / transform:: functions (include/measurementh5.h).  Golden data: tests/golden/measure.npz,
produced by the reference's own header compiled unmodified over an HDF5 recorder shim
(tests/golden/make_golden.py: measure_case).  Tolerance 1e-10 relative (north_star); names, dims exact."""
import os

import numpy as np
import pytest

from tests.common import CONFIGS, make_ctx, make_oracle
from oracle import dqmc_oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-10


def rel(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if np.iscomplexobj(a) or np.iscomplexobj(b):
        a, b = np.stack([a.real, a.imag]), np.stack([b.real, b.imag])
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def test_transforms_against_reference_golden():
    from dqmc_b200 import capi

    g = np.load(os.path.join(GOLD, "measure.npz"))
    for (a, b) in ((6, 6), (5, 5), (4, 6)):
        r = capi.chi_site_to_chi_r(g[f"tf_{a}x{b}_site"], a, b)
        assert r.shape == (a, b, 3)
        assert rel(r, g[f"tf_{a}x{b}_r"]) < TOL
        if a == b:
            k = capi.chi_r_to_chi_k(g[f"tf_{a}x{b}_r"], a, b)
            assert k.shape == (a, b, 3) and rel(k, g[f"tf_{a}x{b}_k"]) < TOL
    tab = g["pbc_shortest_L4567"]
    for li, L in enumerate((4, 5, 6, 7)):
        for di, d in enumerate(range(-6, 7)):
            if abs(d) < L:
                assert capi.pbc_shortest(d, L) == tab[li, di]


@pytest.mark.parametrize("L1,L2,nm", [(8, 8, 2), (12, 12, 1), (3, 7, 4), (16, 16, 1), (2, 2, 3), (1, 5, 2)])
def test_transforms_against_oracle_more_shapes(L1, L2, nm):
    """Ragged / odd / degenerate lattices incl. non-square (k-point (n,m) convention, oracle docstring)."""
    from dqmc_b200 import capi

    rng = np.random.default_rng(L1 * 100 + L2)
    c = rng.standard_normal((nm, L1 * L2, L1 * L2))
    r = capi.chi_site_to_chi_r(c, L1, L2)
    assert rel(r, O.chi_site_to_chi_r(c, L1, L2)) < 1e-12
    assert rel(capi.chi_r_to_chi_k(r, L1, L2), O.chi_r_to_chi_k(r, L1, L2)) < 1e-12
    # empty batch is a no-op
    assert capi.chi_site_to_chi_r(np.zeros((0, L1 * L2, L1 * L2)), L1, L2).shape == (L1, L2, 0)


def test_bin_pipeline_reproduces_reference_datasets():
    """8 G's of the reference trajectory held by 4 walkers: two device-side measure() calls per bin are the
    reference's 4 measure() calls per bin up to summation order -> same dataset tree."""
    from dqmc_b200 import measure

    g = np.load(os.path.join(GOLD, "measure.npz"))
    cfg = (6, 6, 4.0, 1.0, -0.1, 2.0, 20, 5)
    m, _ = make_oracle(cfg, 1)
    # bin 0 <- G[0..3], bin 1 <- G[4..7]; walkers w=0,1 hold the two G's of each call
    ctx = make_ctx(cfg, 2, m.expK, m.invexpK)
    mm = measure.MeasurementManager(ctx, 6, 6, rank=2)
    for b in range(2):
        for call in range(2):
            ctx.set_G(g["G"][4 * b + 2 * call: 4 * b + 2 * call + 2])
            mm.measure()
        new = mm.accumulate()
        assert len(new) == 5
    paths = [str(p) for p in g["paths"]]
    assert sorted(paths) == sorted(mm.datasets)
    for p in paths:
        want = g["ds:" + p]
        assert mm.datasets[p].shape == want.shape, p
        assert rel(mm.datasets[p], want) < TOL, p
    assert os.path.basename(mm.file_name()) == "data_2.npz"
    ctx.close()


def test_bin_accumulation_many_walkers_and_reset(tmp_path):
    """W=300 walkers x 3 sweeps on the fast path: bin sums equal the per-walker observables summed on the
    host (dqmcb_measure_equal_time + oracle fold); fetch(reset) clears; save() writes the tree."""
    from dqmc_b200 import capi, measure

    cfg = CONFIGS["C1s"]
    L1, L2 = cfg[0], cfg[1]
    W = 300
    m, _ = make_oracle(cfg, 3)
    ctx = make_ctx(cfg, W, m.expK, m.invexpK)
    ctx.set_fields(np.random.default_rng(4).integers(0, 4, size=(W, cfg[6], L1 * L2), dtype=np.int8))
    ctx.rng_mode(capi.RNG_PHILOX, 9)
    ctx.init_stacks()
    ctx.init_greenfunctions()
    mm = measure.MeasurementManager(ctx, L1, L2, out_dir=str(tmp_path))
    sc_sum, corr_sum, n = np.zeros(3), np.zeros((L1 * L2, L1 * L2)), 0
    for _ in range(3):
        ctx.sweep_0_to_beta()
        ctx.sweep_beta_to_0()
        mm.measure()
        sc, corr = ctx.measure_equal_time(True)
        sc_sum += sc.sum(axis=0)
        corr_sum += corr.sum(axis=0)
        n += W
    new = mm.accumulate()
    assert rel([new[f"/bin_0/scalar/{k}"][0] for k in measure.SCALARS], sc_sum / n) < 1e-12
    want_r = O.chi_site_to_chi_r(corr_sum / n, L1, L2)
    assert rel(new["/bin_0/equaltime/densityCorr"], want_r) < 1e-11
    want_k = O.chi_r_to_chi_k(want_r, L1, L2)
    assert rel(new["/binK_0/equaltime/densityCorr"], np.stack([want_k.real, want_k.imag], axis=-1)) < 1e-11
    with pytest.raises(RuntimeError):
        mm.accumulate()  # bin was reset: nothing measured
    path = mm.save()
    z = np.load(path)
    assert sorted(z.files) == sorted(new) and z["/binK_0/equaltime/densityCorr"].shape == (L1, L2, 1, 2)
    # error behaviour: wrong lattice
    with pytest.raises(ValueError):
        measure.MeasurementManager(ctx, L1, L2 + 1)
    with pytest.raises(capi.DqmcbError):
        ctx.bin_configure(3, 5)
    ctx.close()


@pytest.mark.parametrize("name", ["C1s", "R7", "C0"])
def test_kinetic_energy_and_spin_correlation_against_restatement(name):
    """dqmcb_measure_kinetic_spin (no reference counterpart) against oracle.kinetic_spin -- itself pinned on the analytic
    U = 0 result in tests/test_oracle.py -- on the G of real trajectories: square, ragged 7 x 5 and the 6 x 6 example."""
    from dqmc_b200 import capi
    from tests.common import CONFIGS, make_ctx, make_oracle

    cfg = CONFIGS[name]
    L1, L2, t = cfg[0], cfg[1], cfg[3]
    W = 3
    m, _ = make_oracle(cfg, 7)
    ctx = make_ctx(cfg, W, m.expK, m.invexpK)
    rng = np.random.default_rng(5)
    ctx.set_fields(rng.integers(0, 4, size=(W, cfg[6], L1 * L2), dtype=np.int8))
    ctx.rng_mode(capi.RNG_PHILOX, 11)
    ctx.init_stacks()
    ctx.init_greenfunctions()
    ctx.sweep_0_to_beta()
    ctx.sweep_beta_to_0()
    G = ctx.get_G()
    ekin, szz = ctx.measure_kinetic_spin(L1, L2, t)
    for w in range(W):
        e0, s0 = O.kinetic_spin(G[w], L1, L2, t)
        assert abs(ekin[w] - e0) < 1e-12 * max(1.0, abs(e0))
        assert np.max(np.abs(szz[w] - s0)) < 1e-12
    Sq = capi.spin_structure_factor(szz)
    assert Sq.shape == (W, L1, L2) and np.allclose(Sq[:, 0, 0], szz.sum(axis=(1, 2)))
    ctx.close()
