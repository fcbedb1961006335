"""GPU parity at the BASELINE configurations themselves, not their short stand-ins:
C1 (8x8, beta=8, nt=80, n_stab=10; 10 full sweeps), C2 (12x12, beta=10, nt=100; n_stab=10 and 5; 2 sweeps),
C3 (16x16, U=6, beta=16, nt=160; n_stab=5: 2 sweeps; n_stab=10: start-up state only, because the reference's
own check_error reaches O(1) there within one sweep -- SURVEY H2).

The goldens (tests/golden/traj_full_*.npz, written by `python tests/golden/make_golden.py full`) are outputs
of the REFERENCE ITSELF (oracle/_ref = its sources compiled unmodified) for utility::random rng(seed).  The
random stream is not stored: the device regenerates std::mt19937(seed) (DQMCB_RNG_MT19937), so a run needs
the seed, the start fields and the reference's exp(-+dtau K) only.

Checked: fields after EVERY half sweep identical (i.e. every accept decision identical); diag(G) after every
half sweep and the full G of the trailing half sweeps to 1e-10 (relative max-norm; the end of a half sweep is
a stabilisation point, where G is a function of the fields alone); log|det M| to 1e-10 relative; the sorted
scales d of EVERY stack entry to 1e-10 relative, element by element, over their whole e^{+-beta W} range;
acceptance rate to 1e-12.  Run on the fast path (N = 64), on the generic any-N engine, and with delayed
rank-k updates.
"""
import os

import numpy as np
import pytest

from tests.common import make_ctx, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    g = np.load(os.path.join(GOLD, f"traj_full_{name}.npz"))
    L1, L2, U, t, mu, beta, nt, n_stab, seed, n_sweeps = g["cfg"]
    cfg = (int(L1), int(L2), float(U), float(t), float(mu), float(beta), int(nt), int(n_stab))
    return g, cfg, int(seed), int(n_sweeps)


def _start(g, cfg, seed, delay_k):
    ctx = make_ctx(cfg, 1, g["expK"], g["invexpK"], delay_k=delay_k)
    ctx.set_fields(g["fields0"][None].astype(np.int8))
    ctx.rng_seed_mt19937(np.array([seed], dtype=np.uint32))
    ctx.init_stacks()
    ctx.init_greenfunctions()
    return ctx


def _check_d(ctx, want, what, tol=TOL):
    worst = 0.0
    for i in range(ctx.n_stack):
        _, dd, _ = ctx.get_ldr(i)
        worst = max(worst, float(np.max(np.abs(np.sort(dd[0]) / want[i] - 1.0))))
    assert worst < tol, f"{what}: sorted d of a stack entry off by {worst:.2e} (relative, element-wise)"
    return worst


def _check_start(ctx, g, d_tol=TOL):
    G = ctx.get_G()[0]
    scale = max(np.max(np.abs(G)), 1e-300)
    assert np.max(np.abs(np.diag(G) - g["G0_diag"])) / scale < TOL
    if "G0" in g.files:
        assert rel_err(G, g["G0"]) < TOL
    ld = float(g["logdet0"])
    assert abs(ctx.get_logdet()[0] - ld) < TOL * max(1.0, abs(ld))
    _check_d(ctx, g["stack0_d"], "init_stacks", d_tol)


@pytest.mark.parametrize("name,engine,delay_k", [
    ("c1", "fast", 0), ("c1", "generic", 0), ("c1", "generic", 16),
    ("c2", "generic", 0), ("c2", "generic", 16), ("c2n5", "generic", 16),
    ("c3n5", "generic", 32), ("c3n5", "generic", 0),
])
def test_reference_trajectory_at_baseline_config(name, engine, delay_k, monkeypatch):
    g, cfg, seed, n_sweeps = _load(name)
    nt = cfg[6]
    if engine == "generic":
        monkeypatch.setenv("DQMCB_FORCE_GENERIC", "1")
    else:
        monkeypatch.delenv("DQMCB_FORCE_GENERIC", raising=False)
    ctx = _start(g, cfg, seed, delay_k)
    _check_start(ctx, g)
    n_half = 2 * n_sweeps
    tail = g["traj_G_tail"]
    first_tail = n_half - tail.shape[0]
    worst_G = 0.0
    for h in range(n_half):
        (ctx.sweep_0_to_beta if h % 2 == 0 else ctx.sweep_beta_to_0)()
        f = ctx.get_fields()[0]
        ndiff = int(np.sum(f != g["traj_fields"][h]))
        assert ndiff == 0, f"{name}: {ndiff} field values differ from the reference after half sweep {h}"
        G = ctx.get_G()[0]
        scale = np.max(np.abs(G))
        e = float(np.max(np.abs(np.diag(G) - g["traj_G_diag"][h])) / scale)
        if h >= first_tail:
            e = max(e, rel_err(G, tail[h - first_tail]))
        worst_G = max(worst_G, e)
        assert e < TOL, f"{name}: G off by {e:.2e} after half sweep {h}"
        ld = float(g["traj_logdet"][h])
        assert abs(ctx.get_logdet()[0] - ld) < TOL * max(1.0, abs(ld))
    worst_d = _check_d(ctx, g["stack_d"], "after the sweeps")
    acc, emax, emean = ctx.get_stats()
    assert abs(acc[0] - float(g["traj_acc_rate"])) < 1e-12
    # wrap-error envelope (check_error): the reference's own running max / mean; ours must be of the same size
    ref_max, ref_mean = g["traj_err"][-1]
    assert emax[0] < 10.0 * ref_max + 1e-12 and emean[0] < 10.0 * ref_mean + 1e-12
    S = ctx.global_action()[0]
    assert abs(S - float(g["traj_action"])) < 1e-10 * abs(S)
    sc, _ = ctx.measure_equal_time(want_corr=False)
    assert np.max(np.abs(sc[0, :3] - g["traj_obs"])) < 1e-10
    print(f"[full-config parity] {name} {engine} delay_k={delay_k}: worst G {worst_G:.2e}, worst d {worst_d:.2e}, "
          f"check_error max {emax[0]:.2e} (reference {ref_max:.2e})")
    ctx.close()


@pytest.mark.parametrize("delay_k", [0])
def test_start_up_state_at_c3_n_stab_10(delay_k, monkeypatch):
    """C3 as stated (n_stab = 10): init_stacks + init_greenfunctions only (see the module docstring)."""
    g, cfg, seed, _ = _load("c3")
    ctx = _start(g, cfg, seed, delay_k)
    # a block product of 10 slices at U = 6 is so ill-conditioned that the smallest scales carry 1e-7 relative noise even
    # between the reference and its numpy restatement, which call the same dgeqp3 (tests/test_oracle.py): 1e-6 here
    _check_start(ctx, g, d_tol=1e-6)
    ctx.close()
