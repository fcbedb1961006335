"""The driver (dqmc_b200/run.py = the run loop of source/main.cpp:36-214) on a small parameter file: output tree,
info file, summary, checkpoint/resume."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PARAMS = """[Lattice]
L1 = 4
L2 = 4
[hubbard]
U = 4.0
t = 1.0
mu = -0.1
[simulation]
beta = 1.0
nt = 10
n_therms = 3
n_sweeps = 2
n_bins = 3
n_stab = 5
symmetric = true
isMeasureUnequalTime = true
[ParallelTempering]
enabled = false
sweep_steps = 20
betas = 5.0, 4.5
"""


def _run(tmp_path, *extra, walkers=8):
    pf = tmp_path / "parameters.in"
    pf.write_text(PARAMS)
    out = tmp_path / "results"
    r = subprocess.run([sys.executable, "-m", "dqmc_b200.run", str(pf), "--walkers", str(walkers), "--seed", "11", "--out", str(out), *extra],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    return out, json.loads(r.stdout.strip().splitlines()[-1]), r.stdout


def test_driver_writes_reference_tree(tmp_path):
    out, summary, stdout = _run(tmp_path)
    assert "Standard DQMC run (Parallel Tempering disabled)." in stdout
    assert (out / "info").read_text() == "L1 4\nL2 4\nn_orb 1\na1_x 1\na1_y 0\na2_x 0\na2_y 1\n"  # include/lattice.h:125-131
    z = np.load(out / "data_0.npz")
    want = set()
    for b in range(3):
        want |= {f"/bin_{b}/scalar/{k}" for k in ("density", "doubleOcc", "swave")}
        want |= {f"/bin_{b}/equaltime/densityCorr", f"/binK_{b}/equaltime/densityCorr"}
        for k in ("greenTau", "doublonTau", "currxxTau"):
            want |= {f"/bin_{b}/unequaltime/{k}", f"/binK_{b}/unequaltime/{k}"}
    assert set(z.files) == want
    assert z["/bin_2/equaltime/densityCorr"].shape == (4, 4, 1) and z["/binK_0/unequaltime/greenTau"].shape == (4, 4, 11, 2)
    for b in range(3):
        assert 0.5 < z[f"/bin_{b}/scalar/density"][0] < 1.5 and 0.0 < z[f"/bin_{b}/scalar/doubleOcc"][0] < 1.0
        # sum rule: the k = 0 component of the connected density correlation is its sum over displacements
        r, k = z[f"/bin_{b}/equaltime/densityCorr"], z[f"/binK_{b}/equaltime/densityCorr"]
        assert abs(k[1, 1, 0, 0] - r.sum()) < 1e-12 * max(1.0, abs(r.sum())) and abs(k[1, 1, 0, 1]) < 1e-12
        # G(tau=0, 0) summed over spin at zero displacement = 2 (1 - n/2) = 2 - density
        g0 = z[f"/bin_{b}/unequaltime/greenTau"][1, 1, 0]
        assert abs(g0 - (2.0 - z[f"/bin_{b}/scalar/density"][0])) < 1e-10
    assert 0.0 < summary["acceptance_rate"] < 1.0 and summary["max_precision_error"] < 1e-6
    assert summary["bins"] == 3 and summary["walkers"] == 8


def test_driver_checkpoint_and_resume(tmp_path):
    out, s1, _ = _run(tmp_path, "--checkpoint-every", "1")
    ck = np.load(out / "checkpoint_0.npz")
    assert int(ck["next_bin"]) == 2 and ck["fields"].shape == (8, 10, 16)
    assert sum(k.startswith("ds:/bin_1/") for k in ck.files) == 3 + 1 + 3  # bins 0 and 1 are in the checkpoint
    first = np.load(out / "data_0.npz")
    keep = {k: first[k] for k in first.files}
    out2, s2, _ = _run(tmp_path, "--checkpoint-every", "1", "--resume")
    z = np.load(out2 / "data_0.npz")
    assert set(z.files) == set(keep)
    for k in keep:  # bins 0 and 1 come from the checkpoint unchanged, bin 2 is re-run
        if "/bin_2/" not in k and "/binK_2/" not in k:
            assert np.array_equal(z[k], keep[k]), k
    assert s2["thermalisation_s"] < 1e-3  # no second thermalisation


def test_driver_mt19937_reproduces_the_reference_program(tmp_path):
    """`--rng mt19937 --walkers 1 --seed 11`: the whole program of source/main.cpp (random start from a copy of
    rng(11), 3 thermalisation sweeps, 3 bins of 2 sweeps with sweep_unequalTime + measure, accumulate) must write
    the datasets the reference's own DQMC + MeasurementManager classes recorded for the same parameter file
    (tests/golden/program.npz, made by tests/golden/make_golden.py:program_case).  No host random numbers: the
    libstdc++ stream is regenerated on the device.  Walker 1 of a 2-walker run is rng(12): its chain differs, and
    the bins are then the mean of two chains."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "program.npz"))
    c = dict(zip([str(k) for k in g["cfg_keys"]], g["cfg"]))
    assert (c["L1"], c["nt"], c["n_therms"], c["n_sweeps"], c["n_bins"], c["seed"]) == (4, 10, 3, 2, 3, 11)
    out, summary, _ = _run(tmp_path, "--rng", "mt19937", walkers=1)
    with np.load(out / "data_0.npz") as f:
        z = {k: f[k] for k in f.files}
    paths = sorted(k[3:] for k in g.files if k.startswith("ds:"))
    assert sorted(z) == paths
    for p in paths:
        want, got = g["ds:" + p], z[p]
        assert got.shape == want.shape, p
        tol = 1e-10 if "/scalar/" in p or "/equaltime/" in p else 1e-7  # tau > 0 goes through naive propagation on both sides
        assert np.max(np.abs(got - want)) <= tol * max(1.0, np.max(np.abs(want))), p
    assert abs(summary["acceptance_rate"] - float(g["acc_rate"]) / 18.0) < 1e-12  # 2 * (3 + 6) half sweeps
    out2, s2, _ = _run(tmp_path, "--rng", "mt19937", walkers=2)
    z2 = np.load(out2 / "data_0.npz")
    assert abs(z2["/bin_0/scalar/doubleOcc"][0] - z["/bin_0/scalar/doubleOcc"][0]) > 1e-6
