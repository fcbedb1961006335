"""The JSON line `bench.py` prints is a contract with the driver: these tests run it as the driver does (a subprocess, one
line on stdout) and check the keys and their meaning, not the numbers.  The reference arm runs on host cores only and is
part of the CPU suite; the product arm needs the GPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config", "e2e"]


def _run(args, timeout):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, f"bench.py must print exactly one line on stdout, got {len(lines)}"
    return json.loads(lines[0])


def _check_base(d, steps, warmup):
    for k in BASE_KEYS:
        assert k in d, k
    assert d["metric"] == "site_flips_per_sec" and d["unit"] == "site-flips/s" and d["higher_is_better"] is True
    assert d["steps"] == steps and d["warmup"] == warmup and d["n_gpus"] == 1
    assert d["scaling"] == "weak" and d["dtype"] == "f64" and d["data"] == "synthetic" and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("C1") and "model" not in d["config"]
    assert d["value"] > 0 and d["ms_per_step"] > 0
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in d["e2e"], k


def test_reference_arm_line():
    """`bench.py --impl reference`: the reference's own CPU sweep loop (oracle/_ref) on the host cores, same metric and
    config as the product arm, no GPU involved."""
    ref = os.path.join(ROOT, "oracle", "_ref", "libdqmc_ref.so")
    if not os.path.exists(ref) and not os.path.isdir("/root/reference"):
        pytest.skip("oracle/_ref not built and the reference sources are not on this machine")
    d = _run(["--impl", "reference", "--steps", "1", "--warmup", "1"], 600)
    _check_base(d, 1, 1)
    assert d["impl"] == "reference"
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_reference_arm_under_torchrun_prints_one_line_from_rank_0():
    """N > 1: the driver launches the reference arm under torchrun like the product arm; rank 0 alone runs and prints, the
    other ranks exit 0 without work."""
    ref = os.path.join(ROOT, "oracle", "_ref", "libdqmc_ref.so")
    if not os.path.exists(ref) and not os.path.isdir("/root/reference"):
        pytest.skip("oracle/_ref not built and the reference sources are not on this machine")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip().startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0


@pytest.mark.gpu
def test_product_arm_line():
    """`bench.py` on one GPU (short: 3 steps, 3 warm-up, 2 thermalisation sweeps, no CPU leg)."""
    d = _run(["--steps", "3", "--warmup", "3", "--therm", "2", "--no-cpu-baseline"], 900)
    _check_base(d, 3, 3)
    assert "impl" not in d or d["impl"] != "reference"
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert r["bound"] in ("hbm", "tensor") and r["unit"] in ("GB/s", "TFLOP/s")
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0.0 < r["frac"] < 1.0
    assert d["gpu_launches"] > 0
    # end to end: host buffers in and out every step, and a number of its own (not a copy of the device-timed value)
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0 and d["e2e"]["value"] != d["value"]
    c = d["clocks"]
    assert c["sm_mhz"] > 0 and c["sm_max_mhz"] >= c["sm_mhz"] and isinstance(c["reasons"], list)
    bad = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert not (bad & set(c["reasons"])), c
