"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/dqmc_b200.h declares, and fails loudly (no CPU fallback) when there is no CUDA device.
No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "dqmc_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dqmcb_[a-zA-Z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    from dqmc_b200 import capi

    L = capi.lib()
    decl = _declared_symbols()
    assert len(decl) >= 45
    for s in decl:
        assert hasattr(L, s), f"{s} declared in include/dqmc_b200.h but not exported by libdqmc_b200.so"
    assert sorted(capi.SYMBOLS) == decl, "dqmc_b200/capi.py SYMBOLS out of date with the header"
    assert b"sm_100a" in L.dqmcb_version()


def test_partner_rank_host_arithmetic():
    """update::partner_rank (source/update.cpp:34-45) against the reference-generated table."""
    from dqmc_b200 import capi

    g = np.load(os.path.join(ROOT, "tests", "golden", "misc.npz"))
    tab = g["partner_rank_ws248_r8_a1to4"]
    L = capi.lib()
    for wi, ws in enumerate((2, 4, 8)):
        for r in range(ws):
            for ai, a in enumerate(range(1, 5)):
                assert L.dqmcb_partner_rank(r, ws, a) == tab[wi, r, ai]


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from dqmc_b200 import capi, host

    n = 16
    e = np.eye(n)
    with pytest.raises(capi.DqmcbError) as ei:
        capi.Context(n, 8, 4, 2, 0.3, host.ETA, host.GAMMA, e, e)
    assert ei.value.code == capi.ECUDA
    with pytest.raises(capi.DqmcbError):
        capi.to_LDR(np.stack([e]))


def test_bad_arguments_rejected_before_cuda():
    from dqmc_b200 import capi, host

    e = np.eye(4)
    with pytest.raises(capi.DqmcbError) as ei:
        capi.Context(300 * 300, 8, 4, 2, 0.3, host.ETA, host.GAMMA, np.eye(2), np.eye(2))  # N > 256
    assert ei.value.code == capi.EINVAL


def test_product_does_not_import_oracle():
    """The product package must never route through the oracle (test infrastructure)."""
    for dp, _, fs in os.walk(os.path.join(ROOT, "dqmc_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"(import\s+oracle|from\s+oracle|#include\s*[<\"][^>\"]*oracle|libdqmc_ref|dlopen)", src), \
                    f"{f} reaches into the oracle"
