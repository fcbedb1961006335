"""dqmcb_to_LDR_noalloc (include/dqmc_b200.h): the allocation-free variant of dqmcb_to_LDR, whose kernel scratch is a
static device array.  No counterpart in the reference, so its oracle is the
existing entry point dqmcb_to_LDR on the same inputs: same kernel, same arithmetic -> compared at the 1e-10 of the
other parity tests (observed: bit-identical).  Two calls in a row on different inputs (the static array keeps the
previous call's contents), at a size whose work tiles live in that array (n = 144) and at one that keeps them in
shared memory (n = 36), then the size limit.  Collected last on purpose.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _graded(rng, n, span):
    Q1, _ = np.linalg.qr(rng.standard_normal((n, n)))
    Q2, _ = np.linalg.qr(rng.standard_normal((n, n)))
    return (Q1 * np.exp(np.linspace(span, -span, n))) @ Q2


def _rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@pytest.mark.parametrize("n,batch", [(144, 3), (36, 5)])
def test_to_LDR_noalloc_matches_to_LDR_on_two_calls_in_a_row(n, batch):
    import torch

    from dqmc_b200 import capi

    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(1000 + n)
    # caller-owned device buffers (allocated by the TEST, once; the entry point itself allocates nothing)
    dM = torch.empty((batch, n, n), dtype=torch.float64, device=dev)
    dL = torch.empty((batch, n, n), dtype=torch.float64, device=dev)
    dR = torch.empty((batch, n, n), dtype=torch.float64, device=dev)
    dd = torch.empty((batch, n), dtype=torch.float64, device=dev)
    for call in range(2):
        M = np.stack([_graded(rng, n, 4.0 + call) for _ in range(batch)])
        Lr, dr, Rr = capi.to_LDR(M)  # the existing entry point (host buffers, cudaMalloc inside)
        # column-major matrices = the transposes in torch's row-major layout
        dM.copy_(torch.from_numpy(np.ascontiguousarray(np.swapaxes(M, -1, -2))))
        dL.fill_(float("nan")); dR.fill_(float("nan")); dd.fill_(float("nan"))
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        capi.to_LDR_noalloc(dM.data_ptr(), dL.data_ptr(), dd.data_ptr(), dR.data_ptr(), n, batch, device=0, stream=0)
        torch.cuda.synchronize()
        if call == 1:  # the first call may load the module lazily; the second must leave device memory untouched
            assert torch.cuda.mem_get_info(0)[0] == free0
        Lg = np.swapaxes(dL.cpu().numpy(), -1, -2)
        Rg = np.swapaxes(dR.cpu().numpy(), -1, -2)
        dg = dd.cpu().numpy()
        assert np.all(np.isfinite(Lg)) and np.all(np.isfinite(Rg)) and np.all(np.isfinite(dg))
        for w in range(batch):
            assert _rel(Lg[w], Lr[w]) < TOL and _rel(Rg[w], Rr[w]) < TOL, f"call {call}, matrix {w}"
            assert np.max(np.abs(dg[w] / dr[w] - 1.0)) < TOL
            assert _rel((Lg[w] * dg[w]) @ Rg[w], M[w]) < 1e-12


def test_to_LDR_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    from dqmc_b200 import capi

    x = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    # n = 256: 3 tiles of 256 x 257 doubles = 1.5 MiB per matrix -> 6 matrices exceed the 8 MiB array
    with pytest.raises(capi.DqmcbError) as ei:
        capi.to_LDR_noalloc(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), 256, 6)
    assert ei.value.code == capi.EINVAL and "DQMCB_NOALLOC_SCRATCH_BYTES" in str(ei.value)
