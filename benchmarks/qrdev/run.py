"""Development harness for the generic pivoted QR: times qr_mini on graded matrices and checks L diag(d) R = M. usage: run.py n batch"""
import ctypes as C, os, sys
import numpy as np, torch
here = os.path.dirname(os.path.abspath(__file__))
L = C.CDLL(os.path.join(here, "libqr_mini.so"))
L.qr_mini.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 5
n, b = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(1)
def graded():
    Q1, _ = np.linalg.qr(rng.standard_normal((n, n))); Q2, _ = np.linalg.qr(rng.standard_normal((n, n)))
    return (Q1 * np.exp(np.linspace(8, -8, n))) @ Q2
Ms = np.stack([graded() for _ in range(min(b, 4))])
Mall = np.concatenate([Ms] * ((b + len(Ms) - 1) // len(Ms)))[:b]
dev = "cuda"
dM = torch.from_numpy(np.ascontiguousarray(np.swapaxes(Mall, -1, -2))).to(dev)
dL = torch.empty_like(dM); dR = torch.empty_like(dM); dd = torch.empty((b, n), dtype=torch.float64, device=dev)
LD = n + 1 if n % 2 == 0 else n
ws = torch.empty(b * 3 * n * LD, dtype=torch.float64, device=dev)
args = (n, b, dM.data_ptr(), dL.data_ptr(), dd.data_ptr(), dR.data_ptr(), ws.data_ptr())
assert L.qr_mini(*args) == 0
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): L.qr_mini(*args)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
Lg = np.swapaxes(dL.cpu().numpy(), -1, -2); Rg = np.swapaxes(dR.cpu().numpy(), -1, -2); dg = dd.cpu().numpy()
err = max(np.max(np.abs((Lg[i] * dg[i]) @ Rg[i] - Mall[i])) / np.max(np.abs(Mall[i])) for i in range(min(b, 4)))
orth = np.max(np.abs(Lg[0].T @ Lg[0] - np.eye(n)))
import scipy.linalg as sl
_, Rref, _ = sl.qr(Mall[0], pivoting=True)
dref = np.abs(np.diag(Rref))
print(f"n={n} batch={b}: {ms:.3f} ms per launch ({ms*1e-3*1.965e9/1e6:.2f} M cycles), recon {err:.1e}, orth {orth:.1e}, d vs dgeqp3 {np.max(np.abs(dg[0]/dref-1)):.1e}")
