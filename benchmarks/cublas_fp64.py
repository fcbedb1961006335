"""Yardstick only (never on the product path): cuBLAS DGEMM 8192^3 through torch.matmul, the FP64 cross-check SURVEY.md
section 6 asks for next to benchmarks/peaks.cu's DMMA / DFMA microbenchmarks.  Prints one JSON object."""
import json
import torch

n = 8192
a = torch.randn(n, n, dtype=torch.float64, device="cuda")
b = torch.randn(n, n, dtype=torch.float64, device="cuda")
c = torch.empty_like(a)
for _ in range(2):
    torch.matmul(a, b, out=c)
best = 1e9
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    torch.matmul(a, b, out=c)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(json.dumps({"cublas_dgemm_8192_tflops": 2.0 * n ** 3 / (best * 1e-3) / 1e12, "ms": best, "how": "torch.matmul float64 8192^3, best of 5, CUDA events"}))
