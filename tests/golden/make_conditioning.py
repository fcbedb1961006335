"""tests/golden/make_conditioning.py -- how well-determined are the quantities the full-config parity tests compare?

For every tests/golden/traj_full_*.npz case, the start-up state (init_stacks + init_greenfunctions, a pure
function of the fields) is recomputed by the numpy restatement (oracle/dqmc_oracle.py, same dgeqp3/dgetrf as
the reference) with exp(-dtau K) perturbed by relative noise of ONE ULP (1.1e-16, three draws) -- a stand-in
for a different rounding order of the matrix products, which is all that separates two correct implementations.
The largest resulting change of G (relative max-norm), of log|det M| and of the sorted LDR scales d (relative,
element-wise) is the conditioning floor below which agreement between independent implementations cannot be
demanded; tests/test_gpu_full_configs.py uses max(1e-10, 30 x floor) as its tolerance and prints what it measured.

    python tests/golden/make_conditioning.py        ->  tests/golden/conditioning.json
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import dqmc_oracle as O  # noqa: E402

out = {}
for name in ("c1", "c2", "c2n5", "c3n5", "c3"):
    g = np.load(os.path.join(HERE, f"traj_full_{name}.npz"))
    L1, L2, U, t, mu, beta, nt, n_stab, seed, _ = g["cfg"]
    L1, L2, nt, n_stab = int(L1), int(L2), int(nt), int(n_stab)
    n = L1 * L2
    rng = np.random.default_rng(20261105)
    runs = []
    for trial in range(4):
        m = O.Model(L1, L2, float(U), float(t), float(mu), float(beta), nt, fields=g["fields0"].astype(np.int64).copy())
        noise = 0.0 if trial == 0 else 1.1e-16
        m.expK = g["expK"] * (1.0 + noise * rng.standard_normal((n, n)))
        m.invexpK = g["invexpK"] * (1.0 + noise * rng.standard_normal((n, n)))
        d = O.DQMC(m, n_stab)
        d.init_stacks()
        d.init_greenfunctions()
        runs.append(([np.sort(s.d) for s in d.stack], d.G.copy(), d.log_det_M))
    base = runs[0]
    sd = max(float(np.max(np.abs(a / b - 1.0))) for r in runs[1:] for a, b in zip(r[0], base[0]))
    sg = max(float(np.max(np.abs(r[1] - base[1])) / np.max(np.abs(base[1]))) for r in runs[1:])
    sl = max(abs(r[2] - base[2]) / max(1.0, abs(base[2])) for r in runs[1:])
    out[name] = {"d_rel": sd, "G_rel": sg, "logdet_rel": sl, "noise": 1.1e-16, "draws": 3}
    print(name, out[name], flush=True)
with open(os.path.join(HERE, "conditioning.json"), "w") as f:
    json.dump(out, f, indent=1)
