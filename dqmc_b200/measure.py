"""Equal-time measurement pipeline: the reference's MeasurementManager
(include/measurementh5.h:120-361) over the batched device engine.

Reference flow (source/main.cpp:114-119,165-170): after every sweep `measure(greens, lat)` adds the
three scalars and the N x N density correlation of the one chain to host accumulators; every
`n_sweeps` sweeps `accumulate(lat)` divides by the count, folds site pairs onto displacements
(transform::chi_site_to_chi_r), Fourier-transforms (chi_r_to_chi_k) and appends
`/bin_<n>/{scalar,equaltime}` and `/binK_<n>/equaltime` to results/data_<rank>.h5.

Here `measure()` is one asynchronous device call for all W walkers of the context
(dqmcb_bin_measure: W samples per call, correlation accumulated already folded), `accumulate()` fetches
3 + L1*L2 doubles, optionally pools them over the ranks of a process group (sum of sums and of counts:
every walker of every GPU weighs the same), transforms on the device and records the same dataset
tree.  HDF5 does not exist in this image: `save()` writes the tree to an .npz whose keys are the HDF5
paths and whose arrays have the dims the reference's datasets have ([L1, L2, 1] and [L1, L2, 1, 2]).
"""
from __future__ import annotations

import os

import numpy as np

from . import capi

SCALARS = ("density", "doubleOcc", "swave")  # source/main.cpp:116-118
EQUAL_TIME = ("densityCorr",)                # source/main.cpp:119
UNEQUAL_TIME = ("greenTau", "doublonTau", "currxxTau")  # source/main.cpp:120-122 (isMeasureUnequalTime)


class MeasurementManager:
    def __init__(self, ctx: "capi.Context", L1: int, L2: int, rank: int = 0, group=None, pool_ranks: bool = False,
                 comm_device=None, out_dir: str = "results", device: int = 0, unequal_time: bool = False):
        if L1 * L2 != ctx.N:
            raise ValueError("L1*L2 must equal the number of sites of the context")
        self.ctx, self.L1, self.L2, self.rank = ctx, L1, L2, rank
        self.group, self.pool, self.comm_device = group, pool_ranks, comm_device
        self.out_dir, self.device = out_dir, device
        self.unequal_time = unequal_time  # [simulation] isMeasureUnequalTime
        self.current_bin = 0
        self.datasets: dict[str, np.ndarray] = {}
        ctx.bin_configure(L1, L2)

    def measure(self) -> None:
        """MeasurementManager::measure (include/measurementh5.h:189-229) for every walker of the context.
        With unequal_time the context must hold the result of sweep_unequal_time() of this sweep
        (source/main.cpp:158-165 calls sweep_unequalTime before measure)."""
        self.ctx.bin_measure()
        if self.unequal_time:
            self.ctx.bin_measure_unequal()

    def accumulate(self) -> dict:
        """MeasurementManager::accumulate (:231-274): close the bin, record its datasets, reset."""
        sc, chi, n = self.ctx.bin_fetch(reset=True)
        if self.pool:
            sc, chi, n = self._pool(sc, chi, n)
        if n == 0:
            raise RuntimeError("accumulate() without a measurement in the bin")
        b = self.current_bin
        new = {}
        for name, v in zip(SCALARS, sc / n):
            new[f"/bin_{b}/scalar/{name}"] = np.array([v])
        chi_r = (chi / n)[:, :, None]                                   # arma::cube (L1, L2, 1)
        chi_k = capi.chi_r_to_chi_k(chi_r, self.L1, self.L2, self.device)
        new[f"/bin_{b}/equaltime/{EQUAL_TIME[0]}"] = chi_r
        new[f"/binK_{b}/equaltime/{EQUAL_TIME[0]}"] = np.stack([chi_k.real, chi_k.imag], axis=-1)
        if self.unequal_time:
            sums, nu = self.ctx.bin_fetch_unequal(reset=True)      # [3, L1, L2, nt+1]
            if self.pool:
                sums, nu = self._pool_generic(sums, nu)
            for k, name in enumerate(UNEQUAL_TIME):
                cube = np.ascontiguousarray(sums[k] / nu)            # arma::cube (L1, L2, nt+1)
                ck = capi.chi_r_to_chi_k(cube, self.L1, self.L2, self.device)
                new[f"/bin_{b}/unequaltime/{name}"] = cube
                new[f"/binK_{b}/unequaltime/{name}"] = np.stack([ck.real, ck.imag], axis=-1)
        self.datasets.update(new)
        self.current_bin += 1
        return new

    def _pool_generic(self, sums, n):
        import torch
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return sums, n
        dev = self.comm_device if self.comm_device is not None else torch.device("cpu")
        buf = torch.from_numpy(np.concatenate([[float(n)], sums.ravel()])).to(dev)
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
        h = buf.cpu().numpy()
        return h[1:].reshape(sums.shape), int(round(h[0]))

    def _pool(self, sc, chi, n):
        import torch
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return sc, chi, n
        dev = self.comm_device if self.comm_device is not None else torch.device("cpu")
        buf = torch.from_numpy(np.concatenate([[float(n)], sc, chi.ravel(order="F")])).to(dev)
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)  # a few hundred bytes per bin
        h = buf.cpu().numpy()
        return h[1:4], h[4:].reshape((self.L1, self.L2), order="F"), int(round(h[0]))

    def file_name(self) -> str:
        return os.path.join(self.out_dir, f"data_{self.rank}.npz")  # results/data_<rank>.h5 in the reference

    def save(self, path: str | None = None) -> str:
        path = path or self.file_name()
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        np.savez(path, **{k: v for k, v in self.datasets.items()})
        return path
