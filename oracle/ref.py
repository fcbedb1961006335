"""oracle/ref.py -- TEST INFRASTRUCTURE, not product code.

ctypes binding of oracle/_ref/libdqmc_ref.so: the reference's own hot-path sources compiled
unmodified (oracle/Makefile) behind the C harness oracle/ref_harness.cpp.  Used to pin
oracle/dqmc_oracle.py, to generate tests/golden/*.npz, and as bench.py's CPU baseline
(`cpu_baseline.kind == "reference"`).  The .so is git-ignored but travels to the GPU box.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# DQMCB_REF_VARIANT=fast selects the build with the reference's own Release flags (-O3 -ffast-math, CMakeLists.txt:7;
# `make -C oracle FAST=1` -> oracle/_ref_fast/): used by bench.py to report the CPU arm both ways
VARIANT = "fast" if os.environ.get("DQMCB_REF_VARIANT", "") == "fast" else "O3"
LIB_PATH = os.path.join(_HERE, "_ref_fast" if VARIANT == "fast" else "_ref", "libdqmc_ref.so")
_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_llp = C.POINTER(C.c_longlong)


def build(reference_root: str = "/root/reference", force: bool = False) -> bool:
    """Compile oracle/_ref when the reference tree is present (this container only)."""
    if not os.path.isdir(reference_root):  # GPU box: the prebuilt .so travels with the snapshot
        return os.path.exists(LIB_PATH)
    if force:
        subprocess.check_call(["make", "-C", _HERE, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", _HERE, f"REF={reference_root}"], stdout=subprocess.DEVNULL)  # incremental
    subprocess.check_call(["make", "-C", _HERE, f"REF={reference_root}", "FAST=1"], stdout=subprocess.DEVNULL)
    return os.path.exists(LIB_PATH)


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libdqmc_ref.so missing: run `make -C oracle` where /root/reference exists")
        import scipy.linalg  # noqa: F401  (makes sure libscipy_openblas is resolvable/loaded)

        L = C.CDLL(LIB_PATH)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_uint]
        for name in ("ref_g", "ref_logdet", "ref_acc_rate", "ref_max_err", "ref_mean_err", "ref_global_action"):
            getattr(L, name).restype = C.c_double
            getattr(L, name).argtypes = [C.c_void_p]
        L.ref_local_update.restype = C.c_double
        L.ref_rng_peek_canonical.restype = C.c_double
        L.ref_rng_peek_canonical.argtypes = [C.c_void_p]
        L.ref_local_update.argtypes = [C.c_void_p, C.c_int]
        L.ref_inv_I_plus_ldr.restype = C.c_double
        L.ref_run_sweeps.restype = C.c_double
        L.ref_run_sweeps.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int]
        L.ref_replica_exchange.argtypes = [C.POINTER(C.c_void_p), C.c_int]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _colmajor(a: np.ndarray) -> np.ndarray:
    """numpy [row, col] -> contiguous buffer in column-major order."""
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).T)


def _from_colmajor(buf: np.ndarray) -> np.ndarray:
    return buf.T.copy()


class RefSim:
    """One reference chain: AttractiveHubbard + DQMC + GF + LDRStack built exactly as
    source/main.cpp:85-106 does, with an injected seed."""

    def __init__(self, L1, L2, U, t, mu, beta, nt, n_stab, seed, unequal=False):
        self.L = lib()
        self.L.ref_create_ex.restype = C.c_void_p
        self.L.ref_create_ex.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_uint, C.c_int]
        self.h = C.c_void_p(self.L.ref_create_ex(L1, L2, U, t, mu, beta, nt, n_stab, seed, int(bool(unequal))))
        if not self.h:
            raise RuntimeError("ref_create failed")
        self.ns = self.L.ref_ns(self.h)
        self.nt = self.L.ref_nt(self.h)
        self.n_stack = self.L.ref_n_stack(self.h)
        self.n_stab = n_stab

    def __del__(self):
        try:
            if self.h:
                self.L.ref_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # --- state ---
    @property
    def g(self):
        return self.L.ref_g(self.h)

    def fields(self) -> np.ndarray:
        """fields[l, site] (int64)."""
        buf = np.empty((self.ns, self.nt), dtype=np.int64)  # column-major nt x N
        self.L.ref_get_fields(self.h, buf.ctypes.data_as(_llp))
        return buf.T.copy()

    def set_fields(self, f: np.ndarray, reinit: bool = True):
        buf = np.ascontiguousarray(np.asarray(f, dtype=np.int64).T)
        self.L.ref_set_fields(self.h, buf.ctypes.data_as(_llp))
        if reinit:
            self.L.ref_init(self.h)

    def reseed(self, seed: int):
        self.L.ref_reseed(self.h, seed)

    def _mat(self, fn, *args) -> np.ndarray:
        buf = np.empty((self.ns, self.ns))
        fn(self.h, *args, _d(buf))
        return _from_colmajor(buf)

    def expK(self):
        return self._mat(self.L.ref_get_expK)

    def invexpK(self):
        return self._mat(self.L.ref_get_invexpK)

    def expK_half(self):
        return self._mat(self.L.ref_get_expK_half)

    def invexpK_half(self):
        return self._mat(self.L.ref_get_invexpK_half)

    def G(self, l: int) -> np.ndarray:
        buf = np.empty((self.ns, self.ns))
        if self.L.ref_get_G(self.h, l, _d(buf)) != 0:
            raise IndexError(f"Gtt[{l}] not populated")
        return _from_colmajor(buf)

    def set_G(self, l: int, G: np.ndarray):
        self.L.ref_set_G(self.h, l, _d(_colmajor(G)))

    def logdet(self):
        return self.L.ref_logdet(self.h)

    def stack(self, i: int):
        Lb = np.empty((self.ns, self.ns))
        d = np.empty(self.ns)
        Rb = np.empty((self.ns, self.ns))
        self.L.ref_get_stack(self.h, i, _d(Lb), _d(d), _d(Rb))
        return _from_colmajor(Lb), d, _from_colmajor(Rb)

    def set_stack(self, i, Lm, d, Rm):
        self.L.ref_set_stack(self.h, i, _d(_colmajor(Lm)), _d(np.ascontiguousarray(d, dtype=np.float64)), _d(_colmajor(Rm)))

    # --- steps ---
    def init(self):
        self.L.ref_init(self.h)

    def sweep_fwd(self):
        self.L.ref_sweep_fwd(self.h)

    def sweep_bwd(self):
        self.L.ref_sweep_bwd(self.h)

    def local_update(self, l: int) -> float:
        return self.L.ref_local_update(self.h, l)

    def local_update_ratio(self, l, i, new):
        R = C.c_double()
        dl = C.c_double()
        self.L.ref_local_update_ratio(self.h, l, i, new, C.byref(R), C.byref(dl))
        return R.value, dl.value

    def update_greens_local(self, delta, l, i):
        self.L.ref_update_greens_local(self.h, C.c_double(delta), l, i)

    def wrap_fwd(self, l):
        self.L.ref_wrap_fwd(self.h, l)

    def wrap_bwd(self, l):
        self.L.ref_wrap_bwd(self.h, l)

    def Bbar(self, i_stack):
        return self._mat(self.L.ref_Bbar, i_stack)

    def expV(self, l):
        out = np.empty(self.ns)
        self.L.ref_expV(self.h, l, _d(out))
        return out

    def update_stack_fwd(self, i, Bbar):
        self.L.ref_update_stack_fwd(self.h, i, _d(_colmajor(Bbar)))

    def update_stack_bwd(self, i, Bbar):
        self.L.ref_update_stack_bwd(self.h, i, _d(_colmajor(Bbar)))

    def stabilize_fwd(self, l):
        self.L.ref_stabilize_fwd(self.h, l)

    def stabilize_bwd(self, l):
        self.L.ref_stabilize_bwd(self.h, l)

    def half_wrap(self):
        return self._mat(self.L.ref_half_wrap)

    # --- unequal time (source/dqmc.cpp:223-280,458-514; needs unequal=True at construction) ---
    def sweep_unequal(self):
        self.L.ref_sweep_unequal(self.h)

    def G_unequal(self, which: int, l: int) -> np.ndarray:
        """which: 0 Gtt[l], 1 Gt0[l], 2 G0t[l]."""
        buf = np.empty((self.ns, self.ns))
        if self.L.ref_get_G_unequal(self.h, which, l, _d(buf)) != 0:
            raise IndexError(f"G[{which}][{l}] not populated")
        return _from_colmajor(buf)

    def half_wrap_unequal(self) -> np.ndarray:
        """[3, nt, N, N]: half-warped Gtt, Gt0, G0t for t < nt."""
        buf = np.empty((3, self.nt, self.ns, self.ns))
        self.L.ref_half_wrap_unequal(self.h, _d(buf))
        return np.ascontiguousarray(np.swapaxes(buf, -1, -2))

    def observables_tau(self):
        """greenTau, doublonTau, currxxTau: [nt+1, N, N] each (source/model.cpp:290-392)."""
        out = [np.empty((self.nt + 1, self.ns, self.ns)) for _ in range(3)]
        self.L.ref_observables_tau(self.h, *[_d(o) for o in out])
        return tuple(np.ascontiguousarray(np.swapaxes(o, -1, -2)) for o in out)

    def global_action(self):
        return self.L.ref_global_action(self.h)

    def observables(self):
        out3 = np.empty(3)
        corr = np.empty((self.ns, self.ns))
        self.L.ref_observables(self.h, _d(out3), _d(corr))
        return out3[0], out3[1], out3[2], _from_colmajor(corr)

    def acc_rate(self):
        return self.L.ref_acc_rate(self.h)

    def max_err(self):
        return self.L.ref_max_err(self.h)

    def mean_err(self):
        return self.L.ref_mean_err(self.h)

    def peek_canonical(self) -> float:
        """u of the next rng.bernoulli() call (replica-exchange decider), generator not advanced."""
        return self.L.ref_rng_peek_canonical(self.h)

    def rng_tape(self, n_calls: int, advance: bool = False):
        """(perm, k, u), each [n_calls, N]: the draws the next n_calls local_update calls make."""
        perm = np.empty((n_calls, self.ns), dtype=np.int32)
        k = np.empty((n_calls, self.ns), dtype=np.int32)
        u = np.empty((n_calls, self.ns), dtype=np.float64)
        self.L.ref_rng_tape(self.h, n_calls, int(advance), perm.ctypes.data_as(_ip), k.ctypes.data_as(_ip), _d(u))
        return perm, k, u


def mt19937_raw(seed: int, n: int) -> np.ndarray:
    out = np.empty(n, dtype=np.uint32)
    lib().ref_mt19937_raw(C.c_uint(seed), n, out.ctypes.data_as(C.POINTER(C.c_uint)))
    return out


def _ldr_bufs(n):
    return np.empty((n, n)), np.empty(n), np.empty((n, n))


def to_LDR(M):
    n = M.shape[0]
    Lb, d, Rb = _ldr_bufs(n)
    if lib().ref_to_LDR(n, _d(_colmajor(M)), _d(Lb), _d(d), _d(Rb)) != 0:
        raise RuntimeError("QR decomposition failed in to_LDR")
    return _from_colmajor(Lb), d, _from_colmajor(Rb)


def _in(F):
    Lm, d, Rm = F
    return _d(_colmajor(Lm)), _d(np.ascontiguousarray(d, dtype=np.float64)), _d(_colmajor(Rm))


def mat_mul_ldr(M, F):
    n = M.shape[0]
    Lb, d, Rb = _ldr_bufs(n)
    keep = (_colmajor(M), _colmajor(F[0]), np.ascontiguousarray(F[1], dtype=np.float64), _colmajor(F[2]))
    lib().ref_mat_mul_ldr(n, _d(keep[0]), _d(keep[1]), _d(keep[2]), _d(keep[3]), _d(Lb), _d(d), _d(Rb))
    return _from_colmajor(Lb), d, _from_colmajor(Rb)


def ldr_mul_mat(F, M):
    n = M.shape[0]
    Lb, d, Rb = _ldr_bufs(n)
    keep = (_colmajor(F[0]), np.ascontiguousarray(F[1], dtype=np.float64), _colmajor(F[2]), _colmajor(M))
    lib().ref_ldr_mul_mat(n, _d(keep[0]), _d(keep[1]), _d(keep[2]), _d(keep[3]), _d(Lb), _d(d), _d(Rb))
    return _from_colmajor(Lb), d, _from_colmajor(Rb)


def ldr_mul_ldr(F1, F2):
    n = F1[0].shape[0]
    Lb, d, Rb = _ldr_bufs(n)
    k = (_colmajor(F1[0]), np.ascontiguousarray(F1[1], dtype=np.float64), _colmajor(F1[2]),
         _colmajor(F2[0]), np.ascontiguousarray(F2[1], dtype=np.float64), _colmajor(F2[2]))
    lib().ref_ldr_mul_ldr(n, *[_d(x) for x in k], _d(Lb), _d(d), _d(Rb))
    return _from_colmajor(Lb), d, _from_colmajor(Rb)


def inv_I_plus_ldr(F):
    n = F[0].shape[0]
    G = np.empty((n, n))
    k = (_colmajor(F[0]), np.ascontiguousarray(F[1], dtype=np.float64), _colmajor(F[2]))
    logdet = lib().ref_inv_I_plus_ldr(n, *[_d(x) for x in k], _d(G))
    return _from_colmajor(G), logdet


def inv_I_plus_ldr_mul_ldr(F1, F2):
    n = F1[0].shape[0]
    G = np.empty((n, n))
    k = (_colmajor(F1[0]), np.ascontiguousarray(F1[1], dtype=np.float64), _colmajor(F1[2]),
         _colmajor(F2[0]), np.ascontiguousarray(F2[1], dtype=np.float64), _colmajor(F2[2]))
    lib().ref_inv_I_plus_ldr_mul_ldr(n, *[_d(x) for x in k], _d(G))
    return _from_colmajor(G)


def inv_invldr_plus_ldr(F1, F2):
    n = F1[0].shape[0]
    G = np.empty((n, n))
    k = (_colmajor(F1[0]), np.ascontiguousarray(F1[1], dtype=np.float64), _colmajor(F1[2]),
         _colmajor(F2[0]), np.ascontiguousarray(F2[1], dtype=np.float64), _colmajor(F2[2]))
    lib().ref_inv_invldr_plus_ldr(n, *[_d(x) for x in k], _d(G))
    return _from_colmajor(G)


def partner_rank(rank, world_size, attempt):
    return lib().ref_partner_rank(rank, world_size, attempt)


def replica_exchange(sims: list[RefSim]):
    arr = (C.c_void_p * len(sims))(*[s.h for s in sims])
    lib().ref_replica_exchange(arr, len(sims))


def run_sweeps(sims: list[RefSim], n_sweeps: int) -> float:
    """Times n_sweeps full sweeps on every chain, one host thread per chain.  Returns seconds."""
    arr = (C.c_void_p * len(sims))(*[s.h for s in sims])
    return lib().ref_run_sweeps(arr, len(sims), n_sweeps)


class RefMeasure:
    """The reference's MeasurementManager + transform:: functions (include/measurementh5.h), compiled
    unmodified over the HDF5 recorder shim (oracle/ref_measure.cpp).  `datasets()` returns what
    results/data_<rank>.h5 would contain: {path: ndarray with the dataset's dims, C order}."""

    def __init__(self, L1: int, L2: int, rank: int = 0, unequal: bool = False):
        L = lib()
        L.refm_create_ex.restype = C.c_void_p
        L.refm_create_ex.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
        L.refm_measure_unequal.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp]
        for name in ("refm_destroy", "refm_accumulate"):
            getattr(L, name).argtypes = [C.c_void_p]
        L.refm_measure.argtypes = [C.c_void_p, _dp]
        L.refm_dataset_info.argtypes = [C.c_int, C.c_char_p, C.c_int, _llp, _ip]
        L.refm_dataset_data.argtypes = [C.c_int, _dp]
        L.refm_file_name.argtypes = [C.c_char_p, C.c_int]
        L.refm_chi_site_to_chi_r.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
        L.refm_chi_r_to_chi_k.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
        self.L, self.L1, self.L2, self.N = L, L1, L2, L1 * L2
        self.h = L.refm_create_ex(L1, L2, rank, int(bool(unequal)))
        if not self.h:
            raise RuntimeError("refm_create failed")

    def close(self):
        if self.h:
            self.L.refm_destroy(self.h)
            self.h = None

    def measure(self, G: np.ndarray) -> None:
        g = np.asfortranarray(G, dtype=np.float64)  # arma::mat is column-major
        self.L.refm_measure(self.h, g.ctypes.data_as(_dp))

    def measure_unequal(self, Gtt, Gt0, G0t) -> None:
        """measure() with the full GF of an unequal-time run: arrays [n_tau, N, N]."""
        bufs = [np.ascontiguousarray(np.swapaxes(np.asarray(x, dtype=np.float64), -1, -2)) for x in (Gtt, Gt0, G0t)]
        self.L.refm_measure_unequal(self.h, bufs[0].shape[0], *[b.ctypes.data_as(_dp) for b in bufs])

    def accumulate(self) -> None:
        self.L.refm_accumulate(self.h)

    def file_name(self) -> str:
        buf = C.create_string_buffer(512)
        self.L.refm_file_name(buf, 512)
        return buf.value.decode()

    def datasets(self) -> dict:
        out = {}
        for i in range(self.L.refm_n_datasets()):
            path = C.create_string_buffer(512)
            dims = (C.c_longlong * 4)()
            rank = C.c_int()
            n = self.L.refm_dataset_info(i, path, 512, dims, C.byref(rank))
            data = np.empty(n, dtype=np.float64)
            self.L.refm_dataset_data(i, data.ctypes.data_as(_dp))
            out[path.value.decode()] = data.reshape([int(dims[k]) for k in range(rank.value)])
        return out

    def chi_site_to_chi_r(self, chi_site: np.ndarray) -> np.ndarray:
        """chi_site [n_tau, N, N] (row i, col j) -> arma cube (L1, L2, n_tau) returned as [dx, dy, slice]."""
        c = np.asarray(chi_site, dtype=np.float64)
        if c.ndim == 2:
            c = c[None]
        nt = c.shape[0]
        buf = np.ascontiguousarray(np.stack([np.asfortranarray(s).ravel(order="F") for s in c]))
        out = np.empty(self.L1 * self.L2 * nt, dtype=np.float64)
        self.L.refm_chi_site_to_chi_r(self.h, nt, buf.ctypes.data_as(_dp), out.ctypes.data_as(_dp))
        return out.reshape((self.L1, self.L2, nt), order="F")

    def chi_r_to_chi_k(self, chi_r: np.ndarray) -> np.ndarray:
        """chi_r [L1, L2, n_slices] -> complex [L1, L2, n_slices]."""
        c = np.asfortranarray(chi_r, dtype=np.float64)
        nt = c.shape[2]
        out = np.empty(2 * c.size, dtype=np.float64)
        self.L.refm_chi_r_to_chi_k(self.h, nt, c.ravel(order="F").ctypes.data_as(_dp), out.ctypes.data_as(_dp))
        z = out[0::2] + 1j * out[1::2]
        return z.reshape((self.L1, self.L2, nt), order="F")

    def pbc_shortest(self, d: int, L: int) -> int:
        return int(self.L.refm_pbc_shortest(d, L))


def params_query(path: str, section: str, key: str, kind: int):
    """utility::parameters of the reference on `path`: kind 0 string, 1 int, 2 double, 3 bool, 4 vector,
    5 hasSection, 6 hasKey.  Returns the value, or RuntimeError when the reference throws."""
    L = lib()
    L.ref_params_query.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, _dp, C.c_char_p, C.c_int]
    num = np.zeros(32)
    buf = C.create_string_buffer(1024)
    if L.ref_params_query(path.encode(), section.encode(), key.encode(), kind, _d(num), buf, 1024) != 0:
        raise RuntimeError("reference parser threw")
    if kind == 0:
        return buf.value.decode()
    if kind == 4:
        return list(num[1:1 + int(num[0])])
    return {1: int, 2: float, 3: bool, 5: bool, 6: bool}[kind](num[0])
