"""ctypes binding of the C ABI in include/dqmc_b200.h (libdqmc_b200.so, built in-tree by
__graft_entry__.build()).  Thin and literal: one Python method per exported symbol, numpy arrays
in the ABI's layouts (FP64 column-major N x N, walker-major batches).  There is no fallback: if the
shared library is missing, importing `lib()` raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdqmc_b200.so")

OK, EINVAL, ECUDA, ENUMERIC, ESTATE = 0, 1, 2, 3, 4
RNG_TAPE, RNG_PHILOX, RNG_MT19937 = 0, 1, 2
PTR_FIELDS, PTR_G, PTR_LOGDET, PTR_OBS, PTR_STACK_L, PTR_STACK_D, PTR_STACK_R = 0, 1, 2, 3, 4, 5, 6


class Config(C.Structure):
    _fields_ = [
        ("n_sites", C.c_int),
        ("nt", C.c_int),
        ("n_stab", C.c_int),
        ("n_walkers", C.c_int),
        ("device", C.c_int),
        ("delay_k", C.c_int),
        ("g", C.c_double),
        ("alpha", C.c_double),
        ("eta", C.c_double * 4),
        ("gamma", C.c_double * 4),
    ]


_dp = C.POINTER(C.c_double)
_i8p = C.POINTER(C.c_int8)
_u8p = C.POINTER(C.c_uint8)
_lib = None

# every symbol include/dqmc_b200.h declares (tests check that the .so exports all of them)
SYMBOLS = [
    "dqmcb_create", "dqmcb_destroy", "dqmcb_last_error", "dqmcb_set_stream", "dqmcb_synchronize", "dqmcb_n_stack",
    "dqmcb_set_half_expK", "dqmcb_set_fields", "dqmcb_get_fields", "dqmcb_set_fields_i64", "dqmcb_get_fields_i64",
    "dqmcb_rng_mode", "dqmcb_rng_seed_mt19937", "dqmcb_set_tape", "dqmcb_init_stacks", "dqmcb_init_greenfunctions",
    "dqmcb_wrap_forward", "dqmcb_wrap_backward", "dqmcb_local_update", "dqmcb_block_product", "dqmcb_get_block_product",
    "dqmcb_stabilize_forward", "dqmcb_stabilize_backward", "dqmcb_sweep_0_to_beta", "dqmcb_sweep_beta_to_0",
    "dqmcb_get_G", "dqmcb_set_G", "dqmcb_get_ldr", "dqmcb_set_ldr", "dqmcb_get_logdet", "dqmcb_get_stats",
    "dqmcb_reset_stats", "dqmcb_device_ptr", "dqmcb_measure_equal_time", "dqmcb_global_action", "dqmcb_half_wrap",
    "dqmcb_run_sweeps_host", "dqmcb_to_LDR", "dqmcb_mat_mul_ldr", "dqmcb_ldr_mul_mat", "dqmcb_ldr_mul_ldr",
    "dqmcb_inv_I_plus_ldr", "dqmcb_inv_I_plus_ldr_mul_ldr", "dqmcb_partner_rank", "dqmcb_version",
    "dqmcb_timing_enable", "dqmcb_timing_get", "dqmcb_inv_invldr_plus_ldr", "dqmcb_local_update_ratio",
    "dqmcb_update_greens_local", "dqmcb_set_single_field", "dqmcb_launch_count", "dqmcb_n_groups",
    "dqmcb_bin_configure", "dqmcb_bin_measure", "dqmcb_bin_fetch", "dqmcb_chi_site_to_chi_r", "dqmcb_chi_r_to_chi_k",
    "dqmcb_pbc_shortest", "dqmcb_sweep_unequal_time", "dqmcb_get_G_unequal", "dqmcb_get_unequal_err",
    "dqmcb_half_wrap_unequal", "dqmcb_measure_unequal_time", "dqmcb_bin_measure_unequal", "dqmcb_bin_fetch_unequal",
    "dqmcb_set_tape_async", "dqmcb_tape_commit", "dqmcb_to_LDR_noalloc", "dqmcb_measure_kinetic_spin",
]


class DqmcbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"dqmc_b200 error {code}: {msg}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "dqmc_b200 has no CPU fallback."
            )
        L = C.CDLL(LIB_PATH)
        L.dqmcb_last_error.restype = C.c_char_p
        L.dqmcb_last_error.argtypes = [C.c_void_p]
        L.dqmcb_version.restype = C.c_char_p
        L.dqmcb_device_ptr.restype = C.c_void_p
        L.dqmcb_device_ptr.argtypes = [C.c_void_p, C.c_int]
        L.dqmcb_create.argtypes = [C.POINTER(Config), _dp, _dp, C.POINTER(C.c_void_p)]
        L.dqmcb_destroy.argtypes = [C.c_void_p]
        L.dqmcb_rng_mode.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
        L.dqmcb_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.dqmcb_launch_count.restype = C.c_longlong
        L.dqmcb_launch_count.argtypes = [C.c_void_p]
        L.dqmcb_n_groups.argtypes = [C.c_void_p]
        L.dqmcb_update_greens_local.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int]
        L.dqmcb_local_update_ratio.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp, _dp]
        L.dqmcb_set_single_field.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.dqmcb_bin_configure.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.dqmcb_bin_measure.argtypes = [C.c_void_p]
        L.dqmcb_bin_fetch.argtypes = [C.c_void_p, _dp, _dp, C.POINTER(C.c_longlong), C.c_int]
        L.dqmcb_chi_site_to_chi_r.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp]
        L.dqmcb_chi_r_to_chi_k.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp]
        L.dqmcb_pbc_shortest.argtypes = [C.c_int, C.c_int]
        L.dqmcb_to_LDR_noalloc.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.dqmcb_measure_kinetic_spin.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, _dp, _dp]
        L.dqmcb_sweep_unequal_time.argtypes = [C.c_void_p]
        L.dqmcb_get_G_unequal.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        L.dqmcb_get_unequal_err.argtypes = [C.c_void_p, _dp]
        L.dqmcb_half_wrap_unequal.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        L.dqmcb_measure_unequal_time.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _dp]
        L.dqmcb_bin_measure_unequal.argtypes = [C.c_void_p]
        L.dqmcb_bin_fetch_unequal.argtypes = [C.c_void_p, _dp, C.POINTER(C.c_longlong), C.c_int]
        for name in ("dqmcb_synchronize", "dqmcb_init_stacks", "dqmcb_init_greenfunctions", "dqmcb_sweep_0_to_beta",
                     "dqmcb_sweep_beta_to_0", "dqmcb_reset_stats", "dqmcb_n_stack"):
            getattr(L, name).argtypes = [C.c_void_p]
        for name in ("dqmcb_wrap_forward", "dqmcb_wrap_backward", "dqmcb_local_update", "dqmcb_block_product",
                     "dqmcb_stabilize_forward", "dqmcb_stabilize_backward"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_int]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _cm(a: np.ndarray) -> np.ndarray:
    """[..., row, col] numpy -> contiguous buffer holding column-major matrices."""
    return np.ascontiguousarray(np.swapaxes(np.asarray(a, dtype=np.float64), -1, -2))


def _from_cm(buf: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(np.swapaxes(buf, -1, -2))


class Context:
    """One dqmcb_ctx: W walkers of one model on one GPU."""

    def __init__(self, n_sites, nt, n_stab, n_walkers, g, eta, gamma, expK, invexpK, alpha=-1.0, device=0, delay_k=0):
        self.L = lib()
        cfg = Config()
        cfg.n_sites, cfg.nt, cfg.n_stab, cfg.n_walkers = n_sites, nt, n_stab, n_walkers
        cfg.device, cfg.delay_k, cfg.g, cfg.alpha = device, delay_k, g, alpha
        for i in range(4):
            cfg.eta[i] = float(eta[i])
            cfg.gamma[i] = float(gamma[i])
        self.N, self.nt, self.n_stab, self.W = n_sites, nt, n_stab, n_walkers
        h = C.c_void_p()
        ek, iek = _cm(expK), _cm(invexpK)
        rc = self.L.dqmcb_create(C.byref(cfg), _d(ek), _d(iek), C.byref(h))
        if rc != OK:
            raise DqmcbError(rc, self.L.dqmcb_last_error(None).decode())
        self.h = h
        self.n_stack = self.L.dqmcb_n_stack(self.h)

    def close(self):
        if getattr(self, "h", None):
            self.L.dqmcb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != OK:
            raise DqmcbError(rc, self.L.dqmcb_last_error(self.h).decode())

    # ---- fields ----
    def set_fields(self, f: np.ndarray):
        """f[W, nt, N] int8 in {0,1,2,3}."""
        f = np.ascontiguousarray(f, dtype=np.int8)
        assert f.shape == (self.W, self.nt, self.N)
        self._ck(self.L.dqmcb_set_fields(self.h, f.ctypes.data_as(_i8p)))

    def get_fields(self) -> np.ndarray:
        f = np.empty((self.W, self.nt, self.N), dtype=np.int8)
        self._ck(self.L.dqmcb_get_fields(self.h, f.ctypes.data_as(_i8p)))
        return f

    def set_fields_i64(self, walker: int, f_colmajor: np.ndarray):
        f = np.ascontiguousarray(f_colmajor, dtype=np.int64)
        self._ck(self.L.dqmcb_set_fields_i64(self.h, walker, f.ctypes.data_as(C.POINTER(C.c_longlong))))

    def get_fields_i64(self, walker: int) -> np.ndarray:
        f = np.empty(self.nt * self.N, dtype=np.int64)
        self._ck(self.L.dqmcb_get_fields_i64(self.h, walker, f.ctypes.data_as(C.POINTER(C.c_longlong))))
        return f

    # ---- rng ----
    def rng_mode(self, mode: int, seed: int = 0):
        self._ck(self.L.dqmcb_rng_mode(self.h, mode, seed))

    def rng_seed_mt19937(self, seeds):
        """Device-resident std::mt19937(seeds[w]) per walker: the reference's libstdc++ stream, no host tape."""
        s = np.ascontiguousarray(seeds, dtype=np.uint32)
        assert s.shape == (self.W,)
        self._ck(self.L.dqmcb_rng_seed_mt19937(self.h, s.ctypes.data_as(C.POINTER(C.c_uint32))))

    def set_tape(self, perm, k, u):
        """perm/k/u: [W, n_visits, N] in visiting order."""
        perm = np.ascontiguousarray(perm, dtype=np.uint8)
        k = np.ascontiguousarray(k, dtype=np.uint8)
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert perm.shape == k.shape == u.shape and perm.shape[0] == self.W and perm.shape[2] == self.N
        self._ck(self.L.dqmcb_set_tape(self.h, perm.shape[1], perm.ctypes.data_as(_u8p), k.ctypes.data_as(_u8p), _d(u)))

    def set_tape_async(self, perm, k, u):
        """Upload the next tape while the current step runs (arrays must be C-contiguous uint8/uint8/float64, ideally
        views of pinned memory, and stay alive until tape_commit)."""
        for a, dt in ((perm, np.uint8), (k, np.uint8), (u, np.float64)):
            assert a.dtype == dt and a.flags["C_CONTIGUOUS"]
        assert perm.shape == k.shape == u.shape and perm.shape[0] == self.W and perm.shape[2] == self.N
        self._pending_tape = (perm, k, u)
        self._ck(self.L.dqmcb_set_tape_async(self.h, perm.shape[1], perm.ctypes.data_as(_u8p), k.ctypes.data_as(_u8p), _d(u)))

    def tape_commit(self):
        self._ck(self.L.dqmcb_tape_commit(self.h))

    # ---- engine ----
    def set_half_expK(self, a, b):
        self._ck(self.L.dqmcb_set_half_expK(self.h, _d(_cm(a)), _d(_cm(b))))

    def init_stacks(self):
        self._ck(self.L.dqmcb_init_stacks(self.h))

    def init_greenfunctions(self):
        self._ck(self.L.dqmcb_init_greenfunctions(self.h))

    def wrap_forward(self, l):
        self._ck(self.L.dqmcb_wrap_forward(self.h, l))

    def wrap_backward(self, l):
        self._ck(self.L.dqmcb_wrap_backward(self.h, l))

    def local_update(self, l):
        self._ck(self.L.dqmcb_local_update(self.h, l))

    def block_product(self, i) -> np.ndarray:
        self._ck(self.L.dqmcb_block_product(self.h, i))
        out = np.empty((self.W, self.N, self.N))
        self._ck(self.L.dqmcb_get_block_product(self.h, _d(out)))
        return _from_cm(out)

    def stabilize_forward(self, l):
        self._ck(self.L.dqmcb_stabilize_forward(self.h, l))

    def stabilize_backward(self, l):
        self._ck(self.L.dqmcb_stabilize_backward(self.h, l))

    def sweep_0_to_beta(self):
        self._ck(self.L.dqmcb_sweep_0_to_beta(self.h))

    def sweep_beta_to_0(self):
        self._ck(self.L.dqmcb_sweep_beta_to_0(self.h))

    def synchronize(self):
        self._ck(self.L.dqmcb_synchronize(self.h))

    # ---- state ----
    def get_G(self) -> np.ndarray:
        out = np.empty((self.W, self.N, self.N))
        self._ck(self.L.dqmcb_get_G(self.h, _d(out)))
        return _from_cm(out)

    def set_G(self, G: np.ndarray):
        self._ck(self.L.dqmcb_set_G(self.h, _d(_cm(G))))

    def get_ldr(self, i):
        Lb = np.empty((self.W, self.N, self.N))
        d = np.empty((self.W, self.N))
        Rb = np.empty((self.W, self.N, self.N))
        self._ck(self.L.dqmcb_get_ldr(self.h, i, _d(Lb), _d(d), _d(Rb)))
        return _from_cm(Lb), d, _from_cm(Rb)

    def set_ldr(self, i, Lm, d, Rm):
        self._ck(self.L.dqmcb_set_ldr(self.h, i, _d(_cm(Lm)), _d(np.ascontiguousarray(d, dtype=np.float64)), _d(_cm(Rm))))

    def get_logdet(self) -> np.ndarray:
        out = np.empty(self.W)
        self._ck(self.L.dqmcb_get_logdet(self.h, _d(out)))
        return out

    def get_stats(self):
        a, m, e = np.empty(self.W), np.empty(self.W), np.empty(self.W)
        self._ck(self.L.dqmcb_get_stats(self.h, _d(a), _d(m), _d(e)))
        return a, m, e

    def reset_stats(self):
        self._ck(self.L.dqmcb_reset_stats(self.h))

    def timing_enable(self, on=True):
        self._ck(self.L.dqmcb_timing_enable(self.h, int(on)))

    def timing_get(self):
        ms = np.zeros(4)
        n = np.zeros(4, dtype=np.int64)
        self._ck(self.L.dqmcb_timing_get(self.h, _d(ms), n.ctypes.data_as(C.POINTER(C.c_longlong))))
        return ms, n

    def launch_count(self) -> int:
        return int(self.L.dqmcb_launch_count(self.h))

    def n_groups(self) -> int:
        return int(self.L.dqmcb_n_groups(self.h))

    def set_stream(self, cuda_stream_handle: int):
        self._ck(self.L.dqmcb_set_stream(self.h, C.c_void_p(cuda_stream_handle)))

    def device_ptr(self, which: int) -> int:
        return self.L.dqmcb_device_ptr(self.h, which)

    def measure_equal_time(self, want_corr=True):
        sc = np.empty((self.W, 3))
        corr = np.empty((self.W, self.N, self.N)) if want_corr else None
        self._ck(self.L.dqmcb_measure_equal_time(self.h, _d(sc), _d(corr)))
        return sc, (_from_cm(corr) if want_corr else None)

    def measure_kinetic_spin(self, L1: int, L2: int, t: float = 1.0):
        """(ekin[W], szz_r[W, L1, L2]): kinetic energy per site and (1/N) sum_i <Sz_i Sz_{i+(dx,dy)}> of every walker
        (no reference counterpart; see include/dqmc_b200.h)."""
        e, s = np.empty(self.W), np.empty((self.W, L1 * L2))
        self._ck(self.L.dqmcb_measure_kinetic_spin(self.h, L1, L2, float(t), _d(e), _d(s)))
        return e, s.reshape((self.W, L2, L1)).transpose(0, 2, 1)  # [w, dx, dy]

    def bin_configure(self, L1: int, L2: int) -> None:
        self._ck(self.L.dqmcb_bin_configure(self.h, L1, L2))
        self._bin_shape = (L1, L2)

    def bin_measure(self) -> None:
        """MeasurementManager::measure for all W walkers (device-side accumulation, asynchronous)."""
        self._ck(self.L.dqmcb_bin_measure(self.h))

    def bin_fetch(self, reset=True):
        """(scalar_sums[3], chi_r_sum[L1, L2], n_samples): sums since the last reset."""
        L1, L2 = self._bin_shape
        sc, chi = np.empty(3), np.empty(L1 * L2)
        n = C.c_longlong(0)
        self._ck(self.L.dqmcb_bin_fetch(self.h, _d(sc), _d(chi), C.byref(n), int(bool(reset))))
        return sc, chi.reshape((L1, L2), order="F"), int(n.value)

    # --- unequal time (source/dqmc.cpp:223-280,458-514) ---
    def sweep_unequal_time(self) -> None:
        self._ck(self.L.dqmcb_sweep_unequal_time(self.h))

    def get_G_unequal(self, which: int, l: int) -> np.ndarray:
        """which: 0 Gtt[l], 1 Gt0[l], 2 G0t[l] -> [W, N, N]."""
        out = np.empty((self.W, self.N, self.N))
        self._ck(self.L.dqmcb_get_G_unequal(self.h, which, l, _d(out)))
        return _from_cm(out)

    def get_unequal_err(self) -> np.ndarray:
        e = np.empty(self.W)
        self._ck(self.L.dqmcb_get_unequal_err(self.h, _d(e)))
        return e

    def half_wrap_unequal(self, which: int, l: int) -> np.ndarray:
        out = np.empty((self.W, self.N, self.N))
        self._ck(self.L.dqmcb_half_wrap_unequal(self.h, which, l, _d(out)))
        return _from_cm(out)

    def measure_unequal_time(self, L1: int, L2: int):
        """greenTau, doublonTau, currxxTau: [W, nt+1, N, N] each."""
        outs = [np.empty((self.W, self.nt + 1, self.N, self.N)) for _ in range(3)]
        self._ck(self.L.dqmcb_measure_unequal_time(self.h, L1, L2, *[_d(o) for o in outs]))
        return tuple(_from_cm(o) for o in outs)

    def bin_measure_unequal(self) -> None:
        self._ck(self.L.dqmcb_bin_measure_unequal(self.h))

    def bin_fetch_unequal(self, reset=True):
        """(chi_r_sums[3, L1, L2, nt+1] for greenTau/doublonTau/currxxTau, n_samples)."""
        L1, L2 = self._bin_shape
        buf = np.empty((3, self.nt + 1, L2, L1))
        n = C.c_longlong(0)
        self._ck(self.L.dqmcb_bin_fetch_unequal(self.h, _d(buf), C.byref(n), int(bool(reset))))
        return np.ascontiguousarray(buf.transpose(0, 3, 2, 1)), int(n.value)

    def global_action(self) -> np.ndarray:
        S = np.empty(self.W)
        self._ck(self.L.dqmcb_global_action(self.h, _d(S)))
        return S

    def local_update_ratio(self, l, site, new_field):
        R, dl = np.empty(self.W), np.empty(self.W)
        self._ck(self.L.dqmcb_local_update_ratio(self.h, l, site, new_field, _d(R), _d(dl)))
        return R, dl

    def update_greens_local(self, walker, delta, site):
        self._ck(self.L.dqmcb_update_greens_local(self.h, walker, C.c_double(delta), site))

    def set_single_field(self, walker, l, site, value):
        self._ck(self.L.dqmcb_set_single_field(self.h, walker, l, site, value))

    def half_wrap(self) -> np.ndarray:
        out = np.empty((self.W, self.N, self.N))
        self._ck(self.L.dqmcb_half_wrap(self.h, _d(out)))
        return _from_cm(out)

    def run_sweeps_host(self, fields_in, n_sweeps, reinit=False, want_G=False):
        """End-to-end call with HOST buffers (the `e2e` path of bench.py)."""
        fin = None if fields_in is None else np.ascontiguousarray(fields_in, dtype=np.int8)
        fout = np.empty((self.W, self.nt, self.N), dtype=np.int8)
        sc = np.empty((self.W, 3))
        G = np.empty((self.W, self.N, self.N)) if want_G else None
        self._ck(self.L.dqmcb_run_sweeps_host(self.h, int(reinit), None if fin is None else fin.ctypes.data_as(_i8p), n_sweeps,
                                              fout.ctypes.data_as(_i8p), _d(sc), _d(G)))
        return fout, sc, (_from_cm(G) if want_G else None)


# ---- stand-alone batched stablelinalg (host buffers) ----
def _b(a):
    a = np.asarray(a, dtype=np.float64)
    return a[None] if a.ndim == 2 else a


def _ckrc(rc):
    if rc != OK:
        raise DqmcbError(rc, lib().dqmcb_last_error(None).decode())


def to_LDR(M, device=0):
    M = _b(M)
    b, n = M.shape[0], M.shape[1]
    Lo, dO, Ro = np.empty((b, n, n)), np.empty((b, n)), np.empty((b, n, n))
    _ckrc(lib().dqmcb_to_LDR(device, n, b, _d(_cm(M)), _d(Lo), _d(dO), _d(Ro)))
    return _from_cm(Lo), dO, _from_cm(Ro)


def spin_structure_factor(szz_r: np.ndarray) -> np.ndarray:
    """S(q) on the grid q = 2 pi (n / L1, m / L2) from szz_r[..., dx, dy]: S(q) = sum_r cos(q.r) szz_r (real part of the DFT)."""
    return np.real(np.fft.fft2(szz_r, axes=(-2, -1)))


def to_LDR_noalloc(dM_ptr: int, dL_ptr: int, dd_ptr: int, dR_ptr: int, n: int, batch: int, device=0, stream=0):
    """dqmcb_to_LDR_noalloc: DEVICE pointers (ints), column-major matrices; asynchronous on `stream`; allocates nothing."""
    rc = lib().dqmcb_to_LDR_noalloc(device, n, batch, C.c_void_p(dM_ptr), C.c_void_p(dL_ptr), C.c_void_p(dd_ptr), C.c_void_p(dR_ptr),
                                    C.c_void_p(stream))
    if rc != OK:
        raise DqmcbError(rc, lib().dqmcb_last_error(None).decode())


def mat_mul_ldr(M, F, device=0):
    M, Lm, d, Rm = _b(M), _b(F[0]), np.atleast_2d(np.asarray(F[1], dtype=np.float64)), _b(F[2])
    b, n = M.shape[0], M.shape[1]
    Lo, dO, Ro = np.empty((b, n, n)), np.empty((b, n)), np.empty((b, n, n))
    k = (_cm(M), _cm(Lm), np.ascontiguousarray(d), _cm(Rm))
    _ckrc(lib().dqmcb_mat_mul_ldr(device, n, b, *[_d(x) for x in k], _d(Lo), _d(dO), _d(Ro)))
    return _from_cm(Lo), dO, _from_cm(Ro)


def ldr_mul_mat(F, M, device=0):
    M, Lm, d, Rm = _b(M), _b(F[0]), np.atleast_2d(np.asarray(F[1], dtype=np.float64)), _b(F[2])
    b, n = M.shape[0], M.shape[1]
    Lo, dO, Ro = np.empty((b, n, n)), np.empty((b, n)), np.empty((b, n, n))
    k = (_cm(Lm), np.ascontiguousarray(d), _cm(Rm), _cm(M))
    _ckrc(lib().dqmcb_ldr_mul_mat(device, n, b, *[_d(x) for x in k], _d(Lo), _d(dO), _d(Ro)))
    return _from_cm(Lo), dO, _from_cm(Ro)


def ldr_mul_ldr(F1, F2, device=0):
    L1, d1, R1 = _b(F1[0]), np.atleast_2d(np.asarray(F1[1], dtype=np.float64)), _b(F1[2])
    L2, d2, R2 = _b(F2[0]), np.atleast_2d(np.asarray(F2[1], dtype=np.float64)), _b(F2[2])
    b, n = L1.shape[0], L1.shape[1]
    Lo, dO, Ro = np.empty((b, n, n)), np.empty((b, n)), np.empty((b, n, n))
    k = (_cm(L1), np.ascontiguousarray(d1), _cm(R1), _cm(L2), np.ascontiguousarray(d2), _cm(R2))
    _ckrc(lib().dqmcb_ldr_mul_ldr(device, n, b, *[_d(x) for x in k], _d(Lo), _d(dO), _d(Ro)))
    return _from_cm(Lo), dO, _from_cm(Ro)


def inv_I_plus_ldr(F, device=0):
    L1, d1, R1 = _b(F[0]), np.atleast_2d(np.asarray(F[1], dtype=np.float64)), _b(F[2])
    b, n = L1.shape[0], L1.shape[1]
    G, ld = np.empty((b, n, n)), np.empty(b)
    k = (_cm(L1), np.ascontiguousarray(d1), _cm(R1))
    _ckrc(lib().dqmcb_inv_I_plus_ldr(device, n, b, *[_d(x) for x in k], _d(G), _d(ld)))
    return _from_cm(G), ld


def inv_I_plus_ldr_mul_ldr(F1, F2, device=0):
    L1, d1, R1 = _b(F1[0]), np.atleast_2d(np.asarray(F1[1], dtype=np.float64)), _b(F1[2])
    L2, d2, R2 = _b(F2[0]), np.atleast_2d(np.asarray(F2[1], dtype=np.float64)), _b(F2[2])
    b, n = L1.shape[0], L1.shape[1]
    G = np.empty((b, n, n))
    k = (_cm(L1), np.ascontiguousarray(d1), _cm(R1), _cm(L2), np.ascontiguousarray(d2), _cm(R2))
    _ckrc(lib().dqmcb_inv_I_plus_ldr_mul_ldr(device, n, b, *[_d(x) for x in k], _d(G)))
    return _from_cm(G)


def inv_invldr_plus_ldr(F1, F2, device=0):
    L1, d1, R1 = _b(F1[0]), np.atleast_2d(np.asarray(F1[1], dtype=np.float64)), _b(F1[2])
    L2, d2, R2 = _b(F2[0]), np.atleast_2d(np.asarray(F2[1], dtype=np.float64)), _b(F2[2])
    b, n = L1.shape[0], L1.shape[1]
    G = np.empty((b, n, n))
    k = (_cm(L1), np.ascontiguousarray(d1), _cm(R1), _cm(L2), np.ascontiguousarray(d2), _cm(R2))
    _ckrc(lib().dqmcb_inv_invldr_plus_ldr(device, n, b, *[_d(x) for x in k], _d(G)))
    return _from_cm(G)


def chi_site_to_chi_r(chi_site, L1, L2, device=0):
    """transform::chi_site_to_chi_r on the device: chi_site [n_tau, N, N] (or [N, N]) -> [L1, L2, n_tau]."""
    c = _b(chi_site)
    nm = c.shape[0]
    out = np.empty((nm, L1 * L2))
    _ckrc(lib().dqmcb_chi_site_to_chi_r(device, L1, L2, nm, _d(_cm(c)), _d(out)))
    return np.ascontiguousarray(out.reshape((nm, L2, L1)).transpose(2, 1, 0))


def chi_r_to_chi_k(chi_r, L1, L2, device=0):
    """transform::chi_r_to_chi_k on the device: chi_r [L1, L2, n_slices] -> complex [L1, L2, n_slices]."""
    c = np.asarray(chi_r, dtype=np.float64)
    if c.ndim == 2:
        c = c[:, :, None]
    nm = c.shape[2]
    buf = np.ascontiguousarray(c.transpose(2, 1, 0))  # [slice][y][x]: x fastest = arma cube memory order
    out = np.empty((nm, L1 * L2, 2))
    _ckrc(lib().dqmcb_chi_r_to_chi_k(device, L1, L2, nm, _d(buf), _d(out)))
    z = out[..., 0] + 1j * out[..., 1]
    return np.ascontiguousarray(z.reshape((nm, L2, L1)).transpose(2, 1, 0))


def pbc_shortest(d, L):
    return lib().dqmcb_pbc_shortest(d, L)


def partner_rank(rank, world_size, attempt):
    return lib().dqmcb_partner_rank(rank, world_size, attempt)
