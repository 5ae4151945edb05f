"""ctypes binding of libchoreo_b200.so (include/choreo_b200.h).

The product path loads ONLY the CUDA library built in-tree by ``choreographies2_b200.build`` and fails
loudly when it is missing or when no sm_100 device is visible: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libchoreo_b200.so")

CHB_OK = 0
ABI_VERSION = 2  # CHB_ABI_VERSION of include/choreo_b200.h
STATUS_NAMES = {
    1: "CHB_ERR_INVALID_ARGUMENT", 2: "CHB_ERR_INVALID_NINT", 3: "CHB_ERR_UNSUPPORTED", 4: "CHB_ERR_CUDA",
    5: "CHB_ERR_NO_DEVICE", 6: "CHB_ERR_WORKSPACE", 7: "CHB_ERR_NOT_READY",
}
# enum chb_query
Q_NINT, Q_NPARAMS, Q_SEGM_SIZE, Q_SEGM_STORE, Q_NSEGM, Q_GEODIM, Q_NLOOP, Q_NBIN, Q_DEVICE, Q_CHUNK, Q_LAUNCHES, \
    Q_FUSED, Q_SEGMPOS_SIZE, Q_HOST_POINTERS, Q_NPARAMS_INCL_O, Q_STAGED_TRANSFORM, Q_PAIR_KERNEL = range(17)
# enum chb_op
(OP_PARAMS_TO_SEGMPOS, OP_SEGMPOS_TO_PARAMS, OP_SEGMPOS_TO_PARAMS_T, OP_ACTION, OP_ACTION_GRAD, OP_ACTION_HESS,
 OP_SEGMPOS_ACTION, OP_SEGMPOS_ACTION_GRAD, OP_SEGMPOS_ACTION_HESS, OP_KIN, OP_POT, OP_POT_GRAD, OP_POT_HESS,
 OP_HASH, OP_PARAMS_TO_SEGMVEL) = range(15)

_p64 = C.POINTER(C.c_int64)
_p32 = C.POINTER(C.c_int32)
_pd = C.POINTER(C.c_double)


class chb_tables(C.Structure):
    _fields_ = [
        ("geodim", C.c_int32), ("nbody", C.c_int32), ("nloop", C.c_int32), ("nsegm", C.c_int32),
        ("nbin", C.c_int32), ("nint_min", C.c_int32), ("greater_nstore", C.c_int32), ("inter_law", C.c_int32),
        ("inter_law_params", C.c_double * 4),
        ("loopnb", _p64), ("loopmass", _pd), ("ncoeff_min", _p64), ("n_sub_fft", _p64), ("shortcut", _p32),
        ("nnz", _p64), ("nppl", _p64), ("nnz_k", _p64), ("nnz_k_shifts", _p64), ("co_in", _p32),
        ("co_in_shifts", _p64), ("params_basis", _pd), ("params_basis_shifts", _p64),
        ("alg_iint", _p64), ("alg_timerev", _p64), ("alg_spacerot", _pd),
        ("bodyloop", _p64),
        ("gensegm_to_body", _p64), ("gensegm_to_iintrel", _p64), ("inter_timerev", _p64),
        ("inter_spacerot", _pd), ("inter_spacerot_isid", _p32),
        ("bin_source", _p64), ("bin_target", _p64), ("bin_spacerot", _pd), ("bin_spacerot_isid", _p32),
        ("bin_prodchargesum", _pd),
        ("bodysegm", _p64), ("body_spacerot", _pd), ("body_timerev", _p64),  # ABI 2, optional (NULL)
    ]


# every symbol include/choreo_b200.h declares: name -> (restype, argtypes)
_vp = C.c_void_p
_sz = C.c_size_t
_i64 = C.c_int64
_dev_op = [_vp, _i64]  # (plan, batch, ...)
SYMBOLS = {
    "chb_abi_version": (C.c_int, []),
    "chb_last_error_string": (C.c_char_p, []),
    "chb_device_count": (C.c_int, []),
    "chb_plan_create": (C.c_int, [C.POINTER(chb_tables), C.c_int, C.POINTER(_vp)]),
    "chb_plan_set_nint": (C.c_int, [_vp, _i64]),
    "chb_plan_destroy": (C.c_int, [_vp]),
    "chb_plan_query": (C.c_int, [_vp, C.c_int, _p64]),
    "chb_plan_set_chunk": (C.c_int, [_vp, _i64]),
    "chb_plan_set_fused": (C.c_int, [_vp, C.c_int]),
    "chb_plan_set_split": (C.c_int, [_vp, C.c_int]),
    "chb_workspace_bytes": (C.c_int, [_vp, _i64, C.POINTER(_sz)]),
    "chb_params_to_segmpos_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_to_params_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_to_params_T_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_params_to_kin_nrg_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_to_pot_nrg_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_to_hash_dev": (C.c_int, _dev_op + [_vp, C.c_double, _vp, _vp, _sz, _vp]),
    "chb_compute_forces_dev": (C.c_int, [_vp, _i64, _vp, _vp, _vp]),
    "chb_compute_grad_forces_dev": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "chb_compute_forces_host": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "chb_arnoldi_step_dev": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "chb_row_stats_dev": (C.c_int, [_vp, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp]),
    "chb_gmres_solution_dev": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int, _vp, _vp, _vp, _vp, C.c_double, _vp, _vp]),
    "chb_segmpos_dparams_to_action_hess_masked_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_gmres_step_dev": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "chb_segmpos_dparams_to_action_hess_listed_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_to_pot_nrg_grad_dev": (C.c_int, _dev_op + [_vp, C.c_double, _vp, _vp, _sz, _vp]),
    "chb_segmpos_to_pot_nrg_hess_dev": (C.c_int, _dev_op + [_vp, _vp, C.c_double, _vp, _vp, _sz, _vp]),
    "chb_params_to_action_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_params_to_action_grad_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_params_to_action_hess_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_noalloc_scratch_bytes": (C.c_int, [C.POINTER(C.c_size_t)]),
    "chb_params_to_action_hess_noalloc_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp]),
    "chb_segmpos_params_to_action_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_params_to_action_grad_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_segmpos_dparams_to_action_hess_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_params_to_segmvel_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _sz, _vp]),
    "chb_params_changevar_dev": (C.c_int, _dev_op + [_vp, C.c_int, C.c_int, _vp, _vp]),
    "chb_segmpos_to_hash_multi_dev": (C.c_int, _dev_op + [_vp, C.c_int, _vp, _vp, _vp]),
    "chb_center_of_mass_dev": (C.c_int, _dev_op + [_vp, _vp, _vp]),
    "chb_detect_escape_dev": (C.c_int, _dev_op + [_vp, C.c_double, _vp, _vp]),
    "chb_path_stats_dev": (C.c_int, _dev_op + [_vp, _vp, _vp, _vp, _vp]),
    "chb_compute_velocities_dev": (C.c_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "chb_newton_workspace_bytes": (C.c_int, [_vp, _i64, C.c_int, C.POINTER(_sz)]),
    "chb_newton_krylov_dev": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "chb_run_host": (C.c_int, [_vp, C.c_int, _i64, _vp, _vp, C.c_double, _vp, _vp]),
    "chb_malloc_device": (C.c_int, [_vp, _sz, C.POINTER(_vp)]),
    "chb_free_device": (C.c_int, [_vp, _vp]),
    "chb_malloc_pinned": (C.c_int, [_sz, C.POINTER(_vp)]),
    "chb_free_pinned": (C.c_int, [_vp]),
    "chb_memcpy_h2d": (C.c_int, [_vp, _vp, _vp, _sz, _vp]),
    "chb_memcpy_d2h": (C.c_int, [_vp, _vp, _vp, _sz, _vp]),
    "chb_stream_synchronize": (C.c_int, [_vp, _vp]),
}


class chb_newton_opts(C.Structure):
    _fields_ = [("f_tol", C.c_double), ("smin", C.c_double), ("maxiter", C.c_int32), ("restart", C.c_int32),
                ("line_search", C.c_int32), ("tol_norm", C.c_int32)]


class ChoreoB200Error(RuntimeError):
    pass


def bind(handle: C.CDLL) -> C.CDLL:
    """Attach prototypes; raises AttributeError if the library lacks a declared symbol."""
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(handle, name)
        fn.restype = res
        fn.argtypes = args
    return handle


_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """The CUDA library (product path).  No fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise ChoreoB200Error(
                f"{LIB_PATH} is missing: build it with `python -m choreographies2_b200.build` "
                "(nvcc, sm_100a). There is no CPU fallback for this path.")
        _lib = bind(C.CDLL(LIB_PATH))
        if _lib.chb_abi_version() != ABI_VERSION:
            raise ChoreoB200Error("libchoreo_b200.so ABI version mismatch")
    return _lib


def check(lib: C.CDLL, status: int) -> None:
    if status == CHB_OK:
        return
    msg = lib.chb_last_error_string().decode("utf-8", "replace")
    if status == 2:
        raise ValueError(msg)  # same exception type as the reference nint setter (_NBodySyst.pyx:545-546)
    if status == 3:
        raise ValueError(f"unsupported on the B200 path: {msg}")
    raise ChoreoB200Error(f"{STATUS_NAMES.get(status, status)}: {msg}")


def tables_struct(T) -> "tuple[chb_tables, list]":
    """Fill a chb_tables from a PlanTables; returns the struct and the arrays that must stay alive."""
    keep = []

    def ptr(a, dtype, ctype):
        a = np.ascontiguousarray(a, dtype=dtype)
        if a.size == 0:
            a = np.zeros(1, dtype=dtype)
        keep.append(a)
        return a.ctypes.data_as(C.POINTER(ctype))

    s = chb_tables()
    s.geodim, s.nbody, s.nloop, s.nsegm, s.nbin = T.geodim, T.nbody, T.nloop, T.nsegm, T.nbin
    s.nint_min, s.greater_nstore, s.inter_law = T.nint_min, int(T.greater_nstore), int(T.inter_law)
    for i in range(4):
        s.inter_law_params[i] = float(T.inter_law_params[i])
    i64, i32, f64 = np.int64, np.int32, np.float64
    s.loopnb = ptr(T.loopnb, i64, C.c_int64)
    s.loopmass = ptr(T.loopmass, f64, C.c_double)
    s.ncoeff_min = ptr(T.ncoeff_min, i64, C.c_int64)
    s.n_sub_fft = ptr(T.n_sub_fft, i64, C.c_int64)
    s.shortcut = ptr(T.shortcut, i32, C.c_int32)
    s.nnz = ptr(T.nnz, i64, C.c_int64)
    s.nppl = ptr(T.nppl, i64, C.c_int64)
    s.nnz_k = ptr(T.nnz_k, i64, C.c_int64)
    s.nnz_k_shifts = ptr(T.nnz_k_shifts, i64, C.c_int64)
    s.co_in = ptr(T.co_in, i32, C.c_int32)
    s.co_in_shifts = ptr(T.co_in_shifts, i64, C.c_int64)
    s.params_basis = ptr(np.ascontiguousarray(T.params_basis, dtype=np.complex128).view(f64), f64, C.c_double)
    s.params_basis_shifts = ptr(T.params_basis_shifts, i64, C.c_int64)
    s.alg_iint = ptr(T.alg_iint, i64, C.c_int64)
    s.alg_timerev = ptr(T.alg_timerev, i64, C.c_int64)
    s.alg_spacerot = ptr(T.alg_spacerot, f64, C.c_double)
    s.bodyloop = ptr(T.bodyloop, i64, C.c_int64)
    s.gensegm_to_body = ptr(T.gensegm_to_body, i64, C.c_int64)
    s.gensegm_to_iintrel = ptr(T.gensegm_to_iintrel, i64, C.c_int64)
    s.inter_timerev = ptr(T.inter_timerev, i64, C.c_int64)
    s.inter_spacerot = ptr(T.inter_spacerot, f64, C.c_double)
    s.inter_spacerot_isid = ptr(T.inter_spacerot_isid, i32, C.c_int32)
    s.bin_source = ptr(T.bin_source, i64, C.c_int64)
    s.bin_target = ptr(T.bin_target, i64, C.c_int64)
    s.bin_spacerot = ptr(T.bin_spacerot, f64, C.c_double)
    s.bin_spacerot_isid = ptr(T.bin_spacerot_isid, i32, C.c_int32)
    s.bin_prodchargesum = ptr(T.bin_prodchargesum, f64, C.c_double)
    if getattr(T, "bodysegm", None) is not None:
        s.bodysegm = ptr(T.bodysegm, i64, C.c_int64)
        s.body_spacerot = ptr(T.body_spacerot, f64, C.c_double)
        s.body_timerev = ptr(T.body_timerev, i64, C.c_int64)
    return s, keep
