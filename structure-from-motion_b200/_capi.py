"""ctypes binding of libsfmba.so (include/sfmba.h).  This is the stub a maintainer of the reference
would add next to ba_processor.py; nothing here computes anything.

The product path has NO CPU fallback: if the library is missing or there is no CUDA device the
calls raise `SfmbaError` immediately.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SFMBA_LIB: another build of the same library (A/B measurements); default = the in-tree build
LIB_PATH = os.environ.get('SFMBA_LIB') or os.path.join(_HERE, 'libsfmba.so')

SFMBA_MAX_COST_HISTORY = 64
NOALLOC_MAX_CAMS = 16384

OK = 0
ERR_INVALID_ARG, ERR_CUDA, ERR_NCCL, ERR_NUMERIC, ERR_UNSUPPORTED, ERR_NO_DEVICE, ERR_COMM = -1, -2, -3, -4, -5, -6, -7
FLAG_QW_NEAR_ZERO, FLAG_DEPTH_ZERO, FLAG_NONFINITE, FLAG_PCG_MAXITER, FLAG_SINGULAR = 1, 2, 4, 8, 16
FLAG_PEER_TIMEOUT, FLAG_BAD_INDEX = 32, 64
MODE_REFERENCE_FIXED, MODE_ADAPTIVE_LM = 0, 1
SOLVER_PCG, SOLVER_DENSE_CHOLESKY = 0, 1
PRECOND_JACOBI, PRECOND_SCHUR_JACOBI = 0, 1
MATVEC_AUTO, MATVEC_F64, MATVEC_F32 = 0, 1, 2
PROBLEM_CAM_U16 = 1
SOLVER_PCG = 0


class SfmbaError(RuntimeError):
    def __init__(self, code, message):
        super().__init__('libsfmba error %d: %s' % (code, message))
        self.code = code


class Problem(C.Structure):
    _fields_ = [('num_cams', C.c_int32), ('num_pts', C.c_int32), ('num_obs', C.c_int32), ('flags', C.c_int32),
                ('obs_uv', C.c_void_p), ('obs_cam', C.c_void_p), ('obs_pt', C.c_void_p), ('pt_rowptr', C.c_void_p),
                ('cam_kinv', C.c_void_p), ('cams', C.c_void_p), ('pts', C.c_void_p)]


class Options(C.Structure):
    _fields_ = [('lambda_', C.c_double), ('iterations', C.c_int32), ('mode', C.c_int32), ('solver', C.c_int32),
                ('pcg_max_iter', C.c_int32), ('pcg_tol', C.c_double), ('dvdc_sign_quirk', C.c_int32),
                ('force_slow_path', C.c_int32), ('compute_costs', C.c_int32), ('precond', C.c_int32),
                ('lm_lambda_min', C.c_double), ('lm_lambda_max', C.c_double), ('huber_delta', C.c_double),
                ('matvec_precision', C.c_int32), ('reuse_index', C.c_int32)]


class Result(C.Structure):
    _fields_ = [('initial_cost', C.c_double), ('final_cost', C.c_double), ('final_lambda', C.c_double),
                ('iterations_done', C.c_int32), ('iterations_accepted', C.c_int32), ('pcg_iterations', C.c_int32),
                ('flags', C.c_uint32), ('num_costs', C.c_int32), ('pcg_refine_iterations', C.c_int32),
                ('costs', C.c_double * SFMBA_MAX_COST_HISTORY),
                ('ms_total', C.c_float), ('ms_index', C.c_float), ('ms_linearize', C.c_float),
                ('ms_camreduce', C.c_float), ('ms_pcg', C.c_float), ('ms_update', C.c_float),
                ('kernel_launches', C.c_int32), ('matvec_precision', C.c_int32),
                ('relinearized', C.c_int32), ('reserved1', C.c_int32)]


# every symbol include/sfmba.h declares: name -> (restype, argtypes)
_vp, _i32, _f64 = C.c_void_p, C.c_int32, C.c_double
SYMBOLS = {
    'sfmba_create': (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, C.c_int, _vp]),
    'sfmba_destroy': (C.c_int, [_vp]),
    'sfmba_last_error': (C.c_char_p, [_vp]),
    'sfmba_version': (C.c_int, []),
    'sfmba_nccl_unique_id': (C.c_int, [_vp]),
    'sfmba_default_options': (None, [C.POINTER(Options)]),
    'sfmba_ba_solve': (C.c_int, [_vp, C.POINTER(Problem), C.POINTER(Options), C.POINTER(Result), _vp]),
    'sfmba_ba_solve_host': (C.c_int, [_vp, C.POINTER(Problem), C.POINTER(Options), C.POINTER(Result)]),
    'sfmba_residual_jacobian': (C.c_int, [_vp, C.POINTER(Problem), C.c_int, _vp, _vp, _vp, _vp, _vp, _vp]),
    'sfmba_residual_jacobian_f64': (C.c_int, [_vp, C.POINTER(Problem), C.c_int, _vp, _vp, _vp, _vp, _vp]),
    'sfmba_ba_cost': (C.c_int, [_vp, C.POINTER(Problem), C.POINTER(_f64), _vp]),
    'sfmba_ba_cost_noalloc': (C.c_int, [C.c_int, C.POINTER(Problem), _vp, _vp, _vp]),
    'sfmba_triangulate_refine_batched': (C.c_int, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f64, _i32,
                                                   C.POINTER(C.c_uint32), _vp]),
    'sfmba_triangulate_refine_batched_host': (C.c_int, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f64, _i32,
                                                        C.POINTER(C.c_uint32)]),
    'sfmba_triangulate_linear_batched': (C.c_int, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp,
                                                   C.POINTER(C.c_uint32), _vp]),
    'sfmba_triangulate_linear_batched_host': (C.c_int, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp,
                                                        C.POINTER(C.c_uint32)]),
    'sfmba_pnp_refine_batched': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f64, _i32, _i32, _i32,
                                           C.POINTER(C.c_uint32), _vp]),
    'sfmba_pnp_refine_batched_host': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f64, _i32, _i32, _i32,
                                                C.POINTER(C.c_uint32)]),
    'sfmba_pnp_ransac': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f64, _vp, _vp, _vp, _vp, _vp, _vp,
                                   C.POINTER(C.c_uint32), _vp]),
    'sfmba_pnp_ransac_host': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f64, _vp, _vp, _vp, _vp, _vp, _vp,
                                        C.POINTER(C.c_uint32)]),
    'sfmba_fund_ransac': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _f64, _vp, _vp, _vp, _vp, _vp,
                                    C.POINTER(C.c_uint32), _vp]),
    'sfmba_fund_ransac_host': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _f64, _vp, _vp, _vp, _vp, _vp,
                                         C.POINTER(C.c_uint32)]),
    'sfmba_build_visibility': (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(_i32), _vp]),
    'sfmba_profile_reset': (C.c_int, [_vp]),
    'sfmba_profile_enable': (C.c_int, [_vp, C.c_int]),
    'sfmba_profile_get': (C.c_int, [_vp, C.c_char_p, C.POINTER(_f64), C.POINTER(C.c_int64)]),
    'sfmba_total_launches': (C.c_int64, [_vp]),
}

_lib = None


def load_library(path: str | None = None):
    """dlopen libsfmba.so and type every entry point.  Raises SfmbaError if it is not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise SfmbaError(ERR_NO_DEVICE, 'libsfmba.so not built at %s (run __graft_entry__.build()); '
                                        'there is no CPU fallback' % p)
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def last_error(ctx=None) -> str:
    lib = load_library()
    msg = lib.sfmba_last_error(ctx)
    return msg.decode('utf-8', 'replace') if msg else ''


def check(code: int, ctx=None):
    if code != OK:
        raise SfmbaError(code, last_error(ctx))


def default_options() -> Options:
    opt = Options()
    load_library().sfmba_default_options(C.byref(opt))
    return opt
