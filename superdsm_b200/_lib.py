"""ctypes binding of ``include/superdsm_b200.h``.  There is no fallback: if the CUDA library is missing the import of
the solver fails loudly."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('SDSM_LIB') or os.path.join(HERE, 'libsuperdsm_b200.so')      # SDSM_LIB: A/B builds (scripts/)

SDSM_OK = 0
ST_OPTIMAL, ST_FALLBACK, ST_SKIPPED, ST_ERROR, ST_BADREGION = range(5)

#: numpy mirror of ``struct sdsm_cand``
CAND_DTYPE = np.dtype([('r0', '<i4'), ('c0', '<i4'), ('h', '<i4'), ('w', '<i4'), ('fp_off', '<i4'), ('fp_len', '<i4'),
                       ('n', '<i4'), ('cell_cap', '<i4'), ('img', '<i4'), ('reserved', '<i4'), ('pix_off', '<i8'), ('cell_off', '<i8'),
                       ('tab_off', '<i8'), ('par_off', '<i8'), ('fg_off', '<i8'), ('vec_off', '<i8'), ('mat_off', '<i8')], align=True)
assert CAND_DTYPE.itemsize == 96
ABI_VERSION = 5
MAX_IMAGES = 32


class Params(ctypes.Structure):
    _fields_ = [('alpha', ctypes.c_double), ('epsilon', ctypes.c_double), ('scale', ctypes.c_double),
                ('sparsity_tol', ctypes.c_double), ('subsample', ctypes.c_int32), ('radius', ctypes.c_int32),
                ('has_smooth', ctypes.c_int32), ('init_elliptical', ctypes.c_int32), ('h_psf', ctypes.c_void_p),
                ('maxiters', ctypes.c_int32), ('skip_single_fg', ctypes.c_int32), ('abstol', ctypes.c_double),
                ('reltol', ctypes.c_double), ('feastol', ctypes.c_double), ('hessian_sparsity_tol', ctypes.c_double)]


class Plan(ctypes.Structure):
    _fields_ = [('workspace_bytes', ctypes.c_size_t), ('params_len', ctypes.c_int64), ('fgmask_len', ctypes.c_int64),
                ('n_max', ctypes.c_int32), ('cells_max', ctypes.c_int32), ('n_cand', ctypes.c_int32),
                ('reserved', ctypes.c_int32), ('totals', ctypes.c_int64 * 6), ('solo_share', ctypes.c_double)]


class ImageDesc(ctypes.Structure):
    _fields_ = [('d_y', ctypes.c_void_p), ('d_atoms', ctypes.c_void_p), ('H', ctypes.c_int32), ('W', ctypes.c_int32)]


class Out(ctypes.Structure):
    _fields_ = [('d_energy', ctypes.c_void_p), ('d_status', ctypes.c_void_p), ('d_info', ctypes.c_void_p),
                ('d_params', ctypes.c_void_p), ('d_fgbox', ctypes.c_void_p), ('d_onboundary', ctypes.c_void_p),
                ('d_fgmask', ctypes.c_void_p), ('d_solver', ctypes.c_void_p)]


EXPORTS = ('sdsm_abi_version', 'sdsm_create', 'sdsm_destroy', 'sdsm_last_error', 'sdsm_set_image', 'sdsm_set_images', 'sdsm_set_params',
           'sdsm_plan_batch', 'sdsm_solve_batch', 'sdsm_eval_candidate', 'sdsm_export_smooth', 'sdsm_trace_solve',
           'sdsm_dfma_peak', 'sdsm_launch_count', 'sdsm_solver_stats', 'sdsm_sweep_eval', 'sdsm_finish_times', 'sdsm_prepare_image',
           'sdsm_preprocess', 'sdsm_prepare_image_noalloc', 'sdsm_eval_counts')

_lib = None


def load():
    """Loads the shared library (once).  Raises if it has not been built: there is no CPU path."""
    global _lib
    if _lib is not None: return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(f'{LIB_PATH} is missing: build it with `python -m superdsm_b200.build` (needs nvcc); '
                          'superdsm_b200 has no CPU fallback')
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64
    lib.sdsm_abi_version.restype = ctypes.c_int
    lib.sdsm_create.argtypes = [ctypes.POINTER(vp), ctypes.c_int]
    lib.sdsm_destroy.argtypes = [vp]
    lib.sdsm_destroy.restype = None
    lib.sdsm_last_error.argtypes = [vp]
    lib.sdsm_last_error.restype = ctypes.c_char_p
    lib.sdsm_set_image.argtypes = [vp, vp, vp, i32, i32]
    lib.sdsm_set_images.argtypes = [vp, ctypes.POINTER(ImageDesc), i32]
    lib.sdsm_set_params.argtypes = [vp, ctypes.POINTER(Params)]
    lib.sdsm_prepare_image.argtypes = [vp, vp, vp, vp, i32, i32, ctypes.c_double, i32, vp, vp, vp, vp, vp]
    lib.sdsm_prepare_image_noalloc.argtypes = [vp, vp, vp, vp, i32, i32, ctypes.c_double, i32, vp, vp, vp, vp]
    lib.sdsm_preprocess.argtypes = [vp, vp, i32, i32, vp, i32, vp, i32, ctypes.c_double, ctypes.c_double, ctypes.c_double, vp, vp, vp]
    lib.sdsm_plan_batch.argtypes = [vp, vp, i32, ctypes.POINTER(Plan)]
    lib.sdsm_solve_batch.argtypes = [vp, ctypes.POINTER(Plan), vp, i32, vp, vp, vp, ctypes.c_size_t, ctypes.POINTER(Out), vp]
    lib.sdsm_eval_candidate.argtypes = [vp, ctypes.POINTER(Plan), vp, i32, vp, vp, vp, vp, vp, vp]
    lib.sdsm_export_smooth.argtypes = [vp, ctypes.POINTER(Plan), vp, i32, vp, vp, vp, vp, vp, vp]
    lib.sdsm_trace_solve.argtypes = [vp, ctypes.POINTER(Plan), vp, i32, i32, vp, vp, vp, vp, vp, vp]
    lib.sdsm_dfma_peak.argtypes = [vp, vp, i32, ctypes.POINTER(ctypes.c_int64), vp]
    lib.sdsm_launch_count.argtypes = [vp]
    lib.sdsm_launch_count.restype = i64
    lib.sdsm_solver_stats.argtypes = [vp, ctypes.POINTER(Plan), vp, vp, vp]
    lib.sdsm_finish_times.argtypes = [vp, ctypes.POINTER(Plan), vp, vp, vp]
    if hasattr(lib, 'sdsm_eval_counts'): lib.sdsm_eval_counts.argtypes = [vp, ctypes.POINTER(Plan), vp, vp, vp]      # (absent from older builds named by SDSM_LIB)
    lib.sdsm_sweep_eval.argtypes = [vp, ctypes.POINTER(Plan), vp, i32, i32, i32, i32, vp, vp]
    if lib.sdsm_abi_version() != ABI_VERSION:
        raise ImportError('libsuperdsm_b200.so: ABI version mismatch')
    _lib = lib
    return lib
