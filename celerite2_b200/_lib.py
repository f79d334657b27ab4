"""ctypes binding of libcelerite2_b200.so (the C ABI declared in include/celerite2_b200.h).

There is no CPU fallback: if the shared library is missing the import fails, and every compute call fails
when no CUDA device is visible.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcelerite2_b200.so")

_vp = ctypes.c_void_p
_i64 = ctypes.c_int64
_i32 = ctypes.c_int32
_sz = ctypes.c_size_t

OP_IDS = {
    "factor": 0, "factor_rev": 1, "solve_lower": 2, "solve_upper": 3, "matmul_lower": 4, "matmul_upper": 5,
    "solve_lower_rev": 6, "solve_upper_rev": 7, "matmul_lower_rev": 8, "matmul_upper_rev": 9,
    "general_matmul": 10, "assemble_rev": 11, "predict_variance": 12,
}
SHARED_T, SHARED_C, SHARED_Y, ACCUMULATE = 1, 2, 4, 8

# name -> argument types (return type is int unless listed in _RESTYPE); mirrors include/celerite2_b200.h
_SIGS = {
    "c2b_get_celerite_matrices": [_i32, _i64, _i32, _i32, _i64] + [_vp] * 12 + [_i32, _vp],
    "c2b_get_celerite_matrices_rev": [_i32, _i64, _i32, _i32, _i64] + [_vp] * 19 + [_i32, _vp, _sz, _vp],
    "c2b_factor": [_i64, _i32, _i64] + [_vp] * 9 + [_i32, _vp, _sz, _vp],
    "c2b_factor_rev": [_i64, _i32, _i64] + [_vp] * 15 + [_i32, _vp, _sz, _vp],
    "c2b_general_matmul_lower": [_i64, _i64, _i32, _i32] + [_vp] * 8 + [_vp, _sz, _vp],
    "c2b_general_matmul_upper": [_i64, _i64, _i32, _i32] + [_vp] * 8 + [_vp, _sz, _vp],
    "c2b_norm_logdet": [_i64, _i64] + [_vp] * 5 + [_vp, _sz, _vp],
    "c2b_norm_logdet_loglike": [_i64, _i64] + [_vp] * 6 + [_vp, _sz, _vp],
    "c2b_add_inplace": [_i64, _vp, _vp, _vp],
    "c2b_term_coefficients": [_i32, _i64, _vp, ctypes.c_double, _i32, _i32, _i32, _i32, _i32] + [_vp] * 6 + [_vp, _vp],
    "c2b_term_coefficients_rev": [_i32, _i64, _vp, ctypes.c_double, _i32, _i32, _i32, _i32, _i32] + [_vp] * 6 + [_vp, _vp],
    "c2b_term_product": [_i64, _i32, _i32] + [_vp] * 6 + [_i32, _i32] + [_vp] * 6 + [_vp] * 6 + [_vp],
    "c2b_term_product_rev": [_i64, _i32, _i32] + [_vp] * 6 + [_i32, _i32] + [_vp] * 6 + [_vp] * 18 + [_vp],
    "c2b_kernel_matrix": [_i64, _i64, _i32, _i32] + [_vp] * 10,
    "c2b_predict_variance": [_i64, _i64, _i32, _i32] + [_vp] * 10 + [_vp, _sz, _vp],
    "c2b_to_dense": [_i64, _i32] + [_vp] * 7,
    "c2b_apply_inverse": [_i64, _i32, _i32, _i64] + [_vp] * 7 + [_i32, _vp, _sz, _vp],
    "c2b_dot_tril": [_i64, _i32, _i32, _i64] + [_vp] * 7 + [_i32, _vp, _sz, _vp],
    "c2b_solve_lower_noalloc": [_i64, _i32, _i32, _i64] + [_vp] * 7 + [_i32, _vp],
}


class Segment(ctypes.Structure):
    """c2b_segment_t (include/celerite2_b200.h)"""
    _fields_ = [("phase", _i32), ("has_left", _i32), ("has_right", _i32), ("reserved", _i32),
                ("state_in", _vp), ("summary_out", _vp), ("edge_out", _vp)]


SEG_SUMMARISE, SEG_APPLY = 1, 2
_segp = ctypes.POINTER(Segment)
_SIGS["c2b_factor_segment"] = _SIGS["c2b_factor"] + [_segp]
_SIGS["c2b_factor_rev_segment"] = _SIGS["c2b_factor_rev"] + [_segp]
_SIGS["c2b_solve_lower_segment"] = [_i64, _i32, _i32, _i64] + [_vp] * 7 + [_i32, _vp, _sz, _vp, _segp]
_SIGS["c2b_solve_lower_rev_segment"] = [_i64, _i32, _i32, _i64] + [_vp] * 13 + [_i32, _vp, _sz, _vp, _segp]
_SIGS["c2b_solve_upper_segment"] = _SIGS["c2b_solve_lower_segment"]
_SIGS["c2b_solve_upper_rev_segment"] = _SIGS["c2b_solve_lower_rev_segment"]
_SIGS["c2b_segment_sizes"] = [_i32, _i32, _i32, ctypes.POINTER(_i64), ctypes.POINTER(_i64)]
_SIGS["c2b_segment_carry"] = [_i32, _i32, _i32, _i32, _i32, _vp, _vp, _vp]
for _op in ("solve_lower", "solve_upper", "matmul_lower", "matmul_upper"):
    _SIGS["c2b_" + _op] = [_i64, _i32, _i32, _i64] + [_vp] * 7 + [_i32, _vp, _sz, _vp]
    _SIGS["c2b_" + _op + "_rev"] = [_i64, _i32, _i32, _i64] + [_vp] * 13 + [_i32, _vp, _sz, _vp]

EXPORTS = sorted(list(_SIGS) + ["c2b_version", "c2b_last_error", "c2b_device_count", "c2b_scratch_bytes", "c2b_testing_set_chunks_per_series",
                                 "c2b_profile_enable", "c2b_profile_report", "c2b_noalloc_scratch_limit",
                                 "c2b_device_malloc", "c2b_device_free", "c2b_copy_to_device", "c2b_copy_to_host",
                                 "c2b_stream_synchronize"])


class C2BError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "celerite2_b200: %s is missing. Build it with `python -m celerite2_b200.build` (needs nvcc); "
            "there is no CPU fallback." % LIB_PATH
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, args in _SIGS.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = ctypes.c_int
    lib.c2b_version.restype = ctypes.c_char_p
    lib.c2b_version.argtypes = []
    lib.c2b_last_error.restype = ctypes.c_char_p
    lib.c2b_last_error.argtypes = []
    lib.c2b_device_count.restype = _i32
    lib.c2b_device_count.argtypes = []
    lib.c2b_scratch_bytes.restype = _sz
    lib.c2b_scratch_bytes.argtypes = [_i32, _i64, _i64, _i32, _i32, _i64]
    lib.c2b_profile_enable.restype = None
    lib.c2b_profile_enable.argtypes = [_i32]
    lib.c2b_profile_report.restype = ctypes.c_char_p
    lib.c2b_profile_report.argtypes = []
    lib.c2b_noalloc_scratch_limit.restype = _sz
    lib.c2b_noalloc_scratch_limit.argtypes = []
    lib.c2b_testing_set_chunks_per_series.restype = None
    lib.c2b_testing_set_chunks_per_series.argtypes = [_i64]
    return lib


lib = _load()


def check(rc):
    if rc != 0:
        raise C2BError("celerite2_b200 error %d: %s" % (rc, lib.c2b_last_error().decode()))


def version():
    return lib.c2b_version().decode()
