"""ctypes mirror of include/aircraft_b200.h (structs, constants, prototypes) and the loader of
libaircraft_b200.so.  There is no fallback: if the CUDA library is missing or does not load,
importing the compute path raises."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

NX, NU = 13, 7
LAYOUT_SOA, LAYOUT_AOS = 0, 1
FLAG_NORMALISE, FLAG_SHARED_U, FLAG_COOPERATIVE, FLAG_PACKED = 1, 2, 4, 8
OK, EINVAL, ECUDA, ENODEV, ENOMEM, EUNSUPPORTED = 0, -1, -2, -3, -4, -5
MODEL_POLY, MODEL_LINEAR, MODEL_DEFAULT = 1, 2, 3

LIB_NAME = "libaircraft_b200.so"
LIB_PATH = Path(__file__).resolve().parent / LIB_NAME


class AircParams(C.Structure):
    _fields_ = [
        ("mass", C.c_double), ("span", C.c_double), ("chord", C.c_double),
        ("reference_area", C.c_double), ("com", C.c_double * 3),
        ("inertia", C.c_double * 9), ("inertia_inv", C.c_double * 9),
        ("gravity", C.c_double * 3), ("epsilon", C.c_double),
        ("stall_scaling", C.c_int32), ("reserved", C.c_int32),
    ]


class AircNNBlob(C.Structure):
    _fields_ = [
        ("W1", C.c_float * 80), ("b1", C.c_float * 16),
        ("W2", C.c_float * 512), ("b2", C.c_float * 32),
        ("W3", C.c_float * 192), ("b3", C.c_float * 6),
        ("input_mean", C.c_float * 5), ("input_std", C.c_float * 5),
        ("output_mean", C.c_float * 6), ("output_std", C.c_float * 6),
    ]


class AircCostParams(C.Structure):
    _fields_ = [
        ("goal", C.c_float * 2), ("vel_param", C.c_float), ("w_goal", C.c_float),
        ("w_du", C.c_float), ("du_eps", C.c_float), ("w_height", C.c_float),
        ("w_speed", C.c_float), ("w_vx", C.c_float), ("w_vyz", C.c_float),
    ]


class AircSweepParams(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64), ("index_offset", C.c_uint64),
        ("sigma0", C.c_float * 4), ("sigma", C.c_float * 4),
        ("lo", C.c_float * 4), ("hi", C.c_float * 4),
        ("ckpt_interval", C.c_int32), ("reserved", C.c_int32),
    ]


SWEEP_RECORD_HEADER = 4
SWEEP_STATIC_WORKSPACE_BYTES = 32 << 20   # AIRC_SWEEP_STATIC_WORKSPACE_BYTES
NDIAG = 22
NZ = 21          # z = (x 13, u 7, dt): AIRC_NZ
DIAG_SLICES = {"v_frd_rel": slice(0, 3), "airspeed": 3, "alpha": 4, "beta": 5, "qbar": 6,
               "coefficients": slice(7, 13), "forces_frd": slice(13, 16), "moments_frd": slice(16, 19),
               "phi": 19, "theta": 20, "psi": 21}

assert C.sizeof(AircNNBlob) == 860 * 4
assert C.sizeof(AircSweepParams) == 88

_vp, _i64, _i32, _u32, _f = C.c_void_p, C.c_int64, C.c_int, C.c_uint32, C.c_float
_PP = C.POINTER(AircParams)
_NP = C.POINTER(AircNNBlob)
_CP = C.POINTER(AircCostParams)
_SP = C.POINTER(AircSweepParams)

# name -> (restype, argtypes); every symbol include/aircraft_b200.h declares
PROTOTYPES = {
    "airc_version": (C.c_int, []),
    "airc_build_info": (C.c_char_p, []),
    "airc_create": (C.c_int, [_PP, _NP, _i32, C.POINTER(_vp)]),
    "airc_destroy": (C.c_int, [_vp]),
    "airc_last_error": (C.c_char_p, [_vp]),
    "airc_create_model": (C.c_int, [_PP, _i32, _vp, _i32, _i32, C.POINTER(_vp)]),
    "airc_state_derivative": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp, _vp]),
    "airc_diagnostics": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp, _vp]),
    "airc_diagnostics_host": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp]),
    "airc_state_update": (C.c_int, [_vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp]),
    "airc_rollout": (C.c_int, [_vp, _vp, _vp, _f, _i64, _i32, _i32, _u32, _i32, _vp, _vp, _vp,
                               _CP, _vp]),
    "airc_linearize": (C.c_int, [_vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp, _vp,
                                 _vp]),
    "airc_linearize_dt": (C.c_int, [_vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp, _vp,
                                    _vp, _vp]),
    "airc_defects": (C.c_int, [_vp, _vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "airc_hessian": (C.c_int, [_vp, _vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp, _vp, _vp]),
    "airc_hessian_host": (C.c_int, [_vp, _vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp, _vp]),
    "airc_argmin": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp]),
    "airc_sweep_record_floats": (C.c_int64, [_i32]),
    "airc_sweep_workspace_bytes": (C.c_size_t, [_i64, _i32, _i32]),
    "airc_sweep": (C.c_int, [_vp, _vp, _vp, _f, _i64, _i32, _i32, _u32, _SP, _CP, _vp, C.c_size_t,
                             _vp, _vp, _vp]),
    "airc_sweep_static": (C.c_int, [_vp, _vp, _vp, _f, _i64, _i32, _i32, _u32, _SP, _CP, _vp, _vp, _vp]),
    "airc_select_record": (C.c_int, [_vp, _vp, _i32, _i64, _vp, _vp]),
    "airc_sweep_controls": (C.c_int, [_vp, _vp, _i32, _SP, _i64, _i64, _i32, _vp, _vp]),
    "airc_sweep_host": (C.c_int, [_vp, _vp, _vp, _f, _i64, _i32, _i32, _u32, _SP, _CP, _vp, _vp]),
    "airc_measure_fp32_peak": (C.c_int, [_vp, _i32, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "airc_state_derivative_host": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _vp]),
    "airc_state_update_host": (C.c_int, [_vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp]),
    "airc_rollout_host": (C.c_int, [_vp, _vp, _vp, _f, _i64, _i32, _i32, _u32, _i32, _vp, _vp,
                                    _vp, _CP]),
    "airc_linearize_host": (C.c_int, [_vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp,
                                      _vp]),
    "airc_linearize_dt_host": (C.c_int, [_vp, _vp, _vp, _f, _vp, _i64, _i32, _u32, _i32, _vp, _vp,
                                         _vp, _vp]),
}

_lib = None


def load_library(path: os.PathLike | str | None = None) -> C.CDLL:
    """dlopen the in-tree CUDA library and bind every prototype.  Raises if it is absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise ImportError(
            f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  aircraft_b200 has no CPU fallback.")
    lib = C.CDLL(str(p))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib
