"""ctypes loader for ``csrc/libekf_b200.so`` — the only way the host side reaches the CUDA kernels.

Importing this module never touches the GPU and never loads the library; ``lib()`` loads it on first use and
raises ``EkfLibraryError`` if it has not been built (there is no CPU fallback).
Signatures mirror ``include/ekf_b200.h`` one to one.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libekf_b200.so")

EKF_COV_PLAIN, EKF_COV_JOSEPH = 0, 1


class EkfLibraryError(RuntimeError):
    pass


class EkfError(RuntimeError):
    """A C-ABI call returned a negative status; the message is ``ekf_last_error()``."""

    def __init__(self, code, msg):
        super().__init__(f"[ekf_b200 {code}] {msg}")
        self.code = code


_p = C.c_void_p
_i = C.c_int
_l = C.c_int64
_d = C.c_double
_u64 = C.c_uint64
_i32 = C.c_int32

# name -> (restype, argtypes); every symbol include/ekf_b200.h declares
SIGNATURES = {
    "ekf_version": (_i, []),
    "ekf_last_error": (C.c_char_p, []),
    "ekf_device_count": (_i, [C.POINTER(_i)]),
    "ekf_delta_phi": (_d, [_l, _l, _l]),
    "ekf_displacement": (_i, [_d, _d, _d, _d, C.POINTER(_d), C.POINTER(_d)]),
    "ekf_batch_layout": (_i, [_i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_l), C.POINTER(_l)]),
    "ekf_batch_create": (_i, [_i, _l, _i, C.POINTER(_p)]),
    "ekf_batch_destroy": (_i, [_p]),
    "ekf_batch_bind_state": (_i, [_p, _p, _p, _p]),
    "ekf_batch_set_noise": (_i, [_p, _p, _p]),
    "ekf_batch_set_flags": (_i, [_p, _i, _i, _i, _i]),
    "ekf_batch_init": (_i, [_p, _d, _p]),
    "ekf_batch_step": (_i, [_p, _p, _p, _p, _p, _p, _i, _p]),
    "ekf_batch_step_ticks": (_i, [_p, _p, _p, _p, _p, _d, _d, _i, _p, _p, _p, _i, _p]),
    "ekf_batch_step_ticks_common": (_i, [_p, _i32, _i32, _i32, _i32, _d, _d, _i, _p, _p, _p, _i, _p]),
    "ekf_batch_pack": (_i, [_p, _p, _p, _l, _l, _p]),
    "ekf_batch_unpack": (_i, [_p, _p, _p, _l, _l, _p]),
    "ekf_batch_step_host": (_i, [_p, _p, _p, _p, _p, _p, _i, _p, _p]),
    "ekf_batch_step_host_async": (_i, [_p, _p, _p, _p, _p, _p, _i, _p, _p]),
    "ekf_batch_host_sync": (_i, [_p]),
    "ekf_batch_fault_count": (_i, [_p, C.POINTER(_l)]),
    "ekf_batch_nonfinite_count": (_i, [_p, C.POINTER(_l), _p]),
    "ekf_batch_associate": (_i, [_p, _i, _p, _p, _p, _p, _p, _i, _p, _p, _p, _p, _p, _i, _p]),
    "ekf_batch_associate_generated": (_i, [_p, _i, _p, _p, _p, _u64, _u64, _u64, _d, _i, _p, _p, _p, _p, _p, _i, _p]),
    "ekf_batch_associate_overflow_count": (_i, [_p, C.POINTER(_l)]),
    "ekf_batch_accumulate_ate": (_i, [_p, _d, _d, _p, _p]),
    "ekf_batch_moments": (_i, [_p, _p, _p]),
    "ekf_batch_pose_moments": (_i, [_p, _p, _p]),
    "ekf_pose_moments_noalloc": (_i, [_p, _l, _l, _p, _i, _p]),
    "ekf_batch_mc_record_doubles": (_i, [_i, C.POINTER(_l)]),
    "ekf_batch_perturb": (_i, [_p, _u64, _u64, _u64, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    "ekf_batch_step_generated": (_i, [_p, _u64, _u64, _u64, _i, _p, _p, _p]),
    "ekf_batch_detection_noise": (_i, [_p, _u64, _u64, _u64, _i, _d, _p, _p, _p]),
    "ekf_batch_mc_step_host_async": (_i, [_p, _u64, _u64, _u64, _i, _p, _p, _i, _d, _d, _p, _p, _p]),
    "ekf_lm_layout": (_i, [_i, C.POINTER(_i), C.POINTER(_l)]),
    "ekf_lm_create": (_i, [_i, _i, C.POINTER(_p)]),
    "ekf_lm_destroy": (_i, [_p]),
    "ekf_lm_bind_state": (_i, [_p, _p, _p, _l, _i, _i]),
    "ekf_lm_upload": (_i, [_p, _i, _p, _p, _p]),
    "ekf_lm_download": (_i, [_p, _i, _p, _p, _p]),
    "ekf_lm_set_noise": (_i, [_p, _d, _d]),
    "ekf_lm_set_flags": (_i, [_p, _i, _i, _i, _i]),
    "ekf_lm_augment": (_i, [_p, _i, _d, _d, _p]),
    "ekf_lm_predict": (_i, [_p, _d, _d, _p]),
    "ekf_lm_update": (_i, [_p, _i, _d, _d, _p]),
    "ekf_lm_update_multi": (_i, [_p, _i, _p, _p, _p, _p]),
    "ekf_lm_export_rows_multi": (_i, [_p, _i, _p, _p, _p]),
    "ekf_lm_update_multi_from_rows": (_i, [_p, _i, _p, _p, _p, _p, _p]),
    "ekf_lm_export_rows": (_i, [_p, _i, _p, _p]),
    "ekf_lm_update_from_rows": (_i, [_p, _i, _d, _d, _p, _p]),
    "ekf_lm_p2p_init": (_i, [_p, _i, _i, _p, _l]),
    "ekf_lm_p2p_connect": (_i, [_p, _p, _l, _p]),
    "ekf_lm_p2p_area": (_i, [_p, C.POINTER(_p)]),
    "ekf_lm_p2p_update": (_i, [_p, _i, _d, _d, _i, _p]),
    "ekf_lm_p2p_update_multi": (_i, [_p, _i, _p, _p, _p, _p, _p]),
    "ekf_lm_p2p_invalidate": (_i, [_p]),
    "ekf_lm_p2p_error": (_i, [_p, C.POINTER(_l)]),
    "ekf_lm_sync": (_i, [_p]),
    "ekf3_predict": (_i, [_l, _p, _p, _p, _p, _p, _p]),
    "ekf3_update": (_i, [_l, _p, _p, _p, _p, _p, _p, _p]),
}

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise EkfLibraryError(
                f"{LIB_PATH} is missing: build it with `make -C {os.path.dirname(LIB_PATH)}` or "
                "`python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(status: int) -> None:
    if status != 0:
        raise EkfError(status, lib().ekf_last_error().decode("utf-8", "replace"))


def call(name: str, *args):
    """Invoke an int-returning entry point and raise EkfError on a negative status."""
    check(getattr(lib(), name)(*args))
