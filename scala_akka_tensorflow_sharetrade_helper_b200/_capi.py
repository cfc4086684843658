"""ctypes binding of include/qpolicy.h — the same stub a JNI/Panama shim would bind.

There is no CPU path: if libqpolicy.so is missing or no CUDA device is present, calls fail loudly.
"""
from __future__ import annotations

import ctypes as C
import os

from . import _build

QP_OK, QP_EINVAL, QP_ECUDA, QP_ENCCL, QP_ESTATE, QP_EUNSUPPORTED, QP_EIO = range(7)
QP_N_KERNELS = 3


class QpConfig(C.Structure):
    _fields_ = [
        ("window", C.c_int32), ("hidden", C.c_int32), ("epsilon", C.c_float), ("gamma", C.c_float),
        ("lr", C.c_float), ("bias", C.c_float), ("step_div", C.c_float), ("adagrad_init_acc", C.c_float),
        ("bookkeeping", C.c_int32), ("device", C.c_int32), ("seed", C.c_uint64),
    ]


class QpTrace(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("action_sel", "action_eff", "explored", "k_next", "reward",
                                          "q_s", "q_n", "target", "budget", "shares")]


# every symbol include/qpolicy.h declares: name -> (restype, argtypes)
_H = C.c_void_p
_P = C.c_void_p
SIGNATURES = {
    "qp_config_default": (C.c_int, [C.POINTER(QpConfig)]),
    "qp_create": (C.c_int, [C.POINTER(QpConfig), C.POINTER(_H)]),
    "qp_destroy": (C.c_int, [_H]),
    "qp_input_dim": (C.c_int64, [_H]),
    "qp_param_count": (C.c_int64, [_H]),
    "qp_set_weights": (C.c_int, [_H, _P, _P]),
    "qp_get_weights": (C.c_int, [_H, _P, _P]),
    "qp_get_last_grad": (C.c_int, [_H, _P]),
    "qp_select_action": (C.c_int, [_H, _P, C.c_int64, _P, _P, _P, _P, C.c_int64]),
    "qp_update_q": (C.c_int, [_H, _P, C.c_int64, _P, _P, C.c_int64, C.c_int64]),
    "qp_env_create": (C.c_int, [_H, _P, C.c_int64, C.c_int64, _P, C.c_int64, C.c_int64, C.c_double, C.c_int32]),
    "qp_env_reset": (C.c_int, [_H]),
    "qp_env_steps_total": (C.c_int64, [_H]),
    "qp_env_steps_done": (C.c_int64, [_H]),
    "qp_env_step": (C.c_int, [_H, C.c_int32]),
    "qp_env_step_async": (C.c_int, [_H, C.c_int32]),
    "qp_sync": (C.c_int, [_H]),
    "qp_set_option": (C.c_int, [_H, C.c_char_p, C.c_int64]),
    "qp_get_option": (C.c_int, [_H, C.c_char_p, C.POINTER(C.c_int64)]),
    "qp_debug_tc_gemm": (C.c_int, [C.c_int32, C.c_int64, C.c_int64, C.c_int64, _P, _P, _P, C.c_int32, C.c_int32, _P]),
    "qp_debug_umma": (C.c_int, [C.c_int32, _P, C.c_int64, _P, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int32, C.c_uint32,
                                C.c_uint32, C.c_uint32, C.c_uint32, C.c_int32, C.c_float, _P, C.c_int32, C.c_uint32]),
    "qp_debug_phase_clocks": (C.c_int, [_H, C.c_int32, _P]),
    "qp_env_step_trace": (C.c_int, [_H, C.c_int32, C.POINTER(QpTrace)]),
    "qp_env_get_portfolio": (C.c_int, [_H, _P]),
    "qp_env_get_state": (C.c_int, [_H, _P, _P, _P]),
    "qp_env_set_state": (C.c_int, [_H, _P, _P, _P, C.c_int64]),
    "qp_env_portfolio_stats": (C.c_int, [_H, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "qp_train": (C.c_int, [_H, _P, C.c_int64, C.c_int64, _P, C.c_int64, C.c_int64, C.c_double, C.c_int32, _P]),
    "qp_enqueue_select": (C.c_int, [_H, _P, C.c_int64, C.c_float, C.c_int64, C.POINTER(C.c_int64)]),
    "qp_enqueue_update": (C.c_int, [_H, _P, C.c_int64, C.c_float, _P, C.c_int64]),
    "qp_flush": (C.c_int, [_H, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "qp_take_action": (C.c_int, [_H, C.c_int64, C.POINTER(C.c_int32)]),
    "qp_pending": (C.c_int64, [_H]),
    "qp_save_checkpoint": (C.c_int, [_H, C.c_char_p]),
    "qp_load_checkpoint": (C.c_int, [_H, C.c_char_p]),
    "qp_set_snapshot": (C.c_int, [_H, C.c_char_p, C.c_int64]),
    "qp_iteration": (C.c_int64, [_H]),
    "qp_parse_prices": (C.c_int, [C.c_char_p, C.c_size_t, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "qp_dist_unique_id": (C.c_int, [_P, C.c_size_t]),
    "qp_dist_p2p_bytes": (C.c_int64, [_H]),
    "qp_dist_p2p_attach": (C.c_int, [_H, _P, C.c_int32, C.c_uint64]),
    "qp_dist_init": (C.c_int, [_H, _P, C.c_int32, C.c_int32]),
    "qp_dist_p2p_export": (C.c_int, [_H, _P, C.c_size_t]),
    "qp_dist_p2p_import": (C.c_int, [_H, _P, C.c_int32]),
    "qp_dist_nccl_version": (C.c_int, [C.POINTER(C.c_int32)]),
    "qp_td_grad_noalloc": (C.c_int, [C.c_int32, C.POINTER(QpConfig), _P, _P, _P, _P, C.c_int64, _P, C.c_void_p]),
    "qp_stream": (C.c_void_p, [_H]),
    "qp_env_step_timed": (C.c_int, [_H, C.c_int32, _P]),
    "qp_measure_fp32_peak": (C.c_int, [_H, C.POINTER(C.c_double)]),
    "qp_measure_fp32_outer_peak": (C.c_int, [_H, C.POINTER(C.c_double)]),
    "qp_launch_count": (C.c_int64, [_H]),
    "qp_last_error": (C.c_char_p, []),
    "qp_version": (C.c_char_p, []),
}

_lib = None


def lib() -> C.CDLL:
    """Load libqpolicy.so (built in-tree by _build.build / __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_build.LIB_PATH):
            raise RuntimeError(
                f"{_build.LIB_PATH} is missing: run __graft_entry__.build() (nvcc, sm_100a). "
                "There is no CPU or PyTorch fallback for this path.")
        L = C.CDLL(_build.LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error() -> str:
    return lib().qp_last_error().decode()
