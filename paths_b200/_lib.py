"""ctypes binding of libpaths_b200.so (C ABI declared in include/paths_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at import of `lib()`, and every
compute entry point fails when no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpaths_b200.so")

MAX_LEVELS = 8
MAX_LAYERS = 4
P_LSTM_COUNT = 4
P_LEVEL_HEAD_COUNT = 14
P_LAYER_COUNT = 15
LEVEL_STRIDE = P_LEVEL_HEAD_COUNT + MAX_LAYERS * P_LAYER_COUNT
NUM_OFFSETS = P_LSTM_COUNT + MAX_LEVELS * LEVEL_STRIDE

# enum paths_lstm_param / paths_level_param / paths_layer_param
P_WG, P_BG, P_WMO, P_BMO = range(4)
(P_WUV, P_B1, P_W2, P_B2, P_BIN, P_SPECIAL, P_NFW, P_NFB, P_WC, P_BC, P_WH0, P_BH0, P_WH2, P_BH2) = range(14)
(P_WQKV, P_BQKV, P_WO, P_BO, P_BCROSS, P_FW1, P_FB1, P_FW2, P_FB2, P_N1W, P_N1B, P_N2W, P_N2B, P_N3W, P_N3B) = range(15)

# enum paths_buffer
BUF_NAMES = ["COUNT", "NTOK", "SEG", "ROW_SLIDE", "ROW_XY", "ROW_PARENT", "PARENT_LOCAL", "SEL", "SEL_COUNT", "CHILD_OF",
             "TSEG", "X", "G", "CN", "PT", "H", "Y", "UV", "ALPHA", "TOK", "Z", "F", "LOGITS", "HAZARDS", "LOSS", "DLOGITS",
             "OVERFLOW", "DF_IN", "DHROW", "DCROW", "LOSS_PER", "HF0", "RISK"]
BUF = {n: i for i, n in enumerate(BUF_NAMES)}
INT_BUFS = {"COUNT", "NTOK", "SEG", "ROW_SLIDE", "ROW_XY", "ROW_PARENT", "PARENT_LOCAL", "SEL", "SEL_COUNT", "CHILD_OF",
            "TSEG", "OVERFLOW"}


class Dims(C.Structure):
    _fields_ = [("D", C.c_int32), ("C", C.c_int32), ("d", C.c_int32), ("heads", C.c_int32), ("layers", C.c_int32),
                ("I", C.c_int32), ("F", C.c_int32), ("nl", C.c_int32), ("L", C.c_int32), ("K", C.c_int32 * MAX_LEVELS),
                ("imp_mul", C.c_int32), ("pe_mode", C.c_int32), ("residual", C.c_int32), ("task", C.c_int32),
                ("precision", C.c_int32), ("lstm", C.c_int32), ("hier_ctx", C.c_int32)]


class Store(C.Structure):
    _fields_ = [("grid_ptr", C.c_void_p), ("gw", C.c_void_p), ("gh", C.c_void_p), ("occ", C.c_void_p),
                ("occ_off", C.c_void_p), ("nonbg", C.c_void_p), ("bf16", C.c_int32)]


_lib = None

# every symbol include/paths_b200.h declares (tests check that the library exports all of them)
SYMBOLS = ["paths_last_error", "paths_abi_version", "paths_engine_workspace_bytes", "paths_engine_create",
           "paths_engine_destroy", "paths_engine_bind_params", "paths_engine_bind_store", "paths_engine_refresh_weights", "paths_engine_begin",
           "paths_engine_begin_padded", "paths_engine_level_forward", "paths_engine_select", "paths_engine_loss",
           "paths_engine_level_backward", "paths_engine_forward", "paths_engine_forward_padded", "paths_engine_forward_packed", "paths_engine_backward", "paths_engine_wait_grads", "paths_engine_overflow_poll", "paths_engine_buffer",
           "paths_engine_set_precision", "paths_engine_set_dropout", "paths_engine_launch_count", "paths_engine_profile", "paths_engine_profile_read", "paths_build_occupancy", "paths_attention_fwd", "paths_attention_bwd", "paths_attention_bwd_static", "paths_lstm_cell_fwd", "paths_select_expand", "paths_gather_rows", "paths_engine_aggregate", "paths_gemm_nt",
           "paths_gemm_tn", "paths_adamw_step", "paths_heatmap_workspace_bytes", "paths_heatmap_composite"]


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m paths_b200.build` (nvcc, sm_100a). "
            "paths_b200 has no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
    L.paths_last_error.restype = C.c_char_p
    L.paths_abi_version.restype = i32
    L.paths_engine_workspace_bytes.restype = sz
    L.paths_engine_workspace_bytes.argtypes = [C.POINTER(Dims), i32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.paths_engine_create.argtypes = [C.POINTER(vp), C.POINTER(Dims), i32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), vp, sz]
    L.paths_engine_destroy.argtypes = [vp]
    L.paths_engine_destroy.restype = None
    L.paths_engine_bind_params.argtypes = [vp, vp, vp, i64, C.POINTER(C.c_int64), vp]
    L.paths_engine_bind_store.argtypes = [vp, C.POINTER(Store)]
    L.paths_engine_refresh_weights.argtypes = [vp, vp]
    L.paths_engine_begin.argtypes = [vp, vp, vp]
    L.paths_engine_begin_padded.argtypes = [vp, vp, vp, i32, vp]
    L.paths_engine_level_forward.argtypes = [vp, i32, vp]
    L.paths_engine_select.argtypes = [vp, i32, vp]
    L.paths_engine_loss.argtypes = [vp, vp, vp, f32, vp]
    L.paths_engine_level_backward.argtypes = [vp, i32, i32, vp]
    L.paths_engine_forward.argtypes = [vp, vp, vp, vp, f32, i32, vp]
    L.paths_engine_forward_padded.argtypes = [vp, vp, vp, i32, vp, vp, f32, i32, vp]
    L.paths_engine_forward_packed.argtypes = [vp, vp, vp, i32, vp, vp, f32, i32, vp]
    L.paths_engine_backward.argtypes = [vp, i32, vp]
    L.paths_engine_wait_grads.argtypes = [vp, i32, vp]
    L.paths_engine_overflow_poll.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.paths_engine_buffer.argtypes = [vp, i32, i32, C.POINTER(vp), C.POINTER(sz)]
    L.paths_engine_set_precision.argtypes = [vp, i32]
    L.paths_engine_set_dropout.argtypes = [vp, f32, C.c_uint64]
    L.paths_engine_launch_count.argtypes = [vp]
    L.paths_engine_launch_count.restype = i64
    L.paths_engine_profile.argtypes = [vp, i32]
    L.paths_engine_profile_read.argtypes = [vp, i32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.paths_build_occupancy.argtypes = [vp, i32, i32, vp, vp, vp]
    L.paths_attention_fwd.argtypes = [i32, vp, vp, i32, vp, vp, i32, i32, i32, f32, C.c_uint64, vp]
    L.paths_attention_bwd.argtypes = [i32, vp, vp, i32, vp, vp, vp, vp, vp, i32, i32, i32, f32, C.c_uint64, vp]
    L.paths_attention_bwd_static.argtypes = [i32, vp, vp, i32, i32, vp, vp, vp, vp, i32, i32, i32, f32, C.c_uint64, vp]
    L.paths_lstm_cell_fwd.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp]
    L.paths_select_expand.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, C.POINTER(Store), i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp,
                                      vp, i32, i32, vp, vp, vp]
    L.paths_gather_rows.argtypes = [C.POINTER(Store), i32, i32, vp, vp, vp, vp, i32, i32, vp, vp]
    L.paths_engine_aggregate.argtypes = [vp, i32, vp]
    L.paths_gemm_nt.argtypes = [i32, vp, i32, vp, i32, i32, vp, i32, vp, i32, i32, i32, i32, vp]
    L.paths_gemm_tn.argtypes = [i32, vp, i32, vp, i32, vp, i32, i32, i32, i32, vp]
    L.paths_adamw_step.argtypes = [vp, vp, vp, vp, i64, i64, C.POINTER(C.c_int64), i32, f32, f32, f32, f32, f32, i32, vp]
    L.paths_heatmap_workspace_bytes.restype = sz
    L.paths_heatmap_workspace_bytes.argtypes = [i32, i32, i32]
    L.paths_heatmap_composite.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, vp, sz, vp, vp, vp]
    _lib = L
    return L


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().paths_last_error().decode(errors="replace")
        raise RuntimeError(f"paths_b200 {what} failed (rc={rc}): {msg}")
