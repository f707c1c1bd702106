"""ctypes binding of libr2d2_b200.so (the C ABI declared in include/r2d2_b200.h).

The library is the product path; there is NO fallback.  If the shared object is missing or a
call fails, a RuntimeError carrying ``r2d2_last_error()`` is raised.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libr2d2_b200.so")

_P, _I, _L, _F, _D = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_double

# name -> (restype, argtypes); mirrors include/r2d2_b200.h one to one
PROTOTYPES = {
    "r2d2_version": (_I, []),
    "r2d2_last_error": (ctypes.c_char_p, []),
    "r2d2_sm_count": (_I, []),
    "r2d2_launch_count": (ctypes.c_longlong, []),
    "r2d2_prof_enable": (_I, [_I]),
    "r2d2_prof_report": (_I, [ctypes.c_char_p, _I, _I]),
    "r2d2_prof_trace": (_I, [ctypes.c_char_p, _I]),
    "r2d2_tree_layout": (_I, [_L, _I, _P, _P, _P, _P]),
    "r2d2_tree_build": (_I, [_P, _P, _L, _P, _P]),
    "r2d2_tree_sample": (_I, [_P, _P, _L, _P, _P, _I, _P, _P]),
    "r2d2_tree_update": (_I, [_P, _P, _L, _P, _P, _I, _P]),
    "r2d2_replay_gather": (_I, [_P] * 6 + [_L, _I, _P, _I, _I, _I, _I] + [_P] * 9 + [_I, _P]),
    "r2d2_slab_to_f32": (_I, [_P, _I, _I, _I, _I, _P, _P]),
    "r2d2_replay_ingest": (_I, [_P] * 10 + [_L] + [_P] * 10 + [_L, _I, _L, _P]),
    "r2d2_update_priority": (_I, [_P, _L, _P, _P, _L, _P]),
    "r2d2_update_priority_seq": (_I, [_P, _L, _P, _I, _I, _I, _P, _P]),
    "r2d2_update_sequence_priority": (_I, [_P, _P, _P, _L, _P, _I, _I, _F, _F, _I, _P, _P]),
    "r2d2_torso_fwd_u8": (_I, [_P, _I, _I, _I, _I] + [_P] * 9 + [_P]),
    "r2d2_torso_fwd_f32": (_I, [_P, _I] + [_P] * 9 + [_P]),
    "r2d2_linear_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "r2d2_lstm_seq_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "r2d2_head_fwd": (_I, [_P] * 7 + [_I, _I, _P]),
    "r2d2_head_out": (_I, [_P, _P, _P, _P, _I, _I, _P]),
    "r2d2_colsum": (_I, [_P, _P, _I, _I, _P]),
    "r2d2_colsum2": (_I, [_P, _P, _P, _I, _I, _P]),
    "r2d2_td_target": (_I, [_P] * 6 + [_I, _F, _I, _F, _F, _F] + [_P] * 7 + [_P]),
    "r2d2_head_bwd": (_I, [_P] * 6 + [_I, _I, _P]),
    "r2d2_linear_dgrad": (_I, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "r2d2_linear_wgrad_scratch_bytes": (_L, [_I, _I, _I]),
    "r2d2_linear_wgrad": (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _P]),
    "r2d2_linear_wgrad_static": (_I, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "r2d2_lstm_seq_bwd": (_I, [_P] * 6 + [_I, _I, _I, _P]),
    "r2d2_torso_bwd_scratch_bytes": (_L, [_I]),
    "r2d2_torso_bwd_u8": (_I, [_P, _I, _I, _I, _I] + [_P] * 13 + [_P]),
    "r2d2_torso_bwd_f32": (_I, [_P, _I] + [_P] * 13 + [_P]),
    "r2d2_tc_conv1_fwd_dual_u8": (_I, [_P, _I, _I, _I, _P, _P, _P, _P, _P, _P, _I, _P]),
    "r2d2_tc_conv1_fwd_dual_ring": (_I, [_P, _L, _P, _I, _I, _P, _P, _P, _P, _P, _P, _I, _P, _P]),
    "r2d2_tc_conv_dgrad2_masked": (_I, [_P, _P, _P, _P, _I, _P]),
    "r2d2_tc_conv1_wgrad_ring": (_I, [_P, _L, _P, _I, _I, _I, _P, _P, _P, _P, _P]),
    "r2d2_tc_torso_fwd_tail": (_I, [_P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "r2d2_tc_torso_fwd_u8": (_I, [_P, _I, _I, _I, _I] + [_P] * 9 + [_P]),
    "r2d2_tc_torso_fwd_f32": (_I, [_P, _I] + [_P] * 9 + [_P]),
    "r2d2_tc_lstm_set_trace": (_I, [_P]),
    "r2d2_tc_lstm_scratch_bytes": (_L, [_I, _I]),
    "r2d2_tc_lstm_seq_fwd": (_I, [_P, _L, _L, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P, _P]),
    "r2d2_tc_lstm_bwd_scratch_bytes": (_L, [_I, _I]),
    "r2d2_tc_lstm_seq_bwd": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P, _P]),
    "r2d2_torso_bwd_dgrad_u8": (_I, [_P, _I, _I, _I, _I] + [_P] * 10 + [_P]),
    "r2d2_tc_torso_wgrad_u8": (_I, [_P, _I, _I, _I, _I] + [_P] * 9 + [_P]),
    "r2d2_reduce_partials": (_I, [_P, _P, _L, _I, _P]),
    "r2d2_tc_conv_dgrad3": (_I, [_P, _P, _P, _P, _I, _P]),
    "r2d2_tc_conv_dgrad2": (_I, [_P, _P, _P, _P, _I, _P]),
    "r2d2_relu_mask": (_I, [_P, _P, _L, _P]),
    "r2d2_chansum": (_I, [_P, _P, _I, _I, _P, _P]),
    "r2d2_tc_conv_wgrad_layer": (_I, [_I, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P]),
    "r2d2_tc_linear": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "r2d2_tc_linear_masked": (_I, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "r2d2_peer_alloc": (_I, [_L, _P, _P]),
    "r2d2_peer_open": (_I, [_P, _P]),
    "r2d2_peer_close": (_I, [_P]),
    "r2d2_peer_free": (_I, [_P]),
    "r2d2_dp_shard_elems": (_L, [_L, _I]),
    "r2d2_dp_reduce_rmsprop": (_I, [_I, _I, _P, _P, _P, _P, _P, _L, _I, _L, _D, _D, _D, _D, _P]),
    "r2d2_tc_linear_x3": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "r2d2_tc_linear_nn": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "r2d2_tc_linear_tn": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "r2d2_tc_linear_x3_ctas": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "r2d2_split_bf16": (_I, [_P, _P, _P, _L, _P]),
    "r2d2_cast_bf16": (_I, [_P, _P, _L, _P]),
    "r2d2_transpose_cast_bf16": (_I, [_P, _P, _I, _I, _P]),
    "r2d2_rmsprop_centered": (_I, [_P, _P, _P, _P, _L, _D, _D, _D, _D, _P]),
    "r2d2_actor_observe": (_I, [_P, _I, _I, _I, _L, _I, _I] + [_P] * 11 + [_F, _F, _F, _F, _F, _P]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m pytorch_r2d2_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().r2d2_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise RuntimeError(f"r2d2_b200 {what} failed (code {rc}): {last_error()}")


def stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


_DT = {"bf16": torch.bfloat16, "f32": torch.float32, "f64": torch.float64, "u8": torch.uint8, "i8": torch.int8,
       "i64": torch.int64, "i32": torch.int32}


def ptr(t, dtype: str | None = None, allow_none: bool = False) -> ctypes.c_void_p:
    """Device pointer of a contiguous CUDA tensor (with dtype check)."""
    if t is None:
        if allow_none:
            return ctypes.c_void_p(0)
        raise ValueError("null tensor passed to r2d2_b200")
    if not t.is_cuda:
        raise ValueError("r2d2_b200 kernels take CUDA tensors only (no CPU fallback)")
    if not t.is_contiguous():
        raise ValueError("r2d2_b200 kernels take contiguous tensors")
    if dtype is not None and t.dtype != _DT[dtype]:
        raise TypeError(f"expected {dtype} tensor, got {t.dtype}")
    return ctypes.c_void_p(t.data_ptr())


def call(name: str, *args) -> None:
    check(getattr(load(), name)(*args), name)


def launch_count() -> int:
    return int(load().r2d2_launch_count())


def prof_enable(on: bool) -> None:
    call("r2d2_prof_enable", int(on))


def prof_trace() -> list:
    """[(kernel name, stream id, start_ms, end_ms)] of every launch recorded since the last reset."""
    buf = ctypes.create_string_buffer(1 << 20)
    call("r2d2_prof_trace", buf, len(buf))
    out = []
    for line in buf.value.decode().splitlines():
        name, sid, t0, t1 = line.rsplit(" ", 3)
        out.append((name, int(sid), float(t0), float(t1)))
    return out


def prof_report(reset: bool = True) -> dict:
    """{kernel name: (launches, total_ms)} measured with CUDA events on the launch stream."""
    buf = ctypes.create_string_buffer(1 << 16)
    call("r2d2_prof_report", buf, len(buf), int(reset))
    out = {}
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.rsplit(" ", 2)
        out[name] = (int(cnt), float(ms))
    return out
