"""ctypes binding of libdqn_b200.so (the C ABI declared in include/dqn_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at import, and
every compute entry point fails without a CUDA device.  PyTorch is used by the shells only as a
device-buffer handle (``tensor.data_ptr()``) and for the current stream.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdqn_b200.so")

if not os.path.exists(LIB_PATH):
    raise RuntimeError(
        "libdqn_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` at the repo "
        "root; this package has no CPU path." % LIB_PATH)

lib = C.CDLL(LIB_PATH)

c_i32, c_i64, c_f32, c_f64, c_u8, c_vp = C.c_int32, C.c_int64, C.c_float, C.c_double, C.c_uint8, C.c_void_p


class Hyper(C.Structure):
    """struct dqn_hyper (include/dqn_b200.h)."""
    _fields_ = [("gamma", c_f64), ("lr", c_f32), ("reg_param", c_f32), ("momentum", c_f32),
                ("rmsprop_decay", c_f32), ("optimizer", c_i32), ("loss", c_i32), ("grad_scale", c_f32),
                ("reserved", c_i32)]


ARCH = {"simple": 0, "cnn": 1, "nature": 2}
DTYPE = {"u8": 0, "f32": 1, "f64": 2}
OPTIMIZER = {"sgd": 0, "momentum": 1, "rmsprop": 2, "adam": 3, "adagrad": 4, "adadelta": 5, "ftrl": 6}
LOSS = {"mse": 0, "huber": 1}
BUF = {"online": 0, "target": 1, "grad": 2, "slot0": 3, "slot1": 4}

# name -> (restype, argtypes); every symbol include/dqn_b200.h declares
SIGNATURES = {
    "dqn_last_error": (C.c_char_p, []),
    "dqn_version": (c_i32, []),
    "dqn_replay_create": (C.c_int, [C.POINTER(c_vp), c_i64, c_i64, c_i32, c_i32, c_i32, c_i32]),
    "dqn_replay_destroy": (C.c_int, [c_vp]),
    "dqn_replay_append_frames": (C.c_int, [c_vp, c_vp, c_i32, C.POINTER(c_i64)]),
    "dqn_replay_append_frames_device": (C.c_int, [c_vp, c_vp, c_i32, C.POINTER(c_i64), c_vp]),
    "dqn_replay_add": (C.c_int, [c_vp, c_i64, c_i64, c_i32, c_f64, c_u8]),
    "dqn_replay_push": (C.c_int, [c_vp, c_vp, c_i32, c_f64, c_u8]),
    "dqn_replay_flush": (C.c_int, [c_vp, c_vp]),
    "dqn_replay_seal": (C.c_int, [c_vp, C.POINTER(c_i32), c_vp]),
    "dqn_replay_ingest_staged": (C.c_int, [c_vp, c_i32, c_vp, c_vp, c_i32, c_vp]),
    "dqn_replay_set_lanes": (C.c_int, [c_vp, c_i32]),
    "dqn_replay_add_batch": (C.c_int, [c_vp, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dqn_replay_current_states": (C.c_int, [c_vp, c_vp, c_vp]),
    "dqn_replay_size": (C.c_int, [c_vp, C.POINTER(c_i64), C.POINTER(c_i64)]),
    "dqn_replay_next_frame_pos": (C.c_int, [c_vp, C.POINTER(c_i64)]),
    "dqn_replay_gather": (C.c_int, [c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dqn_index_table_next": (C.c_int, [c_vp, c_i64, c_i32, c_vp, c_vp, c_vp]),
    "dqn_net_create": (C.c_int, [C.POINTER(c_vp), c_i32, C.POINTER(c_i32), c_i32, c_i32, c_i32, c_i32, c_i32]),
    "dqn_net_destroy": (C.c_int, [c_vp]),
    "dqn_net_param_count": (C.c_int, [c_vp, C.POINTER(c_i64), C.POINTER(c_i32)]),
    "dqn_net_tensor_info": (C.c_int, [c_vp, c_i32, C.POINTER(c_i64), C.POINTER(c_i64), C.POINTER(c_i32),
                                      C.POINTER(c_i64)]),
    "dqn_net_buffer": (C.c_int, [c_vp, c_i32, C.POINTER(c_vp), C.POINTER(c_i64)]),
    "dqn_net_upload": (C.c_int, [c_vp, c_i32, c_vp, c_i64]),
    "dqn_net_download": (C.c_int, [c_vp, c_i32, c_vp, c_i64]),
    "dqn_net_set_hyper": (C.c_int, [c_vp, C.POINTER(Hyper), c_i32]),
    "dqn_net_global_step": (C.c_int, [c_vp, C.POINTER(c_i64), c_vp]),
    "dqn_net_set_precision": (C.c_int, [c_vp, c_i32]),
    "dqn_net_set_global_step": (C.c_int, [c_vp, c_i64]),
    "dqn_net_forward": (C.c_int, [c_vp, c_i32, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "dqn_net_train_step": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "dqn_net_backward": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_i32, c_vp]),
    "dqn_net_apply": (C.c_int, [c_vp, c_vp]),
    "dqn_net_target_sync": (C.c_int, [c_vp, c_vp]),
    "dqn_net_dp_handle": (C.c_int, [c_vp, c_vp]),
    "dqn_net_dp_connect": (C.c_int, [c_vp, c_i32, c_i32, c_vp]),
    "dqn_net_dp_set_timeout": (C.c_int, [c_vp, c_f64]),
    "dqn_net_dp_status": (C.c_int, [c_vp, C.POINTER(c_i64), c_vp]),
    "dqn_net_refresh_operands": (C.c_int, [c_vp, c_i32, c_vp]),
    "dqn_net_dp_step": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "dqn_train_step_dp": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "dqn_train_step": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "dqn_train_step_table": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_vp, c_i32, c_vp]),
    "dqn_preproc_create": (C.c_int, [C.POINTER(c_vp), c_i32, c_i32, c_i32, c_i32, c_i32]),
    "dqn_preproc_destroy": (C.c_int, [c_vp]),
    "dqn_preproc_run": (C.c_int, [c_vp, c_vp, c_i32, c_vp, c_vp]),
    "dqn_cnn_forward_noalloc": (C.c_int, [c_vp, c_i32, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "dqn_launch_count": (c_i64, []),
    "dqn_net_set_timing": (C.c_int, [c_vp, c_i32]),
    "dqn_net_timing_collect": (C.c_int, [c_vp]),
    "dqn_net_timing_read": (C.c_int, [c_vp, c_i32, C.POINTER(c_f64), C.POINTER(c_i32)]),
    "dqn_net_timing_floor": (C.c_int, [c_vp, c_vp, C.POINTER(c_f64)]),
    "dqn_net_timing_floor_enqueue": (C.c_int, [c_vp, c_vp]),
    "dqn_stage_name": (C.c_char_p, [c_i32]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)          # AttributeError here = the .so does not export a declared symbol
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    return (lib.dqn_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    """The reference's only exception type is RuntimeError (network.py:47, 180; frame_buffer.py:21)."""
    if rc != 0:
        raise RuntimeError("%s%s (dqn_status %d)" % (what + ": " if what else "", last_error(), rc))


def ptr(t) -> int:
    """Device (or host) address of a torch tensor / numpy array / None."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
