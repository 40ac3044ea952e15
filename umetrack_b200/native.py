"""ctypes binding of libumetrack_b200.so (declared in include/umetrack_b200.h).

There is no fallback: if the shared library is missing or a call fails, the error is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_uint8, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libumetrack_b200.so")

MODE_FP32_EQUIV = 0
MODE_BF16 = 1
MODE_FP16 = 2
HEAD_KNOWN = 0
HEAD_UNKNOWN = 1
RESAMPLE_PARAMS = 48
STATUS_MEMORY_INDEX = 1
STATUS_FP16_RANGE = 2

# every symbol the header declares: (name, restype, argtypes)
SIGNATURES = {
    "umt_last_error": (c_char_p, []),
    "umt_abi_version": (c_int, []),
    "umt_resample_crops": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                   c_void_p, c_void_p, c_void_p, c_void_p]),
    "umt_crop_cameras": (c_int, [c_void_p] * 7 + [c_int] * 6 + [c_double] + [c_void_p] * 7),
    "umt_resample_homography": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                        c_void_p]),
    "umt_model_num_tensors": (c_int, []),
    "umt_model_tensor_name": (c_char_p, [c_int]),
    "umt_model_tensor_numel": (c_int64, [c_int]),
    "umt_model_create": (c_int, [POINTER(c_void_p), c_int, POINTER(c_void_p)]),
    "umt_model_destroy": (None, [c_void_p]),
    "umt_model_workspace_bytes": (c_int, [c_void_p, c_int, c_int, c_int, POINTER(c_size_t)]),
    "umt_model_bind_workspace": (c_int, [c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_void_p]),
    "umt_model_forward": (c_int, [c_void_p, c_int, c_int] + [c_void_p] * 7 + [c_int, c_int, c_void_p,
                                  c_void_p, c_int, c_void_p, c_void_p, c_int] + [c_void_p] * 5),
    "umt_model_forward_u8": (c_int, [c_void_p, c_int, c_int] + [c_void_p] * 7 + [c_int, c_int, c_void_p,
                                     c_void_p, c_int, c_void_p, c_void_p, c_int] + [c_void_p] * 5),
    "umt_model_encode_u8": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "umt_model_tail": (c_int, [c_void_p, c_int] + [c_void_p] * 7 + [c_int, c_int, c_void_p,
                               c_void_p, c_int, c_void_p, c_void_p, c_int] + [c_void_p] * 5),
    "umt_model_tail_scan": (c_int, [c_void_p, c_int] + [c_void_p] * 7 + [c_int, c_void_p, c_void_p, c_int,
                                    c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int] + [c_void_p] * 5),
    "umt_model_encode": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "umt_model_encode_static": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "umt_model_last_launch_count": (c_int, [c_void_p]),
    "umt_model_status": (c_int, [c_void_p, POINTER(ctypes.c_uint), c_void_p]),
    "umt_model_profile_layers": (c_int, [c_void_p, c_int, c_void_p, c_int, POINTER(c_float), c_void_p]),
    "umt_model_debug_halo_check": (c_int, [c_void_p, POINTER(ctypes.c_longlong), c_void_p]),
    "umt_model_debug_layer": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                      POINTER(c_int), c_void_p]),
    "umt_fk_landmarks": (c_int, [c_void_p] * 8 + [c_int, c_void_p, c_void_p]),
    "umt_mpjpe_accumulate": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "umt_accel_accumulate": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "umt_keypoint_errors": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "umt_pck_accumulate": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int, c_void_p, c_void_p]),
}

_lib = None


class NativeError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Load the shared library once; raise loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C umetrack_b200/csrc`).  umetrack_b200 has no CPU fallback.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().umt_last_error()
        raise NativeError(f"{what} failed with code {rc}: {msg.decode() if msg else ''}")


def ptr(t) -> int:
    """Device (or host) address of a torch tensor / None."""
    return 0 if t is None else t.data_ptr()


def current_stream_ptr() -> int:
    """cudaStream_t of torch's current stream on the current device (the raw getter: the Stream
    object of torch.cuda.current_stream() costs ~20 us to build, which a per-frame caller pays
    several times per frame)."""
    import torch

    raw = getattr(torch._C, "_cuda_getCurrentRawStream", None)
    if raw is not None:
        return raw(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream
