"""ctypes binding of lib/libsplitq.so (the C ABI declared in include/splitq.h).

There is no fallback: if the shared library is missing or a call fails, an exception is
raised.  torch is used only to own device memory and to name the current stream.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SPLITQ_LIB points at another build of the same library (kernel-variant experiments)
LIB_PATH = os.environ.get("SPLITQ_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libsplitq.so")

OK = 0
SCALE_GROUP, SCALE_TOKEN = 0, 1
SCALE_CHANNEL_BASE = 2  # + CHANNEL_METHODS[name]: per-channel payloads
BITS_RAW = 32
NORM_LAYERNORM, NORM_RMSNORM = 0, 1
FLAG_CAUSAL_HINT = 1
FLAG_DEQUANT_RECIP = 0x100  # OR into quant_pack's scale_mode: receiver computes (q * fl(1/L)) * m like CUDA ATen
MAX_SELECT_S = 32768
NOALLOC_MAX_WINDOWS = 1024
FUSED_F16X3, FUSED_F16X1 = 0, 1
CHANNEL_METHODS = {"channel_8": 0, "channel_4": 1, "channel_1_mean": 2, "channel_1_max": 3}

_vp, _sz, _i, _u = C.c_void_p, C.c_size_t, C.c_int, C.c_uint
_ip = C.POINTER(C.c_int)
_fp = C.POINTER(C.c_float)
_llp = C.POINTER(C.c_longlong)

# name -> (restype, argtypes); must list every symbol include/splitq.h declares
SIGNATURES = {
    "splitq_version": (_i, []),
    "splitq_last_error": (C.c_char_p, []),
    "splitq_importance_colmean_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "splitq_importance_colmean": (_i, [_vp, _i, _i, _i, _i, _fp, _vp, _vp, _vp, _sz, _u, _vp]),
    "splitq_importance_fused_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "splitq_importance_fused_tile_cols": (_i, []),
    "splitq_importance_fused": (_i, [_vp, _llp, _vp, _llp, _i, _i, _i, _i, _i, C.c_float, _i, _i, _fp, _vp, _vp, _vp,
                                     _sz, _vp]),
    "splitq_importance_fused_dual": (_i, [_vp, _llp, _vp, _llp, _i, _i, _i, _i, _i, C.c_float, _i, _i, _fp, _vp, _vp, _fp, _vp,
                                          _vp, _sz, _vp]),
    "splitq_select_bits": (_i, [_vp, _i, _i, _i, _ip, _ip, _i, _vp, _vp, _vp]),
    "splitq_select_bits_per_window": (_i, [_vp, _i, _i, _i, _i, _ip, C.POINTER(C.c_double), _vp, _vp, _vp, _vp]),
    "splitq_payload_bytes": (_sz, [_i, _i, _ip, _ip, _i, _i]),
    "splitq_quant_pack_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "splitq_quant_pack": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _vp, _sz, _vp, _sz, _vp]),
    "splitq_quant_pack_group_noalloc": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "splitq_dequant": (_i, [_vp, _sz, _i, _i, _i, _vp, _vp]),
    "splitq_dequant_norm": (_i, [_vp, _sz, _i, _i, _i, _i, _vp, _vp, C.c_float, _vp, _vp, _vp, _vp, _vp, _vp]),
    "splitq_channel_workspace_bytes": (_sz, [_i, _i, _i]),
    "splitq_channel_fake_quant": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "splitq_layer_js": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "splitq_channel_payload_bytes": (_sz, [_i, _i, _i]),
    "splitq_channel_quant_pack": (_i, [_vp, _i, _i, _i, _i, _vp, _sz, _vp, _sz, _vp]),
    "splitq_transfer_peer": (_i, [_vp, _i, _vp, _i, _sz, _vp]),
    "splitq_ipc_export": (_i, [_vp, _vp, C.POINTER(C.c_size_t)]),
    "splitq_ipc_open": (_i, [_vp, _sz, C.POINTER(C.c_void_p)]),
    "splitq_ipc_close": (_i, [_vp, _sz]),
    "splitq_stream_write_u32": (_i, [_vp, C.c_uint32, _vp]),
    "splitq_stream_wait_geq_u32": (_i, [_vp, C.c_uint32, _vp]),
    "splitq_copy_async": (_i, [_vp, _vp, _sz, _vp]),
}
IPC_HANDLE_BYTES = 64


class SplitqError(RuntimeError):
    pass


_lib = None


def load() -> C.CDLL:
    """Load the library once; raise if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SplitqError(
                f"{LIB_PATH} not found: build it with `make -C llm-inference-in-distributed-edge-networks_b200/csrc` "
                "(or __graft_entry__.build()); there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != OK:
        raise SplitqError(f"splitq error {rc}: {load().splitq_last_error().decode()}")


def int_array(values):
    values = list(values)
    return (C.c_int * max(1, len(values)))(*values) if values else None


def float_array(values):
    values = list(values)
    return (C.c_float * len(values))(*values)
