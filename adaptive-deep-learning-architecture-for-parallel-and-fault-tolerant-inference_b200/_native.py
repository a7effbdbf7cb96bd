"""ctypes binding of libadapt_b200.so (declared in include/adapt_b200.h).

There is deliberately no fallback: if the shared object is missing, or a compute entry point
is called on a machine without a B200, the call raises.  The only thing that works without a
GPU is loading the library and looking at its symbols.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libadapt_b200.so")

# mirror of include/adapt_b200.h -------------------------------------------------------------
OK, EINVAL, ECUDA, ENODEV, ENOMEM, ESTATE = 0, -1, -2, -3, -4, -5
MODE_BF16, MODE_FP32_CHECK = 0, 1
OP_INGEST, OP_CONV, OP_MAXPOOL, OP_GAP, OP_SOFTMAX, OP_AFFINE, OP_ZEROPAD, OP_CONV_CHAIN, OP_STEM_POOL = 0, 1, 2, 3, 4, 5, 6, 7, 8
T_INTERNAL, T_INPUT, T_OUTPUT = 0, 1, 2
DT_WORK, DT_F32, DT_U8 = 0, 1, 2
IN_F32, IN_U8_RGB_CAFFE = 0, 1  # INGEST input formats (include/adapt_b200.h)
INPUT_FORMATS = {"float32": IN_F32, "uint8": IN_U8_RGB_CAFFE}
NOALLOC_MAX_LOGITS = 256 * 1024  # ADAPT_NOALLOC_MAX_LOGITS


class WeightDesc(C.Structure):
    _fields_ = [
        ("kernel", C.c_void_p),
        ("scale", C.c_void_p),
        ("bias", C.c_void_p),
        ("kh", C.c_int32),
        ("kw", C.c_int32),
        ("cin", C.c_int32),
        ("cout", C.c_int32),
    ]


class TensorDesc(C.Structure):
    _fields_ = [("elems_per_image", C.c_int64), ("dtype", C.c_int32), ("role", C.c_int32)]


class OpDesc(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("in_tensor", C.c_int32),
        ("res_tensor", C.c_int32),
        ("out_tensor", C.c_int32),
        ("H", C.c_int32),
        ("W", C.c_int32),
        ("Cin", C.c_int32),
        ("Ho", C.c_int32),
        ("Wo", C.c_int32),
        ("Cout", C.c_int32),
        ("kh", C.c_int32),
        ("kw", C.c_int32),
        ("stride", C.c_int32),
        ("pad", C.c_int32),
        ("relu", C.c_int32),
        ("weight", C.c_int32),
        ("out_f32", C.c_int32),
        ("in_format", C.c_int32),
        ("weight2", C.c_int32),
    ]


# every symbol the header declares: name -> (restype, argtypes)
_VP, _I, _SZ, _U32 = C.c_void_p, C.c_int, C.c_size_t, C.c_uint32
SYMBOLS = {
    "adapt_version": (C.c_char_p, []),
    "adapt_last_error": (C.c_char_p, []),
    "adapt_device_count": (_I, []),
    "adapt_stage_create": (_I, [_I, _I, _I, C.POINTER(WeightDesc), _I, C.POINTER(TensorDesc), _I, C.POINTER(OpDesc), _I,
                                C.POINTER(_VP)]),
    "adapt_stage_reslice": (_I, [_VP, C.POINTER(TensorDesc), _I, C.POINTER(OpDesc), _I]),
    "adapt_stage_bind_io": (_I, [_VP, _I, _VP, _VP]),
    "adapt_stage_prepare": (_I, [_VP, _I, _I]),
    "adapt_stage_enqueue": (_I, [_VP, _I, _I, _VP]),
    "adapt_stage_profile": (_I, [_VP, _I, _I, _VP, _I, C.POINTER(C.c_float), _I]),
    "adapt_stage_time_kind": (_I, [_VP, _I, _I, _VP, _I, _I, C.POINTER(C.c_float), C.POINTER(_I)]),
    "adapt_stage_read_tensor": (_I, [_VP, _I, _VP, _SZ, _VP]),
    "adapt_stage_num_launches": (_I, [_VP, _I]),
    "adapt_stage_destroy": (None, [_VP]),
    "adapt_conv2d_bf16": (_I, [_I, _VP, _VP, _VP, _VP, _VP, _VP, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I,
                               _VP]),
    "adapt_conv_chain_bf16": (_I, [_I, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _I, _I, _I, _I, _VP]),
    "adapt_stem_pool_bf16": (_I, [_I, _VP, _VP, _VP, _VP, _VP, _I, _I, _I, _I, _I, _VP]),
    "adapt_dense_softmax_noalloc": (_I, [_I, _VP, _VP, _VP, _VP, _VP, _I, _I, _I, _VP]),
    "adapt_device_alloc": (_I, [_I, _SZ, C.POINTER(_VP)]),
    "adapt_device_free": (_I, [_I, _VP]),
    "adapt_peer_enable": (_I, [_I, _I]),
    "adapt_ipc_export": (_I, [_I, _VP, _VP]),
    "adapt_ipc_import": (_I, [_I, _VP, C.POINTER(_VP)]),
    "adapt_ipc_close": (_I, [_I, _VP]),
    "adapt_stream_write_flag": (_I, [_I, _VP, _VP, _U32]),
    "adapt_stream_wait_flag_geq": (_I, [_I, _VP, _VP, _U32]),
    "adapt_memcpy_h2d": (_I, [_I, _VP, _VP, _SZ, _VP]),
    "adapt_memcpy_d2h": (_I, [_I, _VP, _VP, _SZ, _VP]),
    "adapt_memcpy_d2d": (_I, [_I, _VP, _VP, _SZ, _VP]),
    "adapt_memset": (_I, [_I, _VP, _I, _SZ, _VP]),
    "adapt_stream_sync": (_I, [_I, _VP]),
    "adapt_heartbeat_start": (_I, [_VP, _I, C.POINTER(_VP)]),
    "adapt_heartbeat_stop": (_I, [_VP]),
    "adapt_stream_query": (_I, [_I, _VP, C.POINTER(_I)]),
    "adapt_host_register": (_I, [_I, _VP, _SZ, C.POINTER(_VP)]),
    "adapt_host_unregister": (_I, [_I, _VP]),
    "adapt_stream_create": (_I, [_I, C.POINTER(_VP)]),
    "adapt_stream_destroy": (_I, [_I, _VP]),
    "adapt_event_create": (_I, [_I, C.POINTER(_VP)]),
    "adapt_event_record": (_I, [_I, _VP, _VP]),
    "adapt_event_sync": (_I, [_I, _VP]),
    "adapt_event_elapsed_ms": (_I, [_I, _VP, _VP, C.POINTER(C.c_float)]),
    "adapt_event_query": (_I, [_I, _VP, C.POINTER(_I)]),
    "adapt_event_destroy": (_I, [_I, _VP]),
    "adapt_debug_spin": (_I, [_I, _VP, _I]),
}


class AdaptError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"adapt_b200 error {code}: {message}")
        self.code = code


_lib = None
_lock = threading.Lock()


def lib() -> C.CDLL:
    """Load (once) and return the shared library with prototypes set.  Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python __graft_entry__.py build` (needs nvcc). "
                "There is no CPU fallback for the stage kernels."
            )
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)  # AttributeError if the .so does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = handle
        return _lib


def check(rc: int) -> None:
    if rc != OK:
        raise AdaptError(rc, lib().adapt_last_error().decode("utf-8", "replace"))


def device_count() -> int:
    return int(lib().adapt_device_count())


def require_gpu() -> None:
    if device_count() <= 0:
        raise AdaptError(ENODEV, "no CUDA device visible; the B200 stage kernels have no CPU fallback")
