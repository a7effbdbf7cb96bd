"""Python handle on one C-ABI stage (`adapt_stage_t`) and the single-device predictor.

`StageExecutor` is what a worker (`Node`) owns in place of the reference's
`model_from_json` + `set_weights` + `model.predict` trio (reference src/node.py:40,45,77-78,177).
It needs a B200: constructing one without a GPU raises (there is no CPU path).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _native as N
from .dag_util import PartModel, StageProgram, build_weight_table, compile_part, construct_model, layer_macs
from .resnet50 import Model

MODES = {"bf16": N.MODE_BF16, "fp32": N.MODE_FP32_CHECK, "fp32-check": N.MODE_FP32_CHECK}


def elem_bytes(dtype: int, mode: int) -> int:
    if dtype == N.DT_F32:
        return 4
    if dtype == N.DT_U8:
        return 1
    return 2 if mode == N.MODE_BF16 else 4


class StageExecutor:
    def __init__(self, model: Model, part: PartModel, device: int = 0, mode: str = "bf16", max_batch: int = 128,
                 reserve_full: bool = False, input_format: str = "float32"):
        """`reserve_full`: size the stage's activation arena for the WHOLE model before installing `part`, so that
        no later re-slice (a fail-over hands this stage more layers) has to allocate device memory."""
        N.require_gpu()
        self.lib = N.lib()
        self.model = model
        self.device = int(device)
        self.mode = MODES[mode] if isinstance(mode, str) else int(mode)
        self.max_batch = int(max_batch)
        if input_format not in N.INPUT_FORMATS:
            raise ValueError(f"input_format must be one of {sorted(N.INPUT_FORMATS)}")
        self.input_format = N.INPUT_FORMATS[input_format]  # of the model input, when this stage takes it
        self.entries, self.windex = build_weight_table(model)
        self._keepalive = self.entries  # host arrays must outlive adapt_stage_create
        whole = None
        if reserve_full and len(part.layer_names) < len(model.layers) - 1:
            from .dag_util import construct_model

            whole = construct_model(model, model.layers[0].name, model.layers[-1].name, "whole")
        self.part = whole or part
        self.program: StageProgram = compile_part(self.part, self.windex, self.mode, self.input_format)
        w = (N.WeightDesc * len(self.entries))()
        for i, e in enumerate(self.entries):
            kh, kw, cin, cout = e.geom
            w[i] = N.WeightDesc(e.kernel.ctypes.data if e.kernel is not None else None, e.scale.ctypes.data,
                                e.bias.ctypes.data, kh, kw, cin, cout)
        t, o = self.program.c_arrays()
        handle = C.c_void_p()
        N.check(self.lib.adapt_stage_create(self.device, self.mode, self.max_batch, w, len(self.entries), t, len(t), o,
                                            len(o), C.byref(handle)))
        self.handle = handle
        if whole is not None:
            self.reslice(part)

    # ---- program / geometry ---------------------------------------------------------------------
    def reslice(self, part: PartModel) -> None:
        """Install another layer range; weights stay resident (metadata-only re-slice)."""
        program = compile_part(part, self.windex, self.mode, self.input_format)
        t, o = program.c_arrays()
        N.check(self.lib.adapt_stage_reslice(self.handle, t, len(t), o, len(o)))
        self.part, self.program = part, program

    @property
    def in_bytes_per_image(self) -> int:
        return self.program.in_elems * elem_bytes(self.program.in_dtype, self.mode)

    @property
    def out_bytes_per_image(self) -> int:
        return self.program.out_elems * elem_bytes(self.program.out_dtype, self.mode)

    @property
    def num_launches(self) -> int:
        return len(self.program.ops)

    # ---- execution ------------------------------------------------------------------------------
    def bind_io(self, slot: int, in_ptr: int, out_ptr: int) -> None:
        N.check(self.lib.adapt_stage_bind_io(self.handle, slot, in_ptr, out_ptr))

    def prepare(self, slot: int, batch: int) -> None:
        """Capture the CUDA graph of (slot, batch).  Configuration-time only (may sync the device)."""
        N.check(self.lib.adapt_stage_prepare(self.handle, slot, batch))

    def enqueue(self, slot: int, batch: int, stream: int = 0) -> None:
        N.check(self.lib.adapt_stage_enqueue(self.handle, slot, batch, stream))

    def profile(self, slot: int, batch: int, stream: int = 0, reps: int = 5) -> List[dict]:
        """Per-op mean device time (CUDA events around each launch) with the op's algorithmic work."""
        n = len(self.program.ops)
        ms = (C.c_float * n)()
        N.check(self.lib.adapt_stage_profile(self.handle, slot, batch, stream, reps, ms, n))
        out = []
        for op, t in zip(self.program.ops, ms):
            # algorithmic work = the Keras layer's own MAC count (not the padded / space-to-depth form it runs in)
            macs = 0
            if op.kind in (N.OP_CONV, N.OP_CONV_CHAIN, N.OP_STEM_POOL):
                macs = batch * sum(layer_macs(self.model, self.model.get_layer(n)) for n in op.layers
                                   if n in self.model._by_name)
            out.append({"layers": op.layers, "kind": op.kind, "ms": float(t), "macs": int(macs),
                        "shape": [op.H, op.W, op.Cin, op.Cout, op.kh, op.stride]})
        return out

    def time_kind(self, slot: int, batch: int, kinds: Sequence[int], stream: int = 0, reps: int = 20) -> Tuple[float, int]:
        """(device ms per step, launches) of all ops of the given kinds, run as their own CUDA graph (PDL overlap kept)."""
        ms, n = C.c_float(0), C.c_int(0)
        mask = sum(1 << int(k) for k in kinds)
        N.check(self.lib.adapt_stage_time_kind(self.handle, slot, batch, stream, mask, reps, C.byref(ms), C.byref(n)))
        return float(ms.value), int(n.value)

    def read_logits(self, batch: int, stream: int = 0) -> np.ndarray:
        if "logits" not in self.program.taps:
            raise KeyError("this stage does not contain the `predictions` layer")
        t, shp, _ = self.program.taps["logits"]
        out = np.empty((batch,) + shp, dtype=np.float32)
        N.check(self.lib.adapt_stage_read_tensor(self.handle, t, out.ctypes.data, out.nbytes, stream))
        return out

    def close(self) -> None:
        if getattr(self, "handle", None):
            self.lib.adapt_stage_destroy(self.handle)
            self.handle = None

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass


class DeviceBuffer:
    """Raw cudaMalloc'ed buffer (IPC-exportable), zero-initialised."""

    def __init__(self, device: int, nbytes: int):
        self.lib = N.lib()
        self.device = device
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        N.check(self.lib.adapt_device_alloc(device, self.nbytes, C.byref(p)))
        self.ptr = p.value

    def free(self) -> None:
        if self.ptr:
            self.lib.adapt_device_free(self.device, self.ptr)
            self.ptr = None


def wire_activation(x_f32: np.ndarray, mode: int) -> np.ndarray:
    """Host encoding of a cut tensor (tests / single-stage runs fed from the host): NHWC, stored
    channels, working dtype (bf16 as uint16 bit patterns, round-to-nearest-even)."""
    x = np.ascontiguousarray(x_f32, dtype=np.float32)
    if mode != N.MODE_BF16:
        return x
    c = x.shape[-1]
    cs = (c + 7) // 8 * 8
    if cs != c:
        pad = [(0, 0)] * (x.ndim - 1) + [(0, cs - c)]
        x = np.pad(x, pad)
    u = x.view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    return ((u + r) >> 16).astype(np.uint16)


def unwire_activation(raw: np.ndarray, logical_shape: Tuple[int, ...], mode: int) -> np.ndarray:
    """Inverse of wire_activation: bytes of a cut tensor -> float32 array of `logical_shape` (B, ..., C)."""
    c = logical_shape[-1]
    if mode != N.MODE_BF16:
        return raw.view(np.float32).reshape(logical_shape).copy()
    cs = (c + 7) // 8 * 8
    u = raw.view(np.uint16).astype(np.uint32) << 16
    x = u.view(np.float32).reshape(logical_shape[:-1] + (cs,))
    return np.ascontiguousarray(x[..., :c])


class LocalPredictor:
    """One stage holding an arbitrary layer range on one GPU, fed from / drained to host numpy
    arrays.  With the full range it is `model.predict` of reference test/local_infer.py:22."""

    def __init__(self, model: Model, mode: str = "bf16", device: int = 0, max_batch: int = 32,
                 start: Optional[str] = None, end: Optional[str] = None, input_format: str = "float32"):
        self.key = (mode, device, model._version)
        self.model = model
        start = start or model.layers[0].name
        end = end or model.layers[-1].name
        self.part = construct_model(model, start, end, "part1")
        self.ex = StageExecutor(model, self.part, device=device, mode=mode, max_batch=max_batch, input_format=input_format)
        self.input_u8 = self.part.takes_model_input and input_format == "uint8"
        self.mode = self.ex.mode
        self.device = device
        self.max_batch = max_batch
        self.lib = N.lib()
        self.d_in = DeviceBuffer(device, self.ex.in_bytes_per_image * max_batch)
        self.d_out = DeviceBuffer(device, self.ex.out_bytes_per_image * max_batch)
        self.ex.bind_io(0, self.d_in.ptr, self.d_out.ptr)
        self.ex.prepare(0, max_batch)

    def run_raw(self, x_wire: np.ndarray, batch: int) -> np.ndarray:
        """x_wire: host array already in wire format for the stage input; returns raw output bytes."""
        x_wire = np.ascontiguousarray(x_wire)
        assert x_wire.nbytes == self.ex.in_bytes_per_image * batch, (x_wire.nbytes, self.ex.in_bytes_per_image, batch)
        out = np.empty(self.ex.out_bytes_per_image * batch, dtype=np.uint8)
        N.check(self.lib.adapt_memcpy_h2d(self.device, self.d_in.ptr, x_wire.ctypes.data, x_wire.nbytes, None))
        self.ex.enqueue(0, batch, 0)
        N.check(self.lib.adapt_memcpy_d2h(self.device, out.ctypes.data, self.d_out.ptr, out.nbytes, None))
        N.check(self.lib.adapt_stream_sync(self.device, None))
        return out

    def run(self, x: np.ndarray) -> np.ndarray:
        """float32 in (model input or cut tensor, logical NHWC shape; raw uint8 RGB images with input_format="uint8")
        -> float32 out (logical shape)."""
        x = np.asarray(x, dtype=np.uint8 if self.input_u8 else np.float32)
        outs = []
        out_shape = self.part.output_shape[1:]
        for i in range(0, x.shape[0], self.max_batch):
            xb = x[i:i + self.max_batch]
            b = xb.shape[0]
            wire = np.ascontiguousarray(xb) if self.part.takes_model_input else wire_activation(xb, self.mode)
            raw = self.run_raw(wire, b)
            if self.ex.program.out_dtype == N.DT_F32:
                outs.append(raw.view(np.float32).reshape((b,) + out_shape).copy())
            else:
                outs.append(unwire_activation(raw, (b,) + out_shape, self.mode))
        return np.concatenate(outs, axis=0)

    def predict(self, x: np.ndarray, batch_size: int = 32) -> np.ndarray:
        return self.run(x)

    def logits(self, batch: int) -> np.ndarray:
        return self.ex.read_logits(batch)

    def close(self) -> None:
        self.ex.close()
        self.d_in.free()
        self.d_out.free()
