"""Graph slicing by layer name, and compilation of a slice into a stage program.

Mirror of reference src/dag_util.py (`get_previous`, `traverse_improved`, `construct_model`):
same names, same argument meaning, same error behaviour (an un-closed cut raises RuntimeError,
reference src/dag_util.py:38-43).  Where the reference rebuilds a `tf.keras.Model` by calling
the original layer objects on new symbolic tensors, this build records the layer range
(start, end] as a `PartModel` and compiles it into the op program the B200 stage executor runs
(include/adapt_b200.h): Conv2D+BatchNormalization(+Add)(+ReLU) groups fused into one
implicit-GEMM launch, ZeroPadding2D folded into the consumer, etc.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _native as N
from .resnet50 import BN_EPSILON, Layer, Model


def get_previous(model: Model, name: str) -> List[str]:
    """Names of the layers feeding `name` (reference src/dag_util.py:3-7)."""
    inbound = model.get_layer(name).inbound_nodes[0].inbound_layers
    if type(inbound) != list:  # noqa: E721  (same normalisation as the reference)
        inbound = [inbound]
    return [layer.name for layer in inbound]


def traverse_improved(model: Model, name: str, start: str, part_name: str, inpt, tensor_cache: Dict[str, object]):
    """Memoised walk from `name` back to `start` (reference src/dag_util.py:10-48).  Instead of
    re-applying Keras layers it records, in `tensor_cache`, every layer the sub-model needs (in
    dependency order) and returns the layer name as the "tensor"."""
    if name == start:
        return start
    if name == part_name:
        return inpt
    if name in tensor_cache:
        return tensor_cache[name]
    layer = model.get_layer(name)
    if not layer.inbound_nodes:
        # the reference gets here by trying to call the InputLayer and wraps the failure
        raise RuntimeError(
            f"Error calling layer {name} with inputs: []. The original model's connection might involve a merge "
            f"layer that needs to be explicitly handled.")
    for n in get_previous(model, name):
        traverse_improved(model, n, start, part_name, inpt, tensor_cache)
    tensor_cache[name] = name
    return name


class PartModel:
    """One pipeline partition: the layers strictly after `start` through `end` inclusive
    (what `construct_model` returns in the reference, a `tf.keras.Model`)."""

    def __init__(self, model: Model, start: str, end: str, layer_names: List[str], name: str):
        self.model = model
        self.start = start
        self.end = end
        self.layer_names = layer_names
        self.name = name
        self.takes_model_input = model.get_layer(start).class_name == "InputLayer"

    @property
    def layers(self) -> List[Layer]:
        return [self.model.get_layer(n) for n in self.layer_names]

    @property
    def input_shape(self) -> Tuple[Optional[int], ...]:
        return self.model.get_layer(self.start).output_shape

    @property
    def output_shape(self) -> Tuple[Optional[int], ...]:
        return self.model.get_layer(self.end).output_shape

    def get_weights(self) -> List[np.ndarray]:
        out: List[np.ndarray] = []
        for l in self.layers:
            out.extend(l.get_weights())
        return out

    def to_json(self) -> str:
        return json.dumps({"class_name": "Functional", "config": {
            "name": self.name, "input": self.start, "output": self.end, "layers": self.layer_names}})

    def macs(self) -> int:
        return sum(layer_macs(self.model, l) for l in self.layers)

    def __repr__(self) -> str:
        return f"<PartModel {self.name}: ({self.start}, {self.end}] {len(self.layer_names)} layers>"


def construct_model(model: Model, start: str, end: str, part_name: str = "part_begin") -> PartModel:
    """Sub-model (start, end] (reference src/dag_util.py:50-62).  Raises RuntimeError when the
    cut is not a single-tensor articulation point, ValueError for unknown layer names."""
    model.get_layer(start)
    model.get_layer(end)
    cache: Dict[str, object] = {}
    traverse_improved(model, end, start, part_name, part_name, cache)
    order = {l.name: i for i, l in enumerate(model.layers)}
    names = sorted((n for n in cache if n != part_name), key=order.__getitem__)
    if not names:
        raise RuntimeError(f"empty partition ({start}, {end}]")
    return PartModel(model, start, end, names, part_name)


def layer_macs(model: Model, l: Layer) -> int:
    if l.class_name == "Conv2D":
        k = l.config["kernel_size"][0]
        cin = model.get_layer(l.inbound_names[0]).output_shape[-1]
        ho, wo, co = l.output_shape[1:]
        return ho * wo * co * k * k * cin
    if l.class_name == "Dense":
        return model.get_layer(l.inbound_names[0]).output_shape[-1] * l.output_shape[-1]
    return 0


def valid_cut_points(model: Model) -> List[str]:
    """Layer names `part_at` may contain: every layer whose output is a single-tensor articulation
    point of the graph (cutting there closes both halves)."""
    out = []
    last = model.layers[-1].name
    first = model.layers[0].name
    for l in model.layers[1:-1]:
        try:
            construct_model(model, l.name, last)
            construct_model(model, first, l.name)
        except RuntimeError:
            continue
        out.append(l.name)
    return out


# ================================================================================================
# weight table: one entry per Conv2D/Dense with its BatchNormalization folded (fp64 math on the host),
# plus the few extra entries odd cuts inside the stem need.
@dataclass
class WeightEntry:
    key: str
    kernel: Optional[np.ndarray]  # HWIO float32 (Dense: (1,1,in,out)) or None for scale/bias only
    scale: np.ndarray
    bias: np.ndarray

    @property
    def geom(self) -> Tuple[int, int, int, int]:
        if self.kernel is None:
            return (0, 0, 0, int(self.scale.shape[0]))
        return tuple(int(v) for v in self.kernel.shape)  # kh, kw, cin, cout


def _bn_affine(bn: Layer) -> Tuple[np.ndarray, np.ndarray]:
    g, b, mu, var = (w.astype(np.float64) for w in bn._weights)
    s = g / np.sqrt(var + BN_EPSILON)
    return s, b - mu * s


def _consumers(model: Model) -> Dict[str, List[str]]:
    cons: Dict[str, List[str]] = {l.name: [] for l in model.layers}
    for l in model.layers:
        for n in l.inbound_names:
            cons[n].append(l.name)
    return cons


def is_s2d_stem(model: Model, l: Layer) -> bool:
    """A 7x7 / stride-2 / VALID Conv2D on a 3-channel image behind ZeroPadding2D(3) whose padded size is
    even: the ResNet stem.  It can run as a 4x4 stride-1 convolution over the space-to-depth image."""
    if l.class_name != "Conv2D" or l.config["kernel_size"] != [7, 7] or l.config["strides"] != [2, 2]:
        return False
    if l.config["padding"] != "valid":
        return False
    src = model.get_layer(l.inbound_names[0])
    if src.class_name != "ZeroPadding2D" or src.config["padding"] != [[3, 3], [3, 3]]:
        return False
    img = model.get_layer(src.inbound_names[0])
    h, w, c = src.output_shape[1:]
    return img.class_name == "InputLayer" and c == 3 and h % 2 == 0 and w % 2 == 0


def s2d_pitch(w: int) -> int:
    """Stored row width (pixels) of the space-to-depth image: whole 128-byte units of four pixels, and exactly 32
    units when the image fits (then a 32-row epilogue slab of the stem kernel is one image row)."""
    return 128 if w <= 128 else (w + 3) // 4 * 4


def s2d_kernel(k: np.ndarray) -> np.ndarray:
    """(7,7,3,Cout) stride-2 kernel -> (4,4,16,Cout) stride-1 kernel over the space-to-depth image whose
    channel (py*2+px)*3+c holds padded pixel (2*hh+py, 2*ww+px, c); taps with r or s == 7 and the four pad
    channels are zero."""
    out = np.zeros((4, 4, 16, k.shape[3]), dtype=np.float32)
    for r in range(7):
        dy, py = divmod(r, 2)
        for s in range(7):
            dx, px = divmod(s, 2)
            out[dy, dx, (py * 2 + px) * 3:(py * 2 + px) * 3 + 3, :] = k[r, s]
    return out


def build_weight_table(model: Model) -> Tuple[List[WeightEntry], Dict[str, int]]:
    cons = _consumers(model)
    entries: List[WeightEntry] = []
    index: Dict[str, int] = {}

    def put(key, kernel, scale, bias):
        index[key] = len(entries)
        entries.append(WeightEntry(key, None if kernel is None else np.ascontiguousarray(kernel, np.float32),
                                   np.ascontiguousarray(scale, np.float32), np.ascontiguousarray(bias, np.float32)))

    for l in model.layers:
        if l.class_name == "Conv2D":
            k, b = l._weights
            nxt = cons[l.name]
            bn = model.get_layer(nxt[0]) if len(nxt) == 1 and model.get_layer(nxt[0]).class_name == "BatchNormalization" else None
            if bn is not None:
                s, t = _bn_affine(bn)
                put(l.name + "+bn", k, s, b.astype(np.float64) * s + t)
                if l.name == "conv1_conv":  # the only conv whose BN a valid cut can split off
                    put(l.name, k, np.ones_like(s), b)
                    put(bn.name, None, s, t)
                if is_s2d_stem(model, l):
                    put(l.name + "+bn:s2d", s2d_kernel(k), s, b.astype(np.float64) * s + t)
            else:
                put(l.name, k, np.ones(k.shape[-1]), b)
        elif l.class_name == "Dense":
            k, b = l._weights
            put(l.name, k.reshape(1, 1, *k.shape), np.ones(k.shape[-1]), b)
    return entries, index


# ================================================================================================
# stage program
@dataclass
class VTensor:
    elems: int          # stored elements per image
    dtype: int          # N.DT_*
    role: int = N.T_INTERNAL
    phys: int = -1
    last_use: int = -1
    name: str = ""


@dataclass
class Op:
    kind: int
    inp: int
    out: int
    res: int = -1
    H: int = 0
    W: int = 0
    Cin: int = 0
    Ho: int = 0
    Wo: int = 0
    Cout: int = 0
    kh: int = 0
    kw: int = 0
    stride: int = 1
    pad: int = 0
    relu: int = 0
    weight: int = -1
    out_f32: int = 0
    in_format: int = 0  # INGEST: N.IN_*
    weight2: int = -1   # CONV_CHAIN: weight-table index of the 1x1
    layers: List[str] = field(default_factory=list)  # Keras layers this op covers


@dataclass
class StageProgram:
    ops: List[Op]
    tensors: List[Tuple[int, int, int]]  # physical buffers: (elems_per_image, dtype, role)
    in_elems: int
    in_dtype: int
    out_elems: int
    out_dtype: int
    taps: Dict[str, Tuple[int, Tuple[int, ...], int]]  # layer name -> (physical tensor, logical shape, dtype) for INTERNAL tensors
    macs: int

    def c_arrays(self):
        t = (N.TensorDesc * len(self.tensors))()
        for i, (e, d, r) in enumerate(self.tensors):
            t[i] = N.TensorDesc(e, d, r)
        o = (N.OpDesc * len(self.ops))()
        for i, p in enumerate(self.ops):
            o[i] = N.OpDesc(p.kind, p.inp, p.res, p.out, p.H, p.W, p.Cin, p.Ho, p.Wo, p.Cout, p.kh, p.kw, p.stride,
                            p.pad, p.relu, p.weight, p.out_f32, p.in_format, p.weight2)
        return t, o


def _stored_c(c: int, mode: int) -> int:
    return (c + 7) // 8 * 8 if mode == N.MODE_BF16 else c


def cut_tensor_elems(model: Model, layer_name: str, mode: int) -> int:
    shp = model.get_layer(layer_name).output_shape[1:]
    if len(shp) == 3:
        return shp[0] * shp[1] * _stored_c(shp[2], mode)
    return _stored_c(shp[0], mode)


def chainable(o: "Op") -> bool:
    """A `_2_conv` group the chain kernel can start from: 3x3 / stride 1 / pad 1, C -> C with C in {64, 128}, BN and ReLU
    fused, no shortcut, on a map at least 28 wide (conv2 and conv3 of the ResNets)."""
    return (o.kind == N.OP_CONV and o.kh == 3 and o.kw == 3 and o.stride == 1 and o.pad == 1 and o.relu == 1 and o.res < 0
            and o.out_f32 == 0 and o.Cin == o.Cout and o.Cout in (64, 128) and o.W >= 28)


def _chain_bottleneck_tails(ops: List["Op"], vt: List["VTensor"]) -> List["Op"]:
    """Fuse every `_2_conv`(3x3)+bn+relu whose ONLY consumer is the block's `_3_conv`(1x1)+bn+add+relu into one
    CONV_CHAIN op (csrc/conv_chain_sm100.cuh): the `_2_relu` tensor stays on chip.  The chained op takes the 1x1's place in
    the program (in a block1 the shortcut convolution `_0_conv` sits between the two and must run first)."""
    uses: Dict[int, List[int]] = {}
    for i, o in enumerate(ops):
        for t in (o.inp, o.res):
            if t >= 0:
                uses.setdefault(t, []).append(i)
    drop, repl = set(), {}
    for i, a in enumerate(ops):
        if not chainable(a) or vt[a.out].role != N.T_INTERNAL or uses.get(a.out, []) == []:
            continue
        cons = uses[a.out]
        if len(cons) != 1:
            continue
        j = cons[0]
        b = ops[j]
        if not (b.kind == N.OP_CONV and b.kh == 1 and b.kw == 1 and b.stride == 1 and b.pad == 0 and b.inp == a.out and b.res >= 0
                and b.res != a.out and b.relu == 1 and b.out_f32 == 0 and b.Cin == a.Cout and b.Cout == 4 * a.Cout
                and (b.H, b.W) == (a.Ho, a.Wo)):
            continue
        drop.add(i)
        repl[j] = Op(N.OP_CONV_CHAIN, a.inp, b.out, res=b.res, H=a.H, W=a.W, Cin=a.Cin, Ho=a.Ho, Wo=a.Wo, Cout=b.Cout, kh=3, kw=3,
                     stride=1, pad=1, relu=1, weight=a.weight, weight2=b.weight, layers=a.layers + b.layers)
    return [repl.get(i, o) for i, o in enumerate(ops) if i not in drop]


def _fuse_stem_pool(ops: List["Op"], vt: List["VTensor"]) -> List["Op"]:
    """Fuse the space-to-depth stem convolution (conv1_conv+bn+relu as 4x4 / stride 1 over 16-channel pixels) with the
    3x3 / stride 2 / pad 1 max-pool that is its only consumer into one STEM_POOL op (csrc/stem_pool_sm100.cuh): the
    112 x 112 x 64 map stays on chip."""
    uses: Dict[int, List[int]] = {}
    for i, o in enumerate(ops):
        for t in (o.inp, o.res):
            if t >= 0:
                uses.setdefault(t, []).append(i)
    for i, a in enumerate(ops):
        if not (a.kind == N.OP_CONV and a.kh == 4 and a.kw == 4 and a.stride == 1 and a.pad == 0 and a.Cin == 16 and a.Cout == 64
                and a.relu == 1 and a.res < 0 and a.out_f32 == 0 and a.W == 128 and a.Ho % 4 == 0 and a.Wo % 4 == 0
                and vt[a.out].role == N.T_INTERNAL and len(uses.get(a.out, [])) == 1):
            continue
        j = uses[a.out][0]
        b = ops[j]
        if not (b.kind == N.OP_MAXPOOL and b.kh == 3 and b.stride == 2 and b.pad == 1 and b.inp == a.out and b.Cin == 64
                and (b.H, b.W) == (a.Ho, a.Wo) and (b.Ho, b.Wo) == (a.Ho // 2, a.Wo // 2) and j == i + 1):
            continue
        fused = Op(N.OP_STEM_POOL, a.inp, b.out, H=a.H, W=a.W, Cin=a.Cin, Ho=a.Ho, Wo=a.Wo, Cout=64, kh=4, kw=4, stride=1, pad=0,
                   relu=1, weight=a.weight, layers=a.layers + b.layers)
        return ops[:i] + [fused] + ops[j + 1:]
    return ops


def compile_part(part: PartModel, windex: Dict[str, int], mode: int = N.MODE_BF16,
                 input_format: int = N.IN_F32) -> StageProgram:
    """Fuse the Keras layers of `part` into stage ops and assign activation buffers.  `input_format` (N.IN_*) only
    matters for the part that takes the model input: fp32 preprocessed images (the reference's contract) or raw
    uint8 RGB images preprocessed inside the ingest kernel."""
    model = part.model
    cons = _consumers(model)
    in_part = set(part.layer_names)
    vt: List[VTensor] = []
    ops: List[Op] = []
    val: Dict[str, int] = {}       # layer name -> virtual tensor holding its output
    pend_pad: Dict[str, Tuple[int, int]] = {}  # ZeroPadding2D layer folded into its consumer: (src vt, pad)
    done = set()

    def new_t(elems, dtype=N.DT_WORK, name="") -> int:
        vt.append(VTensor(int(elems), dtype, name=name))
        return len(vt) - 1

    def shape_of(name) -> Tuple[int, ...]:
        return model.get_layer(name).output_shape[1:]

    def sole_consumer(name: str) -> Optional[Layer]:
        c = cons[name]
        if len(c) == 1 and c[0] in in_part and c[0] not in done:
            return model.get_layer(c[0])
        return None

    # stage input
    s2d_pad: Optional[str] = None  # ZeroPadding2D layer absorbed by the space-to-depth ingest
    if part.takes_model_input:
        h, w, c = shape_of(part.start)
        if input_format == N.IN_U8_RGB_CAFFE and c != 3:
            raise ValueError("uint8 input needs a 3-channel image")
        t_in = new_t(h * w * c, N.DT_U8 if input_format == N.IN_U8_RGB_CAFFE else N.DT_F32, part.start)
        vt[t_in].role = N.T_INPUT
        # bf16 stem fast path: pad -> 7x7/2 conv -> BN all inside this part => space-to-depth ingest
        if mode == N.MODE_BF16 and len(part.layer_names) >= 3:
            l0, l1, l2 = (model.get_layer(n) for n in part.layer_names[:3])
            if (l0.class_name == "ZeroPadding2D" and is_s2d_stem(model, l1) and l1.inbound_names == [l0.name]
                    and l2.class_name == "BatchNormalization" and (l1.name + "+bn:s2d") in windex):
                s2d_pad = l0.name
        if s2d_pad is not None:
            hp, wp = h + 6, w + 6
            # rows are stored S2D_PITCH(wp // 2) pixels wide (zero beyond the image): whole 128-byte units of four
            # 16-channel pixels, which is what the stem's phase-decomposed loads walk
            t0 = new_t((hp // 2) * s2d_pitch(wp // 2) * 16, N.DT_WORK, part.start + ":s2d")
            ops.append(Op(N.OP_INGEST, t_in, t0, H=h, W=w, Cin=c, Ho=hp // 2, Wo=s2d_pitch(wp // 2), Cout=16, pad=3,
                          in_format=input_format, layers=[part.start, s2d_pad]))
            done.add(s2d_pad)
            val[s2d_pad] = t0
        else:
            t0 = new_t(h * w * _stored_c(c, mode), N.DT_WORK, part.start + ":ingest")
            ops.append(Op(N.OP_INGEST, t_in, t0, H=h, W=w, Cin=c, Ho=h, Wo=w, Cout=_stored_c(c, mode),
                          in_format=input_format, layers=[part.start]))
        val[part.start] = t0
    else:
        t_in = new_t(cut_tensor_elems(model, part.start, mode), N.DT_WORK, part.start)
        vt[t_in].role = N.T_INPUT
        val[part.start] = t_in

    for name in part.layer_names:
        if name in done:
            continue
        l = model.get_layer(name)
        done.add(name)
        src_name = l.inbound_names[0]
        cls = l.class_name
        if cls == "ZeroPadding2D":
            p = l.config["padding"][0][0]
            nxt = sole_consumer(name)
            if nxt is not None and nxt.class_name in ("Conv2D", "MaxPooling2D") and nxt.config.get("padding") == "valid":
                pend_pad[name] = (val[src_name], p)
                continue
            h, w, c = shape_of(src_name)
            t = new_t((h + 2 * p) * (w + 2 * p) * _stored_c(c, mode), name=name)
            ops.append(Op(N.OP_ZEROPAD, val[src_name], t, H=h, W=w, Cin=_stored_c(c, mode), Ho=h + 2 * p, Wo=w + 2 * p,
                          Cout=_stored_c(c, mode), pad=p, layers=[name]))
            val[name] = t
        elif cls in ("Conv2D", "Dense"):
            covered = [name]
            if cls == "Dense":
                cin = shape_of(src_name)[0]
                h = w = ho = wo = 1
                k, stride, pad, cout = 1, 1, 0, l.config["units"]
                src_t = val[src_name]
            else:
                k = l.config["kernel_size"][0]
                stride = l.config["strides"][0]
                cout = l.config["filters"]
                if src_name == s2d_pad:
                    # 7x7 stride-2 stem as a 4x4 stride-1 VALID convolution over the space-to-depth image
                    src_t = val[src_name]
                    hp, wp, _ = shape_of(src_name)
                    h, w, cin = hp // 2, s2d_pitch(wp // 2), 16   # (stored width; the output size comes from the layer)
                    k, stride, pad = 4, 1, 0
                elif src_name in pend_pad:
                    src_t, pad = pend_pad[src_name]
                    covered.insert(0, src_name)
                    h, w, cin = shape_of(model.get_layer(src_name).inbound_names[0])
                else:
                    src_t = val[src_name]
                    h, w, cin = shape_of(src_name)
                    pad = (k - 1) // 2 if l.config["padding"] == "same" else 0
                    if l.config["padding"] == "same" and stride != 1:
                        raise NotImplementedError("strided SAME convolution (asymmetric padding) is not in ResNet v1")
                ho, wo = l.output_shape[1:3]
            wkey, relu, res_t, out_f32 = name, 0, -1, 0
            tail = l
            nxt = sole_consumer(tail.name)
            if nxt is not None and nxt.class_name == "BatchNormalization" and (name + "+bn") in windex:
                wkey = name + ("+bn:s2d" if src_name == s2d_pad else "+bn")
                done.add(nxt.name)
                covered.append(nxt.name)
                tail = nxt
                nxt = sole_consumer(tail.name)
                if nxt is not None and nxt.class_name == "Add":
                    other = [n for n in nxt.inbound_names if n != tail.name]
                    if len(other) == 1 and other[0] in val:
                        res_t = val[other[0]]
                        done.add(nxt.name)
                        covered.append(nxt.name)
                        tail = nxt
                        nxt = sole_consumer(tail.name)
            elif cls == "Conv2D" and (name + "+bn") in windex and name not in windex:
                raise RuntimeError(f"cut separates {name} from its BatchNormalization; unsupported split")
            if cls == "Dense":
                out_f32 = 1  # logits stay fp32; the softmax activation is its own op
            elif nxt is not None and nxt.class_name == "Activation":
                relu = 1
                done.add(nxt.name)
                covered.append(nxt.name)
                tail = nxt
            t = new_t(ho * wo * (cout if out_f32 else _stored_c(cout, mode)), N.DT_F32 if out_f32 else N.DT_WORK,
                      tail.name + (":logits" if cls == "Dense" else ""))
            ops.append(Op(N.OP_CONV, src_t, t, res=res_t, H=h, W=w, Cin=_stored_c(cin, mode), Ho=ho, Wo=wo, Cout=cout,
                          kh=k, kw=k, stride=stride, pad=pad, relu=relu, weight=windex[wkey], out_f32=out_f32,
                          layers=covered))
            for n in covered:
                val[n] = t
            if cls == "Dense":
                tp = new_t(cout, N.DT_F32, name)
                ops.append(Op(N.OP_SOFTMAX, t, tp, H=1, W=1, Cin=cout, Ho=1, Wo=1, Cout=cout, layers=[name + ":softmax"]))
                val[name] = tp
        elif cls in ("BatchNormalization", "Activation"):
            covered = [name]
            h, w, c = shape_of(name)
            weight, relu = -1, 0
            tail = l
            if cls == "BatchNormalization":
                if name not in windex:
                    raise RuntimeError(f"stand-alone BatchNormalization {name} is not supported at this cut")
                weight = windex[name]
                nxt = sole_consumer(name)
                if nxt is not None and nxt.class_name == "Activation":
                    relu = 1
                    done.add(nxt.name)
                    covered.append(nxt.name)
                    tail = nxt
            else:
                relu = 1
            t = new_t(h * w * _stored_c(c, mode), name=tail.name)
            ops.append(Op(N.OP_AFFINE, val[src_name], t, H=h, W=w, Cin=_stored_c(c, mode), Ho=h, Wo=w,
                          Cout=_stored_c(c, mode), relu=relu, weight=weight, layers=covered))
            for n in covered:
                val[n] = t
        elif cls == "MaxPooling2D":
            covered = [name]
            k = l.config["pool_size"][0]
            stride = l.config["strides"][0]
            if src_name in pend_pad:
                src_t, pad = pend_pad[src_name]
                covered.insert(0, src_name)
                h, w, c = shape_of(model.get_layer(src_name).inbound_names[0])
            else:
                src_t, pad = val[src_name], 0
                h, w, c = shape_of(src_name)
            ho, wo = l.output_shape[1:3]
            t = new_t(ho * wo * c, name=name)
            ops.append(Op(N.OP_MAXPOOL, src_t, t, H=h, W=w, Cin=c, Ho=ho, Wo=wo, Cout=c, kh=k, kw=k, stride=stride, pad=pad,
                          layers=covered))
            val[name] = t
        elif cls == "GlobalAveragePooling2D":
            h, w, c = shape_of(src_name)
            t = new_t(c, name=name)
            ops.append(Op(N.OP_GAP, val[src_name], t, H=h, W=w, Cin=c, Ho=1, Wo=1, Cout=c, layers=[name]))
            val[name] = t
        elif cls == "Add":
            raise RuntimeError(f"stand-alone Add {name}: a cut may not separate a merge from its inputs")
        else:
            raise NotImplementedError(f"layer class {cls}")

    t_out = val[part.end]
    if vt[t_out].role == N.T_INPUT:
        raise RuntimeError("partition has no compute")
    vt[t_out].role = N.T_OUTPUT
    if mode == N.MODE_BF16 and os.environ.get("ADAPT_CHAIN", "1") != "0":
        ops = _chain_bottleneck_tails(ops, vt)
    if mode == N.MODE_BF16 and os.environ.get("ADAPT_STEM_POOL", "1") != "0":
        ops = _fuse_stem_pool(ops, vt)

    # ---- liveness-based physical buffer assignment (internal tensors only)
    for i, o in enumerate(ops):
        for t in (o.inp, o.res):
            if t >= 0:
                vt[t].last_use = i
    phys: List[Tuple[int, int, int]] = []
    free: List[int] = []
    for v in vt:
        if v.role != N.T_INTERNAL:
            v.phys = len(phys)
            phys.append((v.elems, v.dtype, v.role))
    for i, o in enumerate(ops):
        v = vt[o.out]
        if v.role == N.T_INTERNAL:
            best = None
            for p in free:
                if phys[p][1] == v.dtype and phys[p][0] >= v.elems and (best is None or phys[p][0] < phys[best][0]):
                    best = p
            if best is None:
                # grow the largest free buffer of that dtype instead of adding another one
                cands = [p for p in free if phys[p][1] == v.dtype]
                if cands:
                    best = max(cands, key=lambda p: phys[p][0])
                    phys[best] = (v.elems, v.dtype, N.T_INTERNAL)
                else:
                    best = len(phys)
                    phys.append((v.elems, v.dtype, N.T_INTERNAL))
            if best in free:
                free.remove(best)
            v.phys = best
        for t in {o.inp, o.res}:
            if t >= 0 and vt[t].role == N.T_INTERNAL and vt[t].last_use == i and vt[t].phys not in free:
                free.append(vt[t].phys)
    taps: Dict[str, Tuple[int, Tuple[int, ...], int]] = {}
    for o in ops:
        # the fp32 logits buffer is never recycled after the softmax, so it can be read back
        if o.kind == N.OP_CONV and o.out_f32 and vt[o.out].role == N.T_INTERNAL:
            taps["logits"] = (vt[o.out].phys, (o.Cout,), N.DT_F32)
    for o in ops:
        o.inp = vt[o.inp].phys
        o.res = vt[o.res].phys if o.res >= 0 else -1
        o.out = vt[o.out].phys
    vin, vout = vt[t_in], vt[t_out]
    return StageProgram(ops=ops, tensors=phys, in_elems=vin.elems, in_dtype=vin.dtype, out_elems=vout.elems,
                        out_dtype=vout.dtype, taps=taps, macs=part.macs())
