"""CPU ORACLE -- test infrastructure, NOT part of the product.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module, and only as the checker / the timed CPU baseline.

What it restates
----------------
The reference (ADAPT/DEFER) performs no arithmetic itself: every stage calls
`model.predict` on a slice of a Keras model (reference src/node.py:177; slices built by
`construct_model`, reference src/dag_util.py:50-62, from the `part_at` names handed to
`DEFER._partition`, reference src/dispatcher.py:39-53; the model is
`tf.keras.applications.ResNet50`, reference test/test.py:13, test/local_infer.py:8).
The arithmetic therefore lives in a third-party dependency that is absent from
/root/reference and from this image: TensorFlow 2.15.0 / Keras 2.15 (pinned only in prose,
reference README.md:19,29).  This file restates the published algorithm of
keras/applications/resnet.py `ResNet50` (ResNet *v1*: stride on the first 1x1 conv of
convN_block1, biased convolutions, BatchNormalization epsilon 1.001e-5, explicit
ZeroPadding2D) layer by layer in float32 numpy, un-fused, so that ANY layer range
(start, end] can be evaluated -- which is exactly what a pipeline stage computes.

PARITY UNPINNED: the reference ships no golden vectors or numeric assertions for this path
(reference test/test.py:33 prints a shape) and TensorFlow cannot be installed here, so the
oracle cannot be checked against reference outputs.  What *is* pinned (tests/test_oracle.py):
the topology constants every Keras install reports for ResNet50 -- 177 layers, 25,636,712
parameters, 320 weight arrays, output shapes per layer -- the MAC count 3,857,973,248, and
agreement between two independent convolution formulations (im2col GEMM here vs. torch's
direct convolution).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

BN_EPS = 1.001e-5  # keras.applications.resnet: BatchNormalization(epsilon=1.001e-5)


@dataclass
class LayerSpec:
    name: str
    kind: str  # InputLayer ZeroPadding2D Conv2D BatchNormalization Activation MaxPooling2D Add
    #            GlobalAveragePooling2D Dense
    inbound: List[str] = field(default_factory=list)
    # hyper-parameters (only the ones the kind uses)
    filters: int = 0
    kernel: int = 0
    stride: int = 1
    padding: str = "valid"
    pad: int = 0  # ZeroPadding2D amount on every side
    units: int = 0
    out_shape: Tuple[int, ...] = ()  # per-image (H, W, C) or (C,)


RESNET_BLOCKS = {"resnet50": (3, 4, 6, 3), "resnet101": (3, 4, 23, 3), "resnet152": (3, 8, 36, 3)}


def build_resnet50_layers(blocks_per_stack: Tuple[int, int, int, int] = (3, 4, 6, 3)) -> List[LayerSpec]:
    """Layer list in the order Keras reports for ResNet50(include_top=True) (`model.layers`).  The same v1 builder
    with other stack depths gives Keras' ResNet101 (3, 4, 23, 3) and ResNet152 (3, 8, 36, 3)."""
    L: List[LayerSpec] = []
    L.append(LayerSpec("input_1", "InputLayer", [], out_shape=(224, 224, 3)))
    L.append(LayerSpec("conv1_pad", "ZeroPadding2D", ["input_1"], pad=3))
    L.append(LayerSpec("conv1_conv", "Conv2D", ["conv1_pad"], filters=64, kernel=7, stride=2))
    L.append(LayerSpec("conv1_bn", "BatchNormalization", ["conv1_conv"]))
    L.append(LayerSpec("conv1_relu", "Activation", ["conv1_bn"]))
    L.append(LayerSpec("pool1_pad", "ZeroPadding2D", ["conv1_relu"], pad=1))
    L.append(LayerSpec("pool1_pool", "MaxPooling2D", ["pool1_pad"], kernel=3, stride=2))
    prev = "pool1_pool"
    b2, b3, b4, b5 = blocks_per_stack
    for stack, filters, blocks, stride1 in (("conv2", 64, b2, 1), ("conv3", 128, b3, 2), ("conv4", 256, b4, 2),
                                            ("conv5", 512, b5, 2)):
        for b in range(1, blocks + 1):
            n = f"{stack}_block{b}"
            s = stride1 if b == 1 else 1
            L.append(LayerSpec(f"{n}_1_conv", "Conv2D", [prev], filters=filters, kernel=1, stride=s))
            L.append(LayerSpec(f"{n}_1_bn", "BatchNormalization", [f"{n}_1_conv"]))
            L.append(LayerSpec(f"{n}_1_relu", "Activation", [f"{n}_1_bn"]))
            L.append(LayerSpec(f"{n}_2_conv", "Conv2D", [f"{n}_1_relu"], filters=filters, kernel=3, padding="same"))
            L.append(LayerSpec(f"{n}_2_bn", "BatchNormalization", [f"{n}_2_conv"]))
            L.append(LayerSpec(f"{n}_2_relu", "Activation", [f"{n}_2_bn"]))
            if b == 1:
                L.append(LayerSpec(f"{n}_0_conv", "Conv2D", [prev], filters=4 * filters, kernel=1, stride=s))
            L.append(LayerSpec(f"{n}_3_conv", "Conv2D", [f"{n}_2_relu"], filters=4 * filters, kernel=1))
            if b == 1:
                L.append(LayerSpec(f"{n}_0_bn", "BatchNormalization", [f"{n}_0_conv"]))
            L.append(LayerSpec(f"{n}_3_bn", "BatchNormalization", [f"{n}_3_conv"]))
            shortcut = f"{n}_0_bn" if b == 1 else prev
            L.append(LayerSpec(f"{n}_add", "Add", [shortcut, f"{n}_3_bn"]))
            L.append(LayerSpec(f"{n}_out", "Activation", [f"{n}_add"]))
            prev = f"{n}_out"
    L.append(LayerSpec("avg_pool", "GlobalAveragePooling2D", [prev]))
    L.append(LayerSpec("predictions", "Dense", ["avg_pool"], units=1000))
    _infer_shapes(L)
    return L


def _infer_shapes(layers: List[LayerSpec]) -> None:
    shp: Dict[str, Tuple[int, ...]] = {}
    for l in layers:
        if l.kind == "InputLayer":
            pass
        else:
            i = shp[l.inbound[0]]
            if l.kind == "ZeroPadding2D":
                l.out_shape = (i[0] + 2 * l.pad, i[1] + 2 * l.pad, i[2])
            elif l.kind == "Conv2D":
                if l.padding == "same":
                    ho, wo = -(-i[0] // l.stride), -(-i[1] // l.stride)
                else:
                    ho, wo = (i[0] - l.kernel) // l.stride + 1, (i[1] - l.kernel) // l.stride + 1
                l.out_shape = (ho, wo, l.filters)
            elif l.kind == "MaxPooling2D":
                l.out_shape = ((i[0] - l.kernel) // l.stride + 1, (i[1] - l.kernel) // l.stride + 1, i[2])
            elif l.kind == "GlobalAveragePooling2D":
                l.out_shape = (i[2],)
            elif l.kind == "Dense":
                l.out_shape = (l.units,)
            else:  # BatchNormalization, Activation, Add
                l.out_shape = i
        shp[l.name] = l.out_shape


def weight_shapes(layers: Sequence[LayerSpec]) -> List[Tuple[str, str, Tuple[int, ...]]]:
    """(layer, array name, shape) in `model.get_weights()` order."""
    shp = {l.name: l.out_shape for l in layers}
    out = []
    for l in layers:
        if l.kind == "Conv2D":
            cin = shp[l.inbound[0]][2]
            out.append((l.name, "kernel", (l.kernel, l.kernel, cin, l.filters)))
            out.append((l.name, "bias", (l.filters,)))
        elif l.kind == "BatchNormalization":
            c = l.out_shape[-1]
            for a in ("gamma", "beta", "moving_mean", "moving_variance"):
                out.append((l.name, a, (c,)))
        elif l.kind == "Dense":
            cin = shp[l.inbound[0]][0]
            out.append((l.name, "kernel", (cin, l.units)))
            out.append((l.name, "bias", (l.units,)))
    return out


def init_weights(layers: Sequence[LayerSpec], seed: int = 1234, randomize_bn: bool = False,
                 randomize_bias: bool = False) -> Dict[str, List[np.ndarray]]:
    """Keras `weights=None` initialisation restated: glorot-uniform kernels, zero biases,
    gamma=1, beta=0, moving_mean=0, moving_variance=1 (values drawn from numpy's PCG64, not TF's
    RNG: parity only needs the SAME tensors on both sides).  `randomize_bn` / `randomize_bias`
    give a second set that exercises BatchNormalization folding and the bias path."""
    rng = np.random.default_rng(seed)
    shp = {l.name: l.out_shape for l in layers}
    W: Dict[str, List[np.ndarray]] = {}
    for l in layers:
        if l.kind == "Conv2D":
            cin = shp[l.inbound[0]][2]
            fan_in, fan_out = l.kernel * l.kernel * cin, l.kernel * l.kernel * l.filters
            lim = np.sqrt(6.0 / (fan_in + fan_out))
            k = rng.uniform(-lim, lim, (l.kernel, l.kernel, cin, l.filters)).astype(np.float32)
            b = (rng.normal(0, 0.1, l.filters) if randomize_bias else np.zeros(l.filters)).astype(np.float32)
            W[l.name] = [k, b]
        elif l.kind == "Dense":
            cin = shp[l.inbound[0]][0]
            lim = np.sqrt(6.0 / (cin + l.units))
            k = rng.uniform(-lim, lim, (cin, l.units)).astype(np.float32)
            b = (rng.normal(0, 0.1, l.units) if randomize_bias else np.zeros(l.units)).astype(np.float32)
            W[l.name] = [k, b]
        elif l.kind == "BatchNormalization":
            c = l.out_shape[-1]
            if randomize_bn:
                g = rng.uniform(0.5, 1.5, c)
                be = rng.normal(0, 0.2, c)
                mu = rng.normal(0, 0.2, c)
                var = rng.uniform(0.5, 1.5, c)
            else:
                g, be, mu, var = np.ones(c), np.zeros(c), np.zeros(c), np.ones(c)
            W[l.name] = [a.astype(np.float32) for a in (g, be, mu, var)]
    return W


# ------------------------------------------------------------------------------------ layer math
def zero_pad(x: np.ndarray, p: int) -> np.ndarray:
    return np.pad(x, ((0, 0), (p, p), (p, p), (0, 0)))


def conv2d(x: np.ndarray, k: np.ndarray, b: np.ndarray, stride: int, padding: str) -> np.ndarray:
    """NHWC float32 convolution as im2col + GEMM.  `same` padding follows TensorFlow:
    total = max((ceil(H/s)-1)*s + kh - H, 0), before = total // 2 (odd remainder goes after)."""
    kh, kw, cin, cout = k.shape
    B, H, W, _ = x.shape
    if padding == "same":
        ho, wo = -(-H // stride), -(-W // stride)
        ph = max((ho - 1) * stride + kh - H, 0)
        pw = max((wo - 1) * stride + kw - W, 0)
        x = np.pad(x, ((0, 0), (ph // 2, ph - ph // 2), (pw // 2, pw - pw // 2), (0, 0)))
    else:
        ho, wo = (H - kh) // stride + 1, (W - kw) // stride + 1
    if kh == 1 and kw == 1:
        cols = x[:, ::stride, ::stride, :][:, :ho, :wo, :].reshape(B * ho * wo, cin)
    else:
        win = np.lib.stride_tricks.sliding_window_view(x, (kh, kw), axis=(1, 2))  # B,H',W',C,kh,kw
        win = win[:, ::stride, ::stride][:, :ho, :wo]
        cols = np.ascontiguousarray(win.transpose(0, 1, 2, 4, 5, 3)).reshape(B * ho * wo, kh * kw * cin)
    y = cols.astype(np.float32) @ k.reshape(kh * kw * cin, cout)
    y += b
    return y.reshape(B, ho, wo, cout)


def batch_norm(x: np.ndarray, gamma, beta, mean, var) -> np.ndarray:
    return (gamma * (x - mean) / np.sqrt(var + np.float32(BN_EPS)) + beta).astype(np.float32)


def relu(x: np.ndarray) -> np.ndarray:
    return np.maximum(x, np.float32(0))


def max_pool(x: np.ndarray, k: int, stride: int) -> np.ndarray:
    win = np.lib.stride_tricks.sliding_window_view(x, (k, k), axis=(1, 2))[:, ::stride, ::stride]
    return win.max(axis=(-1, -2))


def global_avg_pool(x: np.ndarray) -> np.ndarray:
    return x.mean(axis=(1, 2), dtype=np.float32)


def softmax(z: np.ndarray) -> np.ndarray:
    z = z - z.max(axis=-1, keepdims=True)
    e = np.exp(z)
    return (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)


def bf16_round(x: np.ndarray) -> np.ndarray:
    """Round float32 to the nearest bfloat16 (ties to even), returned as float32."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    return ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)


# ------------------------------------------------------------------------------------ evaluation
class ResNet50Oracle:
    def __init__(self, weights: Optional[Dict[str, List[np.ndarray]]] = None, seed: int = 1234,
                 blocks_per_stack: Optional[Tuple[int, int, int, int]] = None):
        if blocks_per_stack is None:
            # infer the depth from the weight names when weights are given (conv4_block23_1_conv => ResNet-101)
            blocks_per_stack = (3, 4, 6, 3)
            if weights is not None:
                depth = [0, 0, 0, 0]
                for name in weights:
                    if name.startswith("conv") and "_block" in name:
                        stack = int(name[4]) - 2
                        if 0 <= stack < 4:
                            depth[stack] = max(depth[stack], int(name.split("_block")[1].split("_")[0]))
                if all(depth):
                    blocks_per_stack = tuple(depth)
        self.layers = build_resnet50_layers(blocks_per_stack)
        self.by_name = {l.name: l for l in self.layers}
        self.weights = weights if weights is not None else init_weights(self.layers, seed)
        self.logits: Optional[np.ndarray] = None  # pre-softmax tap of the last evaluated `predictions`

    # -- topology helpers ------------------------------------------------------------------
    def layers_between(self, start: str, end: str) -> List[str]:
        """Names of the layers a sub-model (start, end] contains, found the way the reference's
        `traverse_improved` walks (reference src/dag_util.py:10-48): backwards from `end` through
        inbound layers until `start`.  Raises RuntimeError when some path reaches the model
        input without meeting `start` (an invalid cut, reference src/dag_util.py:38-43)."""
        if start not in self.by_name or end not in self.by_name:
            raise ValueError(f"unknown layer {start!r} or {end!r}")
        needed: List[str] = []
        seen = set()

        def walk(name: str) -> None:
            if name == start or name in seen:
                return
            l = self.by_name[name]
            if not l.inbound:
                raise RuntimeError(f"cut ({start}, {end}] is not closed: path reaches {name}")
            seen.add(name)
            for n in l.inbound:
                walk(n)
            needed.append(name)

        walk(end)
        order = {l.name: i for i, l in enumerate(self.layers)}
        return sorted(needed, key=order.__getitem__)

    # -- arithmetic --------------------------------------------------------------------------
    def _apply(self, l: LayerSpec, ins: List[np.ndarray]) -> np.ndarray:
        w = self.weights.get(l.name)
        if l.kind == "ZeroPadding2D":
            return zero_pad(ins[0], l.pad)
        if l.kind == "Conv2D":
            return conv2d(ins[0], w[0], w[1], l.stride, l.padding)
        if l.kind == "BatchNormalization":
            return batch_norm(ins[0], *w)
        if l.kind == "Activation":
            return relu(ins[0])
        if l.kind == "MaxPooling2D":
            return max_pool(ins[0], l.kernel, l.stride)
        if l.kind == "Add":
            return (ins[0] + ins[1]).astype(np.float32)
        if l.kind == "GlobalAveragePooling2D":
            return global_avg_pool(ins[0])
        if l.kind == "Dense":
            self.logits = (ins[0] @ w[0] + w[1]).astype(np.float32)
            return softmax(self.logits)
        raise ValueError(l.kind)

    def run_range(self, x: np.ndarray, start: str, end: str, taps: Sequence[str] = ()) -> Tuple[np.ndarray, Dict[str, np.ndarray]]:
        """Evaluate the sub-model (start, end] on `x` (= the output of layer `start`)."""
        names = self.layers_between(start, end)
        vals: Dict[str, np.ndarray] = {start: np.asarray(x, dtype=np.float32)}
        tapped: Dict[str, np.ndarray] = {}
        for n in names:
            l = self.by_name[n]
            vals[n] = self._apply(l, [vals[i] for i in l.inbound])
            if n in taps:
                tapped[n] = vals[n]
        return vals[end], tapped

    def predict(self, x: np.ndarray, taps: Sequence[str] = ()) -> Tuple[np.ndarray, Dict[str, np.ndarray]]:
        return self.run_range(x, "input_1", "predictions", taps)

    def pipeline(self, x: np.ndarray, part_at: Sequence[str]) -> np.ndarray:
        """Chain of sub-models exactly as `DEFER._partition` cuts them (reference
        src/dispatcher.py:39-53): part p = (part_at[p-1], part_at[p]]."""
        bounds = ["input_1"] + list(part_at) + ["predictions"]
        for a, b in zip(bounds[:-1], bounds[1:]):
            x, _ = self.run_range(x, a, b)
        return x


def total_macs(layers: Sequence[LayerSpec]) -> int:
    shp = {l.name: l.out_shape for l in layers}
    macs = 0
    for l in layers:
        if l.kind == "Conv2D":
            cin = shp[l.inbound[0]][2]
            macs += l.out_shape[0] * l.out_shape[1] * l.filters * l.kernel * l.kernel * cin
        elif l.kind == "Dense":
            macs += shp[l.inbound[0]][0] * l.units
    return macs


def total_params(layers: Sequence[LayerSpec]) -> int:
    return sum(int(np.prod(s)) for _, _, s in weight_shapes(layers))


def synthetic_images(batch: int, seed: int = 0, kind: str = "caffe") -> np.ndarray:
    """Synthetic inputs of SURVEY 8(d): uniform [0,255) RGB with Keras caffe-mode
    `preprocess_input` (RGB->BGR, subtract ImageNet means; reference test/test.py:23), or N(0,1)."""
    rng = np.random.default_rng(seed)
    if kind == "caffe":
        x = rng.uniform(0, 255, (batch, 224, 224, 3)).astype(np.float32)
        x = x[..., ::-1] - np.array([103.939, 116.779, 123.68], dtype=np.float32)
        return np.ascontiguousarray(x, dtype=np.float32)
    if kind == "normal":
        return rng.standard_normal((batch, 224, 224, 3)).astype(np.float32)
    raise ValueError(kind)
