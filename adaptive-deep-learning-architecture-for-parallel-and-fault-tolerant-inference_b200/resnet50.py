"""Keras-compatible ResNet stand-in: the `model` argument of `DEFER.run_defer`.

The reference hands `run_defer` a `tf.keras.Model` (reference test/test.py:13,43) and touches
exactly this part of the Keras API: `model.input._keras_history[0].name`,
`model.output._keras_history[0].name` (reference src/dispatcher.py:43,48),
`get_layer(name).inbound_nodes[0].inbound_layers`, `.output` (reference src/dag_util.py:4,13),
`to_json()` (src/dispatcher.py:235), `get_weights()` (src/dispatcher.py:243) and, on the worker,
`set_weights` / `predict` (src/node.py:45,78,177).  TensorFlow is not available on the B200
box, so this module supplies an object with that surface: the layer graph of
`keras.applications.ResNet50` (v1) with Keras layer names in Keras order, weights in Keras
layouts (Conv2D [kernel HWIO, bias]; BatchNormalization [gamma, beta, moving_mean,
moving_variance]; Dense [kernel (in, out), bias]) and Keras `weights=None` initialisers.
ResNet101 / ResNet152 share every kernel and are provided by the same builder.
"""
from __future__ import annotations

import json
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

BN_EPSILON = 1.001e-5


class KerasTensor:
    """Symbolic output of a layer; only `_keras_history[0]` (the producing layer) and `shape` are used."""

    def __init__(self, layer: "Layer", shape: Tuple[Optional[int], ...]):
        self._keras_history = (layer, 0, 0)
        self.shape = shape
        self.name = layer.name

    def __repr__(self) -> str:
        return f"<KerasTensor shape={self.shape} from {self._keras_history[0].name!r}>"


class InboundNode:
    def __init__(self, inbound_layers):
        # Keras returns a bare layer for one input and a list for several (the reference
        # normalises both, src/dag_util.py:4-6)
        self.inbound_layers = inbound_layers


class Layer:
    def __init__(self, name: str, class_name: str, config: dict, inbound: List["Layer"], out_shape: Tuple[int, ...]):
        self.name = name
        self.class_name = class_name
        self.config = dict(config, name=name)
        self._inbound = inbound
        self.inbound_nodes = [InboundNode(inbound[0] if len(inbound) == 1 else list(inbound))] if inbound else []
        self.output_shape = (None,) + tuple(out_shape)
        self.output = KerasTensor(self, self.output_shape)
        self._weights: List[np.ndarray] = []

    # Keras API
    def get_weights(self) -> List[np.ndarray]:
        return [w.copy() for w in self._weights]

    def set_weights(self, weights: Sequence[np.ndarray]) -> None:
        if len(weights) != len(self._weights):
            raise ValueError(f"layer {self.name}: expected {len(self._weights)} arrays, got {len(weights)}")
        for cur, new in zip(self._weights, weights):
            if tuple(cur.shape) != tuple(np.shape(new)):
                raise ValueError(f"layer {self.name}: shape {np.shape(new)} != {cur.shape}")
        self._weights = [np.ascontiguousarray(w, dtype=np.float32) for w in weights]

    def count_params(self) -> int:
        return int(sum(w.size for w in self._weights))

    @property
    def inbound_names(self) -> List[str]:
        return [l.name for l in self._inbound]

    def __repr__(self) -> str:
        return f"<{self.class_name} {self.name}>"


class Model:
    """Functional-model stand-in (single input, single output)."""

    def __init__(self, name: str, layers: List[Layer]):
        self.name = name
        self.layers = layers
        self._by_name: Dict[str, Layer] = {l.name: l for l in layers}
        self.input = layers[0].output
        self.output = layers[-1].output
        self.inputs = [self.input]
        self.outputs = [self.output]
        self._version = 0  # bumped by set_weights so executors know to re-upload
        self._predictor = None

    def get_layer(self, name: str) -> Layer:
        try:
            return self._by_name[name]
        except KeyError:
            raise ValueError(f"No such layer: {name}. Existing layers are: {[l.name for l in self.layers][:8]}...") from None

    def get_weights(self) -> List[np.ndarray]:
        out: List[np.ndarray] = []
        for l in self.layers:
            out.extend(l.get_weights())
        return out

    def set_weights(self, weights: Sequence[np.ndarray]) -> None:
        weights = list(weights)
        n = sum(len(l._weights) for l in self.layers)
        if len(weights) != n:
            raise ValueError(f"expected {n} weight arrays, got {len(weights)}")
        i = 0
        for l in self.layers:
            k = len(l._weights)
            if k:
                l.set_weights(weights[i:i + k])
                i += k
        self._version += 1

    def count_params(self) -> int:
        return int(sum(l.count_params() for l in self.layers))

    def to_json(self) -> str:
        cfg = {
            "class_name": "Functional",
            "config": {
                "name": self.name,
                "layers": [
                    {"class_name": l.class_name, "name": l.name, "config": l.config,
                     "inbound_nodes": [[[n, 0, 0, {}] for n in l.inbound_names]] if l.inbound_names else []}
                    for l in self.layers
                ],
                "input_layers": [[self.layers[0].name, 0, 0]],
                "output_layers": [[self.layers[-1].name, 0, 0]],
            },
            "backend": "adapt_b200",
        }
        return json.dumps(cfg)

    def summary(self, print_fn=print) -> None:
        print_fn(f'Model: "{self.name}"')
        for l in self.layers:
            print_fn(f"{l.name:28s} {l.class_name:24s} {str(l.output_shape):22s} {l.count_params():>10d}")
        print_fn(f"Total params: {self.count_params():,}")

    def predict(self, x, batch_size: int = 32, verbose=0, mode: str = "bf16", device: int = 0):
        """Un-partitioned single-device inference (reference test/local_infer.py:22) on one B200
        through the same stage kernels; there is no CPU implementation."""
        from .executor import LocalPredictor

        if self._predictor is None or self._predictor.key != (mode, device, self._version):
            if self._predictor is not None:
                self._predictor.close()
            self._predictor = LocalPredictor(self, mode=mode, device=device, max_batch=max(1, batch_size))
        return self._predictor.predict(np.asarray(x, dtype=np.float32), batch_size)


# ----------------------------------------------------------------------------------------------
def _glorot_uniform(rng: np.random.Generator, shape: Tuple[int, ...]) -> np.ndarray:
    if len(shape) == 4:
        rf = shape[0] * shape[1]
        fan_in, fan_out = rf * shape[2], rf * shape[3]
    else:
        fan_in, fan_out = shape
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, shape).astype(np.float32)


def _build(name: str, blocks_per_stack: Sequence[int], input_shape=(224, 224, 3), classes: int = 1000,
           seed: int = 1234) -> Model:
    rng = np.random.default_rng(seed)
    layers: List[Layer] = []

    def add(name_, cls, cfg, inbound, shape) -> Layer:
        l = Layer(name_, cls, cfg, inbound, shape)
        layers.append(l)
        return l

    def conv(name_, x: Layer, filters, k, stride=1, padding="valid") -> Layer:
        h, w, c = x.output_shape[1:]
        if padding == "same":
            ho, wo = -(-h // stride), -(-w // stride)
        else:
            ho, wo = (h - k) // stride + 1, (w - k) // stride + 1
        l = add(name_, "Conv2D", {"filters": filters, "kernel_size": [k, k], "strides": [stride, stride],
                                  "padding": padding, "use_bias": True}, [x], (ho, wo, filters))
        l._weights = [_glorot_uniform(rng, (k, k, c, filters)), np.zeros(filters, np.float32)]
        return l

    def bn(name_, x: Layer) -> Layer:
        c = x.output_shape[-1]
        l = add(name_, "BatchNormalization", {"axis": [3], "epsilon": BN_EPSILON}, [x], x.output_shape[1:])
        l._weights = [np.ones(c, np.float32), np.zeros(c, np.float32), np.zeros(c, np.float32), np.ones(c, np.float32)]
        return l

    def act(name_, x: Layer) -> Layer:
        return add(name_, "Activation", {"activation": "relu"}, [x], x.output_shape[1:])

    x = add("input_1", "InputLayer", {"batch_input_shape": [None, *input_shape]}, [], tuple(input_shape))
    h, w, c = input_shape
    x = add("conv1_pad", "ZeroPadding2D", {"padding": [[3, 3], [3, 3]]}, [x], (h + 6, w + 6, c))
    x = conv("conv1_conv", x, 64, 7, 2)
    x = bn("conv1_bn", x)
    x = act("conv1_relu", x)
    h, w, c = x.output_shape[1:]
    x = add("pool1_pad", "ZeroPadding2D", {"padding": [[1, 1], [1, 1]]}, [x], (h + 2, w + 2, c))
    h, w, c = x.output_shape[1:]
    x = add("pool1_pool", "MaxPooling2D", {"pool_size": [3, 3], "strides": [2, 2], "padding": "valid"}, [x],
            ((h - 3) // 2 + 1, (w - 3) // 2 + 1, c))
    for si, (filters, nblocks) in enumerate(zip((64, 128, 256, 512), blocks_per_stack)):
        stack = f"conv{si + 2}"
        for b in range(1, nblocks + 1):
            n = f"{stack}_block{b}"
            s = (1 if si == 0 else 2) if b == 1 else 1
            inp = x
            y = conv(f"{n}_1_conv", inp, filters, 1, s)
            y = bn(f"{n}_1_bn", y)
            y = act(f"{n}_1_relu", y)
            y = conv(f"{n}_2_conv", y, filters, 3, 1, "same")
            y = bn(f"{n}_2_bn", y)
            y = act(f"{n}_2_relu", y)
            if b == 1:
                sc = conv(f"{n}_0_conv", inp, 4 * filters, 1, s)
            y = conv(f"{n}_3_conv", y, 4 * filters, 1)
            if b == 1:
                sc = bn(f"{n}_0_bn", sc)
            else:
                sc = inp
            y = bn(f"{n}_3_bn", y)
            y = add(f"{n}_add", "Add", {}, [sc, y], y.output_shape[1:])
            x = act(f"{n}_out", y)
    c = x.output_shape[-1]
    x = add("avg_pool", "GlobalAveragePooling2D", {}, [x], (c,))
    d = add("predictions", "Dense", {"units": classes, "activation": "softmax", "use_bias": True}, [x], (classes,))
    d._weights = [_glorot_uniform(rng, (c, classes)), np.zeros(classes, np.float32)]
    return Model(name, layers)


def _reject_pretrained(weights) -> None:
    if weights is not None:
        raise ValueError(
            "only weights=None (Keras random initialisation) is available: there is no network on the "
            "B200 box to download ImageNet weights.  Load your own with model.set_weights(list_of_arrays)."
        )


def ResNet50(weights=None, include_top: bool = True, input_shape=(224, 224, 3), classes: int = 1000,
             seed: int = 1234) -> Model:
    """`tf.keras.applications.ResNet50(weights=None, include_top=True)` stand-in (177 layers,
    25,636,712 parameters)."""
    _reject_pretrained(weights)
    if not include_top:
        raise ValueError("include_top=False is not supported (the pipeline's result is the softmax vector)")
    return _build("resnet50", (3, 4, 6, 3), input_shape, classes, seed)


def ResNet101(weights=None, include_top: bool = True, input_shape=(224, 224, 3), classes: int = 1000,
              seed: int = 1234) -> Model:
    _reject_pretrained(weights)
    return _build("resnet101", (3, 4, 23, 3), input_shape, classes, seed)


def ResNet152(weights=None, include_top: bool = True, input_shape=(224, 224, 3), classes: int = 1000,
              seed: int = 1234) -> Model:
    _reject_pretrained(weights)
    return _build("resnet152", (3, 8, 36, 3), input_shape, classes, seed)


def preprocess_input(x: np.ndarray) -> np.ndarray:
    """Keras `resnet50.preprocess_input` (caffe mode: RGB->BGR, subtract the ImageNet channel means,
    no scaling), as used by reference test/test.py:23."""
    x = np.asarray(x, dtype=np.float32)[..., ::-1]
    return np.ascontiguousarray(x - np.array([103.939, 116.779, 123.68], dtype=np.float32))


def randomize_batchnorm(model: Model, seed: int = 7, biases: bool = True) -> None:
    """Give every BatchNormalization non-trivial statistics (and convolutions non-zero biases) so
    that BN folding is exercised; `weights=None` alone makes every BN a scalar."""
    rng = np.random.default_rng(seed)
    for l in model.layers:
        if l.class_name == "BatchNormalization":
            c = l._weights[0].shape[0]
            l._weights = [rng.uniform(0.5, 1.5, c).astype(np.float32), rng.normal(0, 0.2, c).astype(np.float32),
                          rng.normal(0, 0.2, c).astype(np.float32), rng.uniform(0.5, 1.5, c).astype(np.float32)]
        elif biases and l.class_name in ("Conv2D", "Dense"):
            l._weights[1] = rng.normal(0, 0.1, l._weights[1].shape[0]).astype(np.float32)
    model._version += 1
