"""Hand-over artefact for pinning the oracle against real TensorFlow: writes
tests/golden/resnet50_seed1234_tf_crosscheck.npz -- inputs, and what the CPU oracle computes for them with the
seed-1234 Keras-initialised weights: probabilities, pre-softmax logits and a strided sample of every cut tensor the
BASELINE partitions hand off.  Run from the repo root, CPU only:

    python tools/export_for_tf.py            # (re)writes the .npz   (~2 minutes of numpy)
    python tools/tf_check.py                 # on a machine WITH TensorFlow 2.15: compares Keras against it

The 25.6 M weights themselves are not stored (102 MB): `tools/tf_check.py` regenerates them with the same pure-numpy
generator (`adapt_b200.ResNet50(weights=None, seed=1234).get_weights()`, Keras `get_weights()` order and layouts:
Conv2D [kernel HWIO, bias], BatchNormalization [gamma, beta, moving_mean, moving_variance], Dense [kernel, bias]) and
checks a digest of them stored here.  Until someone has run tf_check.py, the oracle's parity with TensorFlow
stays UNPINNED."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import adapt_b200  # noqa: E402
from oracle import keras_resnet50_oracle as KO  # noqa: E402

SEED = 1234
TAPS = ["pool1_pool", "conv2_block3_out", "conv3_block4_out", "conv4_block6_out", "conv5_block3_out", "avg_pool"]
STRIDE = 97   # every 97th element of a flattened cut tensor is kept (prime: walks all channels and positions)
OUT = os.path.join(ROOT, "tests", "golden", "resnet50_seed1234_tf_crosscheck.npz")


def weights_digest(weights) -> str:
    h = hashlib.sha256()
    for w in weights:
        h.update(np.ascontiguousarray(w, np.float32).tobytes())
    return h.hexdigest()


def make_inputs() -> np.ndarray:
    """The two inputs of the fixture, regenerated from their seeds (1.2 MB of floats are not worth committing): image 0
    = uniform[0,255) pixels through caffe preprocessing, image 1 = N(0,1) -- numpy `default_rng` streams."""
    return np.concatenate([KO.synthetic_images(1, seed=0, kind="caffe"), KO.synthetic_images(1, seed=1, kind="normal")], 0).astype(np.float32)


def main() -> None:
    model = adapt_b200.ResNet50(weights=None, seed=SEED)
    ws = model.get_weights()
    oracle = KO.ResNet50Oracle(weights={l.name: l.get_weights() for l in model.layers if l.get_weights()})
    # two inputs of each kind the tests use: caffe-preprocessed uniform images (logits O(10)) and N(0,1) images
    x = make_inputs()
    probs, taps = oracle.predict(x, taps=TAPS)
    out = {"seed": np.int64(SEED), "weights_sha256": np.array(weights_digest(ws)), "n_weight_arrays": np.int64(len(ws)),
           "inputs_sum": np.float64(np.float64(x).sum()), "probs": probs.astype(np.float32), "logits": oracle.logits.astype(np.float32),
           "tap_stride": np.int64(STRIDE), "tap_names": np.array(TAPS)}
    for name in TAPS:
        t = taps[name]
        out["tap_" + name] = t.reshape(t.shape[0], -1)[:, ::STRIDE].astype(np.float32)
        out["tapshape_" + name] = np.array(t.shape[1:], np.int64)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes; weights sha256", out["weights_sha256"])


if __name__ == "__main__":
    main()
