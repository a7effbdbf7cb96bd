"""Run on a machine WITH TensorFlow 2.15 (none exists in the build image): Keras' own ResNet50 with the seed-1234
weights against tests/golden/resnet50_seed1234_tf_crosscheck.npz (written by tools/export_for_tf.py from the CPU
oracle).  Same shape as the reference's single-device script (reference test/local_infer.py:8-26): build
`ResNet50(weights=None)`, `set_weights`, `predict`.  Passing pins the oracle -- and with it every parity test of
this repo -- to TensorFlow; exit code 0 = pinned."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import tensorflow as tf  # noqa: E402

import adapt_b200  # noqa: E402  (pure numpy here: the weight generator only)
from export_for_tf import OUT, make_inputs, weights_digest  # noqa: E402

g = np.load(OUT)
ws = adapt_b200.ResNet50(weights=None, seed=int(g["seed"])).get_weights()
assert weights_digest(ws) == str(g["weights_sha256"]), "weight generator drifted from the exported fixture"
x = make_inputs()
assert abs(float(np.float64(x).sum()) - float(g["inputs_sum"])) < 1e-3, "input generator drifted from the exported fixture"
model = tf.keras.applications.ResNet50(weights=None, include_top=True)
assert [tuple(w.shape) for w in model.get_weights()] == [tuple(w.shape) for w in ws], "get_weights() order / layouts differ"
model.set_weights(ws)
names = [str(n) for n in g["tap_names"]]
logits_model = tf.keras.Model(model.input, [model.get_layer(n).output for n in names] + [model.layers[-2].output])
outs = logits_model.predict(x, verbose=0)
probs = model.predict(x, verbose=0)
dense_k, dense_b = model.layers[-1].get_weights()
logits = outs[-1] @ dense_k + dense_b
worst = {"probs": float(np.abs(probs - g["probs"]).max()),
         "logits_rel": float(np.abs(logits - g["logits"]).max() / np.abs(g["logits"]).max())}
for n, t in zip(names, outs[:-1]):
    s = t.reshape(t.shape[0], -1)[:, ::int(g["tap_stride"])]
    worst[n] = float(np.abs(s - g["tap_" + n]).max() / max(np.abs(g["tap_" + n]).max(), 1e-30))
print(worst)
ok = worst["probs"] <= 1e-5 and all(v <= 1e-4 for k, v in worst.items() if k != "probs") and \
    probs.argmax(-1).tolist() == g["probs"].argmax(-1).tolist()
print("ORACLE PINNED AGAINST TENSORFLOW" if ok else "MISMATCH: the oracle does not restate Keras ResNet50")
sys.exit(0 if ok else 1)
