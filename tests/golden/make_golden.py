"""Regenerate tests/golden/resnet50_seed1234_golden.json (run from the repo root, CPU only).

The reference cannot run here (SyntaxErrors + no TensorFlow), so these vectors come
from the oracle restatement evaluated on the product's Keras-initialised weights (seed 1234):
they pin the oracle + weight generator against drift, they do not pin parity with TensorFlow.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import adapt_b200  # noqa: E402
from _util import KO, oracle_for  # noqa: E402

model = adapt_b200.ResNet50(weights=None)
o = oracle_for(model)
batch, seed, kind = 2, 11, "normal"
x = KO.synthetic_images(batch, seed=seed, kind=kind)
taps = ["conv1_relu", "pool1_pool", "conv2_block3_out", "conv3_block4_out", "conv4_block6_out", "conv5_block3_out",
        "avg_pool"]
probs, t = o.predict(x, taps=taps)
gold = {
    "batch": batch, "input_seed": seed, "input_kind": kind, "input_sum": float(np.float64(x).sum()),
    "argmax": probs.argmax(-1).tolist(),
    "top5_probs": np.sort(probs, -1)[:, -5:].tolist(),
    "tap_mean_abs": {k: float(np.abs(v).mean()) for k, v in t.items()},
    "logits_head": o.logits[:, :8].tolist(),
}
out = os.path.join(ROOT, "tests", "golden", "resnet50_seed1234_golden.json")
json.dump(gold, open(out, "w"), indent=1)
print("wrote", out)
