"""Shared helpers for the tests: the oracle is only ever used here as the checker."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import keras_resnet50_oracle as KO  # noqa: E402


def oracle_for(model):
    """Oracle evaluating the SAME weight tensors as `model` (the product's Keras stand-in)."""
    weights = {l.name: l.get_weights() for l in model.layers if l.get_weights()}
    return KO.ResNet50Oracle(weights=weights)


def rel_err(got, ref):
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))
