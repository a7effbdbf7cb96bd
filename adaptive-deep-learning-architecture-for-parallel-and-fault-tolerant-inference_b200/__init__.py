"""adapt-b200: B200-native pipeline-partitioned ResNet-50 inference behind the ADAPT/DEFER API.

The directory name is not a Python identifier; import it with
`importlib.import_module("adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200")`
or through the short alias package `adapt_b200` / the reference-compatible `src` package at the repo root.
"""
from . import _native  # noqa: F401  (ctypes binding; loading the .so is deferred to first use)
from .resnet50 import ResNet50, ResNet101, ResNet152, preprocess_input, randomize_batchnorm  # noqa: F401
from .dag_util import construct_model, get_previous, traverse_improved, valid_cut_points  # noqa: F401

__all__ = ["ResNet50", "ResNet101", "ResNet152", "preprocess_input", "randomize_batchnorm", "construct_model",
           "get_previous", "traverse_improved", "valid_cut_points"]
