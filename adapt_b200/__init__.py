"""Short import alias for the package directory
`adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200/` (whose name is
not a Python identifier).  `import adapt_b200` gives that package; `adapt_b200.dispatcher`,
`from adapt_b200._native import X` etc. resolve to the SAME module objects as under the long name (a
second copy of a submodule would have its own classes, e.g. its own `AdaptError`)."""
import importlib
import os
import pkgutil
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_REAL_NAME = "adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200"
_real = importlib.import_module(_REAL_NAME)
for _m in pkgutil.iter_modules(_real.__path__):
    if _m.ispkg:
        continue
    sys.modules[f"{__name__}.{_m.name}"] = importlib.import_module(f"{_REAL_NAME}.{_m.name}")
sys.modules[__name__] = _real
