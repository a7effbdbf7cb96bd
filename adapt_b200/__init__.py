"""Short import alias for the package directory
`adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200/` (whose name is
not a Python identifier).  `import adapt_b200` gives that package; `adapt_b200.dispatcher` etc. work."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_real = importlib.import_module("adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200")
sys.modules[__name__] = _real
