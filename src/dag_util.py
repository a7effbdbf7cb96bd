"""reference src/dag_util.py surface."""
import adapt_b200  # noqa: F401
from adapt_b200.dag_util import construct_model, get_previous, traverse_improved  # noqa: F401
