"""`from src.dispatcher import DEFER` (reference test/test.py:1) -> the B200 dispatcher."""
import adapt_b200  # noqa: F401  (registers the real package)
from adapt_b200.dispatcher import CONFIG_PORT, DATA_PORT, DEFER, RESULT_PORT, balanced_part_at  # noqa: F401
