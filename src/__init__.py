"""Drop-in import surface of the reference (`from src.dispatcher import DEFER`, reference
test/test.py:1): thin re-exports of the B200 package."""
