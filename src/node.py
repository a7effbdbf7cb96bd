"""`python -m src.node` (reference README.md:44): worker role for one GPU."""
import adapt_b200  # noqa: F401
from adapt_b200.node import Node, NodeState, StageWorker  # noqa: F401

if __name__ == "__main__":
    Node().run()
