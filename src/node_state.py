"""reference src/node_state.py surface (NodeState, StateEnum).  The TCP helpers socket_send /
socket_recv have no counterpart on a single NVSwitch box."""
import adapt_b200  # noqa: F401
from adapt_b200.fabric import StateEnum  # noqa: F401
from adapt_b200.node import NodeState  # noqa: F401
