"""GPU: a B200 stage behind the reference's wire protocol (`Node(listen=True)`).  The test plays the
unmodified remote dispatcher: configuration handshake on the config port (model JSON, partition index, weights, ACK
0x06), data messages (4-byte index + framed payload), results collected on the result port -- and checks every
result against the un-partitioned GPU stage on THAT input (bit for bit: the cut is a block output) and the oracle."""
import socket
import threading

import numpy as np
import pytest

from _util import KO, StreamChecker, distinct_images, oracle_for
from test_wire import _free_port, drive_wire_node

pytestmark = pytest.mark.gpu


def test_b200_stage_behind_the_reference_wire_protocol(pkg):
    import adapt_b200
    from adapt_b200 import wire
    from adapt_b200.executor import LocalPredictor
    from adapt_b200.node import Node

    model = adapt_b200.ResNet50(weights=None, seed=3)       # its own instance: the handshake re-installs weights
    codec = wire.RawCodec()
    rp = _free_port()
    node = Node(model, device=0, listen=True, data_port=_free_port(), config_port=_free_port(), dispatcher_addr=("127.0.0.1", rp),
                codec=codec, bind_host="127.0.0.1", max_batch=4)
    t = threading.Thread(target=node.run, daemon=True)
    t.start()
    assert node.ready.wait(10)
    xs = [distinct_images(b, 900, i) for i, b in enumerate([4, 2, 3])]
    first = LocalPredictor(model, mode="bf16", max_batch=4, end="conv3_block4_out")
    cuts = [first.run(x) for x in xs]                       # what a first stage elsewhere would send (fp32 on the wire)
    first.close()
    try:
        cases = [(2, "conv3_block4_out", "predictions", c) for c in cuts]
        cases.append((1, "input_1", "conv3_block4_out", xs[0]))        # re-configured with another partition
        outs = drive_wire_node(node, model, cases, codec, rp)
    finally:
        node.stop()
        t.join(20)
    assert node.errors == [] and node.served == 4
    chk = StreamChecker(model, ["conv3_block4_out"], max_batch=4)
    try:
        chk.check(xs, outs[:3])                             # bit-identical to the whole model on one stage + oracle
    finally:
        chk.close()
    np.testing.assert_array_equal(outs[3], cuts[0])         # partition 1 through the wire == the stage run directly
    o = oracle_for(model)
    ref = o.run_range(xs[0], "input_1", "conv3_block4_out")[0]
    assert np.abs(outs[3] - ref).max() <= 2e-2 * np.abs(ref).max()
