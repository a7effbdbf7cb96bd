"""CPU: the reference's wire protocol -- `adapt_b200.wire` against byte strings built by hand from the
reference's framing rules, against the reference's OWN `socket_send` / `socket_recv` where /root/reference is
present, and `Node(listen=True)` driven the way an unmodified remote dispatcher would drive it (the stage compute
is the oracle here; tests/test_wire_gpu.py does the same with a B200 stage)."""
import ast
import json
import os
import select
import socket
import threading
import time

import numpy as np
import pytest

from _util import KO, oracle_for

REF_NODE_STATE = "/root/reference/src/node_state.py"


def _pair():
    a, b = socket.socketpair()
    return a, b


def _read_exact(sock, n):
    buf = b""
    while len(buf) < n:
        r = sock.recv(n - len(buf))
        if not r:
            break
        buf += r
    return buf


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


# ------------------------------------------------------------------------------------------------ framing
def test_socket_send_emits_the_reference_framing(pkg):
    from adapt_b200 import wire

    a, b = _pair()
    payload = bytes(range(256)) * 9                       # 2304 bytes
    wire.socket_send(payload, a, chunk_size=1000)         # 3 pieces: 1000 + 1000 + 304
    a.close()
    raw = _read_exact(b, 10 ** 6)
    assert raw == (2304).to_bytes(8, "big") + payload     # src/node_state.py:50-52: 8-byte big-endian size, then the body
    b.close()
    a, b = _pair()
    wire.socket_send(b"", a, 512000)                      # empty payload: header only
    a.close()
    assert _read_exact(b, 100) == bytes(8)
    b.close()


def test_socket_recv_parses_hand_built_bytes(pkg):
    from adapt_b200 import wire

    a, b = _pair()
    body = os.urandom(70000)
    threading.Thread(target=lambda: (a.sendall((70000).to_bytes(8, "big")), [a.sendall(body[i:i + 777]) for i in range(0, 70000, 777)]),
                     daemon=True).start()
    assert wire.socket_recv(b, chunk_size=512) == body    # read in pieces of at most chunk_size
    a.sendall((3).to_bytes(8, "big") + b"ab")             # truncated body
    a.close()
    with pytest.raises(socket.error, match="Incomplete data"):
        wire.socket_recv(b, 512000)
    b.close()
    a, b = _pair()
    a.close()
    assert wire.socket_recv(b, 512000) == b""             # clean close before any header byte (src/node_state.py:118-121)
    b.close()
    a, b = _pair()
    a.sendall(b"\x00\x00\x00")                            # closed inside the header
    a.close()
    with pytest.raises(socket.error, match="Incomplete header"):
        wire.socket_recv(b, 512000)
    b.close()


def test_framing_on_non_blocking_sockets(pkg):
    """The reference runs these helpers on non-blocking sockets (EAGAIN -> select, src/node_state.py:60-66)."""
    from adapt_b200 import wire

    a, b = _pair()
    a.setblocking(False)
    b.setblocking(False)
    big = os.urandom(3_000_000)                           # far more than the socket buffers hold
    out = {}
    t = threading.Thread(target=lambda: out.setdefault("got", wire.socket_recv(b, 512000)), daemon=True)
    t.start()
    wire.socket_send(big, a, 512000)
    t.join(30)
    assert out.get("got") == big
    a.close()
    b.close()


def _reference_functions():
    """`socket_send` / `socket_recv` exactly as the reference ships them, lifted out of src/node_state.py by name (the
    module itself cannot be imported: etcd3 is absent)."""
    src = open(REF_NODE_STATE).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("socket_send", "socket_recv")]
    ns = {"socket": socket, "select": select}
    exec(compile(ast.Module(body=keep, type_ignores=[]), REF_NODE_STATE, "exec"), ns)
    return ns["socket_send"], ns["socket_recv"]


@pytest.mark.skipif(not os.path.exists(REF_NODE_STATE), reason="the reference tree is only present in the build container")
def test_interoperates_with_the_reference_own_framing(pkg):
    """Our sender -> the reference's receiver and the reference's sender -> our receiver, blocking and non-blocking,
    chunked bodies, chunk_size=1 (how the partition index is framed, src/dispatcher.py:241)."""
    from adapt_b200 import wire

    ref_send, ref_recv = _reference_functions()
    for blocking in (True, False):
        for payload, chunk in ((os.urandom(1_300_000), 512000), (b"7", 1), (b"", 512000), (os.urandom(4097), 100)):
            if payload == b"":
                continue  # (the reference's receiver reports an empty body and a clean close alike; not a data case)
            a, b = _pair()
            a.setblocking(blocking)
            b.setblocking(blocking)
            got = {}
            t = threading.Thread(target=lambda: got.setdefault("v", ref_recv(b, chunk)), daemon=True)
            t.start()
            wire.socket_send(payload, a, chunk)
            t.join(30)
            assert got.get("v") == payload, (blocking, len(payload), chunk)
            t = threading.Thread(target=lambda: got.setdefault("w", wire.socket_recv(b, chunk)), daemon=True)
            t.start()
            ref_send(payload, a, chunk)
            t.join(30)
            assert got.get("w") == payload, (blocking, len(payload), chunk)
            a.close()
            b.close()


# ------------------------------------------------------------------------------------------------ messages
def test_data_message_bytes(pkg):
    """4-byte big-endian partition index, framed payload, then end of stream (src/dispatcher.py:204-220)."""
    from adapt_b200 import wire

    srv = socket.socket()
    srv.bind(("127.0.0.1", 0))
    srv.listen(1)
    port = srv.getsockname()[1]
    payload = os.urandom(12345)
    threading.Thread(target=wire.forward_data_to_worker, args=(("127.0.0.1", port), payload, 3, 4000), daemon=True).start()
    cli, _ = srv.accept()
    raw = _read_exact(cli, 10 ** 6)
    assert raw == (3).to_bytes(4, "big") + (12345).to_bytes(8, "big") + payload
    cli.close()
    # and the other direction: hand-built bytes into the worker-side parser
    a, b = _pair()
    a.sendall((2).to_bytes(4, "big") + (5).to_bytes(8, "big") + b"hello")
    a.shutdown(socket.SHUT_WR)
    assert wire.recv_data_message(b, 512000) == (2, b"hello")
    assert wire.recv_data_message(b, 512000) is None
    a.close()
    b.close()
    srv.close()


def test_config_handshake_bytes_both_directions(pkg):
    """JSON framed, index as ASCII framed, 8-byte array count, framed arrays, 1-byte ACK 0x06
    (src/dispatcher.py:223-264, 76-89; src/node.py:65-98, 101-119)."""
    from adapt_b200 import wire

    codec = wire.RawCodec()
    weights = [np.arange(6, dtype=np.float32).reshape(2, 3), np.ones(4, np.float32)]
    blobs = [codec.encode(w) for w in weights]
    model_json = json.dumps({"class_name": "Functional", "config": {"name": "part2", "layers": ["a"]}})
    expect = ((len(model_json)).to_bytes(8, "big") + model_json.encode() + (2).to_bytes(8, "big") + b"12"
              + (2).to_bytes(8, "big") + b"".join(len(x).to_bytes(8, "big") + x for x in blobs))
    # dispatcher side (ours) against a hand-written worker that reads raw bytes and ACKs
    srv = socket.socket()
    srv.bind(("127.0.0.1", 0))
    srv.listen(1)
    port = srv.getsockname()[1]
    seen = {}

    def fake_worker():
        cli, _ = srv.accept()
        seen["raw"] = _read_exact(cli, len(expect))
        cli.sendall(b"\x06")
        time.sleep(0.05)
        cli.close()

    t = threading.Thread(target=fake_worker, daemon=True)
    t.start()
    ack = wire.send_full_configuration(("127.0.0.1", port), model_json, 12, weights, codec=codec)
    t.join(10)
    assert ack == wire.ACK == b"\x06" and seen["raw"] == expect
    srv.close()
    # worker side (ours) fed with the hand-built byte string
    a, b = _pair()
    a.sendall(expect)
    js, idx, ws = wire.recv_full_configuration(b, 512000, codec)
    assert js == model_json and idx == 12 and len(ws) == 2
    np.testing.assert_array_equal(ws[0], weights[0])
    np.testing.assert_array_equal(ws[1], weights[1])
    a.close()
    b.close()


def test_codec_and_partition_json(pkg, model50):
    from adapt_b200 import wire
    from adapt_b200.dag_util import construct_model

    c = wire.RawCodec()
    for arr in (np.random.default_rng(0).standard_normal((2, 7, 7, 5)).astype(np.float32), np.arange(10, dtype=np.uint8)):
        back = c.decode(c.encode(arr))
        assert back.dtype == arr.dtype
        np.testing.assert_array_equal(back, arr)
    with pytest.raises(ValueError):
        c.decode(b"not an array")
    assert wire.default_codec().name in ("raw", "lz4+zfp")
    order = [l.name for l in model50.layers]
    part = construct_model(model50, "conv3_block4_out", "predictions", "part2")
    assert wire.partition_bounds_from_json(part.to_json(), order) == ("conv3_block4_out", "predictions")
    # a Keras functional JSON of the same sub-model, as `tf.keras.Model.to_json()` writes it: an InputLayer named after
    # the part, then the layers by name with their inbound nodes
    keras_layers = [{"class_name": "InputLayer", "name": "part2", "config": {"name": "part2"}, "inbound_nodes": []}]
    for n in part.layer_names:
        inbound = [[[("part2" if i == "conv3_block4_out" else i), 0, 0, {}] for i in model50.get_layer(n).inbound_names]]
        keras_layers.append({"class_name": model50.get_layer(n).class_name, "name": n, "config": {"name": n}, "inbound_nodes": inbound})
    kj = json.dumps({"class_name": "Functional", "config": {"name": "model", "layers": keras_layers, "input_layers": [["part2", 0, 0]],
                                                            "output_layers": [["predictions", 0, 0]]}})
    assert wire.partition_bounds_from_json(kj, order) == ("conv3_block4_out", "predictions")
    first = construct_model(model50, "input_1", "conv2_block3_out", "part1")
    assert wire.partition_bounds_from_json(first.to_json(), order) == ("input_1", "conv2_block3_out")


# ------------------------------------------------------------------------------------------------ Node(listen=True)
class _OraclePredictor:
    def __init__(self, model, start, end):
        self.o, self.start, self.end = oracle_for(model), start, end

    def run(self, x):
        return self.o.run_range(np.asarray(x, np.float32), self.start, self.end)[0]

    def close(self):
        pass


def drive_wire_node(node, model, cases, codec, result_port, timeout=120.0):
    """Act as the remote dispatcher: for each (index, start, end, x) configure the partition (expect ACK), send one
    data message, and collect the result message on `result_port`.  Returns the decoded results."""
    from adapt_b200 import wire
    from adapt_b200.dag_util import construct_model

    rs = socket.socket()
    rs.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    rs.bind(("127.0.0.1", result_port))
    rs.listen(4)
    rs.settimeout(timeout)
    outs = []
    try:
        for index, start, end, x in cases:
            part = construct_model(model, start, end, f"part{index}")
            ack = wire.send_full_configuration(("127.0.0.1", node.config_port), part.to_json(), index, part.get_weights(), codec=codec,
                                               ack_timeout=timeout)
            assert ack == wire.ACK, (ack, node.errors)
            wire.forward_data_to_worker(("127.0.0.1", node.data_port), codec.encode(x), index)
            cli, _ = rs.accept()
            cli.settimeout(timeout)
            msg = wire.recv_data_message(cli, wire.CHUNK_SIZE)
            cli.close()
            assert msg is not None and msg[0] == index, (msg and msg[0], node.errors)
            outs.append(codec.decode(msg[1]))
    finally:
        rs.close()
    return outs


def test_node_listen_serves_partitions_over_the_reference_protocol(pkg, model50):
    from adapt_b200 import wire
    from adapt_b200.node import Node

    codec = wire.RawCodec()
    rp = _free_port()
    node = Node(model50, listen=True, data_port=_free_port(), config_port=_free_port(), dispatcher_addr=("127.0.0.1", rp),
                predictor_factory=_OraclePredictor, codec=codec, bind_host="127.0.0.1")
    t = threading.Thread(target=node.run, daemon=True)
    t.start()
    assert node.ready.wait(10)
    o = oracle_for(model50)
    x = KO.synthetic_images(2, seed=5, kind="normal")
    cut = o.run_range(x, "input_1", "conv4_block6_out")[0]
    try:
        outs = drive_wire_node(node, model50, [(2, "conv4_block6_out", "predictions", cut), (1, "input_1", "conv4_block6_out", x)],
                               codec, rp)
    finally:
        node.stop()
        t.join(10)
    assert node.errors == [] and node.served == 2
    np.testing.assert_allclose(outs[0], o.predict(x)[0], rtol=1e-5, atol=1e-7)      # partition 2 on the cut tensor
    np.testing.assert_allclose(outs[1], cut, rtol=1e-5, atol=1e-6)                   # partition 1 on the images
