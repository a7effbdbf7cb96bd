"""The reference's wire protocol, byte for byte, so that a B200 stage can sit behind an UNMODIFIED remote
dispatcher / edge client.

On one NVSwitch box nothing of this is on the hot path (activations move over NVLink, node.py); it is the
compatibility surface for deployments where the dispatcher stays where it is and only a worker moves to a B200.

What is restated (reference file:line):
  framing       8-byte big-endian length, then the payload in pieces of at most `chunk_size` bytes, on blocking
                or non-blocking sockets (EAGAIN -> select)                     src/node_state.py:39-89, 92-161
  data message  4-byte big-endian partition index, framed payload, SHUT_WR      src/dispatcher.py:204-220
  config        framed model JSON, framed ASCII partition index (sent with chunk_size=1), 8-byte big-endian
                array count, one framed array per weight, then wait for the 1-byte ACK 0x06
                                                     src/dispatcher.py:223-264, 76-89; src/node.py:65-98, 101-119
  ports         6000 data, 6001 config, 6003 results                           src/dispatcher.py:15-17, src/node.py:20-22
  payload codec `lz4.frame.compress(zfpy.compress_numpy(arr))`                  src/dispatcher.py:92-98, src/node.py:122-125

lz4 and zfpy are not installed in this image, so the codec is pluggable: `Lz4ZfpCodec` is used when both import,
`RawCodec` (a self-describing uncompressed array) otherwise -- both ends of a connection must agree, exactly as
both ends of the reference must have the same zfpy.  [inferred] marks what the reference's surviving fragments do
not spell out (the message a new-generation worker sends back to port 6003: assumed symmetric to the data message).
"""
from __future__ import annotations

import errno
import json
import select
import socket
import struct
from typing import Iterable, List, Optional, Tuple

import numpy as np

DATA_PORT = 6000     # receive input data           (src/node.py:20)
CONFIG_PORT = 6001   # receive model config         (src/node.py:21)
RESULT_PORT = 6003   # send results                 (src/node.py:22)
CHUNK_SIZE = 512 * 1000   # src/dispatcher.py:24, src/node.py:182
ACK = b"\x06"             # src/node.py:90, src/dispatcher.py:254

_WOULD_BLOCK = (errno.EAGAIN, errno.EWOULDBLOCK)


# ------------------------------------------------------------------------------------------------ framing
def _send_all(view, sock: socket.socket) -> None:
    while len(view) > 0:
        try:
            sent = sock.send(view)
            view = view[sent:]
        except (BlockingIOError, InterruptedError):
            select.select([], [sock], [])
        except socket.error as e:
            if e.errno not in _WOULD_BLOCK:
                raise
            select.select([], [sock], [])   # buffer full on a non-blocking socket: wait until writable


def socket_send(bytes_to_send, sock: socket.socket, chunk_size: int = CHUNK_SIZE) -> None:
    """8-byte big-endian size header, then the body in chunks of at most `chunk_size` bytes
    (reference src/node_state.py:39-89; same name and argument order)."""
    view = memoryview(bytes_to_send).cast("B")
    _send_all(memoryview(len(view).to_bytes(8, "big")), sock)
    step = max(1, int(chunk_size))
    for i in range(0, len(view), step):
        _send_all(view[i:i + step], sock)


def _recv_into(view, sock: socket.socket, chunk_size: int, what: str, allow_clean_close: bool = False) -> bool:
    got = 0
    n = len(view)
    while got < n:
        try:
            r = sock.recv_into(view[got:], min(n - got, max(1, int(chunk_size))))
        except (BlockingIOError, InterruptedError):
            select.select([sock], [], [])
            continue
        except socket.error as e:
            if e.errno not in _WOULD_BLOCK:
                raise
            select.select([sock], [], [])
            continue
        if r == 0:
            if got == 0 and allow_clean_close:
                return False
            raise socket.error(f"Incomplete {what} received, connection closed.")
        got += r
    return True


def socket_recv(sock: socket.socket, chunk_size: int = CHUNK_SIZE) -> bytes:
    """Counterpart of `socket_send` (reference src/node_state.py:92-161): b'' when the peer closed cleanly before
    the first header byte, socket.error when the connection ends inside the header or the body."""
    header = bytearray(8)
    if not _recv_into(memoryview(header), sock, 8, "header", allow_clean_close=True):
        return b""
    size = int.from_bytes(header, "big")
    body = bytearray(size)
    _recv_into(memoryview(body), sock, chunk_size, "data")
    return bytes(body)


# ------------------------------------------------------------------------------------------------ payload codec
class RawCodec:
    """Self-describing uncompressed array: b"ADPT", dtype code, ndim, shape (uint32 big-endian each), C-order bytes.
    Stand-in for lz4(zfp(arr)) when those libraries are absent; lossless like the reference's codec (zfpy is called
    without tolerance / rate / precision, i.e. in its reversible mode)."""

    name = "raw"
    _DTYPES = {0: np.float32, 1: np.float64, 2: np.uint8, 3: np.int32, 4: np.int64, 5: np.float16}

    def encode(self, arr) -> bytes:
        a = np.ascontiguousarray(arr)
        code = next((k for k, v in self._DTYPES.items() if a.dtype == v), None)
        if code is None:
            a, code = a.astype(np.float32), 0
        return b"ADPT" + struct.pack(">II", code, a.ndim) + struct.pack(">" + "I" * a.ndim, *a.shape) + a.tobytes()

    def decode(self, byts) -> np.ndarray:
        if bytes(byts[:4]) != b"ADPT":
            raise ValueError("payload is not a raw-codec array (both ends of a connection must use the same codec)")
        code, nd = struct.unpack(">II", bytes(byts[4:12]))
        shape = struct.unpack(">" + "I" * nd, bytes(byts[12:12 + 4 * nd]))
        return np.frombuffer(byts, dtype=self._DTYPES[code], offset=12 + 4 * nd).reshape(shape)


class Lz4ZfpCodec:
    """The reference's `_comp` / `_decomp` (src/dispatcher.py:92-98, src/node.py:122-125)."""

    name = "lz4+zfp"

    def __init__(self):
        import lz4.frame  # noqa: F401  (ImportError when absent: use RawCodec)
        import zfpy  # noqa: F401

        self._lz4, self._zfpy = lz4.frame, zfpy

    def encode(self, arr) -> bytes:
        return self._lz4.compress(self._zfpy.compress_numpy(np.ascontiguousarray(arr)))

    def decode(self, byts) -> np.ndarray:
        return self._zfpy.decompress_numpy(self._lz4.decompress(bytes(byts)))


def default_codec():
    try:
        return Lz4ZfpCodec()
    except ImportError:
        return RawCodec()


# ------------------------------------------------------------------------------------------------ config handshake
def send_weights(weights: Iterable[np.ndarray], sock: socket.socket, chunk_size: int = CHUNK_SIZE, codec=None) -> None:
    """8-byte big-endian array count, then every array framed (reference src/dispatcher.py:76-89)."""
    codec = codec or default_codec()
    weights = list(weights)
    _send_all(memoryview(len(weights).to_bytes(8, "big")), sock)
    for w in weights:
        socket_send(codec.encode(w), sock, chunk_size)


def recv_weights(sock: socket.socket, chunk_size: int = CHUNK_SIZE, codec=None) -> List[np.ndarray]:
    """Counterpart on the worker (reference src/node.py:101-119)."""
    codec = codec or default_codec()
    header = bytearray(8)
    _recv_into(memoryview(header), sock, 8, "weight count")
    return [codec.decode(socket_recv(sock, chunk_size)) for _ in range(int.from_bytes(header, "big"))]


def send_full_configuration(worker_addr: Tuple[str, int], model_json: str, worker_index: int, weights, chunk_size: int = CHUNK_SIZE,
                            codec=None, connect_timeout: float = 5.0, ack_timeout: float = 60.0) -> bytes:
    """Dispatcher side of the configuration handshake (reference src/dispatcher.py:223-264): JSON, partition index
    as ASCII digits framed with chunk_size=1, weights, then the 1-byte acknowledgement.  Returns the byte received
    (b'\\x06' from a healthy worker)."""
    s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    s.settimeout(connect_timeout)
    try:
        s.connect(worker_addr)
        s.setblocking(True)
        socket_send(model_json.encode(), s, chunk_size)
        socket_send(str(int(worker_index)).encode(), s, 1)
        send_weights(weights, s, chunk_size, codec)
        ready = select.select([s], [], [], ack_timeout)
        if not ready[0]:
            raise socket.timeout("Timed out waiting for Worker ACK")
        return s.recv(1)
    finally:
        s.close()


def recv_full_configuration(client_sock: socket.socket, chunk_size: int = CHUNK_SIZE, codec=None):
    """Worker side (reference src/node.py:65-80): returns (model_json, worker_index, weights); the caller loads the
    partition and then answers with `ACK`."""
    model_json = socket_recv(client_sock, chunk_size)
    worker_index_bytes = socket_recv(client_sock, 1)
    if not model_json or not worker_index_bytes:
        raise socket.error("Incomplete config received.")
    weights = recv_weights(client_sock, chunk_size, codec)
    return model_json.decode(), int(worker_index_bytes.decode()), weights


# ------------------------------------------------------------------------------------------------ data messages
def forward_data_to_worker(worker_addr: Tuple[str, int], data: bytes, target_index: int, chunk_size: int = CHUNK_SIZE) -> None:
    """One data message on a connection of its own: 4-byte big-endian partition index, framed payload, SHUT_WR
    (reference src/dispatcher.py:204-220)."""
    s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    try:
        s.connect(worker_addr)
        s.sendall(int(target_index).to_bytes(4, "big"))
        socket_send(data, s, chunk_size)
        s.shutdown(socket.SHUT_WR)
    finally:
        s.close()


def recv_data_message(client_sock: socket.socket, chunk_size: int = CHUNK_SIZE) -> Optional[Tuple[int, bytes]]:
    """Counterpart of `forward_data_to_worker` (reference src/node.py:127-136, docstring): (index, payload), or None
    when the peer closed without sending anything."""
    head = bytearray(4)
    if not _recv_into(memoryview(head), client_sock, 4, "partition index", allow_clean_close=True):
        return None
    return int.from_bytes(head, "big"), socket_recv(client_sock, chunk_size)


# ------------------------------------------------------------------------------------------------ model JSON
def partition_bounds_from_json(model_json: str, layer_order: List[str]) -> Tuple[str, str]:
    """(start, end] of the partition a configuration message describes.  Accepts this package's own
    `PartModel.to_json()` and a Keras functional-model JSON (`tf.keras.Model.to_json()`, what the reference ships,
    src/dispatcher.py:235): there the partition's layers are listed by name; its InputLayer stands for the output of
    the layer preceding the first of them in the full model (`construct_model`, src/dag_util.py:50-62)."""
    cfg = json.loads(model_json)["config"]
    if "input" in cfg and "output" in cfg:
        return cfg["input"], cfg["output"]
    known = set(layer_order)
    names = [l["name"] if isinstance(l, dict) else l for l in cfg["layers"]]
    body = [n for n in names if n in known]
    if not body:
        raise ValueError("configuration JSON names no layer of the served model")
    idx = {n: i for i, n in enumerate(layer_order)}
    body.sort(key=idx.__getitem__)
    outs = cfg.get("output_layers") or []
    end = outs[0][0] if outs and outs[0][0] in known else body[-1]
    first = idx[body[0]]
    if first == 0:
        return layer_order[0], end
    # the producer of the partition's input: the inbound layer of its first layer that is NOT part of the partition
    inbound = None
    for l in cfg["layers"]:
        if isinstance(l, dict) and l.get("name") == body[0] and l.get("inbound_nodes"):
            try:
                cand = l["inbound_nodes"][0][0][0]
                inbound = cand if cand in known else None
            except (IndexError, KeyError, TypeError):
                inbound = None
    return inbound or layer_order[first - 1], end
