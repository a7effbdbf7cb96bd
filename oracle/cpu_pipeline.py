"""CPU ORACLE (timed baseline) -- test infrastructure, NOT part of the product.

"Restated reference CPU pipeline (TensorFlow unavailable on this box)": what the reference's
dispatcher + localhost workers do for one stream of inputs, restated so that it can run here.
Only `bench.py` (cpu_baseline leg, `--impl reference`) and `tests/` may import this module.

Follows, per hop, the reference's old-generation linear chain (the only fully specified one,
SURVEY 3.2):
  dispatcher `_startDistEdgeInference`  input.get -> _comp -> socket_send      (src/dispatcher.py:99-107)
  worker     `_data_server`             socket_recv -> _decomp -> queue         (src/node.py:137-161)
  worker     `_data_client`             queue.get -> model.predict -> _comp -> socket_send   (src/node.py:163-179)
  dispatcher `_result_server`           socket_recv -> _decomp -> output.put    (src/dispatcher.py:109-119)
with the reference's framing restated from src/node_state.py:39-161 (8-byte big-endian length,
payload sent in chunks of at most chunk_size = 512,000 bytes).  Differences, all stated in
BASELINE.md: (1) `model.predict` is this repo's float32 restatement of Keras ResNet-50 evaluated
with torch CPU kernels (oneDNN), all host threads; (2) the zfp+lz4 codec is omitted -- it is
lossless (zfpy called without tolerance/rate/precision, src/dispatcher.py:92-93), cannot change
results, and the libraries are absent; (3) every worker listens on its own localhost port instead of
the fixed 6000/6003 pair (one worker per host in the reference).
"""
from __future__ import annotations

import multiprocessing as mp
import os
import socket
import struct
import time
from typing import Dict, List, Sequence

import numpy as np

CHUNK = 512 * 1000  # reference src/dispatcher.py:24


# ------------------------------------------------------------------------------------ framing
def socket_send(data: bytes, sock: socket.socket, chunk_size: int = CHUNK) -> None:
    """8-byte big-endian length, then the payload in <= chunk_size pieces (src/node_state.py:39-89)."""
    sock.sendall(struct.pack(">Q", len(data)))
    view = memoryview(data)
    for off in range(0, len(data), chunk_size):
        sock.sendall(view[off:off + chunk_size])


def _recv_exact(sock: socket.socket, n: int) -> bytearray:
    buf = bytearray(n)
    view = memoryview(buf)
    got = 0
    while got < n:
        r = sock.recv_into(view[got:], min(n - got, CHUNK))
        if r == 0:
            raise ConnectionError("peer closed")
        got += r
    return buf


def socket_recv(sock: socket.socket, chunk_size: int = CHUNK) -> bytearray:
    (n,) = struct.unpack(">Q", bytes(_recv_exact(sock, 8)))
    return _recv_exact(sock, n)


def _pack(arr: np.ndarray) -> bytes:
    arr = np.ascontiguousarray(arr, dtype=np.float32)
    hdr = struct.pack(">I", arr.ndim) + b"".join(struct.pack(">I", d) for d in arr.shape)
    return hdr + arr.tobytes()


def _unpack(buf) -> np.ndarray:
    (nd,) = struct.unpack(">I", bytes(buf[:4]))
    shape = struct.unpack(">" + "I" * nd, bytes(buf[4:4 + 4 * nd]))
    return np.frombuffer(buf, dtype=np.float32, offset=4 + 4 * nd).reshape(shape)


# ------------------------------------------------------------------------------------ stage compute
class TorchCpuStage:
    """Layers (start, end] of the oracle's ResNet-50, evaluated un-fused with torch CPU ops."""

    def __init__(self, weights: Dict[str, List[np.ndarray]], start: str, end: str, threads: int):
        import torch

        from . import keras_resnet50_oracle as KO

        torch.set_num_threads(max(1, threads))
        self.torch = torch
        self.KO = KO
        o = KO.ResNet50Oracle(weights=weights)
        self.by = o.by_name
        self.start = start
        self.names = o.layers_between(start, end)
        self.end = end
        self.w = {}
        for n in self.names:
            l = self.by[n]
            if l.kind == "Conv2D":
                k, b = weights[n]
                self.w[n] = (torch.from_numpy(np.ascontiguousarray(k.transpose(3, 2, 0, 1))).contiguous(
                    memory_format=torch.channels_last), torch.from_numpy(b))
            elif l.kind == "BatchNormalization":
                self.w[n] = tuple(torch.from_numpy(a) for a in weights[n])
            elif l.kind == "Dense":
                k, b = weights[n]
                self.w[n] = (torch.from_numpy(np.ascontiguousarray(k.T)), torch.from_numpy(b))

    def __call__(self, x: np.ndarray) -> np.ndarray:
        torch, F, KO = self.torch, self.torch.nn.functional, self.KO
        t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        if t.ndim == 4:
            t = t.permute(0, 3, 1, 2)  # NHWC -> logical NCHW (channels_last memory)
        vals = {self.start: t}
        with torch.no_grad():
            for n in self.names:
                l = self.by[n]
                a = vals[l.inbound[0]]
                if l.kind == "ZeroPadding2D":
                    y = F.pad(a, (l.pad, l.pad, l.pad, l.pad))
                elif l.kind == "Conv2D":
                    k, b = self.w[n]
                    pad = (l.kernel - 1) // 2 if l.padding == "same" else 0
                    y = F.conv2d(a, k, b, stride=l.stride, padding=pad)
                elif l.kind == "BatchNormalization":
                    g, be, mu, var = self.w[n]
                    y = F.batch_norm(a, mu, var, g, be, training=False, eps=KO.BN_EPS)
                elif l.kind == "Activation":
                    y = F.relu(a)
                elif l.kind == "MaxPooling2D":
                    y = F.max_pool2d(a, l.kernel, l.stride)
                elif l.kind == "Add":
                    y = a + vals[l.inbound[1]]
                elif l.kind == "GlobalAveragePooling2D":
                    y = a.mean(dim=(2, 3))
                elif l.kind == "Dense":
                    k, b = self.w[n]
                    y = F.softmax(F.linear(a, k, b), dim=-1)
                else:
                    raise ValueError(l.kind)
                vals[n] = y
        out = vals[self.end]
        if out.ndim == 4:
            out = out.permute(0, 2, 3, 1)
        return np.ascontiguousarray(out.numpy())


# ------------------------------------------------------------------------------------ processes
def _worker_main(idx: int, port: int, next_port: int, weights, start: str, end: str, threads: int, ready) -> None:
    stage = TorchCpuStage(weights, start, end, threads)
    srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("127.0.0.1", port))
    srv.listen(1)
    ready.set()
    cli, _ = srv.accept()
    nxt = None
    try:
        while True:
            try:
                data = socket_recv(cli)
            except ConnectionError:
                break
            if len(data) == 0:
                break
            if nxt is None:  # connect lazily, like the reference's _data_client (src/node.py:163-172)
                nxt = socket.create_connection(("127.0.0.1", next_port))
            y = stage(_unpack(data))
            socket_send(_pack(y), nxt)
    finally:
        if nxt is not None:
            try:
                socket_send(b"", nxt)
            except OSError:
                pass
            nxt.close()
        cli.close()
        srv.close()


class CpuPipeline:
    """Dispatcher side: spawns one localhost worker process per partition, streams inputs in,
    collects results (FIFO)."""

    def __init__(self, weights: Dict[str, List[np.ndarray]], part_at: Sequence[str], base_port: int = 0):
        n = len(part_at) + 1
        self.n = n
        bounds = ["input_1"] + list(part_at) + ["predictions"]
        cores = os.cpu_count() or 1
        self.threads_per_stage = max(1, cores // n)
        self.cores = cores
        ports = self._free_ports(n + 1)
        self.result_port = ports[n]
        ctx = mp.get_context("spawn")
        self.procs = []
        readies = []
        self.result_srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.result_srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self.result_srv.bind(("127.0.0.1", self.result_port))
        self.result_srv.listen(1)
        for i in range(n):
            ev = ctx.Event()
            p = ctx.Process(target=_worker_main, args=(i, ports[i], ports[i + 1], weights, bounds[i], bounds[i + 1],
                                                       self.threads_per_stage, ev), daemon=True)
            p.start()
            self.procs.append(p)
            readies.append(ev)
        for ev in readies:
            if not ev.wait(120):
                raise RuntimeError("CPU pipeline worker did not start")
        self.data_sock = socket.create_connection(("127.0.0.1", ports[0]))
        self.result_cli = None

    @staticmethod
    def _free_ports(n: int) -> List[int]:
        socks = [socket.socket() for _ in range(n)]
        for s in socks:
            s.bind(("127.0.0.1", 0))
        ports = [s.getsockname()[1] for s in socks]
        for s in socks:
            s.close()
        return ports

    def send(self, x: np.ndarray) -> None:
        socket_send(_pack(x), self.data_sock)

    def recv(self) -> np.ndarray:
        if self.result_cli is None:
            self.result_cli, _ = self.result_srv.accept()
        return _unpack(socket_recv(self.result_cli)).copy()

    def run(self, batches: Sequence[np.ndarray]) -> List[np.ndarray]:
        """Feed every batch, then drain (the sockets buffer the in-flight ones, as in the reference)."""
        import threading

        outs: List[np.ndarray] = []
        t = threading.Thread(target=lambda: [self.send(b) for b in batches], daemon=True)
        t.start()
        for _ in batches:
            outs.append(self.recv())
        t.join()
        return outs

    def close(self) -> None:
        try:
            socket_send(b"", self.data_sock)
        except OSError:
            pass
        self.data_sock.close()
        for p in self.procs:
            p.join(timeout=10)
            if p.is_alive():
                p.kill()
        if self.result_cli is not None:
            self.result_cli.close()
        self.result_srv.close()


def time_cpu_pipeline(weights, part_at: Sequence[str], batch: int, steps: int, warmup: int, seed: int = 0) -> dict:
    """images/s of the restated CPU pipeline on `steps` micro-batches of `batch` synthetic images."""
    from . import keras_resnet50_oracle as KO

    x = KO.synthetic_images(batch, seed=seed, kind="caffe")
    pipe = CpuPipeline(weights, part_at)
    try:
        pipe.run([x] * max(1, warmup))
        t0 = time.perf_counter()
        outs = pipe.run([x] * steps)
        dt = time.perf_counter() - t0
    finally:
        pipe.close()
    return {"value": steps * batch / dt, "seconds": dt, "cores": pipe.cores, "stages": pipe.n,
            "threads_per_stage": pipe.threads_per_stage, "batch": batch, "steps": steps,
            "checksum": float(np.float64(outs[-1]).sum())}
