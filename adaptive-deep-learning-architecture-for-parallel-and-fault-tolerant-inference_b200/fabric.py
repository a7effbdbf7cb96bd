"""Host-side plumbing between the dispatcher and the stage workers on one NVSwitch box.

Replaces, for a single host, the reference's TCP framing (`socket_send`/`socket_recv`,
reference src/node_state.py:39-161), its per-hop routing through the dispatcher
(src/dispatcher.py:176-220) and the etcd membership store (src/node_state.py:9-23,
src/start_etcd.sh).  Nothing here touches tensor data: activations move GPU-to-GPU over NVLink
(the producing kernel stores into the consumer's buffer), and the only things that travel
through host shared memory are 32-byte micro-batch headers, heartbeats, the partition table and
the final (B, 1000) result rows.

  ShmRegion     named shared-memory segment (works across threads and processes)
  ControlRing   single-producer single-consumer ring of fixed-size int64 records
  HealthTable   in-dispatcher health/heartbeat table + partition table ("etcd" replacement)
  ResultRing    slots the last stage DMA-writes probabilities into and the dispatcher drains
  CudaBackend   device operations of a worker through the C ABI (include/adapt_b200.h)
"""
from __future__ import annotations

import ctypes as C
import json
import os
import time
import uuid
from multiprocessing import shared_memory
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _native as N

# Flag values are `flag_base + k + 1` ("micro-batch k of this epoch done"), compared cyclically on 32 bits
# (cuStreamWaitValue32 GEQ: (int32)(*flag - value) >= 0).  Every reconfiguration moves the base past anything the old
# epoch can have written (dispatcher._reconfigure), so values of different epochs never alias.
FLAG_BASE_STEP = 0x10000     # added on top of the old epoch's highest value
POISON_OFFSET = 0x40000000   # base + this satisfies every outstanding ">= base + k" wait of the epoch


def poison_value(flag_base: int) -> int:
    return (int(flag_base) + POISON_OFFSET) & 0xFFFFFFFF


# ------------------------------------------------------------------------------------------------
class ShmRegion:
    def __init__(self, name: Optional[str] = None, size: int = 0, create: bool = False):
        if create:
            name = name or ("adapt_" + uuid.uuid4().hex[:12])
            self.shm = shared_memory.SharedMemory(name=name, create=True, size=size)
            self.buf_view = np.frombuffer(self.shm.buf, dtype=np.uint8)
            self.buf_view[:] = 0
        else:
            self.shm = shared_memory.SharedMemory(name=name, create=False)
            try:  # attaching must not make this process's resource tracker unlink the segment
                from multiprocessing import resource_tracker

                resource_tracker.unregister(self.shm._name, "shared_memory")  # type: ignore[attr-defined]
            except Exception:
                pass
            self.buf_view = np.frombuffer(self.shm.buf, dtype=np.uint8)
        self.name = self.shm.name
        self.size = self.shm.size
        self.owner = create
        self.addr = self.buf_view.ctypes.data

    def view(self, dtype, offset: int, count: int) -> np.ndarray:
        return np.frombuffer(self.shm.buf, dtype=dtype, count=count, offset=offset)

    def close(self) -> None:
        self.buf_view = None
        try:
            self.shm.close()
        except BufferError:
            # numpy views are still alive somewhere (a worker thread winding down): leave the mapping
            # to the OS and stop SharedMemory.__del__ from retrying the close
            self.shm.close = lambda: None  # type: ignore[method-assign]
        if self.owner:
            try:
                self.shm.unlink()
            except FileNotFoundError:
                pass


# ------------------------------------------------------------------------------------------------
REC_WORDS = 6  # kind, epoch, k (dense index within the epoch), seq (global id), batch, aux
KIND_DATA, KIND_STOP = 1, 2
KIND_PINNED = 0x100  # header flag: `aux` points at pinned host memory (DMA straight from it)
KIND_RESIDENT = 0x200  # header flag: the first stage's input slots were preloaded (device-resident bench leg)
KIND_FED = 0x400  # header flag: the dispatcher process itself copies this micro-batch's input into the first stage's
#                   slot (the first stage lives in another process and cannot read the dispatcher's host memory)


class ControlRing:
    """SPSC ring of int64[REC_WORDS] records.  Layout: int64 head, int64 tail, records."""

    CAP = 256
    BYTES = 16 + CAP * REC_WORDS * 8

    def __init__(self, region: ShmRegion, offset: int):
        self.ctr = region.view(np.int64, offset, 2)
        self.rec = region.view(np.int64, offset + 16, self.CAP * REC_WORDS).reshape(self.CAP, REC_WORDS)

    def push(self, rec: Sequence[int], timeout: float = 5.0) -> bool:
        t0 = time.monotonic()
        while self.ctr[0] - self.ctr[1] >= self.CAP:
            if time.monotonic() - t0 > timeout:
                return False
            time.sleep(50e-6)
        h = int(self.ctr[0])
        self.rec[h % self.CAP, :] = rec
        self.ctr[0] = h + 1  # publish after the record is written (x86 keeps store order)
        return True

    def pop(self, timeout: float = 0.01) -> Optional[Tuple[int, ...]]:
        t0 = time.monotonic()
        spins = 0
        while self.ctr[0] == self.ctr[1]:
            spins += 1
            if spins > 200:
                if time.monotonic() - t0 > timeout:
                    return None
                time.sleep(20e-6)
        t = int(self.ctr[1])
        rec = tuple(int(v) for v in self.rec[t % self.CAP])
        self.ctr[1] = t + 1
        return rec

    def clear(self) -> None:
        self.ctr[1] = self.ctr[0]


# ------------------------------------------------------------------------------------------------
class StateEnum:
    """Worker states (reference src/node_state.py:164-168) plus the ones fail-over needs."""

    IDLE = 0
    BUSY = 1
    UNKNOWN = 2
    PARSE_ERROR = 3
    DEAD = 4
    ERROR = 5


PHASE_INIT, PHASE_RUN, PHASE_QUIESCE, PHASE_CONFIGURE, PHASE_SHUTDOWN = 0, 1, 2, 3, 4
MAX_WORKERS = 16
_G_WORDS = 16          # global header
_W_WORDS = 16          # per worker
_TABLE_BYTES = 8192    # JSON partition table

# global words
G_EPOCH, G_PHASE, G_NWORKERS, G_TABLE_LEN, G_T0, G_T1, G_TIMING_REQ = 0, 1, 2, 3, 4, 5, 6
# per-worker words
(W_HEARTBEAT, W_STATE, W_KILL, W_ACK_QUIESCE, W_ACK_READY, W_LAST_K, W_PID, W_DEVICE, W_ERRORS, W_ELAPSED_US,
 W_TIMING_ACK, W_BUSY_US, W_BUSY_N, W_BUSY_MIN_US, W_XFER_NS, W_XFER_BYTES) = range(16)
# W_BUSY_US: moving average of the device time this worker's stage spends per micro-batch (microseconds, this epoch);
# W_BUSY_N: micro-batches folded into it; W_BUSY_MIN_US: the fastest single micro-batch of the epoch (the stage's own
# healthy baseline -- kept by the worker, so that it does not depend on when the dispatcher happens to look).
# Read by the dispatcher's straggler watchdog.
# W_XFER_NS / W_XFER_BYTES: duration (CUDA events on the transfer stream, waits excluded) and size of the copy-engine
# hand-off of the last micro-batch of the timing window; 0 for a stage whose last kernel stores into the peer's slot.


class HealthTable:
    """The dispatcher's in-process health table (shared with the workers through shared memory):
    heartbeats, states, kill-injection flags, epoch/phase of a reconfiguration and the current
    partition table.  Replaces the etcd `/workers/<ip>` keys (reference src/node_state.py:18-20)
    and the undefined `_worker_monitor` / `_task_watchdog` (reference src/dispatcher.py:276,303)."""

    BYTES = (_G_WORDS + MAX_WORKERS * _W_WORDS) * 8 + _TABLE_BYTES

    def __init__(self, region: ShmRegion, offset: int = 0):
        self.g = region.view(np.int64, offset, _G_WORDS)
        self.w = region.view(np.int64, offset + _G_WORDS * 8, MAX_WORKERS * _W_WORDS).reshape(MAX_WORKERS, _W_WORDS)
        self.table = region.view(np.uint8, offset + (_G_WORDS + MAX_WORKERS * _W_WORDS) * 8, _TABLE_BYTES)

    # -- partition table -------------------------------------------------------------------
    def write_table(self, table: dict) -> None:
        raw = json.dumps(table).encode()
        if len(raw) > _TABLE_BYTES:
            raise ValueError("partition table too large")
        self.g[G_TABLE_LEN] = 0
        self.table[:len(raw)] = np.frombuffer(raw, dtype=np.uint8)
        self.g[G_TABLE_LEN] = len(raw)

    def read_table(self) -> dict:
        n = int(self.g[G_TABLE_LEN])
        return json.loads(bytes(self.table[:n]).decode()) if n else {}

    # -- worker side -------------------------------------------------------------------------
    def beat(self, wid: int) -> None:
        self.w[wid, W_HEARTBEAT] = time.monotonic_ns()

    def set_state(self, wid: int, state: int) -> None:
        self.w[wid, W_STATE] = state

    # -- dispatcher side ---------------------------------------------------------------------
    def age_ms(self, wid: int) -> float:
        return (time.monotonic_ns() - int(self.w[wid, W_HEARTBEAT])) / 1e6

    @property
    def epoch(self) -> int:
        return int(self.g[G_EPOCH])

    @property
    def phase(self) -> int:
        return int(self.g[G_PHASE])


# ------------------------------------------------------------------------------------------------
class ResultRing:
    """R slots of [int64 header x8 | payload].  The last stage DMA-writes the payload, then a
    stream-ordered 32-bit write publishes `ready = flag_base + k + 1`; the dispatcher polls, copies out and
    sets `free = k + 1`.  header words: 0 ready(u32 in low half), 1 free, 2 epoch, 3 seq, 4 batch."""

    SLOTS = 8
    HDR = 64

    def __init__(self, region: ShmRegion, offset: int, payload_bytes: int):
        self.region = region
        self.offset = offset
        self.payload_bytes = int(payload_bytes)
        self.stride = self.HDR + (self.payload_bytes + 63) // 64 * 64
        self.hdr = [region.view(np.int64, offset + i * self.stride, 8) for i in range(self.SLOTS)]
        self.ready = [region.view(np.uint32, offset + i * self.stride, 1) for i in range(self.SLOTS)]

    @classmethod
    def bytes_needed(cls, payload_bytes: int) -> int:
        return cls.SLOTS * (cls.HDR + (int(payload_bytes) + 63) // 64 * 64)

    def slot_addr(self, i: int) -> int:
        return self.region.addr + self.offset + i * self.stride

    def payload(self, i: int, nbytes: int) -> np.ndarray:
        return self.region.view(np.uint8, self.offset + i * self.stride + self.HDR, nbytes)

    def reset(self) -> None:
        for h in self.hdr:
            h[:] = 0


# ------------------------------------------------------------------------------------------------
class CudaBackend:
    """Device operations of one worker, all through the C ABI.  Needs a B200."""

    name = "cuda"

    def __init__(self, device: int):
        N.require_gpu()
        self.lib = N.lib()
        self.device = int(device)
        self._registered: List[int] = []

    # memory
    def alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        N.check(self.lib.adapt_device_alloc(self.device, int(nbytes), C.byref(p)))
        return int(p.value)

    def free(self, ptr: int) -> None:
        N.check(self.lib.adapt_device_free(self.device, ptr))

    def export(self, ptr: int) -> bytes:
        h = (C.c_ubyte * 64)()
        N.check(self.lib.adapt_ipc_export(self.device, ptr, h))
        return bytes(h)

    def open(self, token: bytes) -> int:
        h = (C.c_ubyte * 64).from_buffer_copy(token)
        p = C.c_void_p()
        N.check(self.lib.adapt_ipc_import(self.device, h, C.byref(p)))
        return int(p.value)

    def peer_enable(self, other_device: int) -> None:
        N.check(self.lib.adapt_peer_enable(self.device, int(other_device)))

    def register_host(self, addr: int, nbytes: int) -> int:
        p = C.c_void_p()
        N.check(self.lib.adapt_host_register(self.device, addr, int(nbytes), C.byref(p)))
        self._registered.append(addr)
        return int(p.value)

    # streams
    def new_stream(self) -> int:
        p = C.c_void_p()
        N.check(self.lib.adapt_stream_create(self.device, C.byref(p)))
        return int(p.value)

    def sync(self, stream: int) -> None:
        N.check(self.lib.adapt_stream_sync(self.device, stream))

    def busy(self, stream: int) -> bool:
        v = C.c_int(0)
        N.check(self.lib.adapt_stream_query(self.device, stream, C.byref(v)))
        return bool(v.value)

    # flags and copies
    def write_flag(self, stream: int, addr: int, value: int) -> None:
        N.check(self.lib.adapt_stream_write_flag(self.device, stream, addr, value & 0xFFFFFFFF))

    def wait_flag(self, stream: int, addr: int, value: int) -> None:
        N.check(self.lib.adapt_stream_wait_flag_geq(self.device, stream, addr, value & 0xFFFFFFFF))

    def h2d(self, dst: int, src_addr: int, nbytes: int, stream: int) -> None:
        N.check(self.lib.adapt_memcpy_h2d(self.device, dst, src_addr, int(nbytes), stream))

    def d2h(self, dst_addr: int, src: int, nbytes: int, stream: int) -> None:
        N.check(self.lib.adapt_memcpy_d2h(self.device, dst_addr, src, int(nbytes), stream))

    def d2d(self, dst: int, src: int, nbytes: int, stream: int) -> None:
        """Device-to-device copy (same GPU, peer GPU or IPC-mapped peer memory) on a copy engine."""
        N.check(self.lib.adapt_memcpy_d2d(self.device, dst, src, int(nbytes), stream))

    def memset(self, dst: int, value: int, nbytes: int, stream: int) -> None:
        N.check(self.lib.adapt_memset(self.device, dst, value, int(nbytes), stream))

    # timing events
    def event_create(self) -> int:
        p = C.c_void_p()
        N.check(self.lib.adapt_event_create(self.device, C.byref(p)))
        return int(p.value)

    def event_record(self, event: int, stream: int) -> None:
        N.check(self.lib.adapt_event_record(self.device, event, stream))

    def event_sync(self, event: int) -> None:
        N.check(self.lib.adapt_event_sync(self.device, event))

    def event_done(self, event: int) -> bool:
        v = C.c_int(0)
        N.check(self.lib.adapt_event_query(self.device, event, C.byref(v)))
        return bool(v.value)

    def spin(self, stream: int, microseconds: int) -> None:
        """Fault injection: keep `stream` busy for `microseconds` (a stage that is alive but slow)."""
        N.check(self.lib.adapt_debug_spin(self.device, stream, int(microseconds)))

    def event_elapsed_ms(self, start: int, stop: int) -> float:
        v = C.c_float(0)
        N.check(self.lib.adapt_event_elapsed_ms(self.device, start, stop, C.byref(v)))
        return float(v.value)

    # liveness
    def start_heartbeat(self, word_addr: int, interval_s: float) -> int:
        p = C.c_void_p()
        N.check(self.lib.adapt_heartbeat_start(word_addr, max(100, int(interval_s * 1e6)), C.byref(p)))
        return int(p.value)

    def stop_heartbeat(self, handle: int) -> None:
        self.lib.adapt_heartbeat_stop(handle)

    # compute
    def make_executor(self, model, part, mode: str, max_batch: int, input_format: str = "float32"):
        from .executor import StageExecutor

        return StageExecutor(model, part, device=self.device, mode=mode, max_batch=max_batch, reserve_full=True,
                             input_format=input_format)

    def close(self) -> None:
        for a in self._registered:
            self.lib.adapt_host_unregister(self.device, a)
        self._registered = []
