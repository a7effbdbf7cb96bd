"""Worker role: one pipeline stage bound to one B200.

Mirror of reference src/node.py `Node` (config server + data server + predict/forward loop,
src/node.py:29-194): a worker receives a partition ("configuration"), then for every
micro-batch takes the incoming activation, runs its model slice and forwards the result to the
next hop (`to_send.get -> model.predict -> _comp -> socket_send`, src/node.py:173-179).  Here
  - the configuration is a row of the dispatcher's partition table (weights are already
    resident on every GPU, so (re)configuration is metadata only);
  - the incoming activation was stored into this worker's double-buffered input arena by the
    previous stage's last kernel (peer stores over NVLink), announced by a 32-bit ready flag;
  - `model.predict` is one CUDA-graph launch of the stage's kernels;
  - forwarding is the epilogue of the stage's last kernel writing into the next worker's arena,
    followed by stream-ordered flag writes; the host never touches activation data and never
    blocks on the GPU except to apply back-pressure on result slots.
"""
from __future__ import annotations

import ctypes
import os
import sys
import threading
import time
import traceback
from typing import Callable, Dict, List, Optional

import numpy as np

from . import _native as N
from .dag_util import construct_model
from .fabric import (G_EPOCH, G_PHASE, G_T0, G_T1, G_TIMING_REQ, KIND_DATA, KIND_FED, KIND_PINNED, KIND_RESIDENT, KIND_STOP, PHASE_CONFIGURE, PHASE_QUIESCE, PHASE_RUN,
                     PHASE_SHUTDOWN, W_ACK_QUIESCE, W_ACK_READY, W_BUSY_MIN_US, W_BUSY_N, W_BUSY_US, W_XFER_BYTES, W_XFER_NS, W_DEVICE, W_ERRORS, W_ELAPSED_US, W_HEARTBEAT, W_KILL, W_LAST_K, W_PID,
                     W_TIMING_ACK,
                     ControlRing, CudaBackend, HealthTable, ResultRing, ShmRegion, StateEnum, poison_value)

# uint32 word indices inside a worker's flag block.  Flags live in the pinned, device-mapped
# control segment (host memory): GPUs write / wait on them with stream memory operations, and the
# host can overwrite them with plain stores -- which is what lets a reconfiguration release a
# blocked stream without having to get a copy scheduled behind it.
F_IN_READY, F_OUT_FREE, F_H2D_DONE, F_IN_CONSUMED, F_OUT_READY, F_OUT_COPIED = 0, 2, 4, 6, 8, 10
F_XFER_PART = 12   # + 2 * j + slot: piece j (1..XFER_STREAMS-1) of a split hand-off copy has landed
BUSY_SAMPLE_EVERY = int(os.environ.get("ADAPT_BUSY_SAMPLE_EVERY", "4"))   # micro-batches per busy-time sample (2 event records)
XFER_STREAMS = 1   # copy streams a large hand-off may be split over.  Measured on B200 (tools/handoff_bench.py, 205 MB
#                    GPU 0 -> GPU 1): ONE copy reaches 737 GB/s, split in 2 / 4 / 8 concurrent pieces 693 / 602 / 488 GB/s
#                    -- several copy engines on one link get in each other's way -- so the default is a single copy;
#                    ADAPT_XFER_STREAMS=n keeps the split path available.
XFER_SPLIT_MIN_BYTES = 8 << 20  # a piece of a split hand-off is at least this big
COPY_HANDOFF_MIN_BYTES = 300 * 1024  # cut tensors at least this big per image go through a copy engine
FLAG_WORDS = 64
IMAGE_BYTES = 224 * 224 * 3 * 4
OUT_ROW_BYTES = 1000 * 4


def arena_slot_bytes(model, mode: str, max_batch: int) -> int:
    """Largest per-slot input a stage can be handed: the fp32 image batch or the biggest cut tensor."""
    from .dag_util import cut_tensor_elems
    from .executor import MODES, elem_bytes

    m = MODES[mode]
    biggest = max(cut_tensor_elems(model, l.name, m) for l in model.layers[1:-1])
    shp = model.layers[0].output_shape[1:]
    return max(int(np.prod(shp)) * 4, biggest * elem_bytes(N.DT_WORK, m)) * max_batch


class NodeState:
    """Per-worker state record (reference src/node_state.py:9-36): chunk_size, dispatcher address
    and the worker's state enum -- backed by a row of the dispatcher's health table instead of an
    etcd key `/workers/<ip>`."""

    def __init__(self, chunk_size: int = 512 * 1000, dispatcher_ip: str = "in-process", health: Optional[HealthTable] = None,
                 wid: int = 0):
        self.chunk_size = chunk_size
        self.dispatcher_ip = dispatcher_ip
        self._health = health
        self._wid = wid
        self._state_enum = StateEnum.IDLE
        self._lock = threading.Lock()
        self.etcd_key = f"/workers/{wid}"

    @property
    def state(self) -> int:
        return self._state_enum

    @state.setter
    def state(self, v: int) -> None:
        with self._lock:
            self._state_enum = v
            if self._health is not None:
                self._health.set_state(self._wid, v)


class StageWorker:
    """Engine of one worker.  `session` is the dict the dispatcher publishes (fabric layout)."""

    def __init__(self, wid: int, device: int, model, session: dict, backend=None):
        self.wid = wid
        self.device = device
        self.model = model
        self.session = session
        self.mode = session["mode"]
        self.max_batch = session["max_batch"]
        self.backend = backend if backend is not None else CudaBackend(device)
        self.region = ShmRegion(session["shm"], create=False) if isinstance(session["shm"], str) else session["shm"]
        self.health = HealthTable(self.region, session["health_off"])
        self.ring = ControlRing(self.region, session["ring_off"] + wid * ControlRing.BYTES)
        self.rings = [ControlRing(self.region, session["ring_off"] + i * ControlRing.BYTES) for i in range(session["n_workers"])]
        self.results = ResultRing(self.region, session["result_off"], self.max_batch * OUT_ROW_BYTES)
        self.state = NodeState(health=self.health, wid=wid)
        self.slot_bytes = session["slot_bytes"]
        self.arena = 0
        self.peers: Dict[int, int] = {}
        self.executor = None
        self.epoch = -1
        self.fb = 0  # flag base of the current epoch (partition table, `flag_base`)
        self.stage_index = -1
        self.n_stages = 0
        self.prev_wid = self.next_wid = -1
        self.thread: Optional[threading.Thread] = None
        self.dead = False
        self.error: Optional[str] = None
        self._stop = threading.Event()
        self._hb_thread: Optional[threading.Thread] = None
        self.launches = 0
        self._busy_seen = 0
        self._timing_events = None
        self._xfer_events = None
        self._xfer_timed_bytes = 0
        self._timing_armed = False
        self.handoff_copy = False
        self.rails: List[dict] = []          # ingest rails ever created (other GPUs' PCIe links)
        self._active_rails: List[dict] = []  # ... of the GPUs in the current stage order
        self.on_kill: Optional[Callable[[], None]] = None
        # Fault injection for the tests (like `DEFER.inject_kill`): ADAPT_FAULT_INJECT="stale_slot:<k>" makes the LAST
        # stage of a multi-stage pipeline run micro-batch k of every epoch on the OTHER input slot, i.e. on the
        # tensor of micro-batch k-1 -- the hand-off bug a value check of the delivered results has to catch.
        self._fault_stale_k = -1
        # "slow_worker:<worker>:<microseconds>:<from k>": from micro-batch k on, that worker's stream is kept busy for the
        # given time on top of every micro-batch (a stage that is alive but slow: what the straggler watchdog is for)
        self._fault_slow = None
        fault = os.environ.get("ADAPT_FAULT_INJECT", "")
        if fault.startswith("stale_slot:"):
            self._fault_stale_k = int(fault.split(":")[1])
        elif fault.startswith("slow_worker:"):
            st, us, k0 = (int(v) for v in fault.split(":")[1:4])
            self._fault_slow = (st, us, k0)
        # per-micro-batch busy time of this stage (device time between two events around the stage's work), folded
        # into a moving average in the health table: the straggler watchdog's input (reference src/dispatcher.py:193)
        self._busy_events: List[tuple] = []
        self._busy_pending: List[tuple] = []
        self._busy_ewma = 0.0
        self._busy_min = 0.0
        self._busy_n = 0
        if hasattr(self.backend, "abort_check"):  # synchronous test backends must be able to give up a wait
            self.backend.abort_check = lambda: (self.health.phase != PHASE_RUN or bool(self.health.w[self.wid, W_KILL])
                                                or self._stop.is_set())

    # ------------------------------------------------------------------ set-up
    def allocate(self):
        """Allocate this worker's arena [in_buf 0 | in_buf 1], its result scratch, its pinned input
        staging and its streams; returns the token peers use to map the arena."""
        b = self.backend
        self.arena = b.alloc(4 * self.slot_bytes)  # [in 0 | in 1 | local out 0 | local out 1]
        self.d_out = b.alloc(self.max_batch * OUT_ROW_BYTES)
        self.stream = b.new_stream()
        self.copy_stream = b.new_stream()
        # copy-engine hand-off of large cut tensors (optionally split over several streams, see XFER_STREAMS)
        n_x = max(1, min(4, int(os.environ.get("ADAPT_XFER_STREAMS", str(XFER_STREAMS)))))
        self.xfer_streams = [b.new_stream() for _ in range(n_x)]
        self.xfer_stream = self.xfer_streams[0]
        # pinned staging for pageable inputs (only the first stage of a session uses it)
        stage_bytes = self.max_batch * self.session["in_row_bytes"]
        self._stage_np = np.zeros(2 * stage_bytes + 4096, dtype=np.uint8)
        base = (self._stage_np.ctypes.data + 4095) // 4096 * 4096
        b.register_host(base, 2 * stage_bytes)
        self.stage_addr = [base, base + stage_bytes]
        if hasattr(b, "event_create"):
            self._timing_events = (b.event_create(), b.event_create())
            self._xfer_events = (b.event_create(), b.event_create())
        # the whole control segment (flags + result ring) is pinned and mapped into this device
        self.ctl_dev = b.register_host(self.region.addr, self.region.size)
        self.results_dev = self.ctl_dev + self.session["result_off"]
        self.flags_host = self.region.view(np.uint32, self.session["flag_off"],
                                           self.session["n_workers"] * FLAG_WORDS).reshape(-1, FLAG_WORDS)
        self.health.w[self.wid, W_PID] = os.getpid()
        self.health.w[self.wid, W_DEVICE] = self.device
        if self.session.get("ipc", False):
            return b.export(self.arena)
        return self.arena

    def connect(self, tokens: Dict[int, object], devices: Dict[int, int]) -> None:
        for w, tok in tokens.items():
            if w == self.wid:
                self.peers[w] = self.arena
            elif self.session.get("ipc", False):
                self.peers[w] = self.backend.open(tok)
            else:
                if devices[w] != self.device:
                    self.backend.peer_enable(devices[w])
                self.peers[w] = int(tok)

    def flag(self, wid: int, word: int) -> int:
        """Device-side address of flag `word` of worker `wid` (in the mapped control segment)."""
        return self.ctl_dev + self.session["flag_off"] + 4 * (wid * FLAG_WORDS + word)

    def in_buf(self, wid: int, slot: int) -> int:
        return self.peers[wid] + slot * self.slot_bytes

    def out_buf(self, slot: int) -> int:
        """Local staging slot for this stage's output when the hand-off goes through a copy engine."""
        return self.arena + (2 + slot) * self.slot_bytes

    # ------------------------------------------------------------------ reconfiguration
    def _write_own_flags(self, value: int) -> None:
        self.flags_host[self.wid, :] = value

    def _drain(self) -> None:
        """Release every GPU-side wait that will never be satisfied and let the queued work run
        out.  A stale flag write from a draining peer can overwrite the poison, so it is re-applied
        until both streams are idle."""
        t0 = time.monotonic()
        while True:
            self._write_own_flags(poison_value(self.fb))
            if not (self.backend.busy(self.copy_stream) or self.backend.busy(self.stream)
                    or any(self.backend.busy(x) for x in self.xfer_streams)
                    or any(r["backend"].busy(r["stream"]) for r in getattr(self, "rails", []))):
                return
            if time.monotonic() - t0 > 20:
                raise TimeoutError("stage streams did not drain")
            time.sleep(200e-6)

    def _quiesce(self, epoch: int) -> None:
        self._drain()
        self.health.w[self.wid, W_ACK_QUIESCE] = epoch

    def _configure(self, epoch: int) -> None:
        table = self.health.read_table()
        assert table["epoch"] == epoch, (table, epoch)
        order: List[int] = table["order"]
        self.fb = int(table.get("flag_base", 0)) & 0xFFFFFFFF
        self._write_own_flags(self.fb)   # "nothing of this epoch done yet"
        self.ring.clear()
        self.k_expect = 0
        if self.wid not in order:
            self.stage_index, self.n_stages = -1, len(order)
        else:
            i = order.index(self.wid)
            bounds = [self.model.layers[0].name] + list(table["part_at"]) + [self.model.layers[-1].name]
            part = construct_model(self.model, bounds[i], bounds[i + 1], f"part{i + 1}")
            if self.executor is None:
                fmt = self.session.get("input_format", "float32")
                if fmt == "float32":
                    self.executor = self.backend.make_executor(self.model, part, self.mode, self.max_batch)
                else:
                    self.executor = self.backend.make_executor(self.model, part, self.mode, self.max_batch, input_format=fmt)
            else:
                self.executor.reslice(part)
            self.stage_index, self.n_stages = i, len(order)
            self.prev_wid = order[i - 1] if i > 0 else -1
            self.next_wid = order[i + 1] if i + 1 < len(order) else -1
            mode = table.get("handoff", "auto")
            out_bytes = getattr(self.executor, "out_bytes_per_image", 0)
            if self.next_wid >= 0 and self.max_batch * out_bytes > self.slot_bytes:
                raise RuntimeError(f"cut tensor of {self.max_batch} x {out_bytes} bytes does not fit a {self.slot_bytes}-byte "
                                   f"hand-off slot (session mode {self.mode!r})")
            big = out_bytes >= COPY_HANDOFF_MIN_BYTES
            self.handoff_copy = self.next_wid >= 0 and (mode == "copy" or (mode == "auto" and big))
            for s in (0, 1):
                if self.next_wid < 0:
                    out_ptr = self.d_out
                elif self.handoff_copy:
                    out_ptr = self.out_buf(s)
                else:
                    out_ptr = self.in_buf(self.next_wid, s)  # epilogue stores straight into the consumer's slot
                self.executor.bind_io(s, self.in_buf(self.wid, s), out_ptr)
            if table.get("warm_batches"):
                self._warm(table["warm_batches"])
        self._setup_rails(order, table)
        self._busy_reset()
        self.epoch = epoch
        self.health.w[self.wid, W_ACK_READY] = epoch

    def _setup_rails(self, order: List[int], table: dict) -> None:
        """Ingest rails of the first stage: one (context, stream, 2 staging buffers) per OTHER GPU of the
        pipeline.  A pinned input batch k goes host -> GPU j over GPU j's own PCIe link, then over NVLink into
        this stage's input slot; rail 0 is this GPU's own link.  Rails are created once and re-used across
        reconfigurations; only the GPUs of the current stage order are active."""
        self._active_rails = []
        if (table.get("ingest_rails", "auto") != "auto" or getattr(self.backend, "name", "") != "cuda"
                or self.stage_index != 0 or len(order) < 2):
            return
        have = {r["device"]: r for r in self.rails}
        nbytes = self.max_batch * self.session["in_row_bytes"]
        for w in order:
            d = int(self.health.w[w, W_DEVICE])
            if w == self.wid or d == self.device or any(r["device"] == d for r in self._active_rails):
                continue
            r = have.get(d)
            if r is None:
                try:
                    rb = CudaBackend(d)
                    rb.peer_enable(self.device)
                    r = {"device": d, "backend": rb, "stream": rb.new_stream(),
                         "staging": [rb.alloc(nbytes), rb.alloc(nbytes)], "count": 0}
                except Exception as e:  # noqa: BLE001  (no peer access / no memory: that GPU is simply not a rail)
                    print(f"[adapt] worker {self.wid}: GPU {d} not usable as an ingest rail: {e}", file=sys.stderr)
                    continue
                self.rails.append(r)
            self._active_rails.append(r)

    def _warm(self, batches) -> None:
        """Build launch records + graphs for the batch sizes the dispatcher announces, so the
        first real micro-batch does not pay for graph capture.  Output goes to real buffers, but
        no flag is written, so downstream never consumes it."""
        for b in batches:
            for s in (0, 1):
                if hasattr(self.executor, "prepare"):
                    self.executor.prepare(s, int(b))
                self.executor.enqueue(s, int(b), self.stream)
        self.backend.sync(self.stream)

    # ------------------------------------------------------------------ steady state
    def _process(self, rec) -> None:
        kind, epoch, k, seq, batch, aux = rec
        b, ex = self.backend, self.executor
        s = k & 1
        # flag values of micro-batch k in this epoch: "k done" = flag_base + k + 1 (cyclic 32-bit compare).  The base
        # moves past everything the previous epoch can have written, so a late write from a stage that was dropped
        # while still running can never satisfy a wait of this epoch.
        v_done = (self.fb + k + 1) & 0xFFFFFFFF   # this micro-batch
        v_prev = (self.fb + k - 1) & 0xFFFFFFFF   # the micro-batch that used slot s before (k - 2)
        first = self.stage_index == 0
        last = self.stage_index == self.n_stages - 1
        if first and (kind & KIND_RESIDENT):
            # measurement mode: the input batch was preloaded into both input slots (no copy); keep the
            # copy-side flag current so that copied micro-batches can follow resident ones
            b.write_flag(self.stream, self.flag(self.wid, F_H2D_DONE + s), v_done)
        elif first and (kind & KIND_FED):
            # the dispatcher process places the input in this slot itself (DEFER._proxy_feed) and raises the flag
            b.wait_flag(self.stream, self.flag(self.wid, F_H2D_DONE + s), v_done)
        elif first:
            # host -> device copy of the input on the copy stream, double buffered against compute
            nbytes = batch * self.session["in_row_bytes"]
            src = aux
            if not (kind & KIND_PINNED):
                # Pageable input: stage it through this worker's pinned slot so that the copy below is
                # truly asynchronous.  (A pageable cudaMemcpyAsync blocks the host until the stream gets
                # there, i.e. possibly forever behind a flag a dead stage never writes.)
                fl = self.flags_host[self.wid]
                t0 = time.monotonic()
                while k >= 2 and ((int(fl[F_H2D_DONE + s]) - v_prev) & 0xFFFFFFFF) >= 0x80000000:
                    if self.health.phase != PHASE_RUN or self._stop.is_set():
                        return
                    if time.monotonic() - t0 > 30:
                        raise TimeoutError("input staging slot never drained")
                    time.sleep(20e-6)
                ctypes.memmove(self.stage_addr[s], aux, nbytes)
                src = self.stage_addr[s]
            rails = self._active_rails if (kind & KIND_PINNED) else []
            lane = k % (len(rails) + 1)
            if lane == 0:
                if k >= 2:
                    b.wait_flag(self.copy_stream, self.flag(self.wid, F_IN_CONSUMED + s), v_prev)
                b.h2d(self.in_buf(self.wid, s), src, nbytes, self.copy_stream)
                b.write_flag(self.copy_stream, self.flag(self.wid, F_H2D_DONE + s), v_done)
            else:
                # over GPU j's PCIe link into its staging buffer (no dependency: runs ahead), then -- once this
                # stage has consumed what the slot held -- one NVLink copy into the input slot
                rail = rails[lane - 1]
                rb, rs = rail["backend"], rail["stream"]
                stg = rail["staging"][rail["count"] & 1]
                rail["count"] += 1
                rb.h2d(stg, src, nbytes, rs)
                if k >= 2:
                    rb.wait_flag(rs, self.flag(self.wid, F_IN_CONSUMED + s), v_prev)
                rb.d2d(self.in_buf(self.wid, s), stg, nbytes, rs)
                rb.write_flag(rs, self.flag(self.wid, F_H2D_DONE + s), v_done)
            b.wait_flag(self.stream, self.flag(self.wid, F_H2D_DONE + s), v_done)
        else:
            b.wait_flag(self.stream, self.flag(self.wid, F_IN_READY + s), v_done)
        staged = (not last) and self.handoff_copy
        if staged and k >= 2:
            # local output slot s is re-used: its copy to the consumer (micro-batch k-2) must be done
            b.wait_flag(self.stream, self.flag(self.wid, F_OUT_COPIED + s), v_prev)
        elif not last and k >= 2:
            b.wait_flag(self.stream, self.flag(self.wid, F_OUT_FREE + s), v_prev)
        if last:
            r = k % ResultRing.SLOTS
            hdr = self.results.hdr[r]
            t0 = time.monotonic()
            while k >= ResultRing.SLOTS and int(hdr[1]) < k - ResultRing.SLOTS + 1:  # back-pressure
                if self.health.epoch != self.epoch or self._stop.is_set():
                    return
                if time.monotonic() - t0 > 30:
                    raise TimeoutError("result slot never freed")
                time.sleep(20e-6)
        timed = self._timing_events is not None
        if timed and k == int(self.health.g[G_T0]):
            b.event_record(self._timing_events[0], self.stream)
        ev = self._busy_begin()
        t_host = time.perf_counter()
        ex.enqueue(s ^ 1 if (k == self._fault_stale_k and last and not first) else s, batch, self.stream)
        if self._fault_slow is not None and self._fault_slow[0] == self.wid and k >= self._fault_slow[2]:
            if hasattr(b, "spin"):
                b.spin(self.stream, self._fault_slow[1])
            else:
                time.sleep(self._fault_slow[1] * 1e-6)
        self._busy_end(ev, t_host)
        self.launches += ex.num_launches
        if first:  # (also for resident micro-batches, so that a later copied one finds the flag current)
            b.write_flag(self.stream, self.flag(self.wid, F_IN_CONSUMED + s), v_done)
        else:
            b.write_flag(self.stream, self.flag(self.prev_wid, F_OUT_FREE + s), v_done)
        if last:
            hdr[2], hdr[3], hdr[4] = epoch, seq, batch
            off = r * self.results.stride
            b.d2h(self.results_dev + off + ResultRing.HDR, self.d_out, batch * OUT_ROW_BYTES, self.stream)
            b.write_flag(self.stream, self.results_dev + off, v_done)
        elif staged:
            # Large cut tensor: the stage wrote it to a LOCAL slot; a copy engine pushes it over NVLink on
            # the transfer stream while this stage's SMs start the next micro-batch (hand-off overlapped
            # with compute instead of stretching the last kernel to NVLink speed).
            x = self.xfer_stream
            b.write_flag(self.stream, self.flag(self.wid, F_OUT_READY + s), v_done)
            nbytes = batch * self.executor.out_bytes_per_image
            n_parts = max(1, min(len(self.xfer_streams), nbytes // XFER_SPLIT_MIN_BYTES))
            step = (nbytes // n_parts + 4095) // 4096 * 4096
            for j in range(n_parts):
                xs = self.xfer_streams[j]
                lo, hi = j * step, min(nbytes, (j + 1) * step)
                b.wait_flag(xs, self.flag(self.wid, F_OUT_READY + s), v_done)
                if k >= 2:
                    b.wait_flag(xs, self.flag(self.wid, F_OUT_FREE + s), v_prev)  # consumer done with its slot s
                time_it = timed and j == 0 and k == int(self.health.g[G_T1]) - 1
                if time_it:
                    b.event_record(self._xfer_events[0], xs)
                b.d2d(self.in_buf(self.next_wid, s) + lo, self.out_buf(s) + lo, hi - lo, xs)
                if time_it:
                    b.event_record(self._xfer_events[1], xs)
                    self._xfer_timed_bytes = hi - lo
                if j > 0:
                    b.write_flag(xs, self.flag(self.wid, F_XFER_PART + 2 * j + s), v_done)
            for j in range(1, n_parts):  # the first stream publishes once every piece has landed
                b.wait_flag(x, self.flag(self.wid, F_XFER_PART + 2 * j + s), v_done)
            b.write_flag(x, self.flag(self.wid, F_OUT_COPIED + s), v_done)
            b.write_flag(x, self.flag(self.next_wid, F_IN_READY + s), v_done)
            self.rings[self.next_wid].push((kind, epoch, k, seq, batch, 0))
        else:
            # Small cut tensor: the last kernel's epilogue stored it straight into the consumer's slot
            b.write_flag(self.stream, self.flag(self.next_wid, F_IN_READY + s), v_done)
            self.rings[self.next_wid].push((kind, epoch, k, seq, batch, 0))
        if timed and k == int(self.health.g[G_T1]) - 1:
            b.event_record(self._timing_events[1], self.stream)
            self._timing_armed = True
        self.health.w[self.wid, W_LAST_K] = k

    # ------------------------------------------------------------------ ingest on behalf of a remote first stage
    def feed_remote_first_stage(self, target_wid: int, fb: int, k: int, batch: int, host_addr: int, pinned: bool) -> None:
        """Dispatcher side of KIND_FED (called from the dispatcher's feeder thread, on the dispatcher-local worker
        object whatever state its own stage is in): copy micro-batch k's input from this process's host memory into
        input slot k & 1 of worker `target_wid` -- another process's GPU memory, mapped here over CUDA IPC -- on this
        worker's copy stream, double buffered against that stage's compute with ITS flags."""
        b = self.backend
        s = k & 1
        v_done, v_prev = (fb + k + 1) & 0xFFFFFFFF, (fb + k - 1) & 0xFFFFFFFF
        nbytes = batch * self.session["in_row_bytes"]
        src = host_addr
        if not pinned:
            fl = self.flags_host[target_wid]
            t0 = time.monotonic()
            while k >= 2 and ((int(fl[F_H2D_DONE + s]) - v_prev) & 0xFFFFFFFF) >= 0x80000000:
                if self.health.phase != PHASE_RUN or self._stop.is_set():
                    return
                if time.monotonic() - t0 > 30:
                    raise TimeoutError("input staging slot never drained")
                time.sleep(20e-6)
            ctypes.memmove(self.stage_addr[s], host_addr, nbytes)
            src = self.stage_addr[s]
        if k >= 2:
            if hasattr(b, "abort_check"):
                # synchronous (test) backend: its waits give up when THIS worker is killed -- which it is, that is why the
                # dispatcher feeds on its behalf -- so the slot's previous micro-batch is waited for here, on the host
                fl = self.flags_host[target_wid]
                t0 = time.monotonic()
                while ((int(fl[F_IN_CONSUMED + s]) - v_prev) & 0xFFFFFFFF) >= 0x80000000:
                    if self.health.phase != PHASE_RUN or self._stop.is_set():
                        return
                    if time.monotonic() - t0 > 30:
                        raise TimeoutError("first stage never consumed its input slot")
                    time.sleep(20e-6)
            b.wait_flag(self.copy_stream, self.flag(target_wid, F_IN_CONSUMED + s), v_prev)
        b.h2d(self.in_buf(target_wid, s), src, nbytes, self.copy_stream)
        b.write_flag(self.copy_stream, self.flag(target_wid, F_H2D_DONE + s), v_done)

    # ------------------------------------------------------------------ busy time (straggler watchdog input)
    def _busy_begin(self):
        b = self.backend
        if not hasattr(b, "event_done"):
            return None
        self._busy_poll()
        # one micro-batch in BUSY_SAMPLE_EVERY carries the two event records (every one at the start of an epoch, so
        # that the stage's baseline exists at once): the watchdog looks at a moving average over 0.5 s anyway
        self._busy_seen += 1
        if self._busy_seen > 8 and self._busy_seen % BUSY_SAMPLE_EVERY:
            return None
        if self._busy_events:
            pair = self._busy_events.pop()
        elif len(self._busy_pending) < 16:
            pair = (b.event_create(), b.event_create())
        else:
            return None  # (every pair is still in flight: skip this sample)
        b.event_record(pair[0], self.stream)
        return pair

    def _busy_end(self, pair, t_host: float) -> None:
        if pair is None:
            if not hasattr(self.backend, "event_done"):  # synchronous test backend: the call itself is the work
                self._busy_fold((time.perf_counter() - t_host) * 1e6)
            return
        self.backend.event_record(pair[1], self.stream)
        self._busy_pending.append(pair)

    def _busy_poll(self) -> None:
        b = self.backend
        while self._busy_pending and b.event_done(self._busy_pending[0][1]):
            pair = self._busy_pending.pop(0)
            self._busy_fold(b.event_elapsed_ms(pair[0], pair[1]) * 1e3)
            self._busy_events.append(pair)

    def _busy_fold(self, us: float) -> None:
        self._busy_n += 1
        self._busy_ewma = us if self._busy_n == 1 else 0.75 * self._busy_ewma + 0.25 * us
        self._busy_min = us if self._busy_n == 1 else min(self._busy_min, us)
        self.health.w[self.wid, W_BUSY_US] = int(self._busy_ewma)
        self.health.w[self.wid, W_BUSY_MIN_US] = max(1, int(self._busy_min))
        self.health.w[self.wid, W_BUSY_N] = self._busy_n

    def _busy_reset(self) -> None:
        self._busy_events += self._busy_pending  # (streams are drained when this runs)
        self._busy_pending = []
        self._busy_ewma, self._busy_n, self._busy_min = 0.0, 0, 0.0
        self._busy_seen = 0
        self.health.w[self.wid, W_BUSY_US] = 0
        self.health.w[self.wid, W_BUSY_MIN_US] = 0
        self.health.w[self.wid, W_BUSY_N] = 0

    def _report_timing(self) -> None:
        """Device time this stage spent on the micro-batches [G_T0, G_T1) (CUDA events on its stream)."""
        token = int(self.health.g[G_TIMING_REQ])
        if not token or int(self.health.w[self.wid, W_TIMING_ACK]) == token or not self._timing_armed:
            return
        self.backend.event_sync(self._timing_events[1])
        ms = self.backend.event_elapsed_ms(self._timing_events[0], self._timing_events[1])
        self.health.w[self.wid, W_ELAPSED_US] = int(ms * 1000.0)
        if self._xfer_timed_bytes:
            self.backend.event_sync(self._xfer_events[1])
            xms = self.backend.event_elapsed_ms(self._xfer_events[0], self._xfer_events[1])
            self.health.w[self.wid, W_XFER_NS] = int(xms * 1e6)
            self.health.w[self.wid, W_XFER_BYTES] = self._xfer_timed_bytes
            self._xfer_timed_bytes = 0
        else:
            self.health.w[self.wid, W_XFER_NS] = 0
            self.health.w[self.wid, W_XFER_BYTES] = 0
        self._timing_armed = False
        self.health.w[self.wid, W_TIMING_ACK] = token

    def _heartbeat_loop(self) -> None:
        while not self._stop.is_set() and not self.dead:
            self.health.beat(self.wid)
            time.sleep(self.session.get("heartbeat_s", 0.005))

    def _start_heartbeat(self) -> None:
        self.health.beat(self.wid)
        self._hb_native = None
        if hasattr(self.backend, "start_heartbeat"):
            # native beacon thread: a Python thread can be starved by the GIL for tens of ms
            addr = self.health.w[self.wid:self.wid + 1, W_HEARTBEAT:W_HEARTBEAT + 1].ctypes.data
            self._hb_native = self.backend.start_heartbeat(addr, self.session.get("heartbeat_s", 0.005))
        else:
            self._hb_thread = threading.Thread(target=self._heartbeat_loop, daemon=True, name=f"hb{self.wid}")
            self._hb_thread.start()

    def _stop_heartbeat(self) -> None:
        if getattr(self, "_hb_native", None):
            self.backend.stop_heartbeat(self._hb_native)
            self._hb_native = None

    def _die(self) -> None:
        """Injected stage kill: go silent -- no more heartbeats, no more work enqueued; in-flight micro-batches are
        lost.  What is already queued on the GPU is NOT flushed here: releasing its flag waits would let it run on
        inputs that have not arrived and hand garbage to the next stage, whose results the collector would deliver
        until the heartbeat time-out fires (a crashed process's queued work simply disappears).  It finishes on real
        inputs or stays blocked, and is flushed when the dispatcher quiesces the pipeline (`run`, below)."""
        self.dead = True
        self._stop_heartbeat()
        if not self.session.get("silent_kill", False):
            self.state.state = StateEnum.DEAD  # a crash would not announce itself; see `silent_kill`
        if self.on_kill is not None:
            self.on_kill()

    @staticmethod
    def pid_alive(pid: int) -> bool:
        if pid <= 0:
            return False
        if pid == os.getpid():
            return True
        try:
            os.kill(pid, 0)
            return True
        except OSError:
            return False

    def _await_peers_drained(self, timeout: float = 2.0) -> None:
        """Shutdown: nobody frees its arena while a peer whose process is alive may still have GPU work queued that
        writes into it (every worker marks W_ACK_READY = -1 once its streams are idle for good)."""
        t0 = time.monotonic()
        while time.monotonic() - t0 < timeout:
            pending = [w for w in range(self.session["n_workers"]) if w != self.wid and int(self.health.w[w, W_ACK_READY]) != -1
                       and self.pid_alive(int(self.health.w[w, W_PID]))]
            if not pending:
                return
            time.sleep(1e-3)

    def run(self) -> None:
        """Blocking worker loop (the reference's `Node.run`, src/node.py:181-194)."""
        h = self.health
        try:
            self._start_heartbeat()
            while not self._stop.is_set():
                phase, epoch = h.phase, h.epoch
                if phase == PHASE_SHUTDOWN:
                    break
                if int(h.w[self.wid, W_KILL]):
                    self._die()
                    # silent until shut down -- or revived (DEFER.revive clears the kill word): then the heartbeat
                    # resumes and the worker idles without a stage until the dispatcher re-slices it back in
                    while int(h.w[self.wid, W_KILL]) and not self._stop.is_set() and h.phase != PHASE_SHUTDOWN:
                        if h.phase == PHASE_QUIESCE and int(h.w[self.wid, W_ACK_QUIESCE]) != h.epoch:
                            # the pipeline is being re-sliced: flush what this dead stage still has queued, like
                            # everyone else (whatever it writes now carries the old epoch's flag values)
                            try:
                                self._quiesce(h.epoch)
                            except Exception:  # noqa: BLE001
                                pass
                        time.sleep(1e-3)
                    if self._stop.is_set() or h.phase == PHASE_SHUTDOWN:
                        break
                    self.dead = False
                    self.stage_index, self.epoch = -1, -1
                    self.state.state = StateEnum.IDLE
                    self._start_heartbeat()
                    continue
                if phase == PHASE_QUIESCE and int(h.w[self.wid, W_ACK_QUIESCE]) != epoch:
                    self.state.state = StateEnum.IDLE
                    self._quiesce(epoch)
                    continue
                if phase == PHASE_CONFIGURE and int(h.w[self.wid, W_ACK_READY]) != epoch:
                    if int(h.w[self.wid, W_ACK_QUIESCE]) != epoch:
                        self._quiesce(epoch)
                    self._configure(epoch)
                    continue
                if phase != PHASE_RUN or self.epoch != epoch or self.stage_index < 0:
                    time.sleep(100e-6)
                    continue
                rec = self.ring.pop(timeout=0.002)
                if rec is None:
                    self.state.state = StateEnum.IDLE
                    self._report_timing()
                    if self._busy_pending:
                        self._busy_poll()
                    continue
                if (rec[0] & 0xFF) == KIND_STOP:
                    if self.next_wid >= 0:
                        self.rings[self.next_wid].push(rec)
                    continue
                if rec[1] != self.epoch:
                    continue  # stale header from before a reconfiguration
                self.state.state = StateEnum.BUSY
                self._process(rec)
        except Exception as e:  # noqa: BLE001 - surfaced to the dispatcher through the health table
            self.error = f"{type(e).__name__}: {e}\n{traceback.format_exc()}"
            self.state.state = StateEnum.ERROR
            self.health.w[self.wid, W_ERRORS] += 1
            self.dead = True
        finally:
            self._stop_heartbeat()
            try:
                if self.arena:
                    self._drain()  # (a stage killed shortly before shutdown may still have work blocked on its flags)
                self.backend.sync(self.stream)
            except Exception:
                pass
            try:
                self.health.w[self.wid, W_ACK_READY] = -1  # streams idle for good: peers may free what they own
            except Exception:
                pass

    def start(self) -> None:
        self.thread = threading.Thread(target=self.run, daemon=True, name=f"stage-worker-{self.wid}")
        self.thread.start()

    def stop(self) -> None:
        self._stop.set()
        if self.thread is not None:
            self.thread.join(timeout=10)

    def close(self) -> None:
        self.stop()
        try:
            if self.health is not None:
                self.health.w[self.wid, W_ACK_READY] = -1
                self._await_peers_drained()
        except Exception:
            pass
        if self.executor is not None:
            self.executor.close()
            self.executor = None
        for r in self.rails:  # ingest rails: staging buffers on the other GPUs
            for p in r["staging"]:
                try:
                    r["backend"].free(p)
                except Exception:
                    pass
        self.rails, self._active_rails = [], []
        try:
            self.backend.close()
        except Exception:
            pass
        if isinstance(self.session.get("shm"), str):  # this worker attached the control segment itself
            self.health = self.ring = self.results = self.flags_host = None
            self.rings = []
            self.region.close()


class Node:
    """Reference-compatible shell: `Node().run()` (reference src/node.py:181,210-211).

    Default (`listen=False`): this process becomes the worker for one GPU of a torchrun-launched deployment.  Like
    the reference's worker it serves one configuration after another -- here: one DEFER session after another, each
    with its own partition table and micro-batch size -- and returns when the dispatcher (rank 0) calls
    `dispatcher.end_sessions()`.

    `listen=True`: the worker speaks the REFERENCE's wire protocol (wire.py) on its ports -- configuration on 6001,
    data on 6000, results to the dispatcher's 6003 -- so that a B200 stage can sit behind an unmodified remote
    dispatcher: a configuration message (model JSON, partition index, weights) installs a partition as
    one GPU stage, every data message is run through it and the result sent back.  `run()` then blocks until
    `stop()`."""

    def __init__(self, model=None, device: Optional[int] = None, backend_factory=None, listen: bool = False,
                 data_port: Optional[int] = None, config_port: Optional[int] = None, dispatcher_addr=None,
                 predictor_factory=None, codec=None, mode: str = "bf16", max_batch: int = 32, chunk_size: int = 512 * 1000,
                 bind_host: str = "0.0.0.0"):
        self.model = model
        self.device = device
        self.backend_factory = backend_factory
        self.sessions_served = 0
        # ---- wire mode
        from . import wire

        self.listen = listen
        self.data_port = wire.DATA_PORT if data_port is None else int(data_port)
        self.config_port = wire.CONFIG_PORT if config_port is None else int(config_port)
        self.dispatcher_addr = dispatcher_addr      # (host, port) results go to; default: config peer, port 6003
        self.predictor_factory = predictor_factory  # (model, start, end) -> object with .run(x) / .close(); default: one GPU stage
        self.codec = codec or wire.default_codec()
        self.mode, self.max_batch, self.chunk_size, self.bind_host = mode, int(max_batch), int(chunk_size), bind_host
        self.node_state = NodeState(chunk_size=chunk_size)
        self.jobs: Dict[int, object] = {}           # partition index -> predictor (the reference's job_queue entry)
        self._jobs_lock = threading.Lock()
        self._wire_stop = threading.Event()
        self._servers: List[object] = []
        self.ready = threading.Event()
        self.served = 0
        self.errors: List[str] = []

    def run(self) -> None:
        if self.listen:
            return self._serve_wire()
        from .dispatcher import join_session_as_worker

        while join_session_as_worker(self.model, self.device, self.backend_factory):
            self.sessions_served += 1

    # ------------------------------------------------------------------ wire mode
    def _make_predictor(self, start: str, end: str):
        if self.predictor_factory is not None:
            return self.predictor_factory(self.model, start, end)
        from .executor import LocalPredictor

        return LocalPredictor(self.model, mode=self.mode, device=self.device or 0, max_batch=self.max_batch, start=start, end=end)

    def _listen_socket(self, port: int):
        import socket

        srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        srv.bind((self.bind_host, port))
        srv.listen(5)
        srv.settimeout(0.1)
        self._servers.append(srv)
        return srv

    def _config_server(self, srv) -> None:
        """Configuration port (reference src/node.py:65-98): JSON + index + weights -> load the partition -> ACK."""
        import socket

        from . import wire
        from .dag_util import construct_model

        while not self._wire_stop.is_set():
            try:
                cli, peer = srv.accept()
            except socket.timeout:
                continue
            except OSError:
                return
            try:
                cli.setblocking(True)
                model_json, index, weights = wire.recv_full_configuration(cli, self.chunk_size, self.codec)
                order = [l.name for l in self.model.layers]
                start, end = wire.partition_bounds_from_json(model_json, order)
                part = construct_model(self.model, start, end, f"part{index}")
                # the shipped arrays are `part.get_weights()`: per layer, in model order (src/dispatcher.py:243)
                it = iter(weights)
                for l in part.layers:
                    n = len(l.get_weights())
                    if n:
                        l.set_weights([np.asarray(next(it)) for _ in range(n)])
                if next(it, None) is not None:
                    raise ValueError("more weight arrays than the partition's layers hold")
                self.model._version += 1
                pred = self._make_predictor(start, end)
                with self._jobs_lock:
                    old = self.jobs.pop(index, None)
                    self.jobs[index] = pred
                if old is not None:
                    old.close()
                if self.dispatcher_addr is None:
                    self.dispatcher_addr = (peer[0], wire.RESULT_PORT)
                self.node_state.state = StateEnum.BUSY
                cli.sendall(wire.ACK)
                time.sleep(0.05)  # (as the reference: let the ACK leave before closing)
            except Exception as e:  # noqa: BLE001  (the reference prints and goes back to listening)
                self.errors.append(f"config: {type(e).__name__}: {e}")
            finally:
                cli.close()

    def _data_server(self, srv) -> None:
        """Data port (reference src/node.py:127-179): index + payload -> predict -> result to the dispatcher."""
        import socket

        from . import wire

        while not self._wire_stop.is_set():
            try:
                cli, _ = srv.accept()
            except socket.timeout:
                continue
            except OSError:
                return
            try:
                cli.setblocking(True)
                msg = wire.recv_data_message(cli, self.chunk_size)
                if msg is None:
                    continue
                index, payload = msg
                with self._jobs_lock:
                    pred = self.jobs.get(index)
                if pred is None:
                    raise KeyError(f"data for partition {index}, which was never configured")
                out = pred.run(self.codec.decode(payload))
                # [inferred] the result message mirrors the data message: index, framed payload, SHUT_WR
                wire.forward_data_to_worker(self.dispatcher_addr, self.codec.encode(out), index, self.chunk_size)
                self.served += 1
            except Exception as e:  # noqa: BLE001
                self.errors.append(f"data: {type(e).__name__}: {e}")
            finally:
                cli.close()

    def _serve_wire(self) -> None:
        if self.model is None:
            raise ValueError("Node(listen=True) needs the model whose partitions it will be sent")
        threads = [threading.Thread(target=self._config_server, args=(self._listen_socket(self.config_port),), daemon=True),
                   threading.Thread(target=self._data_server, args=(self._listen_socket(self.data_port),), daemon=True)]
        for t in threads:
            t.start()
        self.ready.set()
        while not self._wire_stop.is_set():
            time.sleep(0.02)
        for t in threads:
            t.join(timeout=2)
        for srv in self._servers:
            srv.close()
        with self._jobs_lock:
            for p in self.jobs.values():
                p.close()
            self.jobs = {}

    def stop(self) -> None:
        self._wire_stop.set()
