"""TEST-ONLY stand-in for the device backend so that the dispatcher / worker protocol (ordering,
back-pressure, fail-over, replay, exactly-once delivery) can be exercised on a machine without a
GPU.  "Device memory" is host memory, streams are synchronous, and the stage compute is the CPU
oracle -- which is why this file lives under tests/ and is never importable from the product."""
import ctypes
import time

import numpy as np

from _util import KO, oracle_for

_KEEP = []


def _u32(addr):
    return ctypes.c_uint32.from_address(addr)


class HostExecutor:
    def __init__(self, model, part, max_batch, delay=0.0, input_format="float32"):
        self.input_format = input_format
        self.oracle = oracle_for(model)
        self.model = model
        self.max_batch = max_batch
        self.delay = delay
        self.io = {}
        self.calls = 0
        self.reslice(part)

    def reslice(self, part):
        self.part = part
        self.in_shape = tuple(part.input_shape[1:])
        self.out_shape = tuple(part.output_shape[1:])
        self.io = {}

    @property
    def num_launches(self):
        return 1

    @property
    def out_bytes_per_image(self):
        return int(np.prod(self.out_shape)) * 4

    def bind_io(self, slot, in_ptr, out_ptr):
        self.io[slot] = (in_ptr, out_ptr)

    def enqueue(self, slot, batch, stream=0):
        in_ptr, out_ptr = self.io[slot]
        n_in = batch * int(np.prod(self.in_shape))
        n_out = batch * int(np.prod(self.out_shape))
        if self.input_format == "uint8" and self.part.takes_model_input:
            # raw RGB uint8 image: Keras caffe-mode preprocess_input, then the same oracle
            raw = np.ctypeslib.as_array((ctypes.c_uint8 * n_in).from_address(in_ptr)).reshape((batch,) + self.in_shape)
            x = raw.astype(np.float32)[..., ::-1] - np.array([103.939, 116.779, 123.68], dtype=np.float32)
        else:
            x = np.ctypeslib.as_array((ctypes.c_float * n_in).from_address(in_ptr)).reshape((batch,) + self.in_shape).copy()
        y, _ = self.oracle.run_range(x, self.part.start, self.part.end)
        if self.delay:
            time.sleep(self.delay)
        np.ctypeslib.as_array((ctypes.c_float * n_out).from_address(out_ptr))[:] = np.ascontiguousarray(y, np.float32).ravel()
        self.calls += 1

    def close(self):
        pass


class HostBackend:
    name = "host-test"

    def __init__(self, device, delay=0.0):
        self.device = device
        self.delay = delay
        self.abort_check = lambda: False

    def alloc(self, nbytes):
        a = np.zeros(int(nbytes) + 64, dtype=np.uint8)
        _KEEP.append(a)
        return (a.ctypes.data + 63) // 64 * 64

    def free(self, ptr):
        pass

    def export(self, ptr):
        return ptr

    def open(self, token):
        return int(token)

    def peer_enable(self, other):
        pass

    def register_host(self, addr, nbytes):
        return addr

    def new_stream(self):
        return 0

    def sync(self, stream):
        pass

    def busy(self, stream):
        return False

    def write_flag(self, stream, addr, value):
        _u32(addr).value = value & 0xFFFFFFFF

    def wait_flag(self, stream, addr, value):
        t0 = time.monotonic()
        while ((_u32(addr).value - value) & 0xFFFFFFFF) >= 0x80000000:   # cyclic "not yet >="
            if self.abort_check():
                return
            if time.monotonic() - t0 > 60:
                raise TimeoutError("flag never arrived")
            time.sleep(50e-6)

    def h2d(self, dst, src, nbytes, stream):
        ctypes.memmove(dst, src, int(nbytes))

    d2h = h2d
    d2d = h2d

    def memset(self, dst, value, nbytes, stream):
        ctypes.memset(dst, value, int(nbytes))

    def make_executor(self, model, part, mode, max_batch, input_format="float32"):
        # cut tensors move as fp32 here; the session's hand-off slots are sized for its mode's element size
        if mode not in ("fp32", "fp32-check"):
            raise ValueError("the host test backend moves fp32 activations: run the session with mode='fp32'")
        return HostExecutor(model, part, max_batch, self.delay, input_format)

    def close(self):
        pass


class SharedHostBackend(HostBackend):
    """Same, but "device memory" comes from named shared memory so that it can be mapped by
    another process (stands in for CUDA IPC in the world_size-2 gloo tests)."""

    def __init__(self, device, delay=0.0):
        super().__init__(device, delay)
        self._mine = {}
        self._opened = []

    def alloc(self, nbytes):
        from multiprocessing import shared_memory

        shm = shared_memory.SharedMemory(create=True, size=int(nbytes) + 64)
        arr = np.frombuffer(shm.buf, dtype=np.uint8)
        arr[:] = 0
        addr = arr.ctypes.data
        self._mine[addr] = (shm, arr)
        return addr

    def export(self, ptr):
        return self._mine[ptr][0].name

    def open(self, token):
        from multiprocessing import resource_tracker, shared_memory

        shm = shared_memory.SharedMemory(name=token, create=False)
        try:
            resource_tracker.unregister(shm._name, "shared_memory")
        except Exception:
            pass
        arr = np.frombuffer(shm.buf, dtype=np.uint8)
        self._opened.append((shm, arr))
        return arr.ctypes.data

    def close(self):
        for shm, arr in self._mine.values():
            del arr
            try:
                shm.close()
            except BufferError:
                pass
            try:
                shm.unlink()
            except FileNotFoundError:
                pass
        self._mine = {}
