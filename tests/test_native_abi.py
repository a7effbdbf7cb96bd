"""CPU: the C-ABI library loads without a GPU, exports every symbol include/adapt_b200.h declares, and
fails loudly (no CPU fallback) when a compute entry point is called without a device."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    text = open(os.path.join(ROOT, "include", "adapt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(adapt_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(pkg):
    from adapt_b200 import _native

    declared = _declared_functions()
    assert len(declared) >= 30
    assert sorted(_native.SYMBOLS) == declared


def test_library_loads_and_exports_every_declared_symbol(pkg):
    from adapt_b200 import _native

    lib = _native.lib()
    for name in _declared_functions():
        assert getattr(lib, name) is not None, name
    assert lib.adapt_version().startswith(b"adapt_b200")
    raw = ctypes.CDLL(_native.LIB_PATH)
    assert ctypes.cast(raw.adapt_stage_enqueue, ctypes.c_void_p).value


def test_struct_layouts_match_header(pkg):
    from adapt_b200 import _native

    assert ctypes.sizeof(_native.WeightDesc) == 3 * 8 + 4 * 4
    assert ctypes.sizeof(_native.TensorDesc) == 16
    assert ctypes.sizeof(_native.OpDesc) == 19 * 4
    hdr = open(os.path.join(ROOT, "include", "adapt_b200.h")).read()
    assert f"#define ADAPT_NOALLOC_MAX_LOGITS ({_native.NOALLOC_MAX_LOGITS // 1024} * 1024)" in hdr
    for name, val in (("ADAPT_OP_ZEROPAD", _native.OP_ZEROPAD), ("ADAPT_OP_CONV_CHAIN", _native.OP_CONV_CHAIN), ("ADAPT_OP_SOFTMAX", _native.OP_SOFTMAX),
                      ("ADAPT_MODE_FP32_CHECK", _native.MODE_FP32_CHECK), ("ADAPT_T_OUTPUT", _native.T_OUTPUT)):
        assert re.search(rf"#define {name} {val}\b", hdr), name


def test_no_cpu_fallback(pkg, model50):
    from adapt_b200 import _native
    from adapt_b200.executor import LocalPredictor

    if _native.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(_native.AdaptError):
        LocalPredictor(model50)
    out = ctypes.c_void_p()
    assert _native.lib().adapt_device_alloc(0, 16, ctypes.byref(out)) == _native.ENODEV
    assert b"no CPU fallback" in _native.lib().adapt_last_error()
    with pytest.raises(_native.AdaptError):
        model50.predict([[0.0]])


def test_heartbeat_needs_no_gpu(pkg):
    import time

    from adapt_b200 import _native
    from adapt_b200.fabric import HealthTable, ShmRegion, W_HEARTBEAT

    reg = ShmRegion(create=True, size=HealthTable.BYTES)
    try:
        h = HealthTable(reg, 0)
        addr = h.w[2:3, W_HEARTBEAT:W_HEARTBEAT + 1].ctypes.data
        handle = ctypes.c_void_p()
        assert _native.lib().adapt_heartbeat_start(addr, 500, ctypes.byref(handle)) == 0
        time.sleep(0.02)
        assert h.age_ms(2) < 15
        _native.lib().adapt_heartbeat_stop(handle)
        time.sleep(0.03)
        assert h.age_ms(2) > 20
    finally:
        reg.close()
