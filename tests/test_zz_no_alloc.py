"""GPU: the allocation-free classifier head (`adapt_dense_softmax_noalloc`) against the existing stage
entry points.

Some users of this library cannot allocate device memory at the moment they need a prediction: their
process gave all HBM to its own arenas at start-up, or sits in a loop where an allocation's device-wide
synchronisation is unacceptable, or is probing a GPU that is out of memory.  For them the Dense+softmax
head has a variant whose only intermediate (the fp32 logits) lives in a fixed array inside the library.
This test drives that variant through the C ABI twice in a row -- the second call re-uses the array that
still holds the first call's logits -- and requires both results to match what the ordinary stage API
(`adapt_stage_create` / `adapt_stage_enqueue`, whose logits buffer is allocated) produces for the same
layer range (avg_pool, predictions].  Both paths see the same bf16 operands and accumulate in fp32, so they may
differ only by summation order: the probabilities must agree to a relative 1e-4 (with the oracle's arithmetic on
the bf16-rounded operands as well), far inside the 1e-2 bf16-mode bound used elsewhere.  The second call is also
checked for the property in the entry point's name: free device memory is the same before and after it.

The file name sorts last on purpose: with `pytest -x`, every other test has run before this one.
"""
import numpy as np
import pytest

from _util import KO

pytestmark = pytest.mark.gpu
PROB_RTOL = 1e-4   # same operands, fp32 accumulation on both sides: summation order only


def test_noalloc_head_matches_stage_api_twice_in_a_row(pkg, model50):
    torch = pytest.importorskip("torch")
    from adapt_b200 import _native as nat
    from adapt_b200.executor import DeviceBuffer, LocalPredictor, wire_activation

    nat.require_gpu()
    lib = nat.lib()
    torch.cuda.mem_get_info(0)                               # (creates torch's view of the context up front)
    dense = model50.get_layer("predictions")
    kernel, bias = dense.get_weights()                       # (2048, 1000), (1000,)
    cin, cout = kernel.shape
    w_bits = wire_activation(np.ascontiguousarray(kernel.T), nat.MODE_BF16)     # [cout][cin] bf16 bit patterns
    scale = np.ones(cout, np.float32)

    def up(a):
        buf = DeviceBuffer(0, a.nbytes)
        nat.check(lib.adapt_memcpy_h2d(0, buf.ptr, a.ctypes.data, a.nbytes, None))
        return buf

    d_w, d_scale, d_bias = up(w_bits), up(scale), up(np.ascontiguousarray(bias, np.float32))
    reference = LocalPredictor(model50, mode="bf16", max_batch=8, start="avg_pool", end="predictions")
    rng = np.random.default_rng(2024)
    try:
        for call, batch in enumerate((5, 3)):                # second call: smaller batch over the stale scratch
            x = (rng.standard_normal((batch, cin)) * (1.0 + call)).astype(np.float32)
            want = reference.run(x)                          # existing entry points, allocated logits buffer
            x_bits = wire_activation(x, nat.MODE_BF16)
            d_x = up(x_bits)
            d_out = DeviceBuffer(0, batch * cout * 4)
            nat.check(lib.adapt_stream_sync(0, None))
            free_before = torch.cuda.mem_get_info(0)[0]
            nat.check(lib.adapt_dense_softmax_noalloc(0, d_x.ptr, d_w.ptr, d_scale.ptr, d_bias.ptr, d_out.ptr, batch, cin,
                                                      cout, None))
            nat.check(lib.adapt_stream_sync(0, None))
            if call > 0:   # (the first call on a device may load the library's device code, scratch array included)
                assert torch.cuda.mem_get_info(0)[0] == free_before, "the call changed the amount of free device memory"
            got = np.empty((batch, cout), np.float32)
            nat.check(lib.adapt_memcpy_d2h(0, got.ctypes.data, d_out.ptr, got.nbytes, None))
            nat.check(lib.adapt_stream_sync(0, None))
            d_x.free()
            d_out.free()
            assert np.isfinite(got).all()
            np.testing.assert_allclose(got.sum(-1), 1.0, atol=1e-4)
            np.testing.assert_allclose(got, want, rtol=PROB_RTOL, atol=1e-9, err_msg=f"call {call}")
            assert got.argmax(-1).tolist() == want.argmax(-1).tolist(), f"call {call}"
            ref = KO.softmax(KO.bf16_round(x) @ KO.bf16_round(kernel) + bias)   # and against the oracle's math
            np.testing.assert_allclose(got, ref, rtol=PROB_RTOL, atol=1e-9, err_msg=f"call {call} vs oracle math")
    finally:
        reference.close()
        for b in (d_w, d_scale, d_bias):
            b.free()


def test_noalloc_head_rejects_sizes_beyond_its_scratch(pkg):
    from adapt_b200 import _native as nat
    from adapt_b200.executor import DeviceBuffer

    nat.require_gpu()
    lib = nat.lib()
    dummy = DeviceBuffer(0, 4096)
    too_many = nat.NOALLOC_MAX_LOGITS // 1000 + 1            # batch * 1000 logits would not fit
    rc = lib.adapt_dense_softmax_noalloc(0, dummy.ptr, dummy.ptr, dummy.ptr, dummy.ptr, dummy.ptr, too_many, 2048, 1000, None)
    dummy.free()
    assert rc == nat.EINVAL
    assert b"exceeds the fixed scratch" in lib.adapt_last_error()
    # a caller that carves its operands out of its own arena must keep them 16-byte aligned: refused, not faulted
    buf = DeviceBuffer(0, 1 << 16)
    rc = lib.adapt_dense_softmax_noalloc(0, buf.ptr, buf.ptr, buf.ptr + 4, buf.ptr, buf.ptr, 1, 64, 8, None)
    buf.free()
    assert rc == nat.EINVAL and b"16-byte aligned" in lib.adapt_last_error()
