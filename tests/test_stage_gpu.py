"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle.

Tolerances (north_star): bf16 mode -- identical arg-max and |dprob| <= 1e-2; fp32 check mode --
|dprob| <= 1e-4.  Cut tensors in bf16 mode: max error <= 2% of the tensor's max magnitude
(bf16 has 8 mantissa bits; errors accumulate over up to 53 layers).
"""
import ctypes as C

import numpy as np
import pytest

from _util import KO, oracle_for, rel_err

pytestmark = pytest.mark.gpu

PROB_TOL_BF16 = 1e-2
PROB_TOL_FP32 = 1e-4
CUT_TOL_BF16 = 2e-2


@pytest.fixture(scope="module")
def nat(pkg):
    from adapt_b200 import _native

    _native.require_gpu()
    return _native


@pytest.fixture(scope="module")
def oracle(model50):
    return oracle_for(model50)


def _predictor(model, **kw):
    from adapt_b200.executor import LocalPredictor

    return LocalPredictor(model, **kw)


# ------------------------------------------------------------------ single-layer kernel parity
CONV_CASES = [
    # B, H, W, Cin, Cout, k, stride, pad, residual, relu, a_mode
    (2, 14, 14, 64, 64, 1, 1, 0, False, False, 0),
    (4, 28, 28, 256, 128, 1, 1, 0, True, True, 0),
    (3, 56, 56, 64, 256, 1, 1, 0, True, True, 0),
    (2, 14, 14, 64, 64, 3, 1, 1, False, True, 1),
    (5, 7, 7, 512, 512, 3, 1, 1, False, True, 1),      # tiles straddle images, ragged last tile
    (1, 7, 7, 64, 64, 3, 1, 1, False, False, 1),       # single partial tile
    (2, 56, 56, 256, 128, 1, 2, 0, False, True, 1),    # strided 1x1 (convN_block1_1_conv)
    (3, 28, 28, 512, 1024, 1, 2, 0, False, False, 1),  # strided shortcut (convN_block1_0_conv)
    (2, 224, 224, 8, 64, 7, 2, 3, False, True, 2),     # stem, channel-padded input
    (2, 20, 20, 24, 64, 3, 1, 1, True, True, 2),       # software gather, Cin not a multiple of 64
    (2, 14, 14, 64, 64, 3, 1, 1, False, True, 2),      # gather vs im2col on the same case
    (3, 28, 28, 128, 128, 3, 1, 1, False, True, 3),    # row reuse (A_ROW3) on CTA pairs, rows straddle tiles
    (6, 56, 56, 64, 64, 3, 1, 1, False, True, 3),      # row reuse, 64-wide tiles, resident B (> 148 tiles)
    (5, 14, 14, 256, 256, 3, 1, 1, False, False, 3),   # row reuse on a narrow map, ragged last tile
    (1, 7, 7, 512, 512, 3, 1, 1, False, True, 3),      # row reuse, one image smaller than a tile
    (16, 14, 14, 256, 256, 3, 1, 1, False, True, -1),  # auto: tensor-bound 3x3 -> im2col on CTA pairs
    (8, 28, 28, 128, 512, 1, 1, 0, True, True, -1),    # auto: 1x1 + residual, slabs used in place
    (7, 56, 56, 256, 64, 1, 1, 0, False, True, -1),    # auto: single n-tile, resident B (172 tiles)
    (3, 20, 24, 16, 64, 4, 1, 0, False, True, 4),      # stem by column phase (A_S2DP): tiles cross rows and images
    (2, 115, 116, 16, 64, 4, 1, 0, False, True, -1),   # auto: the space-to-depth stem geometry -> A_S2DP
    (2, 9, 8, 16, 64, 4, 1, 0, False, False, 4),       # A_S2DP, rows narrower than four units
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("block_n", [0, 64])
def test_conv_kernel_matches_oracle(nat, case, block_n):
    from adapt_b200.executor import DeviceBuffer, unwire_activation, wire_activation

    B, H, W, Cin, Cout, k, stride, pad, use_res, relu, a_mode = case
    lib = nat.lib()
    rng = np.random.default_rng(hash(case) % (2 ** 31))
    x = KO.bf16_round(rng.standard_normal((B, H, W, Cin)).astype(np.float32))
    w = KO.bf16_round((rng.standard_normal((k, k, Cin, Cout)) / np.sqrt(k * k * Cin)).astype(np.float32))
    scale = rng.uniform(0.5, 1.5, Cout).astype(np.float32)
    bias = rng.standard_normal(Cout).astype(np.float32)
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    res = KO.bf16_round(rng.standard_normal((B, Ho, Wo, Cout)).astype(np.float32)) if use_res else None
    ref = KO.conv2d(KO.zero_pad(x, pad), w, np.zeros(Cout, np.float32), stride, "valid") * scale + bias
    if use_res:
        ref = ref + res
    if relu:
        ref = np.maximum(ref, 0)

    def up(a):
        buf = DeviceBuffer(0, a.nbytes)
        nat.check(lib.adapt_memcpy_h2d(0, buf.ptr, a.ctypes.data, a.nbytes, None))
        return buf

    w_ohwi = np.ascontiguousarray(w.transpose(3, 0, 1, 2))
    bufs = [up(wire_activation(x, 0)), up(wire_activation(w_ohwi, 0)), up(scale), up(bias)]
    d_res = up(wire_activation(res, 0)) if use_res else None
    d_out = DeviceBuffer(0, B * Ho * Wo * Cout * 2)
    nat.check(lib.adapt_conv2d_bf16(0, bufs[0].ptr, bufs[1].ptr, bufs[2].ptr, bufs[3].ptr, d_res.ptr if use_res else None,
                                    d_out.ptr, 0, B, H, W, Cin, Cout, k, k, stride, pad, int(relu), a_mode, block_n, None))
    raw = np.empty(B * Ho * Wo * Cout * 2, np.uint8)
    nat.check(lib.adapt_memcpy_d2h(0, raw.ctypes.data, d_out.ptr, raw.nbytes, None))
    nat.check(lib.adapt_stream_sync(0, None))
    got = unwire_activation(raw, (B, Ho, Wo, Cout), 0)
    for b in bufs + [d_out] + ([d_res] if use_res else []):
        b.free()
    # inputs are exact bf16 values, accumulation is fp32 on both sides: the only difference is the
    # final rounding of the output to bf16 (<= 2^-8 relative) plus fp32 summation order.
    np.testing.assert_allclose(got, ref, rtol=2 ** -7, atol=2e-3)


# ------------------------------------------------------------------ chained 3x3 -> 1x1 + shortcut block tail
CHAIN_CASES = [
    # B, H, W, C
    (2, 56, 56, 64),      # conv2 geometry: single CTA, resident 3x3 weights
    (6, 56, 56, 64),      # more tiles than SMs: several units per CTA, residual ring across units
    (1, 28, 28, 64),      # fewer tiles than SMs
    (3, 28, 28, 128),     # conv3 geometry: CTA pairs, streamed weights
    (16, 28, 28, 128),    # several units per pair
    (1, 28, 28, 128),     # odd number of tiles: the last pair hangs over the end
    (2, 29, 34, 128),     # rows / images that do not divide the tile
    (3, 31, 30, 64),
]


@pytest.mark.parametrize("case", CHAIN_CASES)
def test_conv_chain_kernel_matches_oracle(nat, case):
    """`adapt_conv_chain_bf16` (convN_blockB_2_conv .. convN_blockB_out in one kernel) against the oracle's layer math
    with the intermediate rounded to bf16 where the two-launch form stores it."""
    from adapt_b200.executor import DeviceBuffer, unwire_activation, wire_activation

    B, H, W, C = case
    lib = nat.lib()
    rng = np.random.default_rng(hash(case) % (2 ** 31))
    x = KO.bf16_round(rng.standard_normal((B, H, W, C)).astype(np.float32))
    w2 = KO.bf16_round((rng.standard_normal((3, 3, C, C)) / np.sqrt(9 * C)).astype(np.float32))
    w3 = KO.bf16_round((rng.standard_normal((1, 1, C, 4 * C)) / np.sqrt(C)).astype(np.float32))
    s1, b1 = rng.uniform(0.5, 1.5, C).astype(np.float32), rng.standard_normal(C).astype(np.float32)
    s2, b2 = rng.uniform(0.5, 1.5, 4 * C).astype(np.float32), rng.standard_normal(4 * C).astype(np.float32)
    res = KO.bf16_round(rng.standard_normal((B, H, W, 4 * C)).astype(np.float32))
    mid = np.maximum(KO.conv2d(KO.zero_pad(x, 1), w2, np.zeros(C, np.float32), 1, "valid") * s1 + b1, 0)
    mid = KO.bf16_round(mid.astype(np.float32))
    ref = np.maximum(KO.conv2d(mid, w3, np.zeros(4 * C, np.float32), 1, "valid") * s2 + b2 + res, 0)

    def up(a):
        buf = DeviceBuffer(0, a.nbytes)
        nat.check(lib.adapt_memcpy_h2d(0, buf.ptr, a.ctypes.data, a.nbytes, None))
        return buf

    bufs = [up(wire_activation(x, 0)), up(wire_activation(np.ascontiguousarray(w2.transpose(3, 0, 1, 2)), 0)), up(s1), up(b1),
            up(wire_activation(np.ascontiguousarray(w3.transpose(3, 0, 1, 2)), 0)), up(s2), up(b2), up(wire_activation(res, 0))]
    d_out = DeviceBuffer(0, B * H * W * 4 * C * 2)
    nat.check(lib.adapt_memset(0, d_out.ptr, 0xFF, B * H * W * 4 * C * 2, None))   # NaN pattern: unwritten pixels show up
    nat.check(lib.adapt_conv_chain_bf16(0, *[b.ptr for b in bufs], d_out.ptr, B, H, W, C, None))
    raw = np.empty(B * H * W * 4 * C * 2, np.uint8)
    nat.check(lib.adapt_memcpy_d2h(0, raw.ctypes.data, d_out.ptr, raw.nbytes, None))
    nat.check(lib.adapt_stream_sync(0, None))
    got = unwire_activation(raw, (B, H, W, 4 * C), 0)
    for b in bufs + [d_out]:
        b.free()
    assert np.isfinite(got).all(), "pixels the kernel never wrote"
    # the intermediate is rounded at the same point on both sides, but a value within half a bf16 ulp of a rounding
    # boundary can land on either side (fp32 summation order): allow a few such flips through the 1x1
    np.testing.assert_allclose(got, ref, rtol=2 ** -6, atol=2e-2)
    assert np.abs(got - ref).mean() <= 2e-3


def test_chained_and_unchained_programs_agree(model50, monkeypatch):
    """The stage program with CONV_CHAIN ops (default) against the same layers as separate 3x3 and 1x1 launches
    (ADAPT_CHAIN=0): same rounding points, so cut tensors agree to bf16 re-rounding and the probabilities to 1e-3
    relative."""
    from adapt_b200 import _native as N

    x = KO.synthetic_images(3, seed=77, kind="normal")
    outs = {}
    for chain in ("1", "0"):
        monkeypatch.setenv("ADAPT_CHAIN", chain)
        p = _predictor(model50, mode="bf16", max_batch=3, end="conv3_block4_out")
        n_chain = sum(o.kind == N.OP_CONV_CHAIN for o in p.ex.program.ops)
        assert n_chain == (7 if chain == "1" else 0)
        q = _predictor(model50, mode="bf16", max_batch=3)
        outs[chain] = (p.run(x), q.run(x))
        p.close()
        q.close()
    assert rel_err(outs["1"][0], outs["0"][0]) <= 2e-2
    assert rel_err(outs["1"][1], outs["0"][1]) <= 1e-3


@pytest.mark.parametrize("batch,kind", [(1, "caffe"), (3, "normal"), (37, "caffe"), (128, "caffe")])
def test_fused_stem_pool_is_bit_identical_to_conv_then_pool(model50, monkeypatch, batch, kind):
    """`stem_pool_kernel` (conv1_conv .. pool1_pool in one launch, default) against the stem convolution and the max-pool
    as two launches (ADAPT_STEM_POOL=0): max commutes with ReLU and with rounding to bf16, so `pool1_pool` must agree bit
    for bit -- for one image, for ranges of tiles that start inside an image (37 images over 148 CTAs: 7 tiles per CTA,
    every range but the first few needs the row handed over from the tile before it) and for the bench's batch."""
    from adapt_b200 import _native as N

    x = KO.synthetic_images(batch, seed=5, kind=kind)
    outs = {}
    for fused in ("1", "0"):
        monkeypatch.setenv("ADAPT_STEM_POOL", fused)
        p = _predictor(model50, mode="bf16", max_batch=batch, end="pool1_pool")
        kinds = [o.kind for o in p.ex.program.ops]
        assert (N.OP_STEM_POOL in kinds) == (fused == "1") and (N.OP_MAXPOOL in kinds) == (fused == "0")
        outs[fused] = p.run(x)
        p.close()
    assert outs["1"].shape == (batch, 56, 56, 64)
    assert np.array_equal(outs["1"], outs["0"])
    assert float(outs["1"].max()) > 0.0


# ------------------------------------------------------------------ whole network
@pytest.mark.parametrize("kind,batch", [("caffe", 4), ("normal", 3), ("normal", 1)])
def test_full_model_bf16(model50, oracle, kind, batch):
    x = KO.synthetic_images(batch, seed=21, kind=kind)
    ref, _ = oracle.predict(x)
    ref_logits = oracle.logits.copy()
    p = _predictor(model50, mode="bf16", max_batch=4)
    got = p.run(x)
    logits = p.logits(batch)
    p.close()
    assert got.shape == (batch, 1000) and got.dtype == np.float32
    np.testing.assert_allclose(got.sum(-1), 1.0, atol=1e-4)
    assert np.abs(got - ref).max() <= PROB_TOL_BF16
    assert got.argmax(-1).tolist() == ref.argmax(-1).tolist()
    assert rel_err(logits, ref_logits) <= 3e-2   # (measured ~4e-3: bf16 storage between 53 layers)
    if kind == "normal":
        # N(0,1) inputs keep logits O(0.1): the 1e-2 bound can be stated on logits too
        assert np.abs(logits - ref_logits).max() <= PROB_TOL_BF16
        top5 = lambda a: [set(r) for r in np.argsort(a, -1)[:, -5:]]  # noqa: E731
        assert sum(len(a & b) for a, b in zip(top5(logits), top5(ref_logits))) >= 5 * batch - batch


def test_full_model_fp32_check_mode(model50, oracle):
    x = KO.synthetic_images(2, seed=22, kind="caffe")
    ref, _ = oracle.predict(x)
    ref_logits = oracle.logits.copy()
    p = _predictor(model50, mode="fp32", max_batch=2)
    got = p.run(x)
    logits = p.logits(2)
    p.close()
    assert np.abs(got - ref).max() <= PROB_TOL_FP32
    assert got.argmax(-1).tolist() == ref.argmax(-1).tolist()
    assert rel_err(logits, ref_logits) <= 1e-4


def test_randomized_batchnorm_and_biases(pkg, oracle):
    m = pkg.ResNet50(weights=None, seed=99)
    pkg.randomize_batchnorm(m, seed=5)
    o = oracle_for(m)
    x = KO.synthetic_images(2, seed=23, kind="normal")
    ref, _ = o.predict(x)
    for mode, tol in (("bf16", PROB_TOL_BF16), ("fp32", PROB_TOL_FP32)):
        p = _predictor(m, mode=mode, max_batch=2)
        got = p.run(x)
        p.close()
        assert np.abs(got - ref).max() <= tol, mode
        assert got.argmax(-1).tolist() == ref.argmax(-1).tolist()


# ------------------------------------------------------------------ partitions (part_at)
PARTS = [
    ["conv3_block4_out"],                                               # BASELINE config 2
    ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"],       # BASELINE config 3
    ["conv2_block2_out", "conv3_block1_out", "conv3_block3_out", "conv4_block1_out", "conv4_block3_out",
     "conv4_block5_out", "conv5_block1_out"],                           # 8-stage suggestion
    ["conv1_pad", "conv1_conv", "conv1_bn", "conv1_relu", "pool1_pad", "pool1_pool"],   # every stem cut
    ["conv2_block1_add", "conv4_block2_add", "avg_pool"],               # cuts at Add outputs and the head
]


@pytest.mark.parametrize("part_at", PARTS)
@pytest.mark.parametrize("mode", ["bf16", "fp32"])
def test_partition_chain_matches_oracle_and_unpartitioned(model50, oracle, part_at, mode):
    batch = 2
    x = KO.synthetic_images(batch, seed=31, kind="normal")
    full = _predictor(model50, mode=mode, max_batch=batch)
    want = full.run(x)
    full.close()
    bounds = ["input_1"] + list(part_at) + ["predictions"]
    cur, cur_ref = x, x
    for a, b in zip(bounds[:-1], bounds[1:]):
        p = _predictor(model50, mode=mode, max_batch=batch, start=a, end=b)
        cur = p.run(cur)
        p.close()
        cur_ref, _ = oracle.run_range(cur_ref, a, b)
        assert cur.shape == cur_ref.shape
        tol = CUT_TOL_BF16 if mode == "bf16" else 1e-4
        assert rel_err(cur, cur_ref) <= tol, (a, b)
    if all(c.endswith("_out") or c in ("pool1_pool", "avg_pool") for c in part_at):
        # cuts on fused-group boundaries: same kernels, same order, same inputs -> bit for bit
        np.testing.assert_array_equal(cur, want)
    else:
        # a cut inside a fused group (conv | bn | relu, add | relu) moves a rounding point
        assert np.abs(cur - want).max() <= (PROB_TOL_BF16 if mode == "bf16" else PROB_TOL_FP32)


def test_invalid_cut_is_rejected(model50):
    from adapt_b200.dag_util import construct_model

    with pytest.raises(RuntimeError):
        construct_model(model50, "conv2_block1_1_relu", "predictions")
    with pytest.raises(ValueError):
        construct_model(model50, "no_such_layer", "predictions")


def test_batch_sizes_and_repeatability(model50, oracle):
    x = KO.synthetic_images(7, seed=41, kind="normal")
    ref, _ = oracle.predict(x)
    p = _predictor(model50, mode="bf16", max_batch=5)   # 7 = 5 + ragged 2
    a = p.run(x)
    b = p.run(x)
    one = p.run(x[3:4])
    p.close()
    np.testing.assert_array_equal(a, b)
    np.testing.assert_array_equal(a[3:4], one)          # batch-size independent results
    assert np.abs(a - ref).max() <= PROB_TOL_BF16


def test_bad_arguments_fail_loudly(model50, nat):
    p = _predictor(model50, mode="bf16", max_batch=2)
    with pytest.raises(nat.AdaptError):
        p.ex.enqueue(0, 3, 0)            # batch > max_batch
    with pytest.raises(nat.AdaptError):
        p.ex.enqueue(1, 1, 0)            # slot 1 never bound
    p.close()


def test_resnet101_shares_every_kernel(pkg):
    """a deeper graph from the same builder runs on the same kernels (23 conv4 blocks)."""
    import adapt_b200

    m = adapt_b200.ResNet101(weights=None)
    o = oracle_for(m)
    x = KO.synthetic_images(2, seed=31, kind="normal")
    ref = o.predict(x)[0]
    got = _predictor(m, mode="bf16", max_batch=2).predict(x)
    assert np.abs(got - ref).max() <= PROB_TOL_BF16
    assert got.argmax(-1).tolist() == ref.argmax(-1).tolist()
    got32 = _predictor(m, mode="fp32", max_batch=2).predict(x)
    assert np.abs(got32 - ref).max() <= PROB_TOL_FP32


@pytest.mark.parametrize("end", [None, "conv1_relu", "conv1_pad"])
def test_uint8_ingest_matches_host_preprocessing(pkg, model50, oracle, end):
    """Raw uint8 RGB images with the caffe preprocessing fused into the ingest kernels (space-to-depth stem,
    plain 3 -> 8 channel ingest when a cut separates the stem, fp32 check mode) against the oracle fed with
    host-preprocessed fp32 images -- and bit-identical to the fp32 input path, which sees the same fp32 values."""
    from adapt_b200 import preprocess_input

    rng = np.random.default_rng(11)
    x8 = rng.integers(0, 256, (3, 224, 224, 3), dtype=np.uint8)
    xf = preprocess_input(x8)
    for mode, tol in (("bf16", PROB_TOL_BF16), ("fp32", PROB_TOL_FP32)):
        p8 = _predictor(model50, mode=mode, max_batch=3, end=end, input_format="uint8")
        pf = _predictor(model50, mode=mode, max_batch=3, end=end)
        got8, gotf = p8.run(x8), pf.run(xf)
        np.testing.assert_array_equal(got8, gotf)
        if end is None:
            ref = oracle.predict(xf)[0]
            assert np.abs(got8 - ref).max() <= tol
            assert got8.argmax(-1).tolist() == ref.argmax(-1).tolist()
        p8.close()
        pf.close()


def test_cuda_path_against_the_committed_fixture(model50):
    """The CUDA path against tests/golden/resnet50_seed1234_tf_crosscheck.npz -- the committed vectors a maintainer
    with TensorFlow checks Keras against (tools/tf_check.py): probabilities, logits and the cut tensors BASELINE's
    partitions hand off, in fp32 check mode (1e-4) and bf16 mode (cut tensors 2 %, logits 3e-2 relative)."""
    import os
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "tools"))
    from export_for_tf import make_inputs

    g = np.load(os.path.join(root, "tests", "golden", "resnet50_seed1234_tf_crosscheck.npz"))
    x = make_inputs()
    stride = int(g["tap_stride"])
    for mode, ptol, ltol, ttol in (("fp32", PROB_TOL_FP32, 1e-4, 1e-4), ("bf16", PROB_TOL_BF16, 3e-2, CUT_TOL_BF16)):
        p = _predictor(model50, mode=mode, max_batch=2)
        probs = p.run(x)
        logits = p.logits(2)
        p.close()
        assert np.abs(probs - g["probs"]).max() <= ptol, mode
        assert probs.argmax(-1).tolist() == g["probs"].argmax(-1).tolist(), mode
        for i in range(2):
            assert rel_err(logits[i], g["logits"][i]) <= ltol, (mode, i)
        for name in [str(n) for n in g["tap_names"]]:
            q = _predictor(model50, mode=mode, max_batch=2, end=name)
            t = q.run(x)
            q.close()
            assert tuple(t.shape[1:]) == tuple(int(v) for v in g["tapshape_" + name])
            s = t.reshape(2, -1)[:, ::stride]
            for i in range(2):
                assert rel_err(s[i], g["tap_" + name][i]) <= ttol, (mode, name, i)
