"""GPU: the reference-facing API (DEFER / part_at / queues) end to end on real B200s, checked
against the oracle.  Stages may share one GPU (computeNodes naming the same device): the
hand-off protocol is identical, the peer pointer just happens to be local."""
import os
import queue
import threading
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # several stages' streams on one GPU must not alias

import numpy as np  # noqa: E402
import pytest  # noqa: E402

from _util import KO, StreamChecker, distinct_images  # noqa: E402

pytestmark = pytest.mark.gpu


def _ngpu():
    from adapt_b200 import _native

    return _native.device_count()


def _run(defer, model, part_at, inputs, kill_after=None, kill_stage=None, timeout=40):
    in_q, out_q = queue.Queue(10), queue.Queue()
    t = threading.Thread(target=defer.run_defer, args=(model, part_at, in_q, out_q), daemon=True)
    t.start()
    defer.wait_ready(120)
    outs = []
    threading.Thread(target=lambda: [in_q.put(x) for x in inputs], daemon=True).start()
    t0, killed = time.time(), False
    while len(outs) < len(inputs):
        try:
            outs.append(out_q.get(timeout=0.2))
        except queue.Empty:
            assert defer.error is None, defer.error
            assert time.time() - t0 < timeout, f"timed out with {len(outs)}/{len(inputs)} results: {defer.debug_state()}"
        if kill_after is not None and not killed and len(outs) >= kill_after:
            defer.inject_kill(kill_stage)
            killed = True
    time.sleep(0.05)
    assert out_q.empty(), "duplicate delivery"
    stats = dict(defer.stats)
    launches = defer.gpu_launches
    defer.shutdown()
    t.join(30)
    return outs, stats, launches


def _check(model, part_at, inputs, outs, mode="bf16", oracle_every=1, **kw):
    """Every delivered micro-batch against the un-partitioned model on that very input (bit for bit on fused-group
    cuts) and against the oracle -- see _util.StreamChecker.  `part_at` may be a list of cut lists (a stream that
    lived through re-slicing): bit-exactness then needs every one of them on fused-group boundaries."""
    cuts = [c for p in part_at for c in p] if part_at and isinstance(part_at[0], (list, tuple)) else list(part_at)
    max_batch = max(int(np.asarray(x).shape[0]) for x in inputs)
    chk = StreamChecker(model, cuts, mode=mode, max_batch=max_batch, **kw)
    try:
        chk.check(inputs, outs, oracle_every=oracle_every)
    finally:
        chk.close()


def test_reference_driver_shape(pkg, model50):
    """The reference's own driver (test/test.py:10-50), restated: 10 x batch-1 through part_at=[]."""
    from src.dispatcher import DEFER

    defer = DEFER(["0.0.0.0"], max_batch=1)
    x = KO.synthetic_images(1, seed=7, kind="caffe")
    outs, stats, launches = _run(defer, model50, [], [x] * 10)
    assert len(outs) == 10 and all(o.shape == (1, 1000) for o in outs)
    _check(model50, [], [x], outs[:1], oracle_tol="absolute")
    for o in outs[1:]:
        np.testing.assert_array_equal(o, outs[0])
    assert launches == 10 * 50          # 58 fused groups; 7 bottleneck tails chained (3x3 -> 1x1), stem + max-pool fused: one launch each


@pytest.mark.parametrize("part_at", [["conv3_block4_out"], ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"]])
def test_pipeline_stages_sharing_one_gpu(pkg, model50, part_at):
    from adapt_b200.dispatcher import DEFER

    n = len(part_at) + 1
    inputs = [distinct_images(b, 300, i) for i, b in enumerate([4, 3, 4, 1, 2, 4, 4, 4])]
    defer = DEFER(["cuda:0"] * n, max_batch=4)
    outs, stats, _ = _run(defer, model50, part_at, inputs)
    assert stats["delivered"] == len(inputs) and stats["failovers"] == 0
    _check(model50, part_at, inputs, outs)
    # the check has teeth: two micro-batches of equal size delivered in each other's place fail it
    swapped = list(outs)
    swapped[5], swapped[6] = swapped[6], swapped[5]
    with pytest.raises(AssertionError):
        _check(model50, part_at, inputs, swapped, oracle_every=0)


def test_value_check_catches_a_stale_input_slot(pkg, model50, monkeypatch):
    """Fault injection (node.StageWorker, ADAPT_FAULT_INJECT): the last stage runs micro-batch 5 on the other input
    slot, i.e. on micro-batch 4's cut tensor.  Delivery counts and order are untouched -- only the values are wrong,
    and the per-micro-batch check must say so (and point at micro-batch 5)."""
    from adapt_b200.dispatcher import DEFER

    part_at = ["conv3_block4_out"]
    inputs = [distinct_images(4, 330, i) for i in range(8)]
    monkeypatch.setenv("ADAPT_FAULT_INJECT", "stale_slot:5")
    defer = DEFER(["cuda:0"] * 2, max_batch=4)
    outs, stats, _ = _run(defer, model50, part_at, inputs)
    monkeypatch.delenv("ADAPT_FAULT_INJECT")
    assert stats["delivered"] == len(inputs)
    with pytest.raises(AssertionError, match="micro-batch 5"):
        _check(model50, part_at, inputs, outs, oracle_every=0)
    _check(model50, part_at, inputs[:5], outs[:5], oracle_every=0)      # everything before it is right
    _check(model50, part_at, inputs[6:], outs[6:], oracle_every=0)      # ... and after


def test_pipeline_across_gpus(pkg, model50):
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    from adapt_b200.dispatcher import DEFER, balanced_part_at

    n = min(_ngpu(), 4)
    part_at = balanced_part_at(model50, n)
    inputs = [distinct_images(4, 400, i) for i in range(12)]
    defer = DEFER([f"cuda:{i}" for i in range(n)], max_batch=4)
    outs, stats, _ = _run(defer, model50, part_at, inputs)
    assert stats["delivered"] == 12
    _check(model50, part_at, inputs, outs, oracle_every=3)


def test_pinned_torch_inputs(pkg, model50):
    torch = pytest.importorskip("torch")
    from adapt_b200.dispatcher import DEFER

    xs = [distinct_images(2, 500, i) for i in range(4)]
    pinned = [torch.from_numpy(x).pin_memory() for x in xs]
    defer = DEFER(["cuda:0"], max_batch=2)
    outs, _, _ = _run(defer, model50, [], pinned)
    _check(model50, [], xs, outs)


def test_ingest_rails_across_gpus(pkg, model50):
    """Pinned inputs spread over every GPU's PCIe link (node.StageWorker._setup_rails): every batch distinct, so a
    batch landing in the wrong slot or overtaking another shows up as a wrong row; a stage dies mid-stream."""
    torch = pytest.importorskip("torch")
    if _ngpu() < 2:
        pytest.skip("needs >= 2 GPUs")
    from adapt_b200.dispatcher import DEFER, balanced_part_at

    n = min(_ngpu(), 4)
    xs = [distinct_images(3, 700, i) for i in range(30)]
    pinned = [torch.from_numpy(x).pin_memory() for x in xs]
    defer = DEFER([f"cuda:{i}" for i in range(n)], max_batch=3, heartbeat_timeout=0.05, ingest_rails="auto")
    outs, stats, _ = _run(defer, model50, balanced_part_at(model50, n), pinned, kill_after=12, kill_stage=n - 1)
    assert stats["failovers"] == 1 and stats["delivered"] == 30
    _check(model50, [balanced_part_at(model50, n), balanced_part_at(model50, n - 1)], xs, outs, oracle_every=6)
    with pytest.raises(ValueError):
        DEFER(["cuda:0"], ingest_rails="sometimes")


def test_failover_resliced_and_replayed(pkg, model50):
    from adapt_b200.dispatcher import DEFER

    devs = [f"cuda:{i % max(1, _ngpu())}" for i in range(3)]
    inputs = [distinct_images(2, 600, i) for i in range(40)]
    defer = DEFER(devs, max_batch=2, heartbeat_timeout=0.05)
    outs, stats, _ = _run(defer, model50, ["conv2_block3_out", "conv4_block1_out"], inputs, kill_after=10, kill_stage=1)
    assert stats["failovers"] == 1 and stats["delivered"] == 40
    assert len(defer.part_at) == 1
    # FIFO, nothing lost, nothing duplicated, and every value that of ITS input (all cuts, before and after the
    # re-slice, are block outputs: bit for bit against the un-partitioned model)
    _check(model50, [["conv2_block3_out", "conv4_block1_out"], defer.part_at], inputs, outs, oracle_every=8)
    assert stats["failover_ms"] and stats["failover_ms"][0] < 1000.0   # sub-second recovery (north_star)


def test_killed_stage_with_queued_work_forwards_nothing_wrong(pkg, model50, monkeypatch):
    """A stage dies while several micro-batches are queued on its stream (the last stage is slowed down so that the
    pipeline backs up).  Its queued GPU work must not be released early: it would run on inputs that have not arrived
    and hand garbage downstream, which the collector would deliver until the heartbeat time-out fires.  Every result
    -- before, during and after the fail-over -- must be that of ITS input."""
    from adapt_b200.dispatcher import DEFER

    monkeypatch.setenv("ADAPT_FAULT_INJECT", "slow_worker:3:3000:0")     # last stage: +3 ms per micro-batch
    devs = [f"cuda:{i % max(1, _ngpu())}" for i in range(4)]
    part_at = ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"]
    inputs = [distinct_images(4, 1100, i) for i in range(36)]
    defer = DEFER(devs, max_batch=4, heartbeat_timeout=0.05, straggler_factor=None)
    outs, stats, _ = _run(defer, model50, part_at, inputs, kill_after=8, kill_stage=1, timeout=60)
    monkeypatch.delenv("ADAPT_FAULT_INJECT")
    assert stats["failovers"] == 1 and stats["delivered"] == len(inputs), repr(sorted(stats.items()))
    _check(model50, [part_at, defer.part_at], inputs, outs, oracle_every=9)


def test_elastic_rejoin_after_failover_gpu(pkg, model50):
    """Kill a stage mid-stream (3 -> 2 stages), revive it, and the watchdog re-slices it back in (2 -> 3) while
    inputs keep flowing: FIFO and exactly once across both reconfigurations, values against the oracle."""
    from adapt_b200.dispatcher import DEFER

    devs = [f"cuda:{i % max(1, _ngpu())}" for i in range(3)]
    inputs = [distinct_images(2, 800, i) for i in range(60)]
    defer = DEFER(devs, max_batch=2, heartbeat_timeout=0.05)
    in_q, out_q = queue.Queue(10), queue.Queue()
    t = threading.Thread(target=defer.run_defer, args=(model50, ["conv2_block3_out", "conv4_block1_out"], in_q, out_q),
                         daemon=True)
    t.start()
    defer.wait_ready(120)
    threading.Thread(target=lambda: [in_q.put(x) for x in inputs], daemon=True).start()
    outs, victim, revived, t0 = [], None, False, time.time()
    while len(outs) < len(inputs):
        try:
            outs.append(out_q.get(timeout=0.2))
        except queue.Empty:
            assert defer.error is None, defer.error
            assert time.time() - t0 < 60, f"timed out with {len(outs)}/{len(inputs)}: {defer.debug_state()}"
        if victim is None and len(outs) >= 10:
            victim = defer.inject_kill(1)
        if victim is not None and not revived and defer.stats["failovers"] == 1 and len(outs) >= 25:
            defer.revive(victim)
            revived = True
    t1 = time.time()
    while defer.stats["rejoins"] == 0 and time.time() - t1 < 10:   # (the stream may end before the watchdog's next look)
        time.sleep(0.01)
    time.sleep(0.05)
    assert out_q.empty(), "duplicate delivery"
    stats, order = dict(defer.stats), list(defer.order)
    defer.shutdown()
    t.join(30)
    assert stats["failovers"] == 1 and stats["rejoins"] == 1 and stats["delivered"] == len(inputs)
    assert len(order) == 3 and victim in order
    _check(model50, [["conv2_block3_out", "conv4_block1_out"]], inputs, outs, oracle_every=12)  # (re-slices cut at block outputs too)


def test_straggler_stage_is_dropped_gpu(pkg, model50, monkeypatch):
    """A stage that is alive (its heartbeat keeps coming) but slow: from micro-batch 12 on, worker 1's stream is kept
    busy for 40 ms on top of every micro-batch (a spin kernel, ADAPT_FAULT_INJECT).  The busy-time watchdog (CUDA
    events around the stage's work -> moving average in the health table) drops it, the model is re-sliced over the
    other two, and every micro-batch is delivered exactly once, in order, with ITS values."""
    from adapt_b200.dispatcher import DEFER

    monkeypatch.setenv("ADAPT_FAULT_INJECT", "slow_worker:1:40000:12")
    devs = [f"cuda:{i % max(1, _ngpu())}" for i in range(3)]
    inputs = [distinct_images(2, 1000, i) for i in range(48)]
    defer = DEFER(devs, max_batch=2, heartbeat_timeout=0.25, straggler_factor=6.0, straggler_window=0.15)
    outs, stats, _ = _run(defer, model50, ["conv2_block3_out", "conv4_block1_out"], inputs, timeout=60)
    monkeypatch.delenv("ADAPT_FAULT_INJECT")
    assert stats["stragglers"] == 1 and stats["failovers"] == 1 and stats["delivered"] == len(inputs), repr(sorted(stats.items()))
    ev = stats["straggler_events"][0]
    assert ev["worker"] == 1 and ev["reason"] == "busy time" and ev["ratio"] >= 6.0, ev
    assert len(defer.part_at) == 1 and 1 not in defer.order
    _check(model50, [["conv2_block3_out", "conv4_block1_out"], defer.part_at], inputs, outs, oracle_every=12)


def test_uint8_inputs_through_defer(pkg, model50):
    """input_format="uint8": pinned and pageable raw RGB batches through a 2-stage pipeline, values against the
    oracle on host-preprocessed images."""
    torch = pytest.importorskip("torch")
    from adapt_b200 import preprocess_input
    from adapt_b200.dispatcher import DEFER

    rng = np.random.default_rng(21)
    xs = [rng.integers(0, 256, (2, 224, 224, 3), dtype=np.uint8) for _ in range(6)]
    items = [torch.from_numpy(x).pin_memory() if i % 2 else x for i, x in enumerate(xs)]
    devs = [f"cuda:{i % max(1, _ngpu())}" for i in range(2)]
    defer = DEFER(devs, max_batch=2, input_format="uint8")
    outs, _, _ = _run(defer, model50, ["conv3_block4_out"], items)
    _check(model50, ["conv3_block4_out"], xs, outs, input_format="uint8", preprocess=preprocess_input,
           oracle_tol="absolute")


def test_fp32_check_mode_through_defer(pkg, model50):
    from adapt_b200.dispatcher import DEFER

    inputs = [KO.synthetic_images(1, seed=700 + i, kind="caffe") for i in range(2)]
    defer = DEFER(["cuda:0", "cuda:0"], mode="fp32", max_batch=1)
    outs, _, _ = _run(defer, model50, ["conv3_block4_out"], inputs)
    _check(model50, ["conv3_block4_out"], inputs, outs, mode="fp32", oracle_tol="absolute")
