"""CPU: host logic of the dispatcher / worker protocol with a synchronous test backend whose
stage compute is the oracle (tests/_host_backend.py).  No CUDA involved; the GPU versions of the
same scenarios are in tests/test_defer_gpu.py."""
import queue
import threading
import time

import numpy as np
import pytest

from _host_backend import HostBackend
from _util import KO, oracle_for


def _run(defer, model, part_at, inputs, kill_after=None, kill_stage=None, timeout=120):
    in_q, out_q = queue.Queue(10), queue.Queue()
    t = threading.Thread(target=defer.run_defer, args=(model, part_at, in_q, out_q), daemon=True)
    t.start()
    defer.wait_ready(60)
    outs = []

    def feed():
        for x in inputs:
            in_q.put(x)

    f = threading.Thread(target=feed, daemon=True)
    f.start()
    t0 = time.time()
    killed = False
    while len(outs) < len(inputs):
        try:
            outs.append(out_q.get(timeout=0.2))
        except queue.Empty:
            assert defer.error is None, defer.error
            assert time.time() - t0 < timeout, f"timed out with {len(outs)}/{len(inputs)} results"
        if kill_after is not None and not killed and len(outs) >= kill_after:
            defer.inject_kill(kill_stage)
            killed = True
    time.sleep(0.05)
    assert out_q.empty(), "more results than inputs (duplicate delivery)"
    stats = dict(defer.stats)
    defer.shutdown()
    t.join(20)
    return outs, stats


def test_partition_names_and_errors(pkg, model50):
    from adapt_b200.dispatcher import DEFER

    d = DEFER(["127.0.0.1"], backend_factory=HostBackend)
    assert d.computeNodes == ["127.0.0.1"] and d.chunk_size == 512 * 1000 and isinstance(d.dispatchIP, str)
    parts = d._partition(model50, ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"])
    assert [p.name for p in parts] == ["part1", "part2", "part3", "part4"]
    assert parts[0].start == "input_1" and parts[0].end == "conv2_block3_out"
    assert parts[1].layer_names[0] == "conv3_block1_1_conv" and parts[1].layer_names[-1] == "conv3_block4_out"
    assert parts[3].end == "predictions"
    assert sum(p.macs() for p in parts) == 3_857_973_248
    assert [round(p.macs() / 1e9, 3) for p in parts] == [0.786, 0.951, 1.387, 0.734]   # SURVEY 8d, config 3 (rounded there)
    with pytest.raises(RuntimeError):
        d._partition(model50, ["conv3_block1_2_relu"])
    with pytest.raises(ValueError):
        d._partition(model50, ["not_a_layer"])
    assert len(d._partition(model50, [])) == 1


def test_balanced_part_at(pkg, model50):
    from adapt_b200.dag_util import valid_cut_points
    from adapt_b200.dispatcher import DEFER, balanced_part_at

    valid = set(valid_cut_points(model50))
    d = DEFER([0], backend_factory=HostBackend)
    for n, eff_floor in ((2, 0.90), (4, 0.85), (7, 0.80), (8, 0.80)):
        cuts = balanced_part_at(model50, n, cost_model="macs")
        assert len(cuts) == n - 1 and set(cuts) <= valid
        tcuts = balanced_part_at(model50, n)              # default: roofline time model, overlapped hand-off
        assert len(tcuts) == n - 1 and set(tcuts) <= valid and tcuts == sorted(tcuts, key=[l.name for l in model50.layers].index)
        macs = [p.macs() for p in d._partition(model50, cuts)]
        assert sum(macs) / (n * max(macs)) >= eff_floor, (n, cuts, macs)
    assert balanced_part_at(model50, 1) == []


def test_control_ring_and_health_table(pkg):
    from adapt_b200.fabric import ControlRing, HealthTable, ResultRing, ShmRegion

    reg = ShmRegion(create=True, size=HealthTable.BYTES + ControlRing.BYTES + ResultRing.bytes_needed(4000) + 4096)
    try:
        ring = ControlRing(reg, HealthTable.BYTES)
        assert ring.pop(timeout=0.001) is None
        for i in range(300):                      # wraps the 256-entry ring
            assert ring.push((1, 0, i, i, 2, 0))
            assert ring.pop() == (1, 0, i, i, 2, 0)
        for i in range(ControlRing.CAP):
            assert ring.push((1, 0, i, i, 1, 0))
        assert not ring.push((1, 0, 0, 0, 1, 0), timeout=0.01)     # full -> back-pressure
        ring.clear()
        assert ring.pop(timeout=0.001) is None
        h = HealthTable(reg, 0)
        h.write_table({"epoch": 3, "order": [0, 2], "part_at": ["conv3_block4_out"]})
        assert h.read_table()["order"] == [0, 2]
        h.beat(1)
        assert h.age_ms(1) < 50
        other = ShmRegion(reg.name)               # second mapping sees the same bytes
        assert HealthTable(other, 0).read_table()["epoch"] == 3
        other.close()
    finally:
        reg.close()


@pytest.mark.parametrize("part_at", [[], ["conv3_block4_out"], ["pool1_pool", "conv3_block1_out", "conv5_block1_add"]])
def test_pipeline_fifo_and_values(pkg, model50, part_at):
    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(b, seed=100 + i, kind="normal") for i, b in enumerate([1, 2, 1, 1])]   # ragged batches
    d = DEFER([f"10.0.0.{i}" for i in range(len(part_at) + 1)], mode="fp32", max_batch=2, backend_factory=HostBackend,
              heartbeat_timeout=5.0)
    outs, stats = _run(d, model50, part_at, inputs)
    assert stats["delivered"] == len(inputs) and stats["failovers"] == 0
    for x, y in zip(inputs, outs):
        assert y.shape == (x.shape[0], 1000) and y.dtype == np.float32
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_failover_replays_exactly_once_in_order(pkg, model50):
    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(1, seed=200 + i, kind="normal") for i in range(7)]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=lambda dev: HostBackend(dev, delay=0.02),
              heartbeat_timeout=0.25)
    outs, stats = _run(d, model50, ["conv2_block3_out", "conv4_block1_out"], inputs, kill_after=2, kill_stage=1)
    assert stats["failovers"] == 1 and stats["delivered"] == 7
    assert stats["replayed"] >= 1                     # something was in flight when the stage died
    assert d.part_at and len(d.part_at) == 1          # re-sliced onto the two survivors
    for x, y in zip(inputs, outs):                    # order preserved, nothing lost, nothing duplicated
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)
    assert len(stats["failover_ms"]) == 1 and stats["failover_ms"][0] > 0


def test_too_few_compute_nodes(pkg, model50):
    from adapt_b200.dispatcher import DEFER

    d = DEFER([0], backend_factory=HostBackend)
    with pytest.raises(ValueError):
        d.run_defer(model50, ["conv3_block4_out"], queue.Queue(), queue.Queue())


def test_constructor_options_are_validated_and_published(pkg, model50):
    """Build-specific DEFER options: bad values fail loudly; the hand-off / ingest choices reach the workers through
    the partition table (host backends have no PCIe rails: the option must be harmless there)."""
    from adapt_b200.dispatcher import DEFER

    with pytest.raises(ValueError):
        DEFER([0], handoff="teleport", backend_factory=HostBackend)
    with pytest.raises(ValueError):
        DEFER([0], ingest_rails="sometimes", backend_factory=HostBackend)
    d = DEFER([0, 1], mode="fp32", max_batch=1, backend_factory=HostBackend, heartbeat_timeout=5.0, handoff="copy",
              ingest_rails="auto")
    inputs = [KO.synthetic_images(1, seed=900 + i, kind="normal") for i in range(3)]
    in_q, out_q = queue.Queue(10), queue.Queue()
    import threading

    t = threading.Thread(target=d.run_defer, args=(model50, ["conv3_block4_out"], in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    table = d.health.read_table()
    assert table["handoff"] == "copy" and table["ingest_rails"] == "auto" and table["part_at"] == ["conv3_block4_out"]
    assert all(not getattr(w, "_active_rails", []) for w in d.workers)
    for x in inputs:
        in_q.put(x)
    outs = [out_q.get(timeout=60) for _ in inputs]
    d.shutdown()
    t.join(30)
    o = oracle_for(model50)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_split_copy_handoff(pkg, model50, monkeypatch):
    """A hand-off copy can be split over several streams (ADAPT_XFER_STREAMS; off by default: measured slower on one
    NVLink) and is then published only once every piece has landed: forced here on small tensors."""
    from adapt_b200 import node
    from adapt_b200.dispatcher import DEFER

    monkeypatch.setattr(node, "XFER_SPLIT_MIN_BYTES", 4096)
    monkeypatch.setenv("ADAPT_XFER_STREAMS", "4")
    o = oracle_for(model50)
    inputs = [KO.synthetic_images(b, seed=950 + i, kind="normal") for i, b in enumerate([1, 2, 2, 1, 2])]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=2, backend_factory=HostBackend, heartbeat_timeout=5.0, handoff="copy")
    outs, stats = _run(d, model50, ["conv2_block3_out", "conv4_block6_out"], inputs)
    assert stats["delivered"] == len(inputs)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_bench_cut_selection(pkg, model50):
    """bench.py: N = 1 is the single chunk; N > 1 defaults to the time-balanced cuts, `--cuts config` to the split
    points BASELINE.json names for 2 and 4 stages; every choice is a valid partition."""
    import importlib.util
    import os

    from adapt_b200.dag_util import valid_cut_points

    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(os.path.dirname(__file__)), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert bench.part_at_for(model50, 1) == [] and bench.part_at_for(model50, 1, "config") == []
    assert bench.part_at_for(model50, 2, "config") == ["conv3_block4_out"]
    assert bench.part_at_for(model50, 4, "config") == ["conv2_block3_out", "conv3_block4_out", "conv4_block6_out"]
    valid = set(valid_cut_points(model50))
    for n in (2, 3, 4, 8):
        cuts = bench.part_at_for(model50, n)
        assert len(cuts) == n - 1 and all(c in valid for c in cuts), cuts


@pytest.mark.parametrize("ctor_name,n_params,n_layers", [("ResNet101", 44707176, 347), ("ResNet152", 60419944, 517)])
def test_deeper_resnets_share_the_front_end(pkg, ctor_name, n_params, n_layers):
    """`construct_model` is model-agnostic.  The Keras-compatible ResNet-101 / -152 graphs have
    the published parameter counts, partition at the same kind of articulation points, compile to stage programs
    and run through the DEFER pipeline (host executor) with the oracle's values."""
    import adapt_b200
    from adapt_b200.dag_util import build_weight_table, compile_part, construct_model, valid_cut_points
    from adapt_b200.dispatcher import DEFER, balanced_part_at

    m = getattr(adapt_b200, ctor_name)(weights=None)
    assert len(m.layers) == n_layers and sum(int(w.size) for w in m.get_weights()) == n_params
    cuts = valid_cut_points(m)
    assert "conv4_block23_out" in cuts and "conv4_block23_2_relu" not in cuts
    part_at = balanced_part_at(m, 3)
    assert len(part_at) == 2 and all(c in cuts for c in part_at)
    _, windex = build_weight_table(m)
    whole = construct_model(m, m.layers[0].name, m.layers[-1].name, "whole")
    assert len(compile_part(whole, windex, 0).ops) > 100
    with pytest.raises(RuntimeError):
        construct_model(m, "conv4_block23_2_relu", m.layers[-1].name, "bad")
    if ctor_name == "ResNet152":
        return  # (the pipeline run below is the same code path; one model keeps the CPU suite short)
    x = KO.synthetic_images(1, seed=77, kind="normal")
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=HostBackend, heartbeat_timeout=5.0)
    outs, stats = _run(d, m, part_at, [x])
    np.testing.assert_allclose(outs[0], oracle_for(m).predict(x)[0], rtol=1e-5, atol=1e-7)


def test_elastic_rejoin_after_failover(pkg, model50):
    """a stage is killed (3 -> 2 stages), later revived; the watchdog hands it a stage again
    (2 -> 3) and the stream continues -- FIFO, nothing lost, nothing duplicated across both reconfigurations."""
    import threading

    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(1, seed=300 + i, kind="normal") for i in range(12)]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=lambda dev: HostBackend(dev, delay=0.01),
              heartbeat_timeout=1.0)   # (host executors hold the GIL for long stretches: a generous liveness window)
    in_q, out_q = queue.Queue(4), queue.Queue()
    t = threading.Thread(target=d.run_defer, args=(model50, ["conv2_block3_out", "conv4_block1_out"], in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    outs = []

    def pump(items):
        for x in items:
            in_q.put(x)
            outs.append(out_q.get(timeout=60))

    pump(inputs[:3])
    victim = d.inject_kill(1)
    with pytest.raises(RuntimeError):
        d.revive(victim)                      # still holds a stage: the fail-over has not happened yet
    pump(inputs[3:7])                         # stalls until the fail-over, then flows through 2 stages
    assert d.stats["failovers"] == 1 and len(d.order) == 2 and victim not in d.order, (d.stats, d.order)
    d.revive(victim)
    t0 = time.time()
    while d.stats["rejoins"] == 0 and time.time() - t0 < 45:   # (the quarantine is long enough for the six slow items above)
        time.sleep(0.01)
    assert d.stats["rejoins"] == 1 and len(d.order) == 3 and victim in d.order and len(d.part_at) == 2
    pump(inputs[7:])
    assert d.error is None and out_q.empty()
    d.shutdown()
    t.join(30)
    assert d.stats["delivered"] == len(inputs)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_spare_workers_idle_until_a_failover_needs_them(pkg, model50):
    """More computeNodes than partitions: the caller's `part_at` defines the pipeline, the spare worker holds no stage
    (no re-slice, no re-join right after start-up).  When a stage dies the spare steps in, and because the stage
    count is whole again the caller's own cut points are kept."""
    import threading

    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    part_at = ["conv3_block4_out"]
    inputs = [KO.synthetic_images(1, seed=370 + i, kind="normal") for i in range(6)]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=lambda dev: HostBackend(dev, delay=0.01),
              heartbeat_timeout=0.3)
    in_q, out_q = queue.Queue(4), queue.Queue()
    t = threading.Thread(target=d.run_defer, args=(model50, part_at, in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    outs = []

    def pump(items):
        for x in items:
            in_q.put(x)
            outs.append(out_q.get(timeout=60))

    pump(inputs[:2])
    time.sleep(1.0)                          # many watchdog ticks
    assert d.order == [0, 1] and d.part_at == part_at and d.stats["rejoins"] == 0 and d.health.epoch == 1
    d.inject_kill(1)
    pump(inputs[2:])
    assert d.stats["failovers"] == 1 and d.order == [0, 2] and d.part_at == part_at and d.stats["rejoins"] == 0
    assert d.error is None and out_q.empty()
    d.shutdown()
    t.join(30)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_flag_values_of_different_epochs_never_alias(pkg, model50):
    """Flag and result-ready values are `flag_base + k + 1`; every reconfiguration moves the base past whatever the
    old epoch can have written, so a late write of a dropped-but-running stage cannot satisfy a wait of the new
    epoch.  Checked here on the published table and by injecting exactly such stale writes after a fail-over."""
    import threading

    from adapt_b200.dispatcher import DEFER
    from adapt_b200.fabric import FLAG_BASE_STEP
    from adapt_b200.node import F_IN_READY

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(1, seed=380 + i, kind="normal") for i in range(8)]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=lambda dev: HostBackend(dev, delay=0.01),
              heartbeat_timeout=0.3)
    in_q, out_q = queue.Queue(4), queue.Queue()
    t = threading.Thread(target=d.run_defer, args=(model50, ["conv2_block3_out", "conv4_block1_out"], in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    outs = []

    def pump(items):
        for x in items:
            in_q.put(x)
            outs.append(out_q.get(timeout=60))

    base0 = d.health.read_table()["flag_base"]
    pump(inputs[:3])
    d.inject_kill(1)
    pump(inputs[3:5])
    base1 = d.health.read_table()["flag_base"]
    assert base1 >= base0 + 3 + FLAG_BASE_STEP       # beyond base0 + (micro-batches fed in epoch 1) + 1
    # what a zombie of the old epoch would still write: "micro-batch k done" values of the OLD base, into the new
    # last stage's input-ready flags and into a result slot's ready word
    last = d.order[-1]
    delivered = d.stats["delivered"]
    for k in range(8):
        d.flags_host[last, F_IN_READY + (k & 1)] = (base0 + k + 1) & 0xFFFFFFFF
        d.results.ready[k % 8][0] = (base0 + k + 1) & 0xFFFFFFFF
    time.sleep(0.2)
    assert d.stats["delivered"] == delivered and out_q.empty() and d.error is None    # nothing was accepted
    # (the stale values also clobbered live flags of this synchronous test backend: start a clean epoch to go on)
    d.revive(1)
    t0 = time.time()
    while d.stats["rejoins"] == 0 and time.time() - t0 < 20:
        time.sleep(0.01)
    assert d.stats["rejoins"] == 1
    pump(inputs[5:])
    d.shutdown()
    t.join(30)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_straggler_is_dropped_and_may_rejoin(pkg, model50):
    """SURVEY 8(f) rank 1, the `start_time` idea (reference src/dispatcher.py:190-194, 302-304): a stage that keeps
    beating but turns 20x slower is not "dead" -- the busy-time watchdog drops it, the model is re-sliced over the
    others, every micro-batch is still delivered once and in order; after its quarantine (and once it is fast again)
    the worker gets a stage back."""
    import threading

    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(1, seed=390 + i, kind="normal") for i in range(26)]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=lambda dev: HostBackend(dev, delay=0.005),
              heartbeat_timeout=5.0, straggler_factor=4.0, straggler_window=0.3, straggler_quarantine=12.0)
    in_q, out_q = queue.Queue(4), queue.Queue()
    t = threading.Thread(target=d.run_defer, args=(model50, ["conv2_block3_out", "conv4_block1_out"], in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    outs = []

    def pump(items):
        for x in items:
            in_q.put(x)
            outs.append(out_q.get(timeout=60))

    pump(inputs[:12])                                   # healthy: baselines settle (>= 8 samples per stage)
    assert d.stats["stragglers"] == 0 and d.order == [0, 1, 2]
    slow = next(w for w in d.workers if w.wid == 1)
    base_delay = slow.executor.delay
    slow.executor.delay = 1.0                           # alive, heart beating, 20x+ slower
    pump(inputs[12:18])
    assert d.stats["stragglers"] == 1 and d.stats["failovers"] == 1 and d.order == [0, 2], (d.stats, d.order)
    ev = d.stats["straggler_events"][0]
    assert ev["worker"] == 1 and ev["reason"] == "busy time" and ev["ratio"] >= 4.0
    slow.executor.delay = base_delay                    # recovered
    t0 = time.time()
    while d.stats["rejoins"] == 0 and time.time() - t0 < 20:
        time.sleep(0.02)
    assert d.stats["rejoins"] == 1 and d.order == [0, 1, 2]
    pump(inputs[18:])
    assert d.error is None and out_q.empty() and d.stats["stragglers"] == 1
    d.shutdown()
    t.join(30)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_task_timeout_names_the_stage_that_sits_on_the_backlog(pkg, model50):
    """`task_timeout`: an in-flight micro-batch older than the limit (its `start_time` in the in-flight registry) makes
    the watchdog drop a stage even when no busy-time evidence exists yet (here: the very first micro-batches)."""
    import threading

    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(1, seed=395 + i, kind="normal") for i in range(5)]
    d = DEFER([0, 1], mode="fp32", max_batch=1, backend_factory=HostBackend, heartbeat_timeout=30.0, straggler_factor=None,
              task_timeout=1.0)
    in_q, out_q = queue.Queue(8), queue.Queue()
    t = threading.Thread(target=d.run_defer, args=(model50, ["conv3_block4_out"], in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    stuck = next(w for w in d.workers if w.wid == 1)
    stuck.executor.delay = 4.0                          # the second stage sits on its first micro-batch
    for x in inputs:
        in_q.put(x)
    outs = [out_q.get(timeout=60) for _ in inputs]
    assert d.stats["stragglers"] == 1 and d.order == [0]
    assert d.stats["straggler_events"][0]["reason"].startswith("task timeout")
    d.shutdown()
    t.join(30)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_cascading_kills(pkg, model50):
    """Two stages die one after the other (3 -> 2 -> 1 stages): each death is detected and re-sliced on its own,
    the stream stays FIFO and exactly-once, and the last survivor runs the whole model."""
    import threading

    from adapt_b200.dispatcher import DEFER

    o = oracle_for(model50)
    inputs = [KO.synthetic_images(1, seed=350 + i, kind="normal") for i in range(9)]
    d = DEFER([0, 1, 2], mode="fp32", max_batch=1, backend_factory=lambda dev: HostBackend(dev, delay=0.01),
              heartbeat_timeout=1.0)
    in_q, out_q = queue.Queue(4), queue.Queue()
    t = threading.Thread(target=d.run_defer, args=(model50, ["conv2_block3_out", "conv4_block1_out"], in_q, out_q), daemon=True)
    t.start()
    d.wait_ready(60)
    outs = []

    def pump(items):
        for x in items:
            in_q.put(x)
            outs.append(out_q.get(timeout=60))

    pump(inputs[:2])
    d.inject_kill(2)
    pump(inputs[2:5])
    assert d.stats["failovers"] == 1 and len(d.order) == 2
    d.inject_kill(1)
    pump(inputs[5:])
    assert d.stats["failovers"] == 2 and len(d.order) == 1 and d.part_at == []
    assert d.error is None and out_q.empty()
    d.shutdown()
    t.join(30)
    for x, y in zip(inputs, outs):
        np.testing.assert_allclose(y, o.predict(x)[0], rtol=1e-5, atol=1e-7)


def test_uint8_input_format(pkg, model50):
    """`input_format="uint8"` -- queue items are raw RGB uint8 images, a quarter of the bytes;
    preprocessing (RGB -> BGR, ImageNet means) moves behind the queue.  Same values as preprocessing on the host."""
    from adapt_b200 import preprocess_input
    from adapt_b200._native import DT_U8, IN_U8_RGB_CAFFE, OP_INGEST
    from adapt_b200.dag_util import build_weight_table, compile_part, construct_model
    from adapt_b200.dispatcher import DEFER

    with pytest.raises(ValueError):
        DEFER([0], input_format="jpeg", backend_factory=HostBackend)
    # the stage program: only the part that takes the model input changes (1-byte input elements, flagged ingest)
    _, windex = build_weight_table(model50)
    first = construct_model(model50, model50.layers[0].name, "conv3_block4_out", "part1")
    second = construct_model(model50, "conv3_block4_out", model50.layers[-1].name, "part2")
    for mode in (0, 1):
        p8, p32 = compile_part(first, windex, mode, IN_U8_RGB_CAFFE), compile_part(first, windex, mode)
        assert p8.in_dtype == DT_U8 and p8.in_elems == 224 * 224 * 3
        assert [o.in_format for o in p8.ops if o.kind == OP_INGEST] == [IN_U8_RGB_CAFFE]
        assert [o.in_format for o in p32.ops if o.kind == OP_INGEST] == [0] and len(p8.ops) == len(p32.ops)
        assert not any(o.kind == OP_INGEST for o in compile_part(second, windex, mode, IN_U8_RGB_CAFFE).ops)
    rng = np.random.default_rng(5)
    imgs = [rng.integers(0, 256, (b, 224, 224, 3), dtype=np.uint8) for b in (1, 2)]
    d = DEFER([0, 1], mode="fp32", max_batch=2, backend_factory=HostBackend, heartbeat_timeout=5.0, input_format="uint8")
    outs, stats = _run(d, model50, ["conv3_block4_out"], imgs)
    o = oracle_for(model50)
    for x, y in zip(imgs, outs):
        np.testing.assert_allclose(y, o.predict(preprocess_input(x))[0], rtol=1e-5, atol=1e-7)


def test_fused_ops_in_the_stage_program(pkg, model50, monkeypatch):
    """What the bf16 stage program of the whole network looks like (host logic, no GPU): 58 fused layer groups, of which
    the seven conv2 / conv3 bottleneck tails run as one CONV_CHAIN each and conv1_conv .. pool1_pool as one STEM_POOL;
    a cut between the stem's convolution and its pool, the check mode and the switches leave the ops apart."""
    from adapt_b200._native import MODE_BF16, OP_CONV, OP_CONV_CHAIN, OP_MAXPOOL, OP_STEM_POOL
    from adapt_b200.dag_util import build_weight_table, compile_part, construct_model

    _, windex = build_weight_table(model50)
    first, last = model50.layers[0].name, model50.layers[-1].name
    whole = construct_model(model50, first, last, "all")
    ops = compile_part(whole, windex, MODE_BF16).ops
    kinds = [o.kind for o in ops]
    assert len(ops) == 50 and kinds.count(OP_CONV_CHAIN) == 7 and kinds.count(OP_STEM_POOL) == 1 and OP_MAXPOOL not in kinds
    stem = ops[kinds.index(OP_STEM_POOL)]
    assert stem.layers[0] == "conv1_conv" and stem.layers[-1] == "pool1_pool"
    assert (stem.H, stem.W, stem.Cin, stem.Ho, stem.Wo, stem.Cout, stem.kh, stem.stride) == (115, 128, 16, 112, 112, 64, 4, 1)
    # every layer is covered exactly once
    covered = [n for o in ops for n in o.layers if not n.endswith(":softmax")]
    assert len(covered) == len(set(covered)) and {"conv1_bn", "pool1_pad", "conv2_block1_2_conv", "conv2_block1_out"} <= set(covered)
    # fp32 check mode: no fused kernels
    kinds32 = [o.kind for o in compile_part(whole, windex, 1).ops]
    assert OP_STEM_POOL not in kinds32 and OP_CONV_CHAIN not in kinds32 and OP_MAXPOOL in kinds32
    # a cut at conv1_relu separates the convolution from its pool
    head = construct_model(model50, first, "conv1_relu", "p1")
    tail = construct_model(model50, "conv1_relu", "pool1_pool", "p2")
    assert OP_STEM_POOL not in [o.kind for o in compile_part(head, windex, MODE_BF16).ops]
    assert [o.kind for o in compile_part(tail, windex, MODE_BF16).ops] == [OP_MAXPOOL]
    # the switches
    monkeypatch.setenv("ADAPT_STEM_POOL", "0")
    monkeypatch.setenv("ADAPT_CHAIN", "0")
    kinds0 = [o.kind for o in compile_part(whole, windex, MODE_BF16).ops]
    assert len(kinds0) == 58 and OP_STEM_POOL not in kinds0 and OP_CONV_CHAIN not in kinds0 and kinds0.count(OP_CONV) == 54
