"""CPU: the one-process-per-stage path (what torchrun launches on the GPU box) with world_size 2 over
gloo: rank 0 = dispatcher + stage 0, rank 1 = `Node().run()` worker.  Device memory is stood in for by
named shared memory and the stage compute by the oracle (tests/_host_backend.py)."""
import os
import queue
import sys
import threading
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch
    import torch.distributed as dist

    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import adapt_b200
    from _host_backend import SharedHostBackend
    from _util import KO, oracle_for
    from adapt_b200.dispatcher import DEFER, end_sessions
    from adapt_b200.node import Node

    model = adapt_b200.ResNet50(weights=None)
    try:
        if rank > 0:
            node = Node(model, device=rank, backend_factory=SharedHostBackend)
            node.run()                       # serves session after session until the dispatcher ends them
            q.put((rank, "ok", node.sessions_served))
            return
        o = oracle_for(model)
        err = 0.0
        # two sessions back to back (what bench.py does for BASELINE's other configs): other cuts, other micro-batch
        for part_at, batch in ((["conv3_block4_out"], 1), (["conv4_block6_out"], 2)):
            inputs = [KO.synthetic_images(batch, seed=900 + i, kind="normal") for i in range(3)]
            defer = DEFER(["10.0.0.1", "10.0.0.2"], mode="fp32", max_batch=batch, backend_factory=SharedHostBackend,
                          heartbeat_timeout=5.0)
            in_q, out_q = queue.Queue(), queue.Queue()
            t = threading.Thread(target=defer.run_defer, args=(model, part_at, in_q, out_q), daemon=True)
            t.start()
            defer.wait_ready(120)
            for x in inputs:
                in_q.put(x)
            outs = [out_q.get(timeout=120) for _ in inputs]
            defer.shutdown()
            t.join(30)
            err = max([err] + [float(np.abs(y - o.predict(x)[0]).max()) for x, y in zip(inputs, outs)])
        # third session: the dispatcher-local FIRST stage is lost.  The surviving stage lives in the other process and
        # cannot read this process's host inputs: the dispatcher copies them into that stage's slot itself (KIND_FED)
        inputs = [KO.synthetic_images(1, seed=930 + i, kind="normal") for i in range(7)]
        defer = DEFER(["10.0.0.1", "10.0.0.2"], mode="fp32", max_batch=1, backend_factory=SharedHostBackend, heartbeat_timeout=2.0)
        in_q, out_q = queue.Queue(), queue.Queue()
        t = threading.Thread(target=defer.run_defer, args=(model, ["conv3_block4_out"], in_q, out_q), daemon=True)
        t.start()
        defer.wait_ready(120)
        stream = []
        for x in inputs[:3]:
            in_q.put(x)
            stream.append(out_q.get(timeout=120))
        defer.inject_kill(0)
        for x in inputs[3:]:                      # fed while the dead stage is being detected and the model re-sliced
            in_q.put(x)
        for _ in inputs[3:]:
            stream.append(out_q.get(timeout=120))
        time.sleep(0.05)
        assert out_q.empty()
        assert defer.stats["failovers"] == 1 and defer.order == [1] and defer.part_at == [], (defer.stats, defer.order)
        defer.shutdown()
        t.join(30)
        err = max([err] + [float(np.abs(y - o.predict(x)[0]).max()) for x, y in zip(inputs, stream)])   # FIFO, exactly once
        end_sessions()
        q.put((0, "ok", err))
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, "fail", f"{e}\n{traceback.format_exc()}"))
    finally:
        try:
            dist.destroy_process_group()
        except Exception:
            pass


def test_two_rank_pipeline_over_gloo():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = {}
    t0 = time.time()
    while len(results) < 2 and time.time() - t0 < 240:
        try:
            r, status, info = q.get(timeout=5)
            results[r] = (status, info)
        except Exception:
            if not any(p.is_alive() for p in procs):
                break
    for p in procs:
        p.join(20)
        if p.is_alive():
            p.kill()
    assert results.get(0, ("missing", None))[0] == "ok", results
    assert results.get(1, ("missing", None)) == ("ok", 3), results      # the worker rank served all three sessions
    assert results[0][1] < 1e-5
