"""What a stage-to-stage hand-off can reach over NVLink: a 205 MB cut tensor (conv2 level, micro-batch 128) pushed
GPU 0 -> GPU 1 by copy engines, as one copy or split over several streams, alone and with the reverse direction busy
(a middle stage receives while it sends).   gpurun --gpus 2 -- 'python tools/handoff_bench.py'"""
import torch

nbytes = 128 * 56 * 56 * 256 * 2
src = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda:1")
rsrc = torch.empty(nbytes, dtype=torch.uint8, device="cuda:1")
rdst = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
for a, b in ((0, 1), (1, 0)):
    torch.cuda.set_device(a)
    assert torch.cuda.can_device_access_peer(a, b)


def run(chunks, reverse, reps=10):
    torch.cuda.set_device(0)
    streams = [torch.cuda.Stream(device=0) for _ in range(chunks)]
    rstream = torch.cuda.Stream(device=1)
    step = (nbytes // chunks + 4095) // 4096 * 4096
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    main = torch.cuda.current_stream(0)
    for it in range(reps + 2):
        if it == 2:
            torch.cuda.synchronize(0)
            torch.cuda.synchronize(1)
            e0.record(main)
        if reverse:
            with torch.cuda.stream(rstream):
                rdst.copy_(rsrc, non_blocking=True)
        for i, s in enumerate(streams):
            s.wait_stream(main)
            with torch.cuda.stream(s):
                lo, hi = i * step, min(nbytes, (i + 1) * step)
                dst[lo:hi].copy_(src[lo:hi], non_blocking=True)
        for s in streams:
            main.wait_stream(s)
    e1.record(main)
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    ms = e0.elapsed_time(e1) / reps
    return nbytes / ms / 1e6


for reverse in (False, True):
    for chunks in (1, 2, 4, 8):
        print(f"chunks={chunks} reverse_busy={reverse}: {run(chunks, reverse):.0f} GB/s", flush=True)
