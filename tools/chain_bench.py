"""Micro-benchmark of the chain kernel (`adapt_conv_chain_bf16`) on the conv2 / conv3 geometries at batch 128:
CUDA events around 20 back-to-back launches, algorithmic HBM bytes and FLOPs.  Also the subject of the ncu captures
(profiles/r2_chain_*):   python tools/chain_bench.py [C ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adapt_b200 import _native as N  # noqa: E402

lib = N.lib()
B = int(os.environ.get("B", "128"))
reps = int(os.environ.get("REPS", "20"))
for C in [int(a) for a in sys.argv[1:]] or [64, 128]:
    H = W = 56 if C == 64 else 28
    dev = "cuda:0"
    x = torch.randn(B, H, W, C, device=dev).bfloat16()
    w2 = (torch.randn(C, 3, 3, C, device=dev) / (3 * C ** 0.5)).bfloat16()
    w3 = (torch.randn(4 * C, C, device=dev) / C ** 0.5).bfloat16()
    s1, b1 = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    s2, b2 = torch.ones(4 * C, device=dev), torch.zeros(4 * C, device=dev)
    res = torch.randn(B, H, W, 4 * C, device=dev).bfloat16()
    out = torch.empty(B, H, W, 4 * C, device=dev, dtype=torch.bfloat16)
    stream = torch.cuda.current_stream().cuda_stream

    def run():
        N.check(lib.adapt_conv_chain_bf16(0, x.data_ptr(), w2.data_ptr(), s1.data_ptr(), b1.data_ptr(), w3.data_ptr(), s2.data_ptr(),
                                          b2.data_ptr(), res.data_ptr(), out.data_ptr(), B, H, W, C, stream))

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    byts = 2 * B * H * W * (C + 8 * C)
    flop = 2 * B * H * W * (9 * C * C + 4 * C * C)
    print(f"chain C={C} {H}x{W} B={B}: {us:.1f} us  {byts / us / 1e3:.0f} GB/s  {flop / us / 1e6:.0f} TFLOP/s", flush=True)
