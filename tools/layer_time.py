"""Time one convolution layer through `adapt_conv2d_bf16` with its operands warm in L2 (the situation inside a stage):
   python tools/layer_time.py H W Cin Cout k stride pad res bn [batch]     (bn = 0: the library's own tile choice)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adapt_b200 import _native as N  # noqa: E402

lib = N.lib()
H, W, Cin, Cout, k, stride, pad, use_res, bn = [int(v) for v in sys.argv[1:10]]
B = int(sys.argv[10]) if len(sys.argv) > 10 else 128
dev = torch.device("cuda:0")
x = torch.randn(B, H, W, Cin, device=dev).to(torch.bfloat16)
w = (torch.randn(Cout, k, k, Cin, device=dev) / (k * k * Cin) ** 0.5).to(torch.bfloat16)
scale = torch.ones(Cout, device=dev)
bias = torch.zeros(Cout, device=dev)
Ho = (H + 2 * pad - k) // stride + 1
Wo = (W + 2 * pad - k) // stride + 1
out = torch.empty(B, Ho, Wo, Cout, device=dev, dtype=torch.bfloat16)
res = torch.randn(B, Ho, Wo, Cout, device=dev).to(torch.bfloat16) if use_res else None
st = torch.cuda.current_stream().cuda_stream


def run():
    N.check(lib.adapt_conv2d_bf16(0, x.data_ptr(), w.data_ptr(), scale.data_ptr(), bias.data_ptr(), res.data_ptr() if use_res else None,
                                  out.data_ptr(), 0, B, H, W, Cin, Cout, k, k, stride, pad, 1, -1, bn, st))


for _ in range(3):
    run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(20):
    run()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 20 * 1e3
flop = 2.0 * B * Ho * Wo * Cout * k * k * Cin
print(f"{H}x{W} {Cin}->{Cout} k{k}s{stride} res={use_res} bn={bn} PAIR={os.environ.get('ADAPT_PAIR', '1')}: {us:.1f} us  {flop / us / 1e6:.0f} TFLOP/s")
