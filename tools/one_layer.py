"""Run one convolution layer a few times (for ncu captures).
   python tools/one_layer.py H W Cin Cout k stride pad res bn [iters] [batch]"""
import importlib.util, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("adapt_native", os.path.join(ROOT, "adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200", "_native.py"))
nat = importlib.util.module_from_spec(spec); spec.loader.exec_module(nat)
lib = nat.lib()
H, W, Cin, Cout, k, stride, pad, use_res, bn = [int(v) for v in sys.argv[1:10]]
iters = int(sys.argv[10]) if len(sys.argv) > 10 else 5
B = int(sys.argv[11]) if len(sys.argv) > 11 else 128
dev = torch.device("cuda:0")
x = torch.randn(B, H, W, Cin, device=dev).to(torch.bfloat16)
w = (torch.randn(Cout, k, k, Cin, device=dev) / (k * k * Cin) ** 0.5).to(torch.bfloat16)
scale = torch.ones(Cout, device=dev); bias = torch.zeros(Cout, device=dev)
Ho = (H + 2 * pad - k) // stride + 1; Wo = (W + 2 * pad - k) // stride + 1
out = torch.empty(B, Ho, Wo, Cout, device=dev, dtype=torch.bfloat16)
res = torch.randn(B, Ho, Wo, Cout, device=dev).to(torch.bfloat16) if use_res else None
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
for i in range(iters):
    flush.zero_()
    rc = lib.adapt_conv2d_bf16(0, x.data_ptr(), w.data_ptr(), scale.data_ptr(), bias.data_ptr(), res.data_ptr() if use_res else None,
                               out.data_ptr(), 0, B, H, W, Cin, Cout, k, k, stride, pad, 1, -1, bn, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.adapt_last_error()
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
