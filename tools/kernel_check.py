"""Developer harness (GPU): exercise adapt_conv2d_bf16 over shapes x A-modes x tile widths and
compare with an fp32 torch convolution of the same bf16-rounded operands.  Each group runs in
its own subprocess so that a faulting kernel cannot poison the other groups.

    python tools/kernel_check.py            # all groups
    python tools/kernel_check.py --group tiled
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200")


def load_native():
    spec = importlib.util.spec_from_file_location("adapt_native", os.path.join(PKG, "_native.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if os.environ.get("KC_LIB"):  # A/B against another build of the library
        mod.LIB_PATH = os.path.abspath(os.environ["KC_LIB"])
    return mod


# (name, B, H, W, Cin, Cout, k, stride, pad, residual, relu, out_f32)
CASES = {
    "tiled": [
        ("t64", 2, 14, 14, 64, 64, 1, 1, 0, 0, 0, 0),
        ("t_res_relu", 2, 14, 14, 64, 64, 1, 1, 0, 1, 1, 0),
        ("t256to128", 4, 28, 28, 256, 128, 1, 1, 0, 1, 1, 0),
        ("t128to512", 4, 28, 28, 128, 512, 1, 1, 0, 1, 1, 0),
        ("t_bigM", 32, 56, 56, 64, 256, 1, 1, 0, 0, 1, 0),
        ("t_k2048", 8, 7, 7, 2048, 512, 1, 1, 0, 0, 1, 0),
        ("fc", 128, 1, 1, 2048, 1000, 1, 1, 0, 0, 0, 1),
        ("fc_b1", 1, 1, 1, 2048, 1000, 1, 1, 0, 0, 0, 1),
    ],
    "im2col": [
        ("c3_14", 2, 14, 14, 64, 64, 3, 1, 1, 0, 1, 0),
        ("c3_28", 3, 28, 28, 128, 128, 3, 1, 1, 0, 1, 0),
        ("c3_7_straddle", 5, 7, 7, 512, 512, 3, 1, 1, 0, 1, 0),
        ("c3_56", 2, 56, 56, 64, 64, 3, 1, 1, 0, 1, 0),
        ("s2_56", 2, 56, 56, 256, 128, 1, 2, 0, 0, 1, 0),
        ("s2_28_512", 3, 28, 28, 512, 1024, 1, 2, 0, 0, 0, 0),
        ("s2_14", 5, 14, 14, 1024, 512, 1, 2, 0, 0, 1, 0),
        ("c7s2", 2, 32, 32, 64, 64, 7, 2, 3, 0, 1, 0),
        ("c3_small_b1", 1, 7, 7, 64, 64, 3, 1, 1, 0, 0, 0),
    ],
    "gather": [
        ("g_c3_14", 2, 14, 14, 64, 64, 3, 1, 1, 0, 1, 0),
        ("g_c3_7_straddle", 5, 7, 7, 512, 512, 3, 1, 1, 0, 1, 0),
        ("g_s2_56", 2, 56, 56, 256, 128, 1, 2, 0, 0, 1, 0),
        ("g_stem", 2, 224, 224, 8, 64, 7, 2, 3, 0, 1, 0),
        ("g_cin24", 2, 20, 20, 24, 64, 3, 1, 1, 1, 1, 0),
    ],
    # CTA-pair mode (run with ADAPT_PAIR=2): odd tile counts, partial last tiles, residuals
    "pair_tiled": [
        ("p_t_even", 8, 28, 28, 256, 256, 1, 1, 0, 0, 1, 0),
        ("p_t_res", 4, 28, 28, 128, 512, 1, 1, 0, 1, 1, 0),
        ("p_t_odd_tiles", 5, 14, 14, 256, 1024, 1, 1, 0, 1, 1, 0),
        ("p_t_k2048", 8, 7, 7, 2048, 512, 1, 1, 0, 0, 1, 0),
        ("p_t_bigM", 32, 56, 56, 64, 256, 1, 1, 0, 1, 1, 0),
    ],
    "pair_im2col": [
        ("p_c3_28", 16, 28, 28, 128, 128, 3, 1, 1, 0, 1, 0),
        ("p_c3_14_odd", 8, 14, 14, 256, 256, 3, 1, 1, 0, 1, 0),
        ("p_c3_7_straddle", 16, 7, 7, 512, 512, 3, 1, 1, 0, 1, 0),
        ("p_c3_7_b5", 5, 7, 7, 512, 512, 3, 1, 1, 1, 1, 0),
        ("p_s2_14", 16, 14, 14, 1024, 2048, 1, 2, 0, 0, 0, 0),
    ],
    # row-reuse mode for 3x3 / stride 1 / pad 1 (A_ROW3): rows straddling tiles and images, partial last tiles
    "row3": [
        ("r3_56_c64", 2, 56, 56, 64, 64, 3, 1, 1, 0, 1, 0),
        ("r3_28_c128", 3, 28, 28, 128, 128, 3, 1, 1, 0, 1, 0),
        ("r3_14_c256", 5, 14, 14, 256, 256, 3, 1, 1, 0, 1, 0),
        ("r3_7_c512", 5, 7, 7, 512, 512, 3, 1, 1, 0, 0, 0),
        ("r3_56_b32", 32, 56, 56, 64, 64, 3, 1, 1, 0, 1, 0),
        ("r3_28_b16", 16, 28, 28, 128, 128, 3, 1, 1, 0, 1, 0),
        ("r3_rect", 3, 20, 36, 64, 128, 3, 1, 1, 0, 1, 0),
        ("r3_b1", 1, 14, 14, 64, 64, 3, 1, 1, 0, 0, 0),
    ],
    # space-to-depth stem by column phase (A_S2DP): rows of whole 4-pixel units; tiles crossing rows and images
    "s2dp": [
        ("sp_small", 5, 20, 24, 16, 64, 4, 1, 0, 0, 1, 0),
        ("sp_b1", 1, 115, 116, 16, 64, 4, 1, 0, 0, 1, 0),
        ("sp_b3", 3, 115, 116, 16, 64, 4, 1, 0, 0, 0, 0),
        ("sp_narrow", 2, 9, 8, 16, 64, 4, 1, 0, 0, 1, 0),
    ],
}
MODE = {"tiled": 0, "im2col": 1, "gather": 2, "pair_tiled": 0, "pair_im2col": 1, "row3": 3, "s2dp": 4}


def run_group(group: str) -> int:
    import torch
    import torch.nn.functional as F

    if group.startswith("pair"):
        os.environ["ADAPT_PAIR"] = "2"
    nat = load_native()
    lib = nat.lib()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    worst = 0.0
    ok_all = True
    for (name, B, H, W, Cin, Cout, k, stride, pad, use_res, relu, out_f32) in CASES[group]:
        for bn in (64, 128, 256):
            if bn > 64 and Cout < bn:
                continue
            x = torch.randn(B, H, W, Cin, device=dev).to(torch.bfloat16)
            w = (torch.randn(Cout, k, k, Cin, device=dev) / (k * k * Cin) ** 0.5).to(torch.bfloat16)
            scale = torch.rand(Cout, device=dev) + 0.5
            bias = torch.randn(Cout, device=dev)
            Ho = (H + 2 * pad - k) // stride + 1
            Wo = (W + 2 * pad - k) // stride + 1
            res = torch.randn(B, Ho, Wo, Cout, device=dev).to(torch.bfloat16) if use_res else None
            out = torch.full((B, Ho, Wo, Cout), float("nan"), device=dev,
                             dtype=torch.float32 if out_f32 else torch.bfloat16)
            torch.cuda.synchronize()
            rc = lib.adapt_conv2d_bf16(0, x.data_ptr(), w.data_ptr(), scale.data_ptr(), bias.data_ptr(),
                                       res.data_ptr() if use_res else None, out.data_ptr(), out_f32, B, H, W, Cin,
                                       Cout, k, k, stride, pad, relu, MODE[group], bn, None)
            if rc != 0:
                print(json.dumps({"case": name, "bn": bn, "error": lib.adapt_last_error().decode()}), flush=True)
                ok_all = False
                continue
            try:
                torch.cuda.synchronize()
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"case": name, "bn": bn, "cuda_error": str(e)[:300]}), flush=True)
                return 2
            ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), stride=stride, padding=pad)
            ref = ref.permute(0, 2, 3, 1) * scale + bias
            if use_res:
                ref = ref + res.float()
            if relu:
                ref = ref.clamp_min(0)
            got = out.float()
            err = (got - ref).abs()
            tol = 2e-2 * ref.abs().clamp_min(1.0) if not out_f32 else 2e-3 * ref.abs().clamp_min(1.0)
            bad = int((~(err <= tol)).sum().item())
            mx = float(err.max().item()) if bad == 0 or not torch.isnan(err).any() else float("nan")
            worst = max(worst, mx if mx == mx else 1e9)
            rec = {"case": name, "bn": bn, "M": B * Ho * Wo, "max_abs_err": mx, "bad": bad,
                   "nan": int(torch.isnan(got).sum().item())}
            if bad:
                ok_all = False
                idx = torch.nonzero(~(err <= tol))[:4].tolist()
                rec["first_bad"] = idx
            print(json.dumps(rec), flush=True)
    print(json.dumps({"group": group, "ok": ok_all, "worst": worst}), flush=True)
    return 0 if ok_all else 1


def run_perf() -> int:
    import torch

    nat = load_native()
    lib = nat.lib()
    dev = torch.device("cuda:0")
    B = 128
    layers = [
        ("conv2_1x1_64_256", 56, 56, 64, 256, 1, 1, 0),
        ("conv2_1x1_256_64", 56, 56, 256, 64, 1, 1, 0),
        ("conv2_3x3_64", 56, 56, 64, 64, 3, 1, 1),
        ("conv3_3x3_128", 28, 28, 128, 128, 3, 1, 1),
        ("conv3_1x1_128_512", 28, 28, 128, 512, 1, 1, 0),
        ("conv4_3x3_256", 14, 14, 256, 256, 3, 1, 1),
        ("conv4_1x1_1024_256", 14, 14, 1024, 256, 1, 1, 0),
        ("conv4_1x1_256_1024", 14, 14, 256, 1024, 1, 1, 0),
        ("conv5_3x3_512", 7, 7, 512, 512, 3, 1, 1),
        ("conv5_1x1_512_2048", 7, 7, 512, 2048, 1, 1, 0),
        ("conv3_s2_256_128", 56, 56, 256, 128, 1, 2, 0),
        ("stem_gather", 224, 224, 8, 64, 7, 2, 3),
        ("stem_s2d", 115, 116, 16, 64, 4, 1, 0),
        ("conv2_3_res", 56, 56, 64, 256, 1, 1, 0, 1),
        ("conv3_3_res", 28, 28, 128, 512, 1, 1, 0, 1),
        ("conv4_3_res", 14, 14, 256, 1024, 1, 1, 0, 1),
        ("conv5_3_res", 7, 7, 512, 2048, 1, 1, 0, 1),
    ]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    only = os.environ.get("KC_ONLY")
    for lay in layers:
        (name, H, W, Cin, Cout, k, stride, pad) = lay[:8]
        if only and only not in name:
            continue
        use_res = len(lay) > 8 and lay[8]
        x = torch.randn(B, H, W, Cin, device=dev).to(torch.bfloat16)
        w = (torch.randn(Cout, k, k, Cin, device=dev) / (k * k * Cin) ** 0.5).to(torch.bfloat16)
        scale = torch.ones(Cout, device=dev)
        bias = torch.zeros(Cout, device=dev)
        Ho = (H + 2 * pad - k) // stride + 1
        Wo = (W + 2 * pad - k) // stride + 1
        out = torch.empty(B, Ho, Wo, Cout, device=dev, dtype=torch.bfloat16)
        res = torch.randn(B, Ho, Wo, Cout, device=dev).to(torch.bfloat16) if use_res else None
        only_bn = int(os.environ.get("KC_BN", "0"))
        iters = int(os.environ.get("KC_ITERS", "6"))
        for bn in (64, 128, 256):
            if bn > 64 and Cout < bn:
                continue
            if only_bn and bn != only_bn:
                continue
            times = []
            for it in range(iters):
                flush.zero_()
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                rc = lib.adapt_conv2d_bf16(0, x.data_ptr(), w.data_ptr(), scale.data_ptr(), bias.data_ptr(),
                                           res.data_ptr() if use_res else None,
                                           out.data_ptr(), 0, B, H, W, Cin, Cout, k, k, stride, pad, 1, -1, bn,
                                           torch.cuda.current_stream().cuda_stream)
                e1.record()
                torch.cuda.synchronize()
                if rc != 0:
                    print(json.dumps({"layer": name, "bn": bn, "error": lib.adapt_last_error().decode()}))
                    break
                if it >= min(2, iters - 1):
                    times.append(e0.elapsed_time(e1))
            if not times:
                continue
            ms = sum(times) / len(times)
            M = B * Ho * Wo
            flops = 2.0 * M * Cout * k * k * Cin
            bytes_ = 2.0 * (M * (stride * stride if k == 1 else 1) * Cin / (stride * stride) + M * Cout * (2 if use_res else 1)
                            + Cout * k * k * Cin)
            print(json.dumps({"layer": name, "bn": bn, "ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1),
                              "gbs": round(bytes_ / ms / 1e6, 1)}), flush=True)
    return 0


def run_trace() -> int:
    """One launch per selected perf layer with the role timeline enabled; prints per-launch medians (SM clocks)."""
    import statistics as st
    import tempfile

    trace_lib = os.path.join(PKG, "csrc", "libadapt_b200_trace.so")
    if not os.path.exists(trace_lib):
        print("build the timeline library first: ADAPT_BUILD_TRACE=1 python <pkg>/csrc/build.py", file=sys.stderr)
        return 2
    os.environ["KC_LIB"] = trace_lib
    tf = tempfile.NamedTemporaryFile(suffix=".jsonl", delete=False)
    tf.close()
    os.environ["ADAPT_TRACE_FILE"] = tf.name
    os.environ["KC_ITERS"] = "1"
    run_perf()
    ev = {"cta": 0, "p0": 1, "p1": 2, "m0": 3, "m1": 4, "m2": 5, "e0": 6, "e1": 7}
    for line in open(tf.name):
        r = json.loads(line)
        T, E, G = r["tiles"], r["events"], r["grid"]
        s = r["stamps"]
        at = lambda c, t, e: s[(c * T + t) * E + e]  # noqa: E731
        agg = {k: [] for k in ("prologue", "first_data", "mma_tile", "mma_gap", "epi_wait", "epi_tile", "span",
                               "prod_tile", "tiles")}
        for c in range(G):
            t0 = at(c, 0, ev["cta"])
            if t0 == 0:
                continue
            n = 0
            while n < T and at(c, n, ev["m2"]) != 0:
                n += 1
            if n == 0:
                continue
            agg["tiles"].append(n)
            agg["prologue"].append(at(c, 0, ev["p0"]) - t0)
            agg["first_data"].append(at(c, 0, ev["m1"]) - at(c, 0, ev["p0"]))
            last = t0
            for t in range(n):
                agg["mma_tile"].append(at(c, t, ev["m2"]) - at(c, t, ev["m1"]))
                agg["prod_tile"].append(at(c, t, ev["p1"]) - at(c, t, ev["p0"]))
                if t:
                    agg["mma_gap"].append(at(c, t, ev["m1"]) - at(c, t - 1, ev["m2"]))
                if at(c, t, ev["e0"]):
                    agg["epi_wait"].append(at(c, t, ev["e0"]) - at(c, t, ev["m2"]))
                    agg["epi_tile"].append(at(c, t, ev["e1"]) - at(c, t, ev["e0"]))
                    last = max(last, at(c, t, ev["e1"]))
            agg["span"].append(last - t0)
        med = {k: (round(st.median(v)) if v else None) for k, v in agg.items()}
        med["span_max"] = max(agg["span"]) if agg["span"] else None
        print(json.dumps({k: r[k] for k in ("grid", "block_n", "a_mode", "pair", "m_tiles", "n_tiles", "k_blocks", "stages")}
                         | med), flush=True)
    os.unlink(tf.name)
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--group", default=None)
    args = ap.parse_args()
    if args.group == "perf":
        return run_perf()
    if args.group == "trace":
        return run_trace()
    if args.group:
        return run_group(args.group)
    rc_all = 0
    groups = os.environ.get("KC_GROUPS", "tiled,im2col,gather,pair_tiled,pair_im2col,row3,s2dp,perf").split(",")
    for g in groups:
        t0 = time.time()
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--group", g], timeout=600)
        print(f"== group {g}: exit {p.returncode} in {time.time() - t0:.1f}s", flush=True)
        rc_all |= p.returncode
    return rc_all


if __name__ == "__main__":
    sys.exit(main())
