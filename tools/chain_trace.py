"""Role timeline of the chain kernel (developer build): one launch per geometry at batch 128, per-unit intervals in SM
clocks.   ADAPT_BUILD_TRACE=1 python <pkg>/csrc/build.py && python tools/chain_trace.py"""
import json
import os
import statistics as st
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from adapt_b200 import _native as N  # noqa: E402

N.LIB_PATH = os.path.join(os.path.dirname(N.LIB_PATH), "libadapt_b200_trace.so")
tf = tempfile.NamedTemporaryFile(suffix=".jsonl", delete=False)
tf.close()
os.environ["ADAPT_TRACE_FILE"] = tf.name
lib = N.lib()
B = int(os.environ.get("B", "128"))
for C in (64, 128):
    H = W = 56 if C == 64 else 28
    dev = "cuda:0"
    x = torch.randn(B, H, W, C, device=dev).bfloat16()
    w2 = (torch.randn(C, 3, 3, C, device=dev) / (3 * C ** 0.5)).bfloat16()
    w3 = (torch.randn(4 * C, C, device=dev) / C ** 0.5).bfloat16()
    s1, b1, s2, b2 = torch.ones(C, device=dev), torch.zeros(C, device=dev), torch.ones(4 * C, device=dev), torch.zeros(4 * C, device=dev)
    res = torch.randn(B, H, W, 4 * C, device=dev).bfloat16()
    out = torch.empty(B, H, W, 4 * C, device=dev, dtype=torch.bfloat16)
    for _ in range(2):
        N.check(lib.adapt_conv_chain_bf16(0, x.data_ptr(), w2.data_ptr(), s1.data_ptr(), b1.data_ptr(), w3.data_ptr(), s2.data_ptr(),
                                          b2.data_ptr(), res.data_ptr(), out.data_ptr(), B, H, W, C, None))
    torch.cuda.synchronize()
EV = dict(cta=0, p0=1, p1=2, m1f=3, m1d=4, a2=5, m2d=6, e1s=7, e1d=8, e2f=9, e2d=10, e2g1=11, c0s=12, c0t=13, c0r=14, c0m=15, c0o=16, c0e=17, c1e=18, c0y=19, c0x=20)
for line in list(open(tf.name))[1::2]:   # second (warm) launch of each geometry
    r = json.loads(line)
    T, E, G, s = r["tiles"], r["events"], r["grid"], r["stamps"]
    at = lambda c, t, e: s[(c * T + t) * E + EV[e]]  # noqa: E731
    agg = {}
    def add(k, v):
        agg.setdefault(k, []).append(v)
    spans = []
    for c in range(0, G, 2 if r["C"] == 128 else 1):   # leader CTAs hold the MMA stamps
        t0 = at(c, 0, "cta")
        n = 0
        while n < T and at(c, n, "e2d"):
            n += 1
        if not n:
            continue
        add("units", n)
        add("prologue_to_first_load", at(c, 0, "p0") - t0)
        add("first_load_to_data", at(c, 0, "m1f") - at(c, 0, "p0"))
        for t in range(n):
            add("mma1_issue", at(c, t, "m1d") - at(c, t, "m1f"))
            add("mma1_done_to_epi1_start", at(c, t, "e1s") - at(c, t, "m1d"))   # (epilogue 1 of unit t runs inside unit t-1's drain)
            add("epi1", at(c, t, "e1d") - at(c, t, "e1s"))
            add("epi1_done_to_a2_ready", at(c, t, "a2") - at(c, t, "e1d"))
            add("mma2_issue", at(c, t, "m2d") - at(c, t, "a2"))
            add("a2_ready_to_epi2_first", at(c, t, "e2f") - at(c, t, "a2"))
            add("epi2_g0", at(c, t, "e2d") - at(c, t, "e2f"))
            add("chunk0_tmem_ld", at(c, t, "c0t") - at(c, t, "c0s"))
            add("chunk0_residual_wait", at(c, t, "c0r") - at(c, t, "c0t"))
            add("chunk0_math", at(c, t, "c0m") - at(c, t, "c0r"))
            add("chunk0_write_out", at(c, t, "c0o") - at(c, t, "c0m"))
            add("chunk0_syncwarp", at(c, t, "c0y") - at(c, t, "c0o"))
            add("chunk0_expect_tx", at(c, t, "c0x") - at(c, t, "c0y"))
            add("chunk0_tma_issue", at(c, t, "c0e") - at(c, t, "c0x"))
            add("chunk1_total", at(c, t, "c1e") - at(c, t, "c0e"))
            add("epi2_g1_end_after_g0", at(c, t, "e2g1") - at(c, t, "e2d"))
            if t:
                add("unit_period", at(c, t, "e2d") - at(c, t - 1, "e2d"))
                add("epi1_start_before_prev_epi2_done", at(c, t - 1, "e2d") - at(c, t, "e1s"))
        spans.append(max(at(c, n - 1, "e2d"), at(c, n - 1, "e2g1")) - t0)
    print(json.dumps({"C": r["C"], "grid": G, "m_tiles": r["m_tiles"], "stages": r["stages"], "ring": r["ring"],
                      **{k: round(st.median(v)) for k, v in agg.items()}, "span_median": round(st.median(spans)), "span_max": max(spans)}))
os.unlink(tf.name)
