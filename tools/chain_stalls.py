"""Summarise the ncu source-page capture of the chain kernel (tools/ncu_chain.sh): where the warps' stall samples fall.
   python tools/chain_stalls.py gpurun_out > profiles/r2_chain_stalls.md"""
import csv
import sys

d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
print("# Chain kernel: warp-stall samples by SASS instruction (`ncu --set full --import-source on`, one launch of "
      "`tools/chain_bench.py` per geometry, batch 128)\n")
for C in (64, 128):
    raw = list(csv.reader(open(f"{d}/r2_chain{C}_raw.csv")))
    m = dict(zip(raw[0], raw[2]))
    keys = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
    print(f"## C = {C}\n")
    print("| metric | value |\n|---|---:|")
    for k in keys:
        if k in m:
            print(f"| `{k}` | {m[k]} |")
    tot = {k[len("smsp__pcsamp_warps_issue_stalled_"):]: float(v) for k, v in m.items()
           if k.startswith("smsp__pcsamp_warps_issue_stalled_") and not k.endswith("_not_issued") and v}
    n = sum(tot.values())
    print("\nStall reasons (all sampled warps, incl. the producer / MMA warps' barrier polls): "
          + ", ".join(f"{k} {100 * v / n:.0f} %" for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:7]) + "\n")
    rows = list(csv.reader(open(f"{d}/r2_chain{C}_source.csv")))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    total = sum(int(r[col["# Samples"]] or 0) for r in data)
    print(f"Top instructions of {total} samples:\n\n| samples | SASS | stall reasons |\n|---:|---|---|")
    for r in sorted(data, key=lambda r: -int(r[col["# Samples"]] or 0))[:22]:
        st = {k[6:]: int(r[col[k]] or 0) for k in hdr if k.startswith("stall_") and "(" not in k}
        st = ", ".join(f"{k} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v)
        print(f"| {r[col['# Samples']]} | `{r[col['Source']].strip()[:80]}` | {st} |")
    print()
