"""Turn the raw gpurun_out/ artefacts into the tracked summaries under profiles/.

  python tools/summarize_profiles.py r2 gpurun_out/launches_r2.csv gpurun_out/r2_full_metrics.csv gpurun_out/r2_ops_b128.json
"""
import collections
import csv
import json
import re
import shutil
import sys

tag, launches_csv, full_csv, ops_json = sys.argv[1:5]
out = [f"# profiles/{tag}: pipelined ResNet-50, 1 x B200, micro-batch 128, bf16\n"]

# ---- launch list (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised)
rows = [r for r in csv.reader(open(launches_csv)) if len(r) > 5]
hdr, data = rows[0], rows[1:]
i_name, i_val = hdr.index("Kernel Name"), hdr.index("Metric Value")
names = [r[i_name] for r in data]
starts = [i for i, n in enumerate(names) if "ingest" in n]
step = data[starts[1]:starts[2]]
agg, cnt = collections.Counter(), collections.Counter()
for r in step:
    n = re.sub(r"\(.*", "", r[i_name]).replace("void adapt::", "").replace("void ", "")
    agg[n] += float(r[i_val].replace(",", ""))
    cnt[n] += 1
tot = sum(agg.values())
out.append(f"## Launch list of one step (`{launches_csv}`, command `python bench.py --steps 4 --warmup 3 --no-cpu-baseline`)\n")
out.append(f"{len(step)} launches per step, {tot / 1e3:.1f} us summed device time (ncu-serialised).\n")
out.append("| kernel | launches | us | share |\n|---|---:|---:|---:|")
for n, v in agg.most_common():
    out.append(f"| `{n}` | {cnt[n]} | {v / 1e3:.1f} | {100 * v / tot:.1f}% |")
conv = sum(v for n, v in agg.items() if "conv_gemm" in n or "conv_chain" in n or "stem_pool" in n)
out.append(f"\n`conv_gemm_kernel` + `conv_chain_kernel` + `stem_pool_kernel` (all instantiations): **{100 * conv / tot:.1f}%** of the step.\n")
shutil.copy(launches_csv, f"profiles/{tag}_launches.csv")

# ---- ncu --set full capture of conv_gemm launches inside the same command
rows = list(csv.reader(open(full_csv)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
out.append(f"## `ncu --set full` capture of the conv_gemm_kernel / conv_chain_kernel / stem_pool_kernel launches of one step (`{full_csv}`)\n")
out.append("| # | instantiation | grid | regs | us | DRAM read MB | DRAM write MB | DRAM % of peak | tensor pipe % |\n"
           "|---|---|---:|---:|---:|---:|---:|---:|---:|")
for r in data:
    inst = re.search(r"conv_(?:gemm|chain)_kernel<[^>]*>|stem_pool_kernel", r[col["Kernel Name"]]).group(0)
    out.append(f"| {r[col['ID']]} | `{inst}` | {r[col['launch__grid_size']]} | {r[col['launch__registers_per_thread']]} | "
               f"{float(r[col['gpu__time_duration.sum']]):.1f} | {float(r[col['dram__bytes_read.sum']]):.1f} | "
               f"{float(r[col['dram__bytes_write.sum']]):.1f} | "
               f"{float(r[col['gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']]):.1f} | "
               f"{float(r[col['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed']]):.1f} |")
shutil.copy(full_csv, f"profiles/{tag}_conv_full_metrics.csv")

# ---- per-launch CUDA-event table from bench.py --dump-ops (not under a profiler)
ops = json.load(open(ops_json))
out.append(f"\n## Per-launch CUDA-event timings (`bench.py --dump-ops`, not under a profiler; batch {ops['batch']})\n")
out.append("| layers | H,W,Cin,Cout,k,s | ms | TFLOP/s (algorithmic) |\n|---|---|---:|---:|")
for o in ops["ops"]:
    tf = 2 * o["macs"] / o["ms"] / 1e9 if o["macs"] else 0
    out.append(f"| {o['layers'][0]} .. {o['layers'][-1]} | {o['shape']} | {o['ms']:.4f} | {tf:.1f} |")
tot_ms = sum(o["ms"] for o in ops["ops"])
conv_ms = sum(o["ms"] for o in ops["ops"] if o["kind"] in (1, 7, 8))
flop = 2 * sum(o["macs"] for o in ops["ops"])
out.append(f"\nsum {tot_ms:.3f} ms (isolated launches: no overlap between consecutive layers); conv kernels {conv_ms:.3f} ms = "
           f"{100 * conv_ms / tot_ms:.1f}% -> {flop / conv_ms / 1e9:.1f} TFLOP/s algorithmic over the conv launches.  Inside "
           f"their own CUDA graph (PDL overlap kept) the same conv launches take {ops.get('conv_ms_in_graph', float('nan')):.3f} ms.\n")
shutil.copy(ops_json, f"profiles/{tag}_ops_b128.json")
open(f"profiles/{tag}_summary.md", "w").write("\n".join(out) + "\n")
print("\n".join(out[:14]))
