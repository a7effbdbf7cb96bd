#!/bin/bash
# Capture recipe behind profiles/r2_*: plain run first, then the ncu launch list, then `ncu --set full` of one step's
# conv-class launches (conv_gemm, conv_chain, stem_pool; report kept in /tmp on the box, only CSV exports come back), then the per-op table.
# Run on a GPU box:  gpurun --timeout 2400 -- 'bash tools/profile.sh' ; then
#   python tools/summarize_profiles.py r2 gpurun_out/launches_r2.csv gpurun_out/r2_full_metrics.csv gpurun_out/r2_ops_b128.json
cd /root/repo
set -x
F="--steps 4 --warmup 3 --ramp-s 0 --no-cpu-baseline --no-uint8-leg --no-baseline-configs --sustained-s 0.01"
python bench.py $F > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
NL=$(python - <<PY
import json
d=json.loads(open("gpurun_out/prof_plain.json").read().strip().splitlines()[-1])
print(d["gpu_launches"]//d["steps"], d["roofline"]["launches_per_step"])
PY
)
set -- $NL
PER_STEP=$1; CONV=$2
echo "launches per step: $PER_STEP, conv launches: $CONV"
export ADAPT_BENCH_DRIVE_TIMEOUT_S=1500 ADAPT_HEARTBEAT_TIMEOUT_S=600 ADAPT_CONFIG_TIMEOUT_S=900
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_r2.csv python bench.py $F > gpurun_out/prof_ncu1.log 2>&1
echo "launch list rc=$?"
# skip the two warm-up passes of configuration (plain run + graph warm-up) and the 3 warm-up steps: conv launches only
SKIP=$((CONV * 5))
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:conv_|stem_pool' -s $SKIP -c $CONV -o /tmp/r2full -f python bench.py $F > gpurun_out/prof_ncu2.log 2>&1
echo "full rc=$?"
ncu -i /tmp/r2full.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,launch__grid_size,launch__registers_per_thread,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,launch__cluster_dim_x > gpurun_out/r2_full_metrics.csv 2> gpurun_out/prof_export.err
wc -l gpurun_out/launches_r2.csv gpurun_out/r2_full_metrics.csv
unset ADAPT_BENCH_DRIVE_TIMEOUT_S ADAPT_HEARTBEAT_TIMEOUT_S ADAPT_CONFIG_TIMEOUT_S
timeout 600 python bench.py --steps 20 --warmup 5 --dump-ops gpurun_out/r2_ops_b128.json > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
tail -c 600 gpurun_out/bench_n1.json
