#!/bin/bash
# Capture recipe behind profiles/r1_*: plain run first, then the ncu launch list, then `ncu --set full` of one step's
# 54 conv_gemm launches (report kept in /tmp on the box, only CSV exports come back), then the per-op table.
# Run on a GPU box:  gpurun --timeout 2400 -- 'bash tools/profile.sh' ; then python tools/summarize_profiles.py r1 ...
cd /root/repo
set -x
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-uint8-leg > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ADAPT_BENCH_DRIVE_TIMEOUT_S=600 ADAPT_HEARTBEAT_TIMEOUT_S=600 ADAPT_CONFIG_TIMEOUT_S=600 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-uint8-leg > gpurun_out/prof_ncu1.log 2>&1
echo "launch list rc=$?"
ADAPT_BENCH_DRIVE_TIMEOUT_S=1500 ADAPT_HEARTBEAT_TIMEOUT_S=600 ADAPT_CONFIG_TIMEOUT_S=900 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s 162 -c 54 -o /tmp/r1full -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-uint8-leg > gpurun_out/prof_ncu2.log 2>&1
echo "full rc=$?"
ncu -i /tmp/r1full.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,launch__grid_size,launch__registers_per_thread,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,launch__cluster_dim_x > gpurun_out/r1_full_metrics.csv 2> gpurun_out/prof_export.err
wc -l gpurun_out/launches_r1.csv gpurun_out/r1_full_metrics.csv
timeout 600 python bench.py --dump-ops gpurun_out/ops_b128.json > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
tail -c 600 gpurun_out/bench_n1.json
