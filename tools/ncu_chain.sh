#!/bin/bash
# ncu full-set capture of ONE launch of each chain geometry (micro-benchmark tools/chain_bench.py), raw + source pages as CSV.
cd /root/repo
for C in 64 128; do
  timeout 60 python tools/chain_bench.py $C > gpurun_out/chain_plain_$C.log 2>&1 || exit 1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_chain -s 3 -c 1 -o /tmp/chain$C -f python tools/chain_bench.py $C > gpurun_out/ncu_chain_$C.log 2>&1
  echo "C=$C ncu rc=$?"
  ncu -i /tmp/chain$C.ncu-rep --page raw --csv > gpurun_out/r2_chain${C}_raw.csv 2>/dev/null
  ncu -i /tmp/chain$C.ncu-rep --page source --csv > gpurun_out/r2_chain${C}_source.csv 2>/dev/null
done
wc -c gpurun_out/r2_chain*_*.csv
