#!/bin/bash
# A/B of the chain kernel's shared-memory split and L2 prefetch:  gpurun -- 'bash tools/chain_sweep.sh'
cd /root/repo
for cfg in "2 0 6" "2 1 6" "3 1 6" "3 0 6" "4 1 6"; do
  set -- $cfg
  ADAPT_CHAIN_RING=$1 ADAPT_CHAIN_L2PF=$2 ADAPT_CHAIN_STAGES=$3 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-uint8-leg --no-baseline-configs --dump-ops gpurun_out/ops_chain_$1_$2.json > gpurun_out/bench_chain_$1_$2.json 2> gpurun_out/bench_chain_$1_$2.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_chain_$1_$2.json").read().strip().splitlines()[-1])
ops=json.load(open("gpurun_out/ops_chain_$1_$2.json"))["ops"]
ch=[round(o["ms"]*1e3,1) for o in ops if o["kind"]==7]
print("ring $1 l2pf $2:", d["value"], d["ms_per_step"], d["roofline"]["conv_ms_in_step"], ch, d["checks"]["mismatched_micro_batches"])
PY
done
