cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for r in 1 2; do timeout 60 python tools/chain_bench.py 2>&1 | tail -2; done
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-uint8-leg --no-baseline-configs --dump-ops gpurun_out/ops_new.json > gpurun_out/b1.json 2> gpurun_out/b1.err; python - <<PY
import json
d=json.loads(open("gpurun_out/b1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["conv_ms_in_step"], d["roofline"]["isolated_launch_sum_ms"], d["checks"]["mismatched_micro_batches"], d["checks"]["value_repeat"])
PY
