cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-uint8-leg --dump-ops gpurun_out/ops_new.json > gpurun_out/b1.json 2> gpurun_out/b1.err; tail -3 gpurun_out/b1.err; python - <<PY
import json
d=json.loads(open("gpurun_out/b1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["conv_ms_in_step"], d["checks"]["mismatched_micro_batches"], d["checks"]["value_repeat"])
c=d["baseline_configs"]["config1_b200"]; print(c["value"], c["e2e"], c["p50_latency_ms"])
o=json.load(open("gpurun_out/ops_new.json"))
for p in o["ops"][-4:]: print(p["layers"][0], p["layers"][-1], round(p["ms"]*1000,1))
PY
