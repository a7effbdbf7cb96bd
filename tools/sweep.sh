#!/bin/bash
# Micro-batch sweep of bench.py: tools/sweep.sh N "B1 B2 ..." [extra bench args]
# Appends one JSON line per micro-batch to gpurun_out/sweep_n$N.jsonl.
N=$1; shift
BATCHES=$1; shift
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for B in $BATCHES; do
  if [ "$N" = "1" ]; then
    timeout 600 python bench.py --gpus 1 --batch $B --steps 30 --warmup 5 --no-cpu-baseline --no-failover "$@" \
      >> gpurun_out/sweep_n$N.jsonl 2> gpurun_out/sweep_n${N}_b$B.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --batch $B --steps 30 --warmup 5 --no-cpu-baseline --no-failover "$@" \
      >> gpurun_out/sweep_n$N.jsonl 2> gpurun_out/sweep_n${N}_b$B.err
  fi
  echo "N=$N B=$B rc=$?"
done
python - <<PY
import json
for l in open("gpurun_out/sweep_n$N.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    e = d["e2e"]
    print(d["n_gpus"], d["config"]["micro_batch"], round(d["value"]), d["ms_per_step"], round(e["value"]), round(e.get("p50_latency_ms", 0), 3),
          d["config"]["part_at"])
PY
