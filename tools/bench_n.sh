#!/bin/bash
# bench.py on N GPUs exactly as the driver launches it:  gpurun --gpus N -- 'bash tools/bench_n.sh N'
cd /root/repo
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"]["part_at"], d["timing"]["stage_ms_per_step"], d["timing"]["stage_busy_ms"], d["checks"]["mismatched_micro_batches"], d.get("failover"))
PY
