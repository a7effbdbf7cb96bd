cd /root/repo
for m in off auto auto; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu-baseline --ingest-rails $m > gpurun_out/b2_$m.json 2> gpurun_out/b2_$m.err
python - <<PY
import json
d=json.loads(open("gpurun_out/b2_$m.json").read().strip().splitlines()[-1])
print("$m", d["value"], d["e2e"]["value"], d["failover"])
PY
done
