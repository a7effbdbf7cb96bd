cd /root/repo
KC_GROUPS=tiled,im2col,pair_im2col,row3 timeout 600 python tools/kernel_check.py 2>&1 | grep -E "group|error|\"bad\": [1-9]"
KC_BN=128 KC_ONLY=conv3 timeout 300 python tools/kernel_check.py --group trace 2>&1 | grep grid | cut -c1-330
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1]); print(d["value"], d["ms_per_step"], d["e2e"]["value"])
PY
