cd /root/repo
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
echo "rc=$?"
tail -c 3000 gpurun_out/bench_n8.json
tail -5 gpurun_out/bench_n8.err
