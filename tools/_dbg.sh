cd /root/repo
python adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200/csrc/build.py --force > /dev/null 2>&1
for v in 2 0; do echo "== L2PF=$v"; B=2 ADAPT_CHAIN_L2PF=$v timeout 120 python tools/chain_bench.py 64 2>&1 | tail -1; done
