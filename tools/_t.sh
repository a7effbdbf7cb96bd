cd /root/repo
for i in 1 2; do
ADAPT_MB_STAGE_CAP=1 timeout 300 python tools/kernel_check.py --group perf > gpurun_out/cap1_$i.log 2>&1
ADAPT_MB_STAGE_CAP=0 timeout 300 python tools/kernel_check.py --group perf > gpurun_out/cap0_$i.log 2>&1
done
