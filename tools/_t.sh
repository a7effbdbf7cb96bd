cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/gpu_tests.log | tail -1
KC_GROUPS=tiled,im2col,gather,pair_tiled,pair_im2col,row3,s2dp timeout 600 python tools/kernel_check.py 2>&1 | grep -E "\"group\"|error"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
