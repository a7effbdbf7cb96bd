cd /root/repo
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/gpu_tests.log
KC_GROUPS=row3 timeout 600 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/kernel_check.py --group row3 > gpurun_out/san_row3.log 2>&1; echo "memcheck row3 rc=$?"; tail -4 gpurun_out/san_row3.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/kernel_check.py --group pair_im2col > gpurun_out/san_pair.log 2>&1; echo "memcheck pair rc=$?"; tail -4 gpurun_out/san_pair.log
