set -x
timeout 900 python -m pytest tests/test_gpu_rtdetr.py -q > gpurun_out/s16l_pytest.log 2>&1; echo rc=$?
tail -40 gpurun_out/s16l_pytest.log
timeout 600 python tools/rtdetr_check.py 2 > gpurun_out/s16l_rtdetr.txt 2>&1
grep -E "overlap|scores max" gpurun_out/s16l_rtdetr.txt
