set -x
T=s18d
timeout 900 python -m pytest tests/test_gpu_yolox.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo rc=$?
tail -15 gpurun_out/${T}_pytest.log
