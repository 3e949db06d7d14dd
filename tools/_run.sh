set -x
T=s16r
timeout 300 python -m pytest tests/test_gpu_motion.py tests/test_gpu_shims.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "rc=$?"
tail -4 gpurun_out/${T}_pytest.log
python tools/aux_probe.py 2>&1 | tail -3
