set -x
T=s18h
timeout 1200 python -m pytest tests/test_gpu_rtdetr_kernels.py tests/test_gpu_rtdetr.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo rc=$?
tail -12 gpurun_out/${T}_pytest.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --latency-iters 0 --profile-iters 1"
TCD_RTDETR_DUMP=gpurun_out/${T}_rtdetr_ops.json timeout 600 $B > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo rc=$?
tail -c 300 gpurun_out/${T}_bench*.err
