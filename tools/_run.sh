set -x
T=s17w
timeout 1200 python -m pytest tests/test_gpu_yolox.py tests/test_gpu_shims.py tests/test_gpu_crop.py tests/test_gpu_motion.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo rc=$?
tail -5 gpurun_out/${T}_pytest.log
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --rtdetr-batch 0 --latency-iters 0"
timeout 300 $B --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo rc=$?
timeout 300 $B --pre stretch --no-aux > gpurun_out/${T}_bench_stretch.json 2> gpurun_out/${T}_bench_stretch.err; echo rc=$?
tail -c 300 gpurun_out/${T}_bench*.err
