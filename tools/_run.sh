set -x
T=s18m
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo rc=$?
tail -4 gpurun_out/${T}_pytest.log
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-aux --latency-iters 0"
timeout 300 $B --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo rc=$?
timeout 300 $B > gpurun_out/${T}_bench_b.json 2> gpurun_out/${T}_bench_b.err; echo rc=$?
tail -c 300 gpurun_out/${T}_bench*.err
