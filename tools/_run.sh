set -x
T=s16i
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
B="python bench.py --steps 100 --warmup 10 --latency-iters 10 --profile-iters 5"
timeout 300 $B --trace gpurun_out/${T}_trace.json --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo rc=$?
TCD_CONV_HALO=2 timeout 200 $B --no-cpu-baseline > gpurun_out/${T}_bench_halo_all.json 2>> gpurun_out/${T}_bench.err
tail -3 gpurun_out/${T}_bench.err
