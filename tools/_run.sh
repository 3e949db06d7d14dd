set -x
T=s17f
B="python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0"
for S in 0 2; do
TCD_CONV_SHARE=$S timeout 300 $B --trace gpurun_out/${T}_trace_share$S.json > gpurun_out/${T}_bench_share$S.json 2> gpurun_out/${T}_bench_share$S.err; echo rc=$?
done
tail -c 400 gpurun_out/${T}_bench*.err
