set -x
T=s18b
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0"
for X in 0 20 8 4; do
TCD_CONV_SHARE_MAX_TILES=$X timeout 300 $B > gpurun_out/${T}_bench_x$X.json 2> gpurun_out/${T}_bench_x$X.err; echo rc=$?
done
tail -c 300 gpurun_out/${T}_bench*.err
