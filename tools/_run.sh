set -x
T=s17o
B="python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0"
for S in 4 6 3; do
timeout 300 $B --slots $S > gpurun_out/${T}_bench_s$S.json 2> gpurun_out/${T}_bench_s$S.err; echo rc=$?
done
tail -c 300 gpurun_out/${T}_bench*.err
