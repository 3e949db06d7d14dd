set -x
T=s17z
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --latency-iters 0 --profile-iters 1"
TCD_RTDETR_SPLIT=2 timeout 600 $B > gpurun_out/${T}_bench_sp2.json 2> gpurun_out/${T}_bench_sp2.err; echo rc=$?
TCD_RTDETR_SPLIT=4 timeout 600 $B > gpurun_out/${T}_bench_sp4.json 2> gpurun_out/${T}_bench_sp4.err; echo rc=$?
TCD_RTDETR_SPLIT=2 TCD_PROGRAM_SHARE=0 timeout 600 $B > gpurun_out/${T}_bench_sp2s0.json 2> gpurun_out/${T}_bench_sp2s0.err; echo rc=$?
tail -c 300 gpurun_out/${T}_bench*.err
