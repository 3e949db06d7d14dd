set -x
T=r01f
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-aux --rtdetr-batch 0 --profile-iters 1 --latency-iters 0"
timeout 300 $B > gpurun_out/${T}_prof_cmd.json 2> gpurun_out/${T}_prof_cmd.err; echo rc=$?
TCD_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_ncu.csv $B > gpurun_out/${T}_ncu1.log 2>&1; echo rc=$?
