set -x
T=s18f
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for N in 8 4 2; do
timeout 600 $TR --nproc-per-node $N --master-port 2951$N bench.py --gpus $N --no-cpu-baseline > gpurun_out/${T}_n$N.json 2> gpurun_out/${T}_n$N.err; echo rc=$?
done
tail -c 300 gpurun_out/${T}_n8.err
