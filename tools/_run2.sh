set -x
T=s18p
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N > gpurun_out/${T}_n$N.json 2> gpurun_out/${T}_n$N.err; echo rc=$?
done
tail -c 200 gpurun_out/${T}_n*.err
