T=s22
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err; echo bench rc=$?
tail -c 400 gpurun_out/${T}_bench_n2.err
T=s21
timeout 300 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/${T}_bench.err
