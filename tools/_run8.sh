T=s23
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29548 bench.py --gpus 8 > gpurun_out/${T}_bench_n8.json 2> gpurun_out/${T}_bench_n8.err; echo n8 rc=$?
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 > gpurun_out/${T}_bench_n4.json 2> gpurun_out/${T}_bench_n4.err; echo n4 rc=$?
tail -c 300 gpurun_out/${T}_bench_n8.err
