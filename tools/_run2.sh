set -x
T=s16n
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err; echo rc=$?
tail -5 gpurun_out/${T}_bench_n2.err
cat gpurun_out/${T}_bench_n2.json | cut -c1-600
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo rc=$?
cat gpurun_out/${T}_bench_ref.json | cut -c1-400
