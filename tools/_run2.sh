set -x
T=s17r
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus 2 --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0 --no-gather > gpurun_out/${T}_nogather.json 2> gpurun_out/${T}_nogather.err; echo rc=$?
timeout 600 $TR --master-port 29512 bench.py --gpus 2 --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0 > gpurun_out/${T}_gather.json 2> gpurun_out/${T}_gather.err; echo rc=$?
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0 > gpurun_out/${T}_n1_a.json 2> gpurun_out/${T}_n1_a.err &
CUDA_VISIBLE_DEVICES=1 timeout 300 python bench.py --no-cpu-baseline --no-aux --rtdetr-batch 0 --latency-iters 0 > gpurun_out/${T}_n1_b.json 2> gpurun_out/${T}_n1_b.err &
wait
nproc; free -g | head -2
