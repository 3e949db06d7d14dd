set -x
T=s18a
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 > gpurun_out/${T}_n8.json 2> gpurun_out/${T}_n8.err; echo rc=$?
timeout 600 $TR --nproc-per-node 4 --master-port 29512 bench.py --gpus 4 > gpurun_out/${T}_n4.json 2> gpurun_out/${T}_n4.err; echo rc=$?
timeout 600 $TR --nproc-per-node 8 --master-port 29513 bench.py --gpus 8 --impl reference --steps 10 --warmup 3 > gpurun_out/${T}_ref_n8.json 2> gpurun_out/${T}_ref_n8.err; echo rc=$?
nproc; free -g | head -2
tail -c 600 gpurun_out/${T}_n8.err
