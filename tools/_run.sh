#!/bin/bash
T=${1:-s57}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_yolox.py tests/test_gpu_bench_config.py tests/test_gpu_layers.py -q -x > gpurun_out/${T}_py.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/${T}_py.log
timeout 300 python bench.py --quick --steps 60 --warmup 10 --dump-ops gpurun_out/${T}_ops.json > gpurun_out/${T}_q.json 2>gpurun_out/${T}.err; cut -c1-120 gpurun_out/${T}_q.json
python tools/show_ops.py gpurun_out/${T}_ops.json | grep -E "preproc|total"
timeout 300 python bench.py --quick --pre stretch --steps 60 --warmup 10 --dump-ops gpurun_out/${T}_ops_s.json > gpurun_out/${T}_qs.json 2>>gpurun_out/${T}.err; cut -c1-120 gpurun_out/${T}_qs.json
python tools/show_ops.py gpurun_out/${T}_ops_s.json | grep -E "preproc|total"
