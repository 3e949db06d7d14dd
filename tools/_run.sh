#!/bin/bash
T=${1:-f08}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/${T}_bench.json 2>/dev/null | sed -n 1,9p
timeout 600 python bench.py --steps 20 --warmup 5 --no-aux > gpurun_out/${T}_bench_k20.json 2>> gpurun_out/${T}_bench.err; echo "k20 rc=$?"
python tools/show_bench.py gpurun_out/${T}_bench_k20.json 2>/dev/null | sed -n 1,5p
