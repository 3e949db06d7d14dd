#!/bin/bash
T=${1:-s38}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|FAILED|Error|confident oracle" gpurun_out/${T}_pytest.log | head -40
: > gpurun_out/${T}_ab.jsonl
run() { echo "== $*" >> gpurun_out/${T}_ab.jsonl; env "$@" timeout 200 python bench.py --quick --steps 120 --warmup 10 ${BARGS} >> gpurun_out/${T}_ab.jsonl 2>> gpurun_out/${T}_ab.err; }
BARGS="" run TCD_ENGINES=3
BARGS="" run TCD_ENGINES=3 TCD_NO_BRANCH=1
BARGS="" run TCD_ENGINES=3 TCD_NO_EMIT=1
BARGS="" run TCD_ENGINES=4
BARGS="" run TCD_ENGINES=2
BARGS="" run TCD_ENGINES=4 TCD_CONV_SHARE=0
cat gpurun_out/${T}_ab.jsonl | cut -c1-200
tail -5 gpurun_out/${T}_ab.err
