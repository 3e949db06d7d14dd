#!/bin/bash
T=${1:-s35}
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/prim_bench tools/prim_bench.cu -lcuda && timeout 60 /tmp/prim_bench > gpurun_out/${T}_prim.txt 2>&1
grep -E "issuing|MMA issue" gpurun_out/${T}_prim.txt
: > gpurun_out/${T}_ab.jsonl
run() { echo "== $*" >> gpurun_out/${T}_ab.jsonl; env "$@" timeout 200 python bench.py --quick --steps 60 --warmup 10 ${BARGS} >> gpurun_out/${T}_ab.jsonl 2>> gpurun_out/${T}_ab.err; }
BARGS="" run TCD_CONV_PAIR=1 TCD_NO_PDL=1
BARGS="" run TCD_CONV_PAIR=1 TCD_PDL_LATE=1
BARGS="" run TCD_CONV_PAIR=0 TCD_PDL_LATE=1
BARGS="" run TCD_CONV_PAIR=0 TCD_NO_PDL=1
BARGS="" run TCD_CONV_PAIR=1 TCD_PDL_LATE=1 TCD_ENGINES=8
BARGS="" run TCD_CONV_PAIR=1 TCD_PDL_LATE=1 TCD_CONV_SHARE=0
BARGS="--batch 32" run TCD_CONV_PAIR=1 TCD_PDL_LATE=1
cat gpurun_out/${T}_ab.jsonl | cut -c1-200
tail -5 gpurun_out/${T}_ab.err
