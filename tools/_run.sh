T=s19e
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo rc=$?
TCD_ENGINES=8 timeout 600 python bench.py --no-aux --rtdetr-batch 0 --no-cpu-baseline --steps 400 > gpurun_out/${T}_bench_e8.json 2> gpurun_out/${T}_bench_e8.err; echo rc=$?
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; echo rc=$?
tail -c 300 gpurun_out/${T}_bench*.err gpurun_out/${T}_smoke.log
