T=s21
timeout 300 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench rc=$?
tail -c 300 gpurun_out/${T}_bench.err
