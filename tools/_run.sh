set -x
timeout 600 python tools/rtdetr_check.py 2 > gpurun_out/s16j_rtdetr.txt 2>&1; echo rc=$?
tail -60 gpurun_out/s16j_rtdetr.txt
