#!/bin/bash
T=${1:-s72}
mkdir -p gpurun_out
timeout 600 python - <<'PY' 2>&1 | grep -v Warning | tail -8
import sys, json
sys.argv=["bench.py"]
import bench, torch, numpy as np
peaks = {"hbm_gbs": 6550.0}
pool = torch.empty((8, bench.FRAME_H, bench.FRAME_W, 3), dtype=torch.uint8)
for i in range(8): pool[i] = torch.from_numpy(bench.synth_frame(i))
dev = pool.cuda()
print(json.dumps(bench.jpeg_side(0, dev, peaks)))
print(json.dumps(bench.enhance_side(0, dev, peaks)))
print(json.dumps(bench.ingest_side(0)))
PY
