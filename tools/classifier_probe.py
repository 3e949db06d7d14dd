"""bench.py's Stage-2 classifier side measurement alone: python tools/classifier_probe.py [batch] > out.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    name = sys.argv[2] if len(sys.argv) > 2 else "timm/eva02_large_patch14_clip_336.merged2b_ft_inat21"
    print(json.dumps(bench.classifier_side(0, batch, bench.measured_peaks(), model_name=name, through_class=len(sys.argv) <= 3)))
