"""Runs the motion-gate / MOG2 / crop / preprocess kernels a few times on 1080p frames (profiling target):
   ncu --set full -k regex:blur_kernel|mog2|roi_count|crop|preproc python tools/aux_probe.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import synth_frame  # noqa: E402
from telescope_cam_detection_b200.engine import EmptyGateEngine, Mog2Engine  # noqa: E402

frames = [torch.from_numpy(synth_frame(i)).cuda() for i in range(6)]
g = EmptyGateEngine(1080, 1920)
m = Mog2Engine(1080, 1920)
for i in range(6):
    g.step(frames[i])
    m.apply(frames[i])
    m.roi_counts(np.array([[100, 100, 500, 400], [0, 0, 1920, 1080]], np.int32))
    print(i, g.last_ms(), m.last_ms())
