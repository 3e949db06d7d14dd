"""Host-side profile of YOLOXDetector.detect_batch over many frames (where does the Python time go?)."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from oracle import synth_weights
from telescope_cam_detection_b200.yolox_detector import YOLOXDetector

sd = synth_weights.make_model(seed=0, obj_bias=-1.4).state_dict()
det = YOLOXDetector(state_dict=sd, max_batch=8, max_out=300)
assert det.load_model()
pool = torch.empty((32, 1080, 1920, 3), dtype=torch.uint8, pin_memory=True)
for i in range(32):
    pool[i] = torch.from_numpy(bench.synth_frame(1000 + i))
hn = pool.numpy()
many = [hn[i % 32] for i in range(320)]
det.detect_batch(many[:192])
t0 = time.perf_counter(); out = det.detect_batch(many); dt = time.perf_counter() - t0
print("detect_batch 320 frames: %.1f ms -> %.0f fps, %d dicts" % (dt * 1e3, 320 / dt, sum(len(o) for o in out)))
del out
import gc
gc.disable()
t0 = time.perf_counter(); out = det.detect_batch(many); dt = time.perf_counter() - t0
print("gc disabled: %.1f ms -> %.0f fps" % (dt * 1e3, 320 / dt))
gc.enable(); del out
pr = cProfile.Profile(); pr.enable(); out = det.detect_batch(many); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(18); print(s.getvalue()[:3500])
