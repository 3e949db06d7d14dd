"""Small pass over every kernel family through the C ABI, for compute-sanitizer:
  compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_target.py
Shapes are the smallest ones that still take every code path (generic / halo / pair / grouped / f32-row convs,
preprocess, decode + sort + NMS + class filter, stream bank, crop gather)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from telescope_cam_detection_b200 import engine
from util import structured_frame, oracle_blob, motion_sequence

dev = "cuda:0"
g = torch.Generator().manual_seed(0)


def conv(N, H, W, cin, cout, k, stride, groups=1, resid=False):
    x = torch.randn(N, H, W, groups * cin, generator=g).to(dev).to(torch.bfloat16)
    w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).float()
    b = torch.zeros(cout)
    y = torch.zeros((N, H // stride, W // stride, cout), device=dev, dtype=torch.bfloat16)
    engine.conv2d_nhwc_bf16(x.data_ptr(), N, H, W, groups * cin, 0, cin, w.numpy(), b.numpy(), k, stride, 1,
                            y.data_ptr(), cout, 0, resid_ptr=y.data_ptr() if resid else 0, r_ctot=cout, r_coff=0,
                            groups=groups)
    torch.cuda.synchronize()
    print("conv", N, H, W, cin, cout, k, stride, groups, float(y.float().abs().mean()))


conv(1, 16, 16, 64, 64, 1, 1)            # generic 1x1
conv(1, 32, 32, 32, 32, 3, 1, resid=True)  # halo form, resident weights, residual
conv(1, 32, 32, 32, 64, 3, 2)            # stride 2 (parity views)
conv(2, 40, 40, 128, 128, 3, 1)          # CTA-pair kernel
conv(1, 40, 40, 128, 256, 3, 1, groups=2)  # CTA-pair kernel, grouped
conv(1, 16, 16, 128, 256, 3, 1, groups=2)  # generic, grouped
eng = engine.DetectorEngine(oracle_blob(0), input_size=(640, 640), max_batch=2, max_out=300, pre_mode=1)
frames = [structured_frame(1, 360, 640), structured_frame(2, 360, 640)]
eng.set_class_filter(2, {None: 0.3, 0: 0.5}, {16: 500.0}, {15: 90000.0}, min_area=100.0, max_det=40, allowed_classes=[0, 14, 15, 16, 21])
rows = eng.detect_batch(frames)
eng.submit(0, frames, [2, 0]); eng.submit(1, frames[:1]); r0 = eng.wait(0); r1 = eng.wait(1)
print("detect", [r.shape for r in rows], [r.shape for r in r0], [r.shape for r in r1])
pred = eng.forward(frames[:1])
print("forward", pred.shape, float(np.abs(pred).mean()))
print("postprocess", [r.shape for r in eng.postprocess(pred)])
eng.close()
bank = engine.StreamBankEngine(2, 480, 640)
seq = motion_sequence(n=4)
for t, (f, boxes) in enumerate(seq):
    bank.upload([f, f if t % 2 else None])
    gate = bank.gate(25)
    rois = np.array([[10, 10, 200, 150], [300, 200, 640, 480], [0, 0, 5, 5]], np.int32)
    offs = np.array([0, 3, 3 + (3 if t % 2 else 0)], np.int32)
    cnt = bank.motion(np.concatenate([rois, rois])[:offs[-1]], offs)
    print("bank", gate.tolist(), cnt.tolist())
bank.close()
m = engine.Mog2Engine(240, 320); e = engine.EmptyGateEngine(240, 320)
f = structured_frame(3, 240, 320)
m.apply(f, repeats=2); print("mog2", m.roi_counts(np.array([[0, 0, 320, 240]], np.int32))[0].tolist(), e.step(f)[0])
m.close(); e.close()
dst = torch.empty((3, 224, 224, 4), dtype=torch.bfloat16, device=dev)
fr = torch.from_numpy(structured_frame(4, 480, 640)).to(dev)
engine.crop_resize(fr.data_ptr(), 480, 640, 640 * 3, np.array([[0, 0, 100, 80], [50, 60, 640, 480], [300, 300, 364, 364]], np.int32), 224, dst.data_ptr(), layout=1)
torch.cuda.synchronize()
print("crop", float(dst.float().abs().mean()))
print("SANITIZE_TARGET_DONE")
