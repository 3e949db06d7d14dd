"""GPU parity of the Stage-2 crop-and-resize gather (K13) against the reference's ops
(src/two_stage_pipeline_yolox.py:289 slice + src/species_classifier.py:311-350 preprocess),
restated with the same torch calls on the CPU.  fp32, tolerance 2e-5 (mul vs div by std)."""
import numpy as np
import pytest

from util import structured_frame

pytestmark = pytest.mark.gpu
MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


def _ref_crop(frame, box, S):
    import torch
    x1, y1, x2, y2 = box
    img = torch.from_numpy(frame[y1:y2, x1:x2])[:, :, [2, 1, 0]]
    if img.shape[0] != S or img.shape[1] != S:
        img = torch.nn.functional.interpolate(img.permute(2, 0, 1)[None].float(), size=(S, S), mode="bilinear",
                                              align_corners=False)[0]
    else:
        img = img.permute(2, 0, 1).float()
    img = img / 255.0
    return (img - torch.tensor(MEAN).view(3, 1, 1)) / torch.tensor(STD).view(3, 1, 1)


@pytest.mark.parametrize("S", [224, 336])
def test_crop_resize_matches_reference_ops(S):
    import torch
    from telescope_cam_detection_b200 import engine
    frame = structured_frame(3)
    rng = np.random.default_rng(S)
    boxes = [[0, 0, 1920, 1080], [100, 200, 100 + S, 200 + S], [1900, 1060, 1920, 1080], [5, 5, 6, 6],
             [300, 100, 364, 900]]
    for _ in range(59):  # log-uniform sizes 64..600 like BASELINE config 5
        w, h = np.exp(rng.uniform(np.log(64), np.log(600), 2)).astype(int)
        x1, y1 = rng.integers(0, 1920 - w), rng.integers(0, 1080 - h)
        boxes.append([x1, y1, x1 + w, y1 + h])
    dev = torch.from_numpy(frame).cuda()
    out = torch.empty(len(boxes), 3, S, S, device="cuda", dtype=torch.float32)
    engine.crop_resize(dev.data_ptr(), 1080, 1920, dev.stride(0), boxes, S, out.data_ptr(), layout=0)
    out16 = torch.empty(len(boxes), S, S, 4, device="cuda", dtype=torch.bfloat16)
    engine.crop_resize(dev.data_ptr(), 1080, 1920, dev.stride(0), boxes, S, out16.data_ptr(), layout=1)
    torch.cuda.synchronize()
    got = out.cpu()
    for i, b in enumerate(boxes):
        ref = _ref_crop(frame, b, S)
        assert (got[i] - ref).abs().max().item() <= 2e-5, "crop %d %s" % (i, b)
    assert torch.equal(out16[..., :3].permute(0, 3, 1, 2).cpu(), got.to(torch.bfloat16))
    assert bool((out16[..., 3] == 0).all())
