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


def test_species_classifier_batched_equals_per_crop():
    """stand-in classifier body (random-init conv net; the timm body is 'next' row N2): the batched
    gather + single forward must give what the reference's one-crop-at-a-time loop gives"""
    import torch
    from telescope_cam_detection_b200.species_classifier import SpeciesClassifier
    from telescope_cam_detection_b200.two_stage_pipeline_yolox import TwoStageDetectionPipeline
    torch.manual_seed(0)
    body = torch.nn.Sequential(torch.nn.Conv2d(3, 16, 7, 4), torch.nn.ReLU(), torch.nn.AdaptiveAvgPool2d(4),
                               torch.nn.Flatten(), torch.nn.Linear(256, 50))
    with torch.no_grad():
        body[-1].weight.mul_(30.0)   # peaky softmax so several confidence bands occur
    clf = SpeciesClassifier(device="cuda:0", input_size=224, model_factory=lambda n: body)
    assert clf.load_model(num_classes=50)
    clf.taxonomy = {str(i): {"common_name": "sp%d" % i, "genus": "g%d" % (i % 7), "family": "f%d" % (i % 3),
                             "order": "o", "class": "c"} for i in range(50)}
    frame = structured_frame(8)
    boxes = [[100, 100, 400, 300], [0, 0, 224, 224], [1500, 700, 1920, 1080], [640, 360, 1000, 900]]
    batched = clf.classify_crops(frame, boxes, top_k=3)
    for b, res in zip(boxes, batched):
        single = clf.classify(frame[b[1]:b[3], b[0]:b[2]], top_k=3)
        assert [(r["class_id"], r["species"], r["taxonomic_level"]) for r in res] == \
               [(r["class_id"], r["species"], r["taxonomic_level"]) for r in single]
        assert all(abs(x["confidence"] - y["confidence"]) < 1e-5 for x, y in zip(res, single))
    assert clf.preprocess(frame[100:300, 100:400]).shape == (1, 3, 224, 224)
    pipe = TwoStageDetectionPipeline(min_crop_size=32)
    pipe.add_species_classifier("bird", clf)
    dets = [{"class_id": 14, "class_name": "bird", "confidence": 0.9,
             "bbox": {"x1": 100.0, "y1": 100.0, "x2": 400.0, "y2": 300.0}},
            {"class_id": 0, "class_name": "person", "confidence": 0.9,
             "bbox": {"x1": 10.0, "y1": 10.0, "x2": 200.0, "y2": 300.0}}]
    out = pipe.process_detections(torch.from_numpy(frame).cuda(), dets)      # GPU-tensor frames are part of the API
    assert out[0]["stage2_category"] == "bird" and "species_confidence" in out[0]
    assert out[1]["species"] is None and "stage2_category" not in out[1]
