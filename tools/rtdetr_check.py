"""Stage-by-stage comparison of the libtcd RT-DETR program against the transformers fp32 oracle
(run on the GPU box):  python tools/rtdetr_check.py [batch]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import rtdetr as oracle  # noqa: E402
from telescope_cam_detection_b200.rtdetr_engine import RTDetrEngine  # noqa: E402


def rel(got, ref):
    ref = np.asarray(ref, np.float32)
    return float(np.abs(got - ref).max() / (np.abs(ref).max() + 1e-12)), float(
        np.sqrt(((got - ref) ** 2).mean()) / (np.sqrt((ref ** 2).mean()) + 1e-12))


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    model = oracle.make_model(0)
    torch.manual_seed(1)
    x = torch.rand(B, 3, 640, 640)
    x = x.to(torch.bfloat16).float()  # the engine's input is bf16: give the oracle the same values
    t0 = time.time()
    ref = oracle.forward(model, x)
    print("oracle forward %.1f s" % (time.time() - t0))
    eng = RTDetrEngine(oracle.state_dict_numpy(model), batch=B, keep_taps=True)
    eng.set_input(x.numpy())
    eng.set_orig_sizes([[1920, 1080]] * B)
    eng.run(graph=False)
    eng.run(graph=True)

    def nchw(t):
        return t.permute(0, 2, 3, 1).numpy()

    for i in range(3):
        print("backbone.%d  max-rel %.4f rms-rel %.4f" % ((i,) + rel(eng.read_tap("backbone.%d" % (i + 1)), nchw(ref["backbone"][i]))))
    print("aifi        max-rel %.4f rms-rel %.4f" % rel(eng.read_tap("aifi"), nchw(ref["aifi"])))
    for i in range(3):
        print("encoder.%d   max-rel %.4f rms-rel %.4f" % ((i,) + rel(eng.read_tap("encoder.%d" % i), nchw(ref["encoder"][i]))))
    logits, boxes, topk = eng.raw_outputs()
    rt = ref["topk"].numpy()
    for b in range(B):
        common = np.intersect1d(topk[b], rt[b])
        print("image %d: top-k overlap %d / %d" % (b, len(common), rt.shape[1]))
        # compare the queries both selected (identified by their anchor index)
        mine = {int(a): i for i, a in enumerate(topk[b])}
        theirs = {int(a): i for i, a in enumerate(rt[b])}
        im = np.array([mine[int(a)] for a in common])
        it = np.array([theirs[int(a)] for a in common])
        for li in range(6):
            h_ref = ref["dec_hidden"][b, li].numpy()[it]
            h_got = eng.read_tap("dec%d" % li)[b, 0][im]
            print("   dec%d hidden rms-rel %.4f" % (li, rel(h_got, h_ref)[1]))
        s_got = 1 / (1 + np.exp(-logits[b][im]))
        s_ref = torch.sigmoid(ref["logits"][b]).numpy()[it]
        print("   scores max |d| %.4f   boxes max |d| %.5f (= %.2f px of 640)" % (
            np.abs(s_got - s_ref).max(), np.abs(boxes[b][im] - ref["pred_boxes"][b].numpy()[it]).max(),
            640 * np.abs(boxes[b][im] - ref["pred_boxes"][b].numpy()[it]).max()))
    labels, bx, sc = eng.outputs()
    rl, rb, rs = oracle.postprocess(ref["logits"], ref["pred_boxes"], torch.tensor([[1920, 1080]] * B))
    print("post: scores[0][:5]", sc[0][:5], "ref", rs[0][:5].numpy())
    print("post: labels[0][:10]", labels[0][:10], "ref", rl[0][:10].numpy())
    prof = eng.prog.profile(3)
    tot = sum(o["ms"] for o in prof)
    fl = sum(o["flops"] for o in prof)
    print("eager per-op sum %.3f ms for batch %d, conv/linear %.1f GFLOP" % (tot, B, fl / 1e9))
    t0 = time.time()
    for _ in range(10):
        eng.prog.run(True)
    eng.prog.sync()
    print("graph replay %.3f ms / batch" % ((time.time() - t0) * 100))
    top = sorted(prof, key=lambda o: -o["ms"])[:12]
    for o in top:
        print("   %-24s %.3f ms" % (o["name"], o["ms"]))


if __name__ == "__main__":
    main()
