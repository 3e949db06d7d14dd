"""Where the host's time goes in the pipelined public call (GPU box): per step, the time in _marshal, in the
tcd_detect_batch_async call, in tcd_detect_batch_wait (blocking) and in building the rows -- for host and device frames.
  python tools/host_probe.py [slots] [steps]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import synth_weights  # noqa: E402  (weight generator only)
from telescope_cam_detection_b200 import _lib, weights  # noqa: E402
from telescope_cam_detection_b200.engine import DetectorEngine, check  # noqa: E402

slots = int(sys.argv[1]) if len(sys.argv) > 1 else 4
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 400
B, H, W, POOL = 8, 1080, 1920, 32
torch.cuda.set_device(0)
blob = weights.pack_yolox(synth_weights.make_model(seed=0).state_dict())
eng = DetectorEngine(blob, input_size=(640, 640), max_batch=B, max_out=300, max_src=(H, W), device=0, pre_mode=1)
rng = np.random.default_rng(0)
host = torch.empty((POOL, H, W, 3), dtype=torch.uint8, pin_memory=True)
host.numpy()[:] = rng.integers(0, 255, (POOL, H, W, 3), dtype=np.uint8)
dev = host.cuda()
hn = host.numpy()
batches = {"host": [[hn[b * B + j] for j in range(B)] for b in range(POOL // B)],
           "device": [[dev[b * B + j] for j in range(B)] for b in range(POOL // B)]}
lib = eng.lib
for kind in ("device", "host"):
    bt = batches[kind]
    acc = {"marshal": 0.0, "async": 0.0, "wait": 0.0, "rows": 0.0, "dev_h2d": 0.0, "dev_pre": 0.0, "dev_net": 0.0}
    for phase in ("warm", "timed"):
        for k in acc:
            acc[k] = 0.0
        torch.cuda.synchronize()
        t_all = time.perf_counter()
        n = steps if phase == "timed" else 20
        for i in range(n):
            sl = i % slots
            if i >= slots:
                t0 = time.perf_counter()
                Bp, _ = eng._pending.pop(sl)
                check(lib.tcd_detect_batch_wait(eng.handle, sl, eng._out.ctypes.data_as(_lib._F32P),
                                                eng._cnt.ctypes.data_as(_lib._I32P)), "wait")
                t1 = time.perf_counter()
                eng._rows(Bp)
                t2 = time.perf_counter()
                acc["wait"] += t1 - t0
                lt = eng.last_timing()
                acc["dev_h2d"] += lt["h2d_ms"] * 1e-3
                acc["dev_pre"] += lt["preprocess_ms"] * 1e-3
                acc["dev_net"] += lt["network_post_ms"] * 1e-3
                acc["rows"] += t2 - t1
            t0 = time.perf_counter()
            m = eng._marshal(bt[i % len(bt)])
            t1 = time.perf_counter()
            check(lib.tcd_detect_batch_async(eng.handle, sl, m[0], m[1], m[2], m[3], m[4], m[5]), "async")
            t2 = time.perf_counter()
            eng._pending[sl] = (m[4], m[6])
            acc["marshal"] += t1 - t0
            acc["async"] += t2 - t1
        for i in range(max(0, n - slots), n):
            eng.wait(i % slots)
        torch.cuda.synchronize()
        t_all = time.perf_counter() - t_all
    print("%-6s slots=%d  step %.3f ms (%.0f fps)  per step: %s" % (
        kind, slots, 1e3 * t_all / n, B * n / t_all,
        "  ".join("%s %.1f us" % (k, 1e6 * v / n) for k, v in acc.items())), flush=True)

# ---- raw pitched H2D of the same shape (8 x 360 rows x 5760 B, source pitch 3 x 5760), idle GPU
from cuda.bindings import runtime as rt  # noqa: E402
dst = torch.empty((4, B, 360, W * 3), dtype=torch.uint8, device="cuda")
sts = [torch.cuda.Stream() for _ in range(4)]
for nst in (1, 4):
    for pitched in (1, 0):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st in sts[:nst]:
            st.wait_event(e0)
        n = 200
        for i in range(n):
            st = sts[i % nst]
            for j in range(B):
                src = host[(i * B + j) % POOL].data_ptr() + W * 3
                d = dst[i % nst, j].data_ptr()
                if pitched:
                    (err,) = rt.cudaMemcpy2DAsync(d, W * 3, src, 3 * W * 3, W * 3, 360, rt.cudaMemcpyKind.cudaMemcpyHostToDevice, st.cuda_stream)
                else:
                    (err,) = rt.cudaMemcpyAsync(d, src, 360 * W * 3, rt.cudaMemcpyKind.cudaMemcpyHostToDevice, st.cuda_stream)
                assert int(err) == 0, err
        for st in sts[:nst]:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print("raw H2D %s, %d stream(s): %.3f ms per step of 8 frames = %.1f GB/s" % (
            "pitched" if pitched else "contiguous", nst, ms, B * 360 * W * 3 / ms / 1e6), flush=True)
