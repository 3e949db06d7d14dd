"""Thin Python handle classes over the libtcd C ABI (ctypes).  numpy for host buffers,
torch only for device pointers handed in by the caller (GPU-tensor frames are part of the
reference API, src/yolox_detector.py:197-203)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check


def device_index(device) -> int:
    """'cuda:1' / 'cuda' / 1 -> ordinal.  Anything else is refused: there is no CPU path."""
    if isinstance(device, int):
        return device
    s = str(device)
    if not s.startswith("cuda"):
        raise RuntimeError("the B200 build needs a CUDA device, got %r" % (device,))
    return int(s.split(":")[1]) if ":" in s else 0


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _frame_desc(frame):
    """-> (ptr, h, w, row_stride_bytes, on_device, keepalive)"""
    if _is_torch(frame):
        t = frame
        if t.dtype != __import__("torch").uint8:
            raise TypeError("frames must be uint8, got %s" % t.dtype)
        if t.ndim != 3 or t.shape[2] != 3:
            raise ValueError("frame tensor must be HxWx3, got %s" % (tuple(t.shape),))
        if t.stride(2) != 1 or t.stride(1) != 3:
            t = t.contiguous()
        on_dev = t.is_cuda
        return t.data_ptr(), t.shape[0], t.shape[1], t.stride(0), on_dev, t
    a = np.asarray(frame)
    if a.dtype != np.uint8:
        raise TypeError("frames must be uint8, got %s" % a.dtype)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("frame must be HxWx3, got %s" % (a.shape,))
    if a.strides[2] != 1 or a.strides[1] != 3:
        a = np.ascontiguousarray(a)
    return a.ctypes.data, a.shape[0], a.shape[1], a.strides[0], False, a


class DetectorEngine:
    """Owns one tcd_detector (weights, activation arenas, CUDA graphs) on one GPU."""

    def __init__(self, weight_blob, input_size=(640, 640), max_batch=8, num_classes=80, max_out=1024,
                 max_src=(1080, 1920), device=0, pre_mode=0, swap_rb=0, scale=1.0,
                 mean=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0), pad_value=114.0):
        self.lib = _lib.load()
        cfg = _lib.DetectorCfg()
        cfg.input_h, cfg.input_w = int(input_size[0]), int(input_size[1])
        cfg.max_batch, cfg.num_classes, cfg.max_out = int(max_batch), int(num_classes), int(max_out)
        cfg.max_src_h, cfg.max_src_w = int(max_src[0]), int(max_src[1])
        cfg.pre = _lib.make_preproc_cfg(input_size[0], input_size[1], pre_mode, swap_rb, scale, mean, std, pad_value)
        self.cfg = cfg
        self.max_batch, self.max_out, self.num_classes = cfg.max_batch, cfg.max_out, cfg.num_classes
        self._blob = bytes(weight_blob)
        h = C.c_void_p()
        check(self.lib.tcd_create_detector(C.byref(cfg), self._blob, len(self._blob), int(device), C.byref(h)),
              "tcd_create_detector")
        self.handle = h
        self.n_anchors = self.lib.tcd_detector_num_anchors(h)
        self._out = np.zeros((self.max_batch, self.max_out, 7), np.float32)
        self._cnt = np.zeros((self.max_batch,), np.int32)
        self._pending = {}

    def close(self):
        if getattr(self, "handle", None):
            self.lib.tcd_destroy_detector(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_thresholds(self, conf, nms):
        check(self.lib.tcd_detector_set_thresholds(self.handle, float(conf), float(nms)), "tcd_detector_set_thresholds")

    def set_class_filter(self, table, cls_conf, cls_min_area=None, cls_max_area=None, min_area=0.0, max_det=1 << 30,
                         allowed_classes=None):
        """Device-side per-camera filter table `table` (1 .. MAX_FILTER_TABLES-1), see tcd_class_filter:
        cls_conf: scalar or {class_id: threshold} over a default given under key None;
        cls_min_area / cls_max_area: {class_id: area}; allowed_classes: iterable of ids (None = all)."""
        f = _lib.ClassFilter()
        default = cls_conf.get(None, float("-inf")) if isinstance(cls_conf, dict) else float(cls_conf)
        for c in range(_lib.MAX_CLASSES):
            f.cls_conf[c] = float(cls_conf.get(c, default)) if isinstance(cls_conf, dict) else default
            f.cls_min_area[c] = float((cls_min_area or {}).get(c, float("-inf")))
            f.cls_max_area[c] = float((cls_max_area or {}).get(c, float("inf")))
        f.min_area = float(min_area)
        f.max_det = int(min(max(int(max_det), -(1 << 30)), 1 << 30))
        mask = [0, 0]
        for c in (range(_lib.MAX_CLASSES) if allowed_classes is None else allowed_classes):
            if 0 <= c < _lib.MAX_CLASSES:
                mask[c >> 6] |= 1 << (c & 63)
        f.class_mask[0], f.class_mask[1] = mask
        check(self.lib.tcd_detector_set_class_filter(self.handle, int(table), C.byref(f)),
              "tcd_detector_set_class_filter")

    def filter_rows(self, table, rows, orig_h, orig_w):
        """Parity tap: class filter `table` on rows [n,7] (network pixels, score order) -> bool keep mask."""
        rows = np.ascontiguousarray(rows, np.float32).reshape(-1, 7)
        keep = np.zeros((rows.shape[0],), np.int32)
        check(self.lib.tcd_detector_filter_rows(self.handle, int(table), rows.ctypes.data_as(_lib._F32P), rows.shape[0],
                                                int(orig_h), int(orig_w), keep.ctypes.data_as(_lib._I32P)),
              "tcd_detector_filter_rows")
        return keep.astype(bool)

    def _marshal(self, frames):
        if not 1 <= len(frames) <= self.max_batch:
            raise ValueError("batch of %d frames outside [1, %d]" % (len(frames), self.max_batch))
        descs = [_frame_desc(f) for f in frames]
        on_dev = {d[4] for d in descs}
        if len(on_dev) != 1:
            raise ValueError("a batch must be all host frames or all device frames")
        return (_lib.ptr_array([d[0] for d in descs]), _lib.i32_array([d[1] for d in descs]),
                _lib.i32_array([d[2] for d in descs]), _lib.i64_array([d[3] for d in descs]),
                len(descs), 1 if on_dev.pop() else 0, [d[5] for d in descs])

    def _rows(self, B):
        return [self._out[b, : self._cnt[b]].copy() for b in range(B)]

    def detect_batch(self, frames):
        """frames: list of HxWx3 uint8 (numpy, or torch CUDA tensors).  -> list of [n,7] float32 arrays
        [x1,y1,x2,y2,obj_conf,class_conf,class_id] in network-input pixels, score-descending."""
        p, h, w, s, B, dev, keep = self._marshal(frames)
        check(self.lib.tcd_detect_batch(self.handle, p, h, w, s, B, dev,
                                        self._out.ctypes.data_as(_lib._F32P), self._cnt.ctypes.data_as(_lib._I32P)),
              "tcd_detect_batch")
        return self._rows(B)

    def submit(self, slot, frames, filter_tables=None):
        """Enqueue a batch on pipeline slot `slot`; filter_tables: per-frame class-filter table ids (None = no filter)."""
        p, h, w, s, B, dev, keep = self._marshal(frames)
        ft = None
        if filter_tables is not None:
            if len(filter_tables) != B:
                raise ValueError("filter_tables has %d entries for %d frames" % (len(filter_tables), B))
            ft = _lib.i32_array(filter_tables)
        check(self.lib.tcd_detector_submit(self.handle, int(slot), p, h, w, s, B, dev, ft), "tcd_detector_submit")
        self._pending[slot] = (B, keep)

    def wait(self, slot):
        B, _ = self._pending.pop(slot)
        check(self.lib.tcd_detect_batch_wait(self.handle, int(slot), self._out.ctypes.data_as(_lib._F32P),
                                             self._cnt.ctypes.data_as(_lib._I32P)), "tcd_detect_batch_wait")
        return self._rows(B)

    def pending(self, slot):
        return slot in self._pending

    def forward(self, frames):
        p, h, w, s, B, dev, keep = self._marshal(frames)
        out = np.empty((B, self.n_anchors, 5 + self.num_classes), np.float32)
        check(self.lib.tcd_forward(self.handle, p, h, w, s, B, dev, out.ctypes.data_as(_lib._F32P)), "tcd_forward")
        return out

    def postprocess(self, pred):
        pred = np.ascontiguousarray(pred, np.float32)
        B = pred.shape[0]
        check(self.lib.tcd_postprocess(self.handle, pred.ctypes.data_as(_lib._F32P), B,
                                       self._out.ctypes.data_as(_lib._F32P), self._cnt.ctypes.data_as(_lib._I32P)),
              "tcd_postprocess")
        return self._rows(B)

    def last_timing(self):
        t = (C.c_float * 5)()
        check(self.lib.tcd_detector_last_timing(self.handle, t), "tcd_detector_last_timing")
        return dict(zip(("h2d_ms", "preprocess_ms", "network_post_ms", "_", "d2h_ms"), list(t)))

    def stream(self, slot=0):
        """cudaStream_t (int) of pipeline slot `slot`."""
        return int(self.lib.tcd_detector_stream(self.handle, int(slot)) or 0)

    def result_device(self, slot):
        """(dets_ptr, counts_ptr): device addresses of slot's [B,max_out,7] f32 table and [B] i32 counts, valid from
        wait(slot) until the next submit(slot)."""
        d, c = C.c_void_p(0), C.c_void_p(0)
        check(self.lib.tcd_detector_result_device(self.handle, int(slot), C.byref(d), C.byref(c)),
              "tcd_detector_result_device")
        return int(d.value), int(c.value)

    @property
    def num_slots(self):
        return int(self.lib.tcd_detector_num_slots(self.handle))

    def trace_op(self, B, idx, reps=3):
        """In-kernel timeline of conv op `idx` (index as in profile()).  -> dict with the clock64
        stamps of the first / last CTA relative to kernel entry, in SM cycles, and the wall ns."""
        buf = (C.c_uint64 * 128)()
        grid, smem = C.c_int32(0), C.c_int32(0)
        check(self.lib.tcd_detector_trace_op(self.handle, int(B), int(idx), int(reps), buf, C.byref(grid),
                                             C.byref(smem)), "tcd_detector_trace_op")
        v = list(buf)
        names = ("setup_done", "first_A_issued", "last_load_issued", "first_operands", "last_mma_commit",
                 "first_acc_ready", "last_store_issued", "stores_drained", "exit")
        out = {"grid": grid.value, "smem": smem.value}
        for sel, tag in ((0, "cta0"), (1, "ctaN")):
            b = v[sel * 16: sel * 16 + 16]
            if b[1] == 0:
                continue
            out[tag] = {n: int(b[2 + i]) - int(b[1]) if b[2 + i] else None for i, n in enumerate(names)}
            for i, n in enumerate(("epi_slab0_loaded", "epi_slab0_staged", "epi_slabs_done", "epi_fence_done")):
                out[tag][n] = int(b[12 + i]) - int(b[1]) if b[12 + i] else None
            out[tag]["wall_ns"] = int(b[11]) - int(b[0])
            out[tag]["start_ns"] = int(b[0])
        t0 = v[1]
        for base, tag in ((32, "tile_prod"), (48, "tile_mma"), (64, "tile_acc"), (80, "tile_store")):
            out[tag] = [int(x) - int(t0) for x in v[base:base + 16] if x]
        out["tile_stage_free"] = [int(x) - int(t0) for x in v[96:112] if x]
        out["tile_staged"] = [int(x) - int(t0) for x in v[112:128] if x]
        if "cta0" in out and "ctaN" in out:
            out["ctaN_start_after_cta0_ns"] = out["ctaN"]["start_ns"] - out["cta0"]["start_ns"]
        return out

    def profile(self, frames, iters=5):
        """Per-kernel device times of one detect_batch (eager pass, CUDA-event pair per kernel).
        -> list of dicts {name, kind ('conv'|'other'), ms, flops}."""
        p, h, w, s, B, dev, keep = self._marshal(frames)
        cap = 512
        ms = (C.c_float * cap)()
        n = C.c_int32(0)
        check(self.lib.tcd_detector_profile(self.handle, p, h, w, s, B, dev, int(iters), ms, cap, C.byref(n)),
              "tcd_detector_profile")
        out = []
        for i in range(n.value):
            out.append({"name": self.lib.tcd_detector_op_name(self.handle, B, i).decode(),
                        "kind": "conv" if self.lib.tcd_detector_op_kind(self.handle, B, i) == 0 else "other",
                        "ms": float(ms[i]), "flops": float(self.lib.tcd_detector_op_flops(self.handle, B, i))})
        return out

    def read_activation(self, name):
        shape = (C.c_int32 * 4)()
        check(self.lib.tcd_detector_read_activation(self.handle, name.encode(), None, 0, shape), "read_activation")
        out = np.empty(tuple(shape), np.float32)
        check(self.lib.tcd_detector_read_activation(self.handle, name.encode(), out.ctypes.data_as(_lib._F32P),
                                                    out.size, shape), "read_activation")
        return out


def preprocess(frames_dev_ptrs, hs, ws, strides, out_ptr, out_size, layout=0, mode=0, swap_rb=0, scale=1.0,
               mean=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0), pad_value=114.0):
    lib = _lib.load()
    cfg = _lib.make_preproc_cfg(out_size[0], out_size[1], mode, swap_rb, scale, mean, std, pad_value)
    check(lib.tcd_preprocess(C.byref(cfg), _lib.ptr_array(frames_dev_ptrs), _lib.i32_array(hs), _lib.i32_array(ws),
                             _lib.i64_array(strides), len(hs), int(layout), C.c_void_p(int(out_ptr)), None),
          "tcd_preprocess")


def conv2d_nhwc_bf16(x_ptr, N, H, W, x_ctot, x_coff, cin, w, bias, k, stride, silu, y_ptr, y_ctot, y_coff,
                     resid_ptr=0, r_ctot=0, r_coff=0, groups=1):
    """cin: input channels PER GROUP; w: [cout, cin, k, k]."""
    lib = _lib.load()
    w = np.ascontiguousarray(w, np.float32)
    bias = np.ascontiguousarray(bias, np.float32)
    cout = w.shape[0]
    check(lib.tcd_conv2d_nhwc_bf16_grouped(C.c_void_p(int(x_ptr)), N, H, W, x_ctot, x_coff, cin,
                                           w.ctypes.data_as(_lib._F32P), bias.ctypes.data_as(_lib._F32P), cout, k,
                                           stride, int(silu), C.c_void_p(int(resid_ptr)) if resid_ptr else None,
                                           r_ctot, r_coff, C.c_void_p(int(y_ptr)), y_ctot, y_coff, int(groups), None),
          "tcd_conv2d_nhwc_bf16_grouped")


def crop_resize(frame_ptr, h, w, stride, boxes, S, out_ptr, layout=0, swap_rb=1, scale=1.0 / 255.0,
                mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)):
    lib = _lib.load()
    boxes = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
    cfg = _lib.make_preproc_cfg(S, S, 0, swap_rb, scale, mean, std, 0.0)
    check(lib.tcd_crop_resize(C.c_void_p(int(frame_ptr)), int(h), int(w), int(stride),
                              boxes.ctypes.data_as(_lib._I32P), boxes.shape[0], int(S), C.byref(cfg), int(layout),
                              C.c_void_p(int(out_ptr)), None), "tcd_crop_resize")


def _plane_desc(frame, channels):
    """-> (ptr, h, w, row_stride_bytes, on_device, keepalive) for an HxWxC / HxW u8 frame."""
    if _is_torch(frame):
        t = frame
        if t.stride(-1) != 1 or (channels == 3 and t.stride(1) != 3):
            t = t.contiguous()
        return t.data_ptr(), t.shape[0], t.shape[1], t.stride(0), t.is_cuda, t
    a = np.asarray(frame)
    if a.dtype != np.uint8:
        raise TypeError("frames must be uint8, got %s" % a.dtype)
    if a.strides[-1] != 1 or (channels == 3 and a.strides[1] != 3):
        a = np.ascontiguousarray(a)
    return a.ctypes.data, a.shape[0], a.shape[1], a.strides[0], False, a


class Mog2Engine:
    """One tcd_mog2 handle: device-resident 5-mode GMM per pixel for one camera stream
    (replaces cv2.createBackgroundSubtractorMOG2, src/motion_filter.py:56-60)."""

    def __init__(self, h, w, history=500, var_threshold=16, detect_shadows=True, device=0):
        self.lib = _lib.load()
        self.h, self.w = int(h), int(w)
        hd = C.c_void_p()
        check(self.lib.tcd_mog2_create(self.h, self.w, int(history), float(var_threshold),
                                       1 if detect_shadows else 0, int(device), C.byref(hd)), "tcd_mog2_create")
        self.handle = hd

    def close(self):
        if getattr(self, "handle", None):
            self.lib.tcd_mog2_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        check(self.lib.tcd_mog2_reset(self.handle), "tcd_mog2_reset")

    def apply(self, frame, repeats=1, want_mask=False):
        ptr, h, w, stride, on_dev, keep = _plane_desc(frame, 3)
        if (h, w) != (self.h, self.w):
            raise ValueError("frame is %dx%d, model is %dx%d" % (h, w, self.h, self.w))
        mask = np.empty((self.h, self.w), np.uint8) if want_mask else None
        check(self.lib.tcd_mog2_apply(self.handle, C.c_void_p(ptr), int(stride), 1 if on_dev else 0, int(repeats),
                                      mask.ctypes.data_as(C.c_void_p) if want_mask else None), "tcd_mog2_apply")
        return mask

    def last_ms(self):
        """Device time (CUDA events) of the last call's kernels: (one MOG2 apply, clean-up + ROI counts), ms."""
        t = (C.c_float * 2)()
        check(self.lib.tcd_mog2_last_ms(self.handle, t), "tcd_mog2_last_ms")
        return float(t[0]), float(t[1])

    def roi_counts(self, rois, shadow_thr=200, ksize=21, blur_thr=25, want_mask=False):
        """rois: int32 [n,4] (x1,y1,x2,y2) already clamped to the frame.  -> (counts[n], clean mask or None)"""
        rois = np.ascontiguousarray(rois, np.int32).reshape(-1, 4)
        n = rois.shape[0]
        counts = np.zeros((max(n, 1),), np.int32)
        mask = np.empty((self.h, self.w), np.uint8) if want_mask else None
        check(self.lib.tcd_mog2_roi_counts(self.handle, int(shadow_thr), int(ksize), int(blur_thr),
                                           rois.ctypes.data_as(_lib._I32P), n, counts.ctypes.data_as(_lib._I32P),
                                           mask.ctypes.data_as(C.c_void_p) if want_mask else None),
              "tcd_mog2_roi_counts")
        return counts[:n], mask


class EmptyGateEngine:
    """One tcd_empty_filter handle (previous blurred gray frame lives on the device)."""

    def __init__(self, h, w, blur_size=21, device=0):
        self.lib = _lib.load()
        self.h, self.w = int(h), int(w)
        hd = C.c_void_p()
        check(self.lib.tcd_empty_filter_create(self.h, self.w, int(blur_size), int(device), C.byref(hd)),
              "tcd_empty_filter_create")
        self.handle = hd

    def close(self):
        if getattr(self, "handle", None):
            self.lib.tcd_empty_filter_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        check(self.lib.tcd_empty_filter_reset(self.handle), "tcd_empty_filter_reset")

    def last_ms(self):
        return float(self.lib.tcd_empty_filter_last_ms(self.handle))

    def step(self, frame, threshold=25, want_blur=False):
        """-> (changed pixel count, or -1 on the first frame; blurred gray plane or None)"""
        channels = 3 if (frame.ndim == 3) else 1
        ptr, h, w, stride, on_dev, keep = _plane_desc(frame, channels)
        if (h, w) != (self.h, self.w):
            raise ValueError("frame is %dx%d, filter is %dx%d" % (h, w, self.h, self.w))
        cnt = C.c_int32(0)
        blur = np.empty((self.h, self.w), np.uint8) if want_blur else None
        check(self.lib.tcd_empty_filter_step(self.handle, C.c_void_p(ptr), int(stride), channels, 1 if on_dev else 0,
                                             int(threshold), C.byref(cnt),
                                             blur.ctypes.data_as(C.c_void_p) if want_blur else None),
              "tcd_empty_filter_step")
        return cnt.value, blur


class StreamBankEngine:
    """One tcd_stream_bank handle: the gate reference planes, MOG2 models and this step's frames of every camera
    stream of one GPU, with batched kernels (include/tcd.h, stream bank)."""

    def __init__(self, n_streams, h, w, history=500, var_threshold=16, detect_shadows=True, blur_size=21, device=0):
        self.lib = _lib.load()
        self.n, self.h, self.w = int(n_streams), int(h), int(w)
        hd = C.c_void_p()
        check(self.lib.tcd_stream_bank_create(self.n, self.h, self.w, int(history), float(var_threshold),
                                              1 if detect_shadows else 0, int(blur_size), device_index(device),
                                              C.byref(hd)), "tcd_stream_bank_create")
        self.handle = hd
        self._keep = None

    def close(self):
        if getattr(self, "handle", None):
            self.lib.tcd_stream_bank_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, stream, model=True, gate=True):
        check(self.lib.tcd_stream_bank_reset(self.handle, int(stream), (1 if model else 0) | (2 if gate else 0)),
              "tcd_stream_bank_reset")

    def upload(self, frames):
        """frames: one HxWx3 uint8 host array (or None) per stream."""
        if len(frames) != self.n:
            raise ValueError("expected %d frames (None for idle streams), got %d" % (self.n, len(frames)))
        ptrs, strides, keep = [], [], []
        for f in frames:
            if f is None:
                ptrs.append(0)
                strides.append(0)
                continue
            ptr, h, w, stride, on_dev, k = _plane_desc(f, 3)
            if on_dev or (h, w) != (self.h, self.w):
                raise ValueError("stream bank frames must be %dx%d host arrays" % (self.h, self.w))
            ptrs.append(ptr)
            strides.append(stride)
            keep.append(k)
        self._keep = keep   # the copies are asynchronous: the arrays stay alive until the next call synchronises
        check(self.lib.tcd_stream_bank_upload(self.handle, _lib.ptr_array(ptrs), _lib.i64_array(strides)),
              "tcd_stream_bank_upload")

    def frame_ptr(self, stream):
        return int(self.lib.tcd_stream_bank_frame(self.handle, int(stream)) or 0)

    def cuda_stream(self):
        return int(self.lib.tcd_stream_bank_cuda_stream(self.handle) or 0)

    def gate(self, threshold=25):
        """-> int32[n]: changed-pixel count per stream, -1 on a stream's first frame, -2 when it had no frame."""
        out = np.zeros((self.n,), np.int32)
        check(self.lib.tcd_stream_bank_gate(self.handle, int(threshold), out.ctypes.data_as(_lib._I32P)),
              "tcd_stream_bank_gate")
        return out

    def motion(self, rois, offsets, shadow_thr=200, ksize=21, blur_thr=25):
        """rois: int32 [total,4] clamped boxes, offsets: int32 [n+1].  -> int32[total] motion-pixel counts."""
        rois = np.ascontiguousarray(rois, np.int32).reshape(-1, 4)
        offsets = np.ascontiguousarray(offsets, np.int32)
        counts = np.zeros((max(rois.shape[0], 1),), np.int32)
        check(self.lib.tcd_stream_bank_motion(self.handle, rois.ctypes.data_as(_lib._I32P),
                                              offsets.ctypes.data_as(_lib._I32P), int(shadow_thr), int(ksize),
                                              int(blur_thr), counts.ctypes.data_as(_lib._I32P)),
              "tcd_stream_bank_motion")
        return counts[:rois.shape[0]]

    def last_ms(self):
        t = (C.c_float * 3)()
        check(self.lib.tcd_stream_bank_last_ms(self.handle, t), "tcd_stream_bank_last_ms")
        return {"gate_ms": float(t[0]), "motion_ms": float(t[1]), "motion_rounds": int(t[2])}
