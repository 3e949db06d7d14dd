"""EmptyFrameFilter -- drop-in for the reference's src/empty_frame_filter.py (:15-113).

Same constructor, methods, counters and verdicts; the BGR->gray, 21x21 fixed-point Gaussian,
absdiff, strict threshold and count run in ONE fused CUDA kernel (csrc/motion.cu blur_kernel),
and the previous blurred frame stays on the device.  All-integer path: bit-exact vs cv2.
"""
import logging

from .engine import EmptyGateEngine, device_index

logger = logging.getLogger(__name__)


class EmptyFrameFilter:
    def __init__(self, min_motion_area: int = 200, threshold: int = 25, blur_size: int = 21, device=0):
        self.min_motion_area = min_motion_area
        self.threshold = threshold
        self.blur_size = blur_size if blur_size % 2 == 1 else blur_size + 1  # odd, like the reference (:37)
        self.device = device_index(device)   # ordinal, "cuda:N" or N
        self._gate = None
        self.prev_gray = None  # kept for attribute compatibility; the plane itself lives on the GPU
        self.total_frames = 0
        self.skipped_frames = 0
        self.motion_frames = 0
        logger.info("EmptyFrameFilter initialized: min_motion_area=%spx², threshold=%s, blur_size=%s",
                    min_motion_area, threshold, self.blur_size)

    def _engine(self, frame):
        h, w = frame.shape[0], frame.shape[1]
        if self._gate is None or (self._gate.h, self._gate.w) != (h, w):
            if self._gate is not None:
                self._gate.close()
            self._gate = EmptyGateEngine(h, w, self.blur_size, self.device)
            self.prev_gray = None
        return self._gate

    def has_motion(self, frame) -> bool:
        """True = run inference, False = frame is empty (reference :51-98)."""
        self.total_frames += 1
        gate = self._engine(frame)
        if self.prev_gray is None:
            gate.reset()
        changed, _ = gate.step(frame, self.threshold)
        self.prev_gray = True
        if changed < 0:  # first frame: nothing to compare with
            self.motion_frames += 1
            return True
        moving = changed >= self.min_motion_area
        if moving:
            self.motion_frames += 1
        else:
            self.skipped_frames += 1
        return moving

    def reset(self):
        self.prev_gray = None

    def get_stats(self) -> dict:
        rate = self.skipped_frames / max(self.total_frames, 1)
        return {"total_frames": self.total_frames, "skipped_frames": self.skipped_frames,
                "motion_frames": self.motion_frames, "skip_rate": rate, "skip_rate_percent": rate * 100}
