"""ORACLE (test infrastructure only) -- CPU restatement of the reference's motion path.

Reference: /root/reference/src/motion_filter.py:56-60,68-127 (MOG2 apply -> threshold 200 ->
GaussianBlur 21x21 -> threshold 25 -> ROI countNonZero -> verdict) and
/root/reference/src/empty_frame_filter.py:51-98 (BGR2GRAY -> GaussianBlur -> absdiff -> threshold
-> countNonZero).  The arithmetic itself lives in the reference's dependency opencv-python
(requirements.txt:12, ">=4.8.0", unpinned; this image has 4.13.0, source not under
/root/reference).  Its published algorithms (Zivkovic MOG2, the u8 fixed-point GaussianBlur and
BGR2GRAY paths) are restated here in NumPy, vectorised over pixels, following SURVEY.md App. C.

Pinning: tests/test_oracle_motion.py checks every function here bit-for-bit against cv2 4.13
itself (the reference's real arithmetic) and against tests/golden/motion_*.npz, which were
produced by importing the reference's own MotionFilter / EmptyFrameFilter classes in the build
container (tests/golden/make_golden.py).  The reference's tests hold no vectors for this path
(SURVEY.md F8), so beyond that: PARITY UNPINNED.
"""
import numpy as np

F = np.float32
NMIX = 5


def gaussian_kernel_q8(n):
    """8.8 fixed-point 1-D Gaussian taps OpenCV uses for u8 images, sigma derived from n
    (cv::getGaussianKernel bit-exact path + error-diffusion quantisation)."""
    small = {1: [1.0], 3: [0.25, 0.5, 0.25], 5: [0.0625, 0.25, 0.375, 0.25, 0.0625],
             7: [0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125]}
    if n in small:
        g = np.array(small[n], np.float64)
    else:
        sigma = ((n - 1) * 0.5 - 1) * 0.3 + 0.8
        x = np.arange(n, dtype=np.float64) - (n - 1) * 0.5
        g = np.exp(-0.5 / (sigma * sigma) * x * x)
        g /= g.sum()
    out = np.zeros(n, np.int64)
    err, total = 0.0, 0
    for i in range(n // 2):
        adj = g[i] * 256.0 + err
        v = int(np.rint(adj))  # half to even
        err = adj - v
        out[i] = out[n - 1 - i] = v
        total += v
    out[n // 2] = 256 - 2 * total
    return out


def _reflect101(idx, n):
    idx = np.abs(idx)
    return np.where(idx >= n, 2 * n - 2 - idx, idx)


def gaussian_blur_u8(img, ksize):
    """cv2.GaussianBlur(img, (k,k), 0) on a u8 plane: exact 8.8 horizontal pass, 16.16 vertical
    pass, round half up, BORDER_REFLECT_101."""
    if ksize <= 1:
        return img.copy()
    k = gaussian_kernel_q8(ksize)
    r = ksize // 2
    h, w = img.shape
    xs = _reflect101(np.arange(-r, w + r), w)
    ys = _reflect101(np.arange(-r, h + r), h)
    src = img.astype(np.int64)[:, xs]
    hor = np.zeros((h, w), np.int64)
    for t in range(ksize):
        hor += k[t] * src[:, t:t + w]
    hor = hor[ys, :]
    ver = np.zeros((h, w), np.int64)
    for t in range(ksize):
        ver += k[t] * hor[t:t + h, :]
    return ((ver + 32768) >> 16).astype(np.uint8)


def bgr_to_gray(frame):
    """cv2.cvtColor(BGR2GRAY) on u8: 15-bit coefficients."""
    f = frame.astype(np.int64)
    return ((f[..., 0] * 3735 + f[..., 1] * 19235 + f[..., 2] * 9798 + 16384) >> 15).astype(np.uint8)


class Mog2:
    """cv2.createBackgroundSubtractorMOG2(history, varThreshold, detectShadows).apply(frame),
    learningRate = -1.  Every intermediate is rounded to fp32 in the order of SURVEY.md App. C."""

    def __init__(self, history=500, var_threshold=16, detect_shadows=True):
        self.history, self.Tb, self.shadows = int(history), F(var_threshold), bool(detect_shadows)
        self.TB, self.Tg, self.var_init, self.var_min, self.var_max = F(0.9), F(9.0), F(15.0), F(4.0), F(75.0)
        self.CT, self.tau = F(0.05), F(0.5)
        self.nframes = 0
        self.n = None

    def _alloc(self, npix):
        self.n = np.zeros(npix, np.int32)
        self.w = np.zeros((NMIX, npix), F)
        self.var = np.zeros((NMIX, npix), F)
        self.mean = np.zeros((NMIX, 3, npix), F)

    def _swap(self, i, j, sel):
        for a in (self.w, self.var):
            t = a[i, sel].copy(); a[i, sel] = a[j, sel]; a[j, sel] = t
        t = self.mean[i][:, sel].copy(); self.mean[i][:, sel] = self.mean[j][:, sel]; self.mean[j][:, sel] = t

    def apply(self, frame):
        h, wd = frame.shape[:2]
        x = frame.reshape(-1, 3).astype(F).T.copy()  # [3, npix], stored channel order
        npix = x.shape[1]
        if self.n is None or self.n.shape[0] != npix:
            self._alloc(npix)
        self.nframes += 1
        alpha = F(1.0 / min(2 * self.nframes, self.history))
        a1 = F(1.0) - alpha
        prune = -(alpha * self.CT)
        w, var, mean = self.w, self.var, self.mean
        bg = np.zeros(npix, bool)
        fits = np.zeros(npix, bool)
        nmodes = self.n.copy()
        tot = np.zeros(npix, F)
        ar = np.arange(npix)
        for mode in range(NMIX):
            act = mode < nmodes  # nmodes may shrink while we walk the slots
            if not act.any():
                break
            wt = (a1 * w[mode]) + prune
            swap = np.zeros(npix, np.int32)
            test = act & ~fits
            d = mean[mode] - x
            dist2 = ((d[0] * d[0]) + (d[1] * d[1])) + (d[2] * d[2])
            v = var[mode]
            bg |= test & (tot < self.TB) & (dist2 < (self.Tb * v))
            fit = test & (dist2 < (self.Tg * v))
            fits |= fit
            wt = np.where(fit, wt + alpha, wt)
            with np.errstate(divide="ignore", invalid="ignore"):
                k = (alpha / wt).astype(F)
            newm = mean[mode] - (k * d)
            vn = v + (k * (dist2 - v))
            vn = np.minimum(np.maximum(vn, self.var_min), self.var_max)
            mean[mode][:, fit] = newm[:, fit]
            var[mode, fit] = vn[fit]
            bubbling = fit.copy()
            for i in range(mode, 0, -1):
                stop = bubbling & (wt < w[i - 1])
                bubbling &= ~stop
                if not bubbling.any():
                    break
                self._swap(i, i - 1, bubbling)
                swap[bubbling] += 1
            dead = act & (wt < -prune)
            wt = np.where(dead, F(0), wt)
            nmodes = nmodes - dead.astype(np.int32)
            dst = mode - swap
            w[dst[act], ar[act]] = wt[act]
            tot = np.where(act, tot + wt, tot)
        inv = np.where(np.abs(tot) > np.finfo(F).eps, F(1.0) / np.where(tot == 0, F(1), tot), F(0)).astype(F)
        for mode in range(NMIX):
            sel = mode < nmodes
            w[mode, sel] = w[mode, sel] * inv[sel]
        new = ~fits if alpha > 0 else np.zeros(npix, bool)
        if new.any():
            full = nmodes == NMIX
            slot = np.where(full, NMIX - 1, nmodes)
            nmodes = np.where(new & ~full, nmodes + 1, nmodes)
            first = new & (nmodes == 1)
            for mode in range(NMIX):
                sel = new & (slot == mode)
                if sel.any():
                    w[mode, sel] = np.where(first[sel], F(1.0), alpha)
                    mean[mode][:, sel] = x[:, sel]
                    var[mode, sel] = self.var_init
                rest = new & ~first & (mode < nmodes - 1) & (slot != mode)
                w[mode, rest] = w[mode, rest] * a1
            moving = new.copy()
            for i in range(NMIX - 1, 0, -1):
                here = moving & (i <= nmodes - 1)
                stop = here & (alpha < w[i - 1])
                moving &= ~stop
                go = here & ~stop
                if go.any():
                    self._swap(i, i - 1, go)
        self.n = nmodes
        out = np.where(bg, 0, 255).astype(np.uint8)
        if self.shadows:
            fg = ~bg
            tw = np.zeros(npix, F)
            done = ~fg
            for mode in range(NMIX):
                live = ~done & (mode < nmodes)
                if not live.any():
                    break
                m = mean[mode]
                num = ((x[0] * m[0]) + (x[1] * m[1])) + (x[2] * m[2])
                den = ((m[0] * m[0]) + (m[1] * m[1])) + (m[2] * m[2])
                zero = live & (den == 0)
                done |= zero
                live &= ~zero
                cand = live & (num <= den) & (num >= (self.tau * den))
                with np.errstate(divide="ignore", invalid="ignore"):
                    a = (num / den).astype(F)
                e = (a * m) - x
                s = ((e[0] * e[0]) + (e[1] * e[1])) + (e[2] * e[2])
                shadow = cand & (s < (((self.Tb * var[mode]) * a) * a))
                out[shadow] = 127
                done |= shadow
                live &= ~shadow
                tw = np.where(live, tw + w[mode], tw)
                done |= live & (tw > self.TB)
        return out.reshape(h, wd)


def clean_mask(fg_mask, blur_size=21, detect_shadows=True):
    """motion_filter.py:91-98."""
    m = np.where(fg_mask > 200, 255, 0).astype(np.uint8) if detect_shadows else fg_mask
    m = gaussian_blur_u8(m, blur_size)
    return np.where(m > 25, 255, 0).astype(np.uint8)


def bbox_verdict(mask, bbox, min_motion_ratio=0.05, min_motion_pixels=10):
    """motion_filter.py:100-127 on an already cleaned mask.  -> (has_motion, ratio)."""
    x1, y1, x2, y2 = bbox["x1"], bbox["y1"], bbox["x2"], bbox["y2"]
    if x1 > x2:
        x1, x2 = x2, x1
    if y1 > y2:
        y1, y2 = y2, y1
    if x2 - x1 < 1:
        x2 = x1 + 1
    if y2 - y1 < 1:
        y2 = y1 + 1
    h, w = mask.shape[:2]
    x1, y1, x2, y2 = max(0, int(x1)), max(0, int(y1)), min(w, int(x2)), min(h, int(y2))
    if x2 <= x1 or y2 <= y1:
        return False, 0.0
    px = int(np.count_nonzero(mask[y1:y2, x1:x2]))
    ratio = px / ((x2 - x1) * (y2 - y1))
    return (px >= min_motion_pixels and ratio > min_motion_ratio), ratio


class EmptyFrame:
    """empty_frame_filter.py:51-98 -> changed-pixel count per frame (-1 for the first)."""

    def __init__(self, blur_size=21):
        self.blur_size = blur_size if blur_size % 2 == 1 else blur_size + 1
        self.prev = None

    def step(self, frame, threshold=25):
        gray = bgr_to_gray(frame) if frame.ndim == 3 else frame
        blurred = gaussian_blur_u8(gray, self.blur_size)
        if self.prev is None:
            self.prev = blurred
            return -1, blurred
        delta = np.abs(self.prev.astype(np.int16) - blurred.astype(np.int16))
        self.prev = blurred
        return int(np.count_nonzero(delta > threshold)), blurred
