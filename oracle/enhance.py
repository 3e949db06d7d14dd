"""ORACLE (test infrastructure only) -- CPU restatement of ImageEnhancer.enhance_clahe_bilateral.

Reference: /root/reference/src/image_enhancement.py:146-184 (BGR -> LAB, CLAHE(clipLimit 2.0, 8x8 tiles) on L,
LAB -> BGR, bilateralFilter(d 9, sigmaColor 75, sigmaSpace 75); a 2-D input takes CLAHE + bilateral directly).
The arithmetic lives in the reference's dependency opencv-python (requirements.txt, ">=4.8.0", unpinned; this image has
4.13.0, source not under /root/reference).  Its published algorithms are restated here in NumPy:

  bgr2lab / lab2bgr   the 8-bit fixed-point CIE L*a*b* conversions (sRGB gamma table x8, cube-root table in 1.15 fixed
                      point, D65 white point; the inverse through the 14-bit L->(y, fy) table, the analytic ab->xz map
                      and the 12-bit inverse-gamma table).  Tables are built the way the library builds them: in
                      float32, the cube root through the Turkowski quartic rational with the result TRUNCATED to
                      float32 -- that truncation decides two table entries (49, 628) and is what makes this bit-exact.
  clahe               Zuiderveld's CLAHE as OpenCV implements it (reflect-101 padding to whole tiles, integer clip
                      limit, batch + strided residual redistribution, float32 LUT scale, float32 bilinear blend of the
                      four neighbouring tile LUTs).
  bilateral           the 8-bit bilateral filter: circular window of radius d/2, exp tables for colour (L1 distance over
                      channels) and space, float32 accumulation in window order with fused multiply-adds, the centre
                      pixel added last with weight 1, division, round-half-even.

Pinning: tests/test_enhance.py checks bgr2lab, lab2bgr and clahe bit-for-bit against cv2 4.13 itself and against
tests/golden/enhance_ref.npz (produced by the reference's own ImageEnhancer class in the build container,
tests/golden/make_golden.py).  bilateral: cv2's own three code paths (IPP, SIMD dispatch, scalar) differ from each
other in 3-5 of 1.44 M output bytes by one LSB (float summation details), so "bit-exact vs cv2" is not defined for it;
this restatement differs from each of them in <= 12 of 1.44 M bytes, always by one LSB, and the tests bound exactly
that.  The reference's tests hold no vectors for this path: beyond the above, PARITY UNPINNED.
"""
import math
import struct

import numpy as np

F = np.float32
LAB_SHIFT, GAMMA_SHIFT = 12, 3
LAB_SHIFT2 = LAB_SHIFT + GAMMA_SHIFT
CBRT_TAB_SIZE = 256 * 3 // 2 * (1 << GAMMA_SHIFT)
BASE = 1 << 14
MIN_AB = -8145
_XYZ = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]])
_RGB = np.array([[3.240479, -1.53715, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]])
_D65 = np.array([0.950456, 1.0, 1.088754])


def _trunc_f32(d):
    b = struct.unpack("<Q", struct.pack("<d", d))[0] & ~((1 << 29) - 1)
    return F(struct.unpack("<d", struct.pack("<Q", b))[0])


def _cbrt_f32(x):
    m, e = math.frexp(float(x))
    shx = e % 3
    if shx > 0:
        shx -= 3
    ex = (e - shx) // 3
    fr = math.ldexp(m, shx)
    fr = ((((45.2548339756803022511987494 * fr + 192.2798368355061050458134625) * fr + 119.1654824285581628956914143) * fr
           + 13.43250139086239872172837314) * fr + 0.1636161226585754240958355063) / \
         ((((14.80884093219134573786480845 * fr + 151.9714051044435648658557668) * fr + 168.5254414101568283957668343) * fr
           + 33.9905941350215598754191872) * fr + 1.0)
    return _trunc_f32(math.ldexp(fr, ex))


_TABLES = None


def tables():
    """(gamma[256], cbrt[3072], fwd coeffs[3,3], LabToYF[256,2], inv coeffs[3,3], inv_gamma[4096])"""
    global _TABLES
    if _TABLES is not None:
        return _TABLES
    gamma = np.zeros(256, np.int64)
    for i in range(256):
        x = F(i) / F(255.0)
        g = x * (F(1.0) / F(12.92)) if x <= F(0.04045) else F(math.pow(float((x + F(0.055)) * (F(1.0) / F(1.055))), 2.4))
        gamma[i] = int(np.rint(F(255.0) * F(1 << GAMMA_SHIFT) * g))
    cbrt = np.zeros(CBRT_TAB_SIZE, np.int64)
    scale = F(1.0) / F(255 * (1 << GAMMA_SHIFT))
    for i in range(CBRT_TAB_SIZE):
        x = scale * F(i)
        v = x * F(7.787) + F(0.13793103448275862) if x < F(0.008856) else _cbrt_f32(x)
        cbrt[i] = int(np.rint(F(1 << LAB_SHIFT2) * v))
    cf = np.rint((1 << LAB_SHIFT) * _XYZ / _D65[:, None]).astype(np.int64)
    yf = np.zeros((256, 2), np.int64)
    for i in range(256):
        fy = F(i * 100 * BASE) / F(255 * 116) + F(16 * BASE) / F(116)
        if i <= 20:
            y = int(np.rint(F(i * BASE * 20 * 9) / F(17 * 29 * 29 * 29)))
        else:
            y = int(np.rint(fy * fy * fy / F(BASE * BASE)))
        yf[i] = (y, int(np.rint(fy)))
    ci = np.rint((1 << LAB_SHIFT) * _RGB * _D65[None, :]).astype(np.int64)
    inv = np.zeros(4096, np.int64)
    for i in range(4096):
        x = F(i) * (F(1.0) / F(4096))
        g = x * F(12.92) if x <= F(0.0031308) else F(1.055) * F(math.pow(float(x), 1.0 / 2.4)) - F(0.055)
        inv[i] = int(np.rint(F(255.0) * g))
    _TABLES = (gamma, cbrt, cf, yf, ci, inv)
    return _TABLES


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def bgr2lab(img):
    """cv2.cvtColor(img, COLOR_BGR2LAB) for uint8 HxWx3."""
    gamma, cbrt, C, _, _, _ = tables()
    B, G, R = (gamma[img[..., k]] for k in range(3))
    fX = cbrt[_descale(R * C[0, 0] + G * C[0, 1] + B * C[0, 2], LAB_SHIFT)]
    fY = cbrt[_descale(R * C[1, 0] + G * C[1, 1] + B * C[1, 2], LAB_SHIFT)]
    fZ = cbrt[_descale(R * C[2, 0] + G * C[2, 1] + B * C[2, 2], LAB_SHIFT)]
    Lscale = (116 * 255 + 50) // 100
    Lshift = -((16 * 255 * (1 << LAB_SHIFT2) + 50) // 100)
    L = _descale(Lscale * fY + Lshift, LAB_SHIFT2)
    a = _descale(500 * (fX - fY) + 128 * (1 << LAB_SHIFT2), LAB_SHIFT2)
    b = _descale(200 * (fY - fZ) + 128 * (1 << LAB_SHIFT2), LAB_SHIFT2)
    return np.clip(np.stack([L, a, b], -1), 0, 255).astype(np.uint8)


def _ab_to_xz(i):
    """the library's abToXZ_b table, computed: C integer division (truncation toward zero) below the knee, cube above."""
    i = np.asarray(i, np.int64)
    lin_num = i * 108
    lin = np.where(lin_num >= 0, lin_num // 841, -((-lin_num) // 841)) - ((BASE * 16 // 116) * 108 // 841)
    cube = (i * i // BASE) * i // BASE
    return np.where(i <= 3390, lin, cube)


def lab2bgr(lab):
    """cv2.cvtColor(lab, COLOR_LAB2BGR) for uint8 HxWx3."""
    _, _, _, yf, C, inv = tables()
    LL, aa, bb = (lab[..., k].astype(np.int64) for k in range(3))
    y, ify = yf[LL, 0], yf[LL, 1]
    adiv = ((5 * aa * 53687 + (1 << 7)) >> 13) - 128 * BASE // 500
    bdiv = ((bb * 41943 + (1 << 4)) >> 9) - 128 * BASE // 200 + 1
    x, z = _ab_to_xz(ify + adiv), _ab_to_xz(ify - bdiv)
    out = []
    for row in (2, 1, 0):     # B, G, R
        v = _descale(C[row, 0] * x + C[row, 1] * y + C[row, 2] * z, 14)
        out.append(inv[np.clip(v, 0, 4095)])
    return np.stack(out, -1).astype(np.uint8)


def reflect101(idx, n):
    idx = np.abs(idx)
    return np.where(idx >= n, 2 * (n - 1) - idx, idx)


def clahe_luts(src, clip=2.0, grid=(8, 8)):
    """per-tile LUTs [ty, tx, 256] (uint8) and the tile size (th, tw) of the padded image."""
    tx, ty = grid
    h, w = src.shape
    eh = h if h % ty == 0 else h + ty - h % ty
    ew = w if w % tx == 0 else w + tx - w % tx
    if (eh, ew) != (h, w):   # the library pads BOTH axes as soon as one does not divide
        eh, ew = h + (ty - h % ty), w + (tx - w % tx)
    ext = src[reflect101(np.arange(eh), h)][:, reflect101(np.arange(ew), w)]
    th, tw = eh // ty, ew // tx
    area = th * tw
    lut_scale = F(255.0) / F(area)
    cl = max(int(clip * area / 256), 1) if clip > 0 else 0
    luts = np.zeros((ty, tx, 256), np.uint8)
    for j in range(ty):
        for i in range(tx):
            hist = np.bincount(ext[j * th:(j + 1) * th, i * tw:(i + 1) * tw].ravel(), minlength=256).astype(np.int64)
            if cl > 0:
                clipped = int(np.maximum(hist - cl, 0).sum())
                hist = np.minimum(hist, cl) + clipped // 256
                resid = clipped % 256
                if resid:
                    step = max(256 // resid, 1)
                    hist[np.arange(resid) * step] += 1
            luts[j, i] = np.clip(np.rint(np.cumsum(hist).astype(np.float32) * lut_scale), 0, 255).astype(np.uint8)
    return luts, (th, tw)


def clahe(src, clip=2.0, grid=(8, 8)):
    """cv2.createCLAHE(clip, grid).apply(src) for uint8 HxW."""
    tx, ty = grid
    h, w = src.shape
    luts, (th, tw) = clahe_luts(src, clip, grid)
    xs = np.arange(w, dtype=np.float32) * (F(1.0) / F(tw)) - F(0.5)
    tx1 = np.floor(xs).astype(np.int32)
    xa = (xs - tx1.astype(np.float32)).astype(np.float32)
    xa1 = F(1.0) - xa
    tx2 = np.minimum(tx1 + 1, tx - 1)
    tx1 = np.maximum(tx1, 0)
    ys = np.arange(h, dtype=np.float32) * (F(1.0) / F(th)) - F(0.5)
    ty1 = np.floor(ys).astype(np.int32)
    ya = (ys - ty1.astype(np.float32)).astype(np.float32)
    ya1 = F(1.0) - ya
    ty2 = np.minimum(ty1 + 1, ty - 1)
    ty1 = np.maximum(ty1, 0)
    out = np.zeros_like(src)
    for y in range(h):
        v = src[y]
        a, b = luts[ty1[y], tx1, v].astype(np.float32), luts[ty1[y], tx2, v].astype(np.float32)
        c, d = luts[ty2[y], tx1, v].astype(np.float32), luts[ty2[y], tx2, v].astype(np.float32)
        res = (a * xa1 + b * xa) * ya1[y] + (c * xa1 + d * xa) * ya[y]
        out[y] = np.clip(np.rint(res), 0, 255).astype(np.uint8)
    return out


def bilateral_tables(d, sigma_color, sigma_space, cn):
    radius = d // 2
    gcc, gsc = -0.5 / (sigma_color * sigma_color), -0.5 / (sigma_space * sigma_space)
    cw = np.array([F(math.exp(i * i * gcc)) for i in range(256 * cn)], np.float32)
    offs, sw = [], []
    for i in range(-radius, radius + 1):
        for j in range(-radius, radius + 1):
            r = math.sqrt(i * i + j * j)
            if r > radius or (i == 0 and j == 0):
                continue
            sw.append(F(math.exp(r * r * gsc)))
            offs.append((i, j))
    return radius, cw, offs, np.array(sw, np.float32)


def bilateral(src, d=9, sigma_color=75.0, sigma_space=75.0):
    """cv2.bilateralFilter(src, d, sigma_color, sigma_space) for uint8 HxW or HxWx3 (see the module docstring for
    what "equal" means here)."""
    squeeze = src.ndim == 2
    img = src[..., None] if squeeze else src
    h, w, cn = img.shape
    radius, cw, offs, sw = bilateral_tables(d, sigma_color, sigma_space, cn)
    yy = reflect101(np.arange(-radius, h + radius), h)
    xx = reflect101(np.arange(-radius, w + radius), w)
    tmp = img[yy][:, xx].astype(np.int32)
    c = tmp[radius:radius + h, radius:radius + w]
    wsum = np.zeros((h, w), np.float32)
    acc = np.zeros((h, w, cn), np.float32)
    for (i, j), s in zip(offs, sw):
        nb = tmp[radius + i:radius + i + h, radius + j:radius + j + w]
        wgt = (s * cw[np.abs(nb - c).sum(-1)]).astype(np.float32)
        wsum = (wsum + wgt).astype(np.float32)
        acc = (nb.astype(np.float64) * wgt[..., None].astype(np.float64) + acc.astype(np.float64)).astype(np.float32)  # fma
    wsum = (wsum + F(1)).astype(np.float32)
    acc = (acc + c.astype(np.float32)).astype(np.float32)
    out = np.clip(np.rint((acc / wsum[..., None]).astype(np.float32)), 0, 255).astype(np.uint8)
    return out[..., 0] if squeeze else out


def enhance_clahe_bilateral(image, clip=2.0, grid=(8, 8), d=9, sigma_color=75.0, sigma_space=75.0):
    """image_enhancement.py:146-184."""
    if image.ndim == 3:
        lab = bgr2lab(image)
        lab[..., 0] = clahe(np.ascontiguousarray(lab[..., 0]), clip, grid)
        enhanced = lab2bgr(lab)
    else:
        enhanced = clahe(image, clip, grid)
    return bilateral(enhanced, d, sigma_color, sigma_space)
