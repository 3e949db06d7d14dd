"""ORACLE (test infrastructure only) -- CPU restatement of the JPEG encode on the result path.

Reference: /root/reference/src/web_server.py:923 and src/snapshot_saver.py:160 (`cv2.imencode('.jpg', frame,
[cv2.IMWRITE_JPEG_QUALITY, q])`), src/snapshot_saver.py:358-404 (`cv2.imwrite(path, frame)`, quality 95).  The arithmetic
lives in the reference's dependency opencv-python, which bundles libjpeg-turbo (3.1.2 in this image's cv2 4.13.0; source
not under /root/reference).  What it does for a BGR uint8 frame is the baseline sequential JPEG of the IJG library, restated
here in NumPy:
  colour conversion    16-bit fixed-point RGB -> YCbCr (the 0.5 - epsilon rounding fudge on Cb / Cr included)
  edges                luma: last column / row replicated to whole 16x16 MCUs; chroma: columns replicated at full resolution,
                       rows only to an even count, then copies of the last DOWNSAMPLED row; luma blocks beyond
                       ceil(w/8) x ceil(h/8) are "dummy" blocks (zero AC, the DC of the block coded just before them)
  4:2:0 downsampling   2x2 sums with the alternating 1, 2 rounding bias
  forward DCT          the "islow" integer DCT (13-bit constants, 2 extra bits after the row pass), output scaled by 8
  quantisation         Annex K tables scaled by the IJG quality rule, round-half-up division with the sign taken out
  entropy coding       Annex K Huffman tables, interleaved MCUs (Y00 Y01 Y10 Y11 Cb Cr), DC prediction per component,
                       0xFF byte stuffing, final byte padded with ones
  headers              SOI, JFIF 1.01 APP0 (density 1:1, no units), two DQT, SOF0, four DHT, SOS, EOI

Pinning: tests/test_jpeg.py requires `encode()` to return the SAME BYTES as cv2.imencode for every test image (sizes with
and without whole MCUs, even / odd, qualities 20-100) -- the reference's own encoder, run in place.  Nothing else to pin.
"""
import struct

import numpy as np

STD_LUMA = np.array([16,11,10,16,24,40,51,61,12,12,14,19,26,58,60,55,14,13,16,24,40,57,69,56,14,17,22,29,51,87,80,62,
                     18,22,37,56,68,109,103,77,24,35,55,64,81,104,113,92,49,64,78,87,103,121,120,101,72,92,95,98,112,100,103,99])
STD_CHROMA = np.array([17,18,24,47,99,99,99,99,18,21,26,66,99,99,99,99,24,26,56,99,99,99,99,99,47,66,99,99,99,99,99,99,
                       99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99,99])
ZIGZAG = np.array([0,1,8,16,9,2,3,10,17,24,32,25,18,11,4,5,12,19,26,33,40,48,41,34,27,20,13,6,7,14,21,28,
                   35,42,49,56,57,50,43,36,29,22,15,23,30,37,44,51,58,59,52,45,38,31,39,46,53,60,61,54,47,55,62,63])
DC_L_BITS = [0,1,5,1,1,1,1,1,1,0,0,0,0,0,0,0]; DC_L_VALS = list(range(12))
DC_C_BITS = [0,3,1,1,1,1,1,1,1,1,1,0,0,0,0,0]; DC_C_VALS = list(range(12))
AC_L_BITS = [0,2,1,3,3,2,4,3,5,5,4,4,0,0,1,0x7d]
AC_L_VALS = [0x01,0x02,0x03,0x00,0x04,0x11,0x05,0x12,0x21,0x31,0x41,0x06,0x13,0x51,0x61,0x07,0x22,0x71,0x14,0x32,0x81,0x91,0xa1,0x08,
 0x23,0x42,0xb1,0xc1,0x15,0x52,0xd1,0xf0,0x24,0x33,0x62,0x72,0x82,0x09,0x0a,0x16,0x17,0x18,0x19,0x1a,0x25,0x26,0x27,0x28,
 0x29,0x2a,0x34,0x35,0x36,0x37,0x38,0x39,0x3a,0x43,0x44,0x45,0x46,0x47,0x48,0x49,0x4a,0x53,0x54,0x55,0x56,0x57,0x58,0x59,
 0x5a,0x63,0x64,0x65,0x66,0x67,0x68,0x69,0x6a,0x73,0x74,0x75,0x76,0x77,0x78,0x79,0x7a,0x83,0x84,0x85,0x86,0x87,0x88,0x89,
 0x8a,0x92,0x93,0x94,0x95,0x96,0x97,0x98,0x99,0x9a,0xa2,0xa3,0xa4,0xa5,0xa6,0xa7,0xa8,0xa9,0xaa,0xb2,0xb3,0xb4,0xb5,0xb6,
 0xb7,0xb8,0xb9,0xba,0xc2,0xc3,0xc4,0xc5,0xc6,0xc7,0xc8,0xc9,0xca,0xd2,0xd3,0xd4,0xd5,0xd6,0xd7,0xd8,0xd9,0xda,0xe1,0xe2,
 0xe3,0xe4,0xe5,0xe6,0xe7,0xe8,0xe9,0xea,0xf1,0xf2,0xf3,0xf4,0xf5,0xf6,0xf7,0xf8,0xf9,0xfa]
AC_C_BITS = [0,2,1,2,4,4,3,4,7,5,4,4,0,1,2,0x77]
AC_C_VALS = [0x00,0x01,0x02,0x03,0x11,0x04,0x05,0x21,0x31,0x06,0x12,0x41,0x51,0x07,0x61,0x71,0x13,0x22,0x32,0x81,0x08,0x14,0x42,0x91,
 0xa1,0xb1,0xc1,0x09,0x23,0x33,0x52,0xf0,0x15,0x62,0x72,0xd1,0x0a,0x16,0x24,0x34,0xe1,0x25,0xf1,0x17,0x18,0x19,0x1a,0x26,
 0x27,0x28,0x29,0x2a,0x35,0x36,0x37,0x38,0x39,0x3a,0x43,0x44,0x45,0x46,0x47,0x48,0x49,0x4a,0x53,0x54,0x55,0x56,0x57,0x58,
 0x59,0x5a,0x63,0x64,0x65,0x66,0x67,0x68,0x69,0x6a,0x73,0x74,0x75,0x76,0x77,0x78,0x79,0x7a,0x82,0x83,0x84,0x85,0x86,0x87,
 0x88,0x89,0x8a,0x92,0x93,0x94,0x95,0x96,0x97,0x98,0x99,0x9a,0xa2,0xa3,0xa4,0xa5,0xa6,0xa7,0xa8,0xa9,0xaa,0xb2,0xb3,0xb4,
 0xb5,0xb6,0xb7,0xb8,0xb9,0xba,0xc2,0xc3,0xc4,0xc5,0xc6,0xc7,0xc8,0xc9,0xca,0xd2,0xd3,0xd4,0xd5,0xd6,0xd7,0xd8,0xd9,0xda,
 0xe2,0xe3,0xe4,0xe5,0xe6,0xe7,0xe8,0xe9,0xea,0xf2,0xf3,0xf4,0xf5,0xf6,0xf7,0xf8,0xf9,0xfa]

def quant_tables(quality):
    quality = min(max(quality, 1), 100)
    scale = 5000 // quality if quality < 50 else 200 - quality * 2
    out = []
    for base in (STD_LUMA, STD_CHROMA):
        t = (base.astype(np.int64) * scale + 50) // 100
        out.append(np.clip(t, 1, 255).astype(np.int64))
    return out

def huff_codes(bits, vals):
    code, k, ehufco, ehufsi = 0, 0, {}, {}
    for l in range(1, 17):
        for _ in range(bits[l - 1]):
            ehufco[vals[k]] = code; ehufsi[vals[k]] = l; code += 1; k += 1
        code <<= 1
    return ehufco, ehufsi

def fix(x): return int(x * 65536 + 0.5)

def color_convert(img):
    b = img[..., 0].astype(np.int64); g = img[..., 1].astype(np.int64); r = img[..., 2].astype(np.int64)
    half, off = 1 << 15, 128 << 16
    y = (fix(0.29900) * r + fix(0.58700) * g + fix(0.11400) * b + half) >> 16
    cb = (-fix(0.16874) * r - fix(0.33126) * g + fix(0.50000) * b + off + half - 1) >> 16
    cr = (fix(0.50000) * r - fix(0.41869) * g - fix(0.08131) * b + off + half - 1) >> 16
    return y, cb, cr

def pad_edge(p, H, W):
    h, w = p.shape
    return np.pad(p, ((0, H - h), (0, W - w)), mode="edge")

def downsample_h2v2(p):
    s = p[0::2, 0::2] + p[0::2, 1::2] + p[1::2, 0::2] + p[1::2, 1::2]
    bias = np.where(np.arange(s.shape[1]) % 2 == 0, 1, 2)[None, :]
    return (s + bias) >> 2

def fdct_islow(block):
    CONST_BITS, PASS1_BITS = 13, 2
    F = lambda x: int(x * (1 << CONST_BITS) + 0.5)
    F_0_298, F_0_390, F_0_541, F_0_765, F_0_899, F_1_175 = F(0.298631336), F(0.390180644), F(0.541196100), F(0.765366865), F(0.899976223), F(1.175875602)
    F_1_501, F_1_847, F_1_961, F_2_053, F_2_562, F_3_072 = F(1.501321110), F(1.847759065), F(1.961570560), F(2.053119869), F(2.562915447), F(3.072711026)
    d = block.astype(np.int64).copy()   # [..., 8, 8]
    def desc(x, n): return (x + (1 << (n - 1))) >> n
    def pass_(d, first):
        t0 = d[..., 0] + d[..., 7]; t7 = d[..., 0] - d[..., 7]
        t1 = d[..., 1] + d[..., 6]; t6 = d[..., 1] - d[..., 6]
        t2 = d[..., 2] + d[..., 5]; t5 = d[..., 2] - d[..., 5]
        t3 = d[..., 3] + d[..., 4]; t4 = d[..., 3] - d[..., 4]
        t10 = t0 + t3; t13 = t0 - t3; t11 = t1 + t2; t12 = t1 - t2
        o = np.empty_like(d)
        if first:
            o[..., 0] = (t10 + t11) << PASS1_BITS; o[..., 4] = (t10 - t11) << PASS1_BITS
        else:
            o[..., 0] = desc(t10 + t11, PASS1_BITS); o[..., 4] = desc(t10 - t11, PASS1_BITS)
        z1 = (t12 + t13) * F_0_541
        n = CONST_BITS - PASS1_BITS if first else CONST_BITS + PASS1_BITS
        o[..., 2] = desc(z1 + t13 * F_0_765, n); o[..., 6] = desc(z1 + t12 * (-F_1_847), n)
        z1 = t4 + t7; z2 = t5 + t6; z3 = t4 + t6; z4 = t5 + t7; z5 = (z3 + z4) * F_1_175
        t4 = t4 * F_0_298; t5 = t5 * F_2_053; t6 = t6 * F_3_072; t7 = t7 * F_1_501
        z1 = z1 * (-F_0_899); z2 = z2 * (-F_2_562); z3 = z3 * (-F_1_961); z4 = z4 * (-F_0_390)
        z3 = z3 + z5; z4 = z4 + z5
        o[..., 7] = desc(t4 + z1 + z3, n); o[..., 5] = desc(t5 + z2 + z4, n)
        o[..., 3] = desc(t6 + z2 + z3, n); o[..., 1] = desc(t7 + z1 + z4, n)
        return o
    d = pass_(d, True)                       # rows
    d = np.swapaxes(pass_(np.swapaxes(d, -1, -2), False), -1, -2)   # columns
    return d

def blocks_of(p):
    H, W = p.shape
    return p.reshape(H // 8, 8, W // 8, 8).swapaxes(1, 2)   # [by, bx, 8, 8]

def quantize(coef, q):
    qv = (q.reshape(8, 8) << 3)
    a = np.abs(coef)
    r = (a + (qv >> 1)) // qv
    return np.where(coef < 0, -r, r)

class BitWriter:
    def __init__(self): self.buf = bytearray(); self.acc = 0; self.n = 0
    def put(self, code, size):
        self.acc = (self.acc << size) | (code & ((1 << size) - 1)); self.n += size
        while self.n >= 8:
            b = (self.acc >> (self.n - 8)) & 0xFF
            self.buf.append(b)
            if b == 0xFF: self.buf.append(0)
            self.n -= 8
        self.acc &= (1 << self.n) - 1
    def flush(self):
        if self.n: self.put(0x7F, 7 if False else (8 - self.n))   # pad with ones
def nbits_of(v):
    v = abs(int(v)); n = 0
    while v: n += 1; v >>= 1
    return n

def encode_block(bw, blk, last_dc, dc, ac):
    zz = blk.reshape(64)[ZIGZAG]
    diff = int(zz[0]) - last_dc
    nb = nbits_of(diff)
    bw.put(dc[0][nb], dc[1][nb])
    if nb: bw.put(diff if diff >= 0 else diff - 1, nb)
    r = 0
    for k in range(1, 64):
        v = int(zz[k])
        if v == 0: r += 1; continue
        while r > 15:
            bw.put(ac[0][0xF0], ac[1][0xF0]); r -= 16
        nb = nbits_of(v)
        sym = (r << 4) | nb
        bw.put(ac[0][sym], ac[1][sym]); bw.put(v if v >= 0 else v - 1, nb)
        r = 0
    if r > 0: bw.put(ac[0][0], ac[1][0])
    return int(zz[0])

def headers(h, w, qt):
    out = bytearray(b"\xff\xd8")
    out += b"\xff\xe0" + struct.pack(">H", 16) + b"JFIF\x00\x01\x01\x00\x00\x01\x00\x01\x00\x00"
    for i, q in enumerate(qt):
        out += b"\xff\xdb" + struct.pack(">H", 67) + bytes([i]) + bytes(int(x) for x in q[ZIGZAG])
    out += b"\xff\xc0" + struct.pack(">HBHHB", 17, 8, h, w, 3) + bytes([1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1])
    for tc_th, bits, vals in ((0x00, DC_L_BITS, DC_L_VALS), (0x10, AC_L_BITS, AC_L_VALS), (0x01, DC_C_BITS, DC_C_VALS), (0x11, AC_C_BITS, AC_C_VALS)):
        out += b"\xff\xc4" + struct.pack(">H", 2 + 1 + 16 + len(vals)) + bytes([tc_th]) + bytes(bits) + bytes(vals)
    out += b"\xff\xda" + struct.pack(">H", 12) + bytes([3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0])
    return out

def encode(img, quality=95):
    h, w = img.shape[:2]
    qt = quant_tables(quality)
    y, cb, cr = color_convert(img)
    H, W = (h + 15) // 16 * 16, (w + 15) // 16 * 16
    # luma: last column / row replicated to whole MCUs.  chroma: columns replicated at full resolution (before the
    # 2x2 average), rows replicated at full resolution only to an even count -- the rest of the MCU row is filled with
    # copies of the last DOWNSAMPLED row
    he = (h + 1) // 2 * 2
    y = pad_edge(y, H, W)
    cb, cr = (pad_edge(downsample_h2v2(pad_edge(c, he, W)), H // 2, W // 2) for c in (cb, cr))
    qy = quantize(fdct_islow(blocks_of(y - 128)), qt[0])
    qcb = quantize(fdct_islow(blocks_of(cb - 128)), qt[1])
    qcr = quantize(fdct_islow(blocks_of(cr - 128)), qt[1])
    dcl, acl = huff_codes(DC_L_BITS, DC_L_VALS), huff_codes(AC_L_BITS, AC_L_VALS)
    dcc, acc = huff_codes(DC_C_BITS, DC_C_VALS), huff_codes(AC_C_BITS, AC_C_VALS)
    bw = BitWriter(); last = [0, 0, 0]
    wb, hb = (w + 7) // 8, (h + 7) // 8          # real Y blocks; the rest of the last MCU column / row are dummies
    for my in range(H // 16):
        for mx in range(W // 16):
            prev = None
            for (dy, dx) in ((0, 0), (0, 1), (1, 0), (1, 1)):
                by, bx = 2 * my + dy, 2 * mx + dx
                if by < hb and bx < wb:
                    blk = qy[by, bx]
                else:                              # dummy block: zero AC, DC of the block coded just before it
                    blk = np.zeros((8, 8), np.int64); blk[0, 0] = prev[0, 0]
                prev = blk
                last[0] = encode_block(bw, blk, last[0], dcl, acl)
            last[1] = encode_block(bw, qcb[my, mx], last[1], dcc, acc)
            last[2] = encode_block(bw, qcr[my, mx], last[2], dcc, acc)
    if bw.n: bw.put((1 << (8 - bw.n)) - 1, 8 - bw.n)
    return bytes(headers(h, w, qt)) + bytes(bw.buf) + b"\xff\xd9"
