"""tcd_enhance_clahe_bilateral_noalloc (operators' addition, no counterpart in the reference): the crop enhancer with its
scratch in a static device array, called through the C ABI and compared with the existing entry point
tcd_enhance_clahe_bilateral on the same device crops.  Tolerance: the enhancement tests' own -- bit-exact (same kernels,
same arithmetic).  Two calls in a row with different crops (the second smaller: stale scratch contents must not show),
then a single-channel crop, a strided view, other parameters, and the size limit.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from util import structured_frame

pytestmark = pytest.mark.gpu


def _cfg(_lib, clip=2.0, grid=(8, 8), d=9, sc=75.0, ss=75.0):
    cfg = _lib.EnhanceCfg()
    cfg.clip_limit, cfg.grid_x, cfg.grid_y, cfg.d = float(clip), int(grid[0]), int(grid[1]), int(d)
    cfg.sigma_color, cfg.sigma_space = float(sc), float(ss)
    return cfg


def _existing(_lib, lib, cfg, t):
    """the handle-based entry point, device in / device out"""
    import torch
    h = C.c_void_p()
    _lib.check(lib.tcd_enhancer_create(C.byref(cfg), 0, C.byref(h)), "tcd_enhancer_create")
    try:
        out = torch.empty(tuple(t.shape), dtype=torch.uint8, device=t.device)
        cn = 3 if t.ndim == 3 else 1
        _lib.check(lib.tcd_enhance_clahe_bilateral(
            h, _lib.ptr_array([t.data_ptr()]), _lib.i32_array([t.shape[0]]), _lib.i32_array([t.shape[1]]),
            _lib.i64_array([t.stride(0)]), 1, cn, 1, _lib.ptr_array([out.data_ptr()]), None, 1), "tcd_enhance_clahe_bilateral")
        return out.cpu().numpy()
    finally:
        lib.tcd_enhancer_destroy(h)


def _noalloc(_lib, lib, cfg, t, stream=None):
    import torch
    out = torch.full(tuple(t.shape), 7, dtype=torch.uint8, device=t.device)
    cn = 3 if t.ndim == 3 else 1
    torch.cuda.synchronize()
    _lib.check(lib.tcd_enhance_clahe_bilateral_noalloc(C.byref(cfg), 0, C.c_void_p(t.data_ptr()), t.shape[0], t.shape[1],
                                                       t.stride(0), cn, C.c_void_p(out.data_ptr()), 0,
                                                       C.c_void_p(stream) if stream else None),
               "tcd_enhance_clahe_bilateral_noalloc")
    return out.cpu().numpy()


def test_noalloc_entry_equals_the_existing_entry_point():
    import torch
    from telescope_cam_detection_b200 import _lib
    lib = _lib.load()
    assert lib.tcd_enhance_noalloc_max_image_bytes() == 512 * 512 * 3
    cfg = _cfg(_lib)
    # two calls in a row: a large colour crop, then a smaller one with sizes the tile grid does not divide
    first = torch.from_numpy(structured_frame(71, 400, 500)).cuda()
    second = torch.from_numpy(structured_frame(72, 131, 77)).cuda()
    ref1, ref2 = _existing(_lib, lib, cfg, first), _existing(_lib, lib, cfg, second)
    got1 = _noalloc(_lib, lib, cfg, first)
    got2 = _noalloc(_lib, lib, cfg, second)
    assert np.array_equal(got1, ref1)
    assert np.array_equal(got2, ref2)
    assert not np.array_equal(got1, first.cpu().numpy())            # it did enhance
    # again in the other order, on a caller's stream: same answers (the scratch keeps nothing that matters)
    s = torch.cuda.Stream()
    assert np.array_equal(_noalloc(_lib, lib, cfg, second, s.cuda_stream), ref2)
    assert np.array_equal(_noalloc(_lib, lib, cfg, first, s.cuda_stream), ref1)
    # a call with every buffer already there leaves the driver's free-memory count where it was
    out = torch.empty_like(first)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(lib.tcd_enhance_clahe_bilateral_noalloc(C.byref(cfg), 0, C.c_void_p(first.data_ptr()), 400, 500, 0, 3,
                                                       C.c_void_p(out.data_ptr()), 0, None), "noalloc")
    assert torch.cuda.mem_get_info()[0] == free0
    assert np.array_equal(out.cpu().numpy(), ref1)
    # single channel, and a strided view into a frame
    gray = torch.from_numpy(np.ascontiguousarray(structured_frame(73, 150, 170)[..., 1])).cuda()
    assert np.array_equal(_noalloc(_lib, lib, cfg, gray), _existing(_lib, lib, cfg, gray))
    frame = torch.from_numpy(structured_frame(74, 540, 960)).cuda()
    view = frame[100:303, 300:427]
    assert np.array_equal(_noalloc(_lib, lib, cfg, view), _existing(_lib, lib, cfg, view))
    # other parameters
    cfg2 = _cfg(_lib, 4.0, (4, 6), 5, 30.0, 10.0)
    assert np.array_equal(_noalloc(_lib, lib, cfg2, second), _existing(_lib, lib, cfg2, second))


def test_noalloc_entry_rejects_what_the_static_scratch_cannot_hold():
    import torch
    from telescope_cam_detection_b200 import _lib
    lib = _lib.load()
    big = torch.zeros((600, 600, 3), dtype=torch.uint8, device="cuda")
    out = torch.zeros_like(big)
    rc = lib.tcd_enhance_clahe_bilateral_noalloc(C.byref(_cfg(_lib)), 0, C.c_void_p(big.data_ptr()), 600, 600, 0, 3,
                                                 C.c_void_p(out.data_ptr()), 0, None)
    assert rc != 0
    msg = lib.tcd_last_error().decode()
    assert "786432" in msg and "exceeds" in msg
    rc = lib.tcd_enhance_clahe_bilateral_noalloc(C.byref(_cfg(_lib, grid=(16, 16))), 0, C.c_void_p(big.data_ptr()), 64, 64, 0, 3,
                                                 C.c_void_p(out.data_ptr()), 0, None)
    assert rc != 0 and "tile grid" in lib.tcd_last_error().decode()
    # 512 x 512 x 3 is exactly the limit and works
    edge = torch.from_numpy(structured_frame(75, 512, 512)).cuda()
    assert np.array_equal(_noalloc(_lib, lib, _cfg(_lib), edge), _existing(_lib, lib, _cfg(_lib), edge))
