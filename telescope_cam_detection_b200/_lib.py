"""ctypes binding of libtcd.so (include/tcd.h).  No fallback: a missing library or a
non-sm_100 device raises, it never silently routes to PyTorch or the CPU oracle."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libtcd.so")


class TcdError(RuntimeError):
    pass


class PreprocCfg(C.Structure):
    _fields_ = [("out_h", C.c_int32), ("out_w", C.c_int32), ("mode", C.c_int32),
                ("swap_rb", C.c_int32), ("scale", C.c_float), ("mean", C.c_float * 3),
                ("inv_std", C.c_float * 3), ("pad_value", C.c_float)]


class DetectorCfg(C.Structure):
    _fields_ = [("input_h", C.c_int32), ("input_w", C.c_int32), ("max_batch", C.c_int32),
                ("num_classes", C.c_int32), ("max_out", C.c_int32), ("max_src_h", C.c_int32),
                ("max_src_w", C.c_int32), ("pre", PreprocCfg)]


MAX_CLASSES = 128          # TCD_MAX_CLASSES
MAX_FILTER_TABLES = 64     # TCD_MAX_FILTER_TABLES


class ClassFilter(C.Structure):
    """tcd_class_filter (include/tcd.h)."""
    _fields_ = [("cls_conf", C.c_double * MAX_CLASSES), ("cls_min_area", C.c_double * MAX_CLASSES),
                ("cls_max_area", C.c_double * MAX_CLASSES), ("min_area", C.c_double),
                ("class_mask", C.c_uint64 * 2), ("max_det", C.c_int32), ("reserved", C.c_int32)]


class ConvDesc(C.Structure):
    """tcd_conv_desc (include/tcd.h)."""
    _fields_ = [("name", C.c_char * 64),
                ("x", C.c_void_p), ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("x_ctot", C.c_int32),
                ("x_coff", C.c_int32), ("cin", C.c_int32), ("x_nstride", C.c_int64),
                ("w_host", C.c_void_p), ("bias_host", C.c_void_p),
                ("cout", C.c_int32), ("k", C.c_int32), ("stride", C.c_int32), ("act", C.c_int32),
                ("y", C.c_void_p), ("y_ctot", C.c_int32), ("y_coff", C.c_int32), ("y_nstride", C.c_int64),
                ("resid", C.c_void_p), ("r_ctot", C.c_int32), ("r_coff", C.c_int32), ("resid_pre", C.c_int32),
                ("r_nstride", C.c_int64),
                ("out_f32", C.c_void_p), ("out_ld", C.c_int32), ("out_col", C.c_int32),
                ("rows_per_img", C.c_int32), ("row_off", C.c_int32),
                ("ln_stats", C.c_void_p), ("ln_colsum_host", C.c_void_p),
                ("ln_parts", C.c_int32), ("ln_count", C.c_int32), ("ln_eps", C.c_float),
                ("stats_out", C.c_void_p), ("stats_parts", C.c_int32), ("stats_slot0", C.c_int32)]


class OpDesc(C.Structure):
    """tcd_op_desc (include/tcd.h)."""
    _fields_ = [("name", C.c_char * 64), ("kind", C.c_int32), ("p", C.c_void_p * 8), ("i", C.c_int32 * 12),
                ("f", C.c_float * 4)]


OP_KINDS = {"layernorm": 1, "add_rows": 2, "mask_rows": 3, "mha": 4, "msda": 5, "rowmax": 6, "topk": 7,
            "gather": 8, "ref_init": 9, "ref_update": 10, "detr_post": 11, "maxpool": 12, "avgpool": 13,
            "upsample": 14, "vit_assemble": 15, "mha_rope": 16, "swiglu_ln": 17, "softmax": 18, "take": 19, "head_gemv": 20, "row_stats": 21}


class EnhanceCfg(C.Structure):
    _fields_ = [("clip_limit", C.c_float), ("grid_x", C.c_int32), ("grid_y", C.c_int32), ("d", C.c_int32),
                ("sigma_color", C.c_float), ("sigma_space", C.c_float)]


def make_preproc_cfg(out_h, out_w, mode=0, swap_rb=0, scale=1.0, mean=(0.0, 0.0, 0.0),
                     std=(1.0, 1.0, 1.0), pad_value=114.0):
    cfg = PreprocCfg()
    cfg.out_h, cfg.out_w, cfg.mode, cfg.swap_rb = int(out_h), int(out_w), int(mode), int(swap_rb)
    cfg.scale = float(scale)
    for i in range(3):
        cfg.mean[i] = float(mean[i])
        cfg.inv_std[i] = 1.0 / float(std[i])
    cfg.pad_value = float(pad_value)
    return cfg


_P = C.c_void_p
_I32P = C.POINTER(C.c_int32)
_I64P = C.POINTER(C.c_int64)
_F32P = C.POINTER(C.c_float)
_U8P = C.POINTER(C.c_uint8)
_PP = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); every symbol include/tcd.h declares
SIGNATURES = {
    "tcd_last_error": (C.c_char_p, []),
    "tcd_version": (C.c_int, []),
    "tcd_launch_count": (C.c_uint64, []),
    "tcd_h2d_byte_count": (C.c_uint64, []),
    "tcd_preprocess": (C.c_int, [C.POINTER(PreprocCfg), _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, _P, _P]),
    "tcd_conv2d_nhwc_bf16": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                       _F32P, _F32P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P,
                                       C.c_int32, C.c_int32, _P, C.c_int32, C.c_int32, _P]),
    "tcd_conv2d_nhwc_bf16_grouped": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                               _F32P, _F32P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P,
                                               C.c_int32, C.c_int32, _P, C.c_int32, C.c_int32, C.c_int32, _P]),
    "tcd_create_detector": (C.c_int, [C.POINTER(DetectorCfg), _P, C.c_size_t, C.c_int32, _PP]),
    "tcd_destroy_detector": (None, [_P]),
    "tcd_detector_set_thresholds": (C.c_int, [_P, C.c_float, C.c_float]),
    "tcd_detector_set_class_filter": (C.c_int, [_P, C.c_int32, C.POINTER(ClassFilter)]),
    "tcd_detector_filter_rows": (C.c_int, [_P, C.c_int32, _F32P, C.c_int32, C.c_int32, C.c_int32, _I32P]),
    "tcd_detector_submit": (C.c_int, [_P, C.c_int32, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, _I32P]),
    "tcd_detect_batch": (C.c_int, [_P, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, _F32P, _I32P]),
    "tcd_detect_batch_async": (C.c_int, [_P, C.c_int32, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32]),
    "tcd_detect_batch_wait": (C.c_int, [_P, C.c_int32, _F32P, _I32P]),
    "tcd_forward": (C.c_int, [_P, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, _F32P]),
    "tcd_postprocess": (C.c_int, [_P, _F32P, C.c_int32, _F32P, _I32P]),
    "tcd_detector_num_anchors": (C.c_int32, [_P]),
    "tcd_detector_last_timing": (C.c_int, [_P, _F32P]),
    "tcd_detector_stream": (C.c_void_p, [_P, C.c_int32]),
    "tcd_detector_num_slots": (C.c_int32, [_P]),
    "tcd_detector_result_device": (C.c_int, [_P, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "tcd_detector_trace_op": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_uint64), _I32P, _I32P]),
    "tcd_detector_profile": (C.c_int, [_P, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, C.c_int32, _F32P,
                                       C.c_int32, _I32P]),
    "tcd_detector_op_name": (C.c_char_p, [_P, C.c_int32, C.c_int32]),
    "tcd_detector_op_kind": (C.c_int32, [_P, C.c_int32, C.c_int32]),
    "tcd_detector_op_flops": (C.c_double, [_P, C.c_int32, C.c_int32]),
    "tcd_detector_read_activation": (C.c_int, [_P, C.c_char_p, _F32P, C.c_size_t, _I32P]),
    "tcd_mog2_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_int32, C.c_int32, _PP]),
    "tcd_mog2_destroy": (None, [_P]),
    "tcd_mog2_reset": (C.c_int, [_P]),
    "tcd_mog2_apply": (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, _P]),
    "tcd_mog2_roi_counts": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, _I32P, C.c_int32, _I32P, _P]),
    "tcd_mog2_last_ms": (C.c_int, [_P, _F32P]),
    "tcd_empty_filter_last_ms": (C.c_float, [_P]),
    "tcd_empty_filter_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, _PP]),
    "tcd_empty_filter_destroy": (None, [_P]),
    "tcd_empty_filter_reset": (C.c_int, [_P]),
    "tcd_empty_filter_step": (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _I32P, _P]),
    "tcd_stream_bank_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_int32, C.c_int32,
                                         C.c_int32, _PP]),
    "tcd_stream_bank_destroy": (None, [_P]),
    "tcd_stream_bank_reset": (C.c_int, [_P, C.c_int32, C.c_int32]),
    "tcd_stream_bank_upload": (C.c_int, [_P, _PP, _I64P]),
    "tcd_stream_bank_frame": (C.c_void_p, [_P, C.c_int32]),
    "tcd_stream_bank_cuda_stream": (C.c_void_p, [_P]),
    "tcd_stream_bank_gate": (C.c_int, [_P, C.c_int32, _I32P]),
    "tcd_stream_bank_motion": (C.c_int, [_P, _I32P, _I32P, C.c_int32, C.c_int32, C.c_int32, _I32P]),
    "tcd_stream_bank_last_ms": (C.c_int, [_P, _F32P]),
    "tcd_enhancer_create": (C.c_int, [C.POINTER(EnhanceCfg), C.c_int32, _PP]),
    "tcd_enhancer_destroy": (None, [_P]),
    "tcd_enhance_clahe_bilateral": (C.c_int, [_P, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, C.c_int32, _PP, _I64P,
                                              C.c_int32]),
    "tcd_enhance_clahe_bilateral_noalloc": (C.c_int, [C.POINTER(EnhanceCfg), C.c_int32, _P, C.c_int32, C.c_int32, C.c_int64,
                                                      C.c_int32, _P, C.c_int64, _P]),
    "tcd_enhance_noalloc_max_image_bytes": (C.c_int64, []),
    "tcd_enhancer_stream": (C.c_int, [_P, _PP]),
    "tcd_enhance_tables": (C.c_int, [C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), _U8P, _I32P, _I32P]),
    "tcd_jpeg_encoder_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, _PP]),
    "tcd_jpeg_encoder_destroy": (None, [_P]),
    "tcd_jpeg_encode": (C.c_int, [_P, _PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, C.c_int32, _PP, _I64P, _I64P]),
    "tcd_jpeg_max_bytes": (C.c_int64, [C.c_int32, C.c_int32]),
    "tcd_frame_ring_create": (C.c_int, [C.c_int32, C.c_int64, C.c_int32, _PP]),
    "tcd_frame_ring_destroy": (None, [_P]),
    "tcd_frame_ring_host": (C.c_void_p, [_P, C.c_int32]),
    "tcd_frame_ring_device": (C.c_void_p, [_P, C.c_int32]),
    "tcd_frame_ring_read_fd": (C.c_int, [_P, C.c_int32, C.c_int32, _I64P]),
    "tcd_frame_ring_publish": (C.c_int, [_P, C.c_int32]),
    "tcd_frame_ring_wait": (C.c_int, [_P, C.c_int32, _P]),
    "tcd_crop_resize": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int64, _I32P, C.c_int32, C.c_int32,
                                  C.POINTER(PreprocCfg), C.c_int32, _P, _P]),
    "tcd_crop_patches": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int64, _I32P, _P, C.c_int32, C.c_int32, C.c_int32,
                                   C.c_int32, C.POINTER(PreprocCfg), _P, _P]),
    "tcd_patchify_f32": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "tcd_program_create": (C.c_int, [C.c_int32, _PP]),
    "tcd_program_destroy": (None, [_P]),
    "tcd_program_alloc": (C.c_int, [_P, C.c_size_t, _PP]),
    "tcd_program_upload": (C.c_int, [_P, _P, C.c_size_t, _PP]),
    "tcd_program_add_conv": (C.c_int, [_P, C.POINTER(ConvDesc)]),
    "tcd_program_add_op": (C.c_int, [_P, C.POINTER(OpDesc)]),
    "tcd_program_run": (C.c_int, [_P, C.c_int32]),
    "tcd_program_sync": (C.c_int, [_P]),
    "tcd_program_stream": (C.c_void_p, [_P]),
    "tcd_program_num_ops": (C.c_int32, [_P]),
    "tcd_program_op_name": (C.c_char_p, [_P, C.c_int32]),
    "tcd_program_op_flops": (C.c_double, [_P, C.c_int32]),
    "tcd_program_profile": (C.c_int, [_P, C.c_int32, _F32P, C.c_int32]),
    "tcd_pil_resize": (C.c_int, [_PP, _I32P, _I32P, _I64P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P]),
    "tcd_malloc_device": (C.c_int, [C.c_size_t, _PP]),
    "tcd_free_device": (C.c_int, [_P]),
    "tcd_malloc_pinned": (C.c_int, [C.c_size_t, _PP]),
    "tcd_free_pinned": (C.c_int, [_P]),
    "tcd_memcpy_h2d": (C.c_int, [_P, _P, C.c_size_t]),
    "tcd_memcpy_d2h": (C.c_int, [_P, _P, C.c_size_t]),
    "tcd_device_synchronize": (C.c_int, []),
}

_lib = None


def load():
    """Load libtcd.so (once).  Raises TcdError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TcdError("libtcd.so is missing (%s): run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "or `python telescope_cam_detection_b200/build.py`; there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, what=""):
    if status != 0:
        msg = load().tcd_last_error().decode("utf-8", "replace")
        raise TcdError("%s failed (status %d): %s" % (what or "libtcd call", status, msg))


def ptr_array(ptrs):
    arr = (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p)) for p in ptrs])
    return arr


def i32_array(vals):
    return (C.c_int32 * len(vals))(*[int(v) for v in vals])


def i64_array(vals):
    return (C.c_int64 * len(vals))(*[int(v) for v in vals])
