"""Bounding-box normalisation helpers, same behaviour as the reference's src/bbox_utils.py
(ensure_valid_bbox :12-59, validate_bbox_coords :62-88, is_valid_bbox :91-122)."""


def _ordered(lo, hi, min_size):
    if lo > hi:
        lo, hi = hi, lo
    if hi - lo < min_size:
        hi = lo + min_size
    return lo, hi


def validate_bbox_coords(x1, y1, x2, y2, min_size=1):
    x1, x2 = _ordered(x1, x2, min_size)
    y1, y2 = _ordered(y1, y2, min_size)
    return (x1, y1, x2, y2)


def ensure_valid_bbox(bbox, min_size=1):
    """Swap inverted corners, enforce the minimum size, recompute width/height/area (float)."""
    x1, y1, x2, y2 = validate_bbox_coords(bbox["x1"], bbox["y1"], bbox["x2"], bbox["y2"], min_size)
    w, h = x2 - x1, y2 - y1
    return {"x1": x1, "y1": y1, "x2": x2, "y2": y2, "width": w, "height": h, "area": w * h}


def is_valid_bbox(bbox, min_size=1):
    try:
        x1, y1, x2, y2 = bbox["x1"], bbox["y1"], bbox["x2"], bbox["y2"]
    except (KeyError, TypeError):
        return False
    try:
        if x1 >= x2 or y1 >= y2 or x2 - x1 < min_size or y2 - y1 < min_size:
            return False
        return min(x1, y1, x2, y2) >= 0
    except TypeError:
        return False
