"""The ctypes mirror of include/tcd.h against the header itself, as a C compiler lays it out.

A binding a maintainer adds on the reference side and this package's own `_lib.py` both depend on the
struct layouts and enum values of include/tcd.h.  gcc compiles a small program that prints sizeof / offsetof of every
field of every struct the header declares with a body, and the value of every op kind; the ctypes classes must agree
field by field.  No GPU, no CUDA: the header is plain C.
"""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest

from telescope_cam_detection_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "tcd.h")

STRUCTS = {
    "tcd_preproc_cfg": _lib.PreprocCfg,
    "tcd_detector_cfg": _lib.DetectorCfg,
    "tcd_class_filter": _lib.ClassFilter,
    "tcd_conv_desc": _lib.ConvDesc,
    "tcd_op_desc": _lib.OpDesc,
    "tcd_enhance_cfg": _lib.EnhanceCfg,
}


def _c_layout(tmp_path):
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "tcd.h"', 'int main(void) {']
    for cname, cls in STRUCTS.items():
        lines.append(f'  printf("S {cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("F {cname} {fname} %zu %zu\\n", offsetof({cname}, {fname}), sizeof((({cname}*)0)->{fname}));')
    for kind in _lib.OP_KINDS:
        lines.append(f'  printf("K {kind} %d\\n", (int)TCD_OP_{kind.upper()});')
    for i, name in enumerate(("TCD_OK", "TCD_ERR_INVALID", "TCD_ERR_CUDA", "TCD_ERR_NOMEM", "TCD_ERR_UNSUPPORTED")):
        lines.append(f'  printf("E {name} %d\\n", (int){name});')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.dirname(HEADER), str(src), "-o", str(exe)], check=True)
    return subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines()


@pytest.fixture(scope="module")
def layout(tmp_path_factory):
    if shutil.which("gcc") is None:
        pytest.skip("gcc not on PATH")
    return _c_layout(tmp_path_factory.mktemp("abi"))


def test_header_is_plain_c_and_every_mirrored_field_exists(layout):
    assert any(l.startswith("S tcd_conv_desc") for l in layout)


def test_struct_sizes_and_field_offsets_match_ctypes(layout):
    sizes = {l.split()[1]: int(l.split()[2]) for l in layout if l.startswith("S ")}
    fields = {(l.split()[1], l.split()[2]): (int(l.split()[3]), int(l.split()[4])) for l in layout if l.startswith("F ")}
    for cname, cls in STRUCTS.items():
        assert C.sizeof(cls) == sizes[cname], cname
        for fname, _ in cls._fields_:
            d = getattr(cls, fname)
            assert (d.offset, d.size) == fields[(cname, fname)], (cname, fname)


def test_no_struct_field_of_the_header_is_missing_from_the_mirror():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    for cname, cls in STRUCTS.items():
        m = re.search(r"typedef struct(?:\s+\w+)?\s*\{([^}]*)\}\s*" + cname + r"\s*;", text)
        assert m, cname
        declared = []
        for stmt in m.group(1).split(";"):
            stmt = stmt.strip()
            if not stmt:
                continue
            # "const void *a, *b" / "int32_t N, H, W" / "float mean[3]"
            first, *rest = stmt.split(",")
            names = [first.split()[-1]] + [r.strip() for r in rest]
            declared += [re.sub(r"\[.*\]", "", n).lstrip("*").strip() for n in names]
        assert declared == [f for f, _ in cls._fields_], cname


def test_op_kinds_and_status_codes_match(layout):
    kinds = {l.split()[1]: int(l.split()[2]) for l in layout if l.startswith("K ")}
    assert kinds == _lib.OP_KINDS
    codes = [int(l.split()[2]) for l in layout if l.startswith("E ")]
    assert codes == [0, 1, 2, 3, 4]


def test_every_op_kind_of_the_header_is_bound():
    text = open(HEADER).read()
    in_header = set(re.findall(r"\bTCD_OP_([A-Z0-9_]+)\b\s*(?:=|,|\n|\})", text))
    in_header = {k.lower() for k in in_header}
    assert in_header == set(_lib.OP_KINDS), in_header ^ set(_lib.OP_KINDS)
