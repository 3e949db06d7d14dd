"""Builds libtcd.so (hand-written sm_100a CUDA + C ABI) in-tree with nvcc.

No torch extension machinery: the library has plain C entry points (include/tcd.h) and is
loaded with ctypes.  The .so lands next to this file so it travels with the repo snapshot.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libtcd.so")
OBJ = os.path.join(HERE, "build")

# TUs whose fp32 arithmetic must round exactly like the scalar reference code (no FMA contraction)
NO_FMAD = {"postproc.cu", "motion.cu", "enhance.cu", "noalloc.cu"}
SOURCES = ["capi_core.cu", "capi_detector.cu", "capi_motion.cu", "capi_ingest.cu", "capi_enhance.cu", "enhance.cu", "capi_jpeg.cu", "jpeg.cu", "noalloc.cu", "conv_umma.cu", "yolox_net.cu", "program.cu",
           "tfm_kernels.cu", "tfm_kernels2.cu", "vit_kernels.cu", "fa_umma.cu", "pil_resize.cu",
           "yolox_plan.cu", "preproc.cu", "postproc.cu", "motion.cu", "crop_resize.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _digest():
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for name in sorted(os.listdir(root)):
            if name.endswith((".cu", ".cuh", ".h", ".c")):
                with open(os.path.join(root, name), "rb") as f:
                    h.update(name.encode())
                    h.update(f.read())
    return h.hexdigest()


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    stamp = os.path.join(OBJ, "digest.txt")
    dig = _digest()
    if not force and os.path.exists(OUT) and os.path.exists(stamp) and open(stamp).read() == dig:
        return OUT
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = [nvcc, *ARCH, *COMMON, "-c", os.path.join(CSRC, src), "-o", obj]
        if src in NO_FMAD:
            cmd.insert(1, "-fmad=false")
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        with open(obj + ".log", "w") as f:
            f.write(" ".join(cmd) + "\n" + log)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, log))
        if verbose:
            print(log)
        return obj

    with ThreadPoolExecutor(max_workers=min(12, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, *ARCH, "-shared", "-o", OUT, *objs, "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    build_fmt()
    with open(stamp, "w") as f:
        f.write(dig)
    return OUT


def fmt_path():
    import sysconfig
    return os.path.join(HERE, "_tcdfmt" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))


def build_fmt():
    """The CPython helper that builds the detection dicts (csrc/tcdfmt.c; host logic, plain gcc)."""
    import sysconfig
    inc = sysconfig.get_paths()["include"]
    if not os.path.exists(os.path.join(inc, "Python.h")):
        return None   # no headers: YOLOXDetector keeps its NumPy formatter
    out = fmt_path()
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-I", inc, os.path.join(CSRC, "tcdfmt.c"), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gcc failed for tcdfmt.c:\n" + r.stdout + r.stderr)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
