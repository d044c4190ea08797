"""Build libwoft_b200.so in-tree with nvcc for sm_100a.

    python -m woft_b200.build [--force] [--verbose]

The library is plain CUDA C++ behind a C ABI (include/woft_b200.h); it does not link against torch.  The
built .so is git-ignored but travels with the working tree, so a GPU box needs no compiler.
"""
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libwoft_b200.so"
STAMP = PKG / ".libwoft_b200.stamp"
SOURCES = ["api.cu", "lookup.cu", "homography.cu", "corr_simt.cu", "corr_tc.cu", "alt_corr.cu", "corr_static.cu",
           "fused_elementwise.cu", "prewarp.cu", "small_convs.cu", "upsample.cu", "select.cu", "preprocess.cu",
           "stem_conv.cu", "weight_head.cu", "conv_tc.cu", "lookup_tma.cu"]
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


def _fingerprint():
    h = hashlib.sha256()
    files = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "woft_b200.h"]
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compile every CUDA source into one shared library.  Returns the library path."""
    fp = _fingerprint()
    if not force and LIB.exists() and STAMP.exists() and STAMP.read_text().strip() == fp:
        return LIB
    nvcc = _nvcc()
    if nvcc is None:
        if LIB.exists():
            return LIB  # GPU box without a toolkit: use the library that travelled with the tree
        raise RuntimeError("nvcc not found and no prebuilt libwoft_b200.so present")
    cmd = [nvcc, "-O3", "-std=c++17", "--threads", "0", *ARCH_FLAGS, "-lineinfo", "-Xcompiler", "-fPIC", "-shared",
           "-I", str(PKG.parent / "include"), "-o", str(LIB)]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [str(CSRC / s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libwoft_b200.so")
    STAMP.write_text(fp)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
