"""Compile the reference's one native op, `alt_cuda_corr`, from its own sources into oracle/_ref/.  TEST
INFRASTRUCTURE ONLY.

    python -m oracle.build_ref        # build container only: needs /root/reference and nvcc

The two source files are compiled where they lie (/root/reference/pytracking/external/RAFT/alt_cuda_corr/
correlation.cpp and correlation_kernel.cu); nothing is copied into this repository and the reference's setup.py is
not run.  The output, oracle/_ref/alt_cuda_corr.so, is git-ignored but travels to the GPU box with the tree, where
tests/test_gpu_parity.py::test_alt_corr_forward_vs_reference_extension imports it as the oracle of
woft_alt_corr_forward (the pure-torch restatement in woft_oracle.alt_corr_forward is pinned by it there as well).
"""
import subprocess
import sys
import sysconfig
from pathlib import Path

SRC = Path("/root/reference/pytracking/external/RAFT/alt_cuda_corr")
OUT = Path(__file__).resolve().parent / "_ref"


def build(verbose=False):
    """Returns the path of the built extension, or None when the reference sources or nvcc are not present."""
    import shutil
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    so = OUT / "alt_cuda_corr.so"
    if not SRC.is_dir() or not Path(nvcc).exists():
        return so if so.exists() else None
    srcs = [SRC / "correlation.cpp", SRC / "correlation_kernel.cu"]
    if so.exists() and all(so.stat().st_mtime >= f.stat().st_mtime for f in srcs):
        return so
    from torch.utils import cpp_extension as ce
    OUT.mkdir(exist_ok=True)
    inc = [f"-I{p}" for p in ce.include_paths("cuda")] + [f"-I{sysconfig.get_paths()['include']}"]
    defs = ["-DTORCH_EXTENSION_NAME=alt_cuda_corr", "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    objs, procs = [], []
    for f in srcs:  # both at once: each takes minutes in torch's headers
        o = OUT / (f.stem + ".o")
        cmd = [nvcc, "-c", "-O2", "-std=c++17", "-gencode", "arch=compute_100,code=sm_100", "-Xcompiler", "-fPIC",
               "-w", "-x", "cu", *defs, *inc, str(f), "-o", str(o)]
        procs.append((f, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(str(o))
    for f, proc in procs:
        out, _ = proc.communicate()
        if verbose or proc.returncode:
            sys.stderr.write(out)
        if proc.returncode:
            raise RuntimeError(f"compiling {f.name} failed")
    libdirs = ce.library_paths("cuda")
    cmd = [nvcc, "-shared", *objs, "-o", str(so), *[f"-L{d}" for d in libdirs],
           *[f"-Xlinker=-rpath={d}" for d in libdirs], "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
           "-ltorch_python", "-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode:
        raise RuntimeError("linking alt_cuda_corr.so failed")
    for o in objs:
        Path(o).unlink()
    return so


def load():
    """Import oracle/_ref/alt_cuda_corr.so as a module (torch must be importable); None if it was never built."""
    import importlib.util
    so = OUT / "alt_cuda_corr.so"
    if not so.exists():
        return None
    import torch  # noqa: F401  (its libraries must be loaded first)
    spec = importlib.util.spec_from_file_location("alt_cuda_corr", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(verbose="--verbose" in sys.argv))
