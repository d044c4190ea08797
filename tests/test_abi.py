"""The C-ABI library: it loads without a GPU, exports exactly what include/woft_b200.h declares, and the host
side refuses to compute anywhere but on CUDA (no fallback)."""
import re
import subprocess
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    from woft_b200 import build, _lib
    build.build()
    return _lib.load()


def header_functions():
    text = (ROOT / "include" / "woft_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(woft_[a-z0-9_]+)\s*\(", text))


def test_header_and_binding_agree(lib):
    from woft_b200 import _lib
    assert header_functions() == set(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol(lib):
    from woft_b200 import _lib
    out = subprocess.run(["nm", "-D", "--defined-only", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (woft_[a-z0-9_]+)", out))
    assert header_functions() <= exported


def test_version_and_sizes_without_gpu(lib):
    assert lib.woft_abi_version() == 1
    simt = lib.woft_corr_workspace_bytes(1, 256, 90, 160, 0)
    tf32 = lib.woft_corr_workspace_bytes(1, 256, 90, 160, 1)
    x3 = lib.woft_corr_workspace_bytes(1, 256, 90, 160, 2)
    bf16 = lib.woft_corr_workspace_bytes(1, 256, 90, 160, 3)
    assert lib.woft_corr_workspace_bytes(1, 256, 90, 160, 4) == bf16
    packed = 14400 * 256 * 4
    assert packed * 2 <= lib.woft_corr_workspace_bytes(1, 256, 90, 160, 5) < packed * 2 + 8192
    assert 0 < simt < 8192
    assert 2 * packed <= tf32 < 2 * packed + 8192
    assert 4 * packed <= x3 < 4 * packed + 8192
    assert packed <= bf16 < packed + 8192
    assert lib.woft_corr_workspace_bytes(0, 256, 90, 160, 1) == 0


def test_argument_errors_are_reported_not_thrown(lib):
    from woft_b200 import _lib
    rc = lib.woft_homography_fit(None, None, None, 1, 500, 0, 0, 1.0, None, None)
    assert rc == _lib.ERR_INVALID_ARGUMENT and "null" in _lib.last_error()
    rc = lib.woft_corr_lookup(None, 4, None, 1, 16, 16, 16, 16, 9, None, 0, None)
    assert rc == _lib.ERR_INVALID_ARGUMENT


def test_no_cpu_fallback():
    from woft_b200 import corr, least_squares_H as lsq
    f = torch.randn(1, 8, 16, 16)
    with pytest.raises(RuntimeError, match="CUDA"):
        corr.CorrBlock(f, f)
    p = torch.rand(1, 10, 2)
    with pytest.raises(AssertionError, match="GPU"):  # the reference's own message (least_squares_H.py:292-293)
        lsq.find_homography_nonhomogeneous_QR(p, p)
    with pytest.raises(AssertionError):
        lsq.find_homography_nonhomogeneous_QR(torch.rand(1, 3, 2), torch.rand(1, 3, 2))


def test_product_never_imports_oracle():
    for py in (ROOT / "woft_b200").rglob("*.py"):
        src = py.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), py
        assert "/root/reference" not in src, py


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from woft_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "libwoft_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()
