"""woft_corr_pyramid_static: the correlation pyramid entry point that brings its own scratch memory.

It exists for callers that hold nothing but the input and output buffers they were given and cannot (or must
not, inside a frame loop) allocate device memory for a workspace: the library keeps a fixed 32 MiB scratch array
on the device instead.  The test drives it through the C ABI exactly as such a caller would (raw pointers via
ctypes) and checks it against the ordinary entry point, woft_corr_pyramid, which gets its workspace from the
caller.  Both run the same kernels, so the results must agree to the tolerance the parity tests use for that
arithmetic mode; two calls in a row on different inputs show that nothing stale survives in the shared scratch.

This file is named to be collected last, so that a `-x` run has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL = {"fp32": 1e-4, "tf32": 6e-3, "bf16": 5e-2}  # same per-mode tolerances as tests/test_gpu_parity.py


def _levels(B, H, W, n):
    return [torch.full((B * H * W, 1, H >> l, W >> l), float("nan"), device=DEV) for l in range(n)]


def _run(lib, _lib, static, f1, f2, mode_id, n_levels=4):
    B, C, H, W = f1.shape
    levels = _levels(B, H, W, n_levels)
    row_mean = torch.full((B, H, W), float("nan"), device=DEV)
    arr, keep = _lib.ptr_array(levels)
    stream = _lib.current_stream(torch.device(DEV))
    if static:
        rc = lib.woft_corr_pyramid_static(_lib.ptr(f1), _lib.ptr(f2), B, C, H, W, n_levels, arr, _lib.ptr(row_mean),
                                          mode_id, stream)
    else:
        nbytes = lib.woft_corr_workspace_bytes(B, C, H, W, mode_id)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
        rc = lib.woft_corr_pyramid(_lib.ptr(f1), _lib.ptr(f2), B, C, H, W, n_levels, arr, _lib.ptr(row_mean),
                                   _lib.ptr(ws), nbytes, mode_id, stream)
    del keep
    if static and rc == _lib.OK:
        # a caller of the static variant has no scratch pointer to give woft_corr_check; this is its counterpart
        assert lib.woft_corr_check_static(stream) == _lib.OK, _lib.last_error()
    torch.cuda.synchronize()
    return rc, levels, row_mean


@pytest.mark.parametrize("mode", ["tf32", "bf16", "fp32"])
def test_static_scratch_entry_point_matches_workspace_entry_point(mode):
    from woft_b200 import _lib
    lib = _lib.load()
    mode_id = _lib.CORR_MODES[mode]
    g = torch.Generator().manual_seed(11)
    shapes = [(1, 256, 48, 64), (2, 64, 45, 80)]  # two calls in a row, different shapes and data
    for shape in shapes:
        f1 = torch.randn(shape, generator=g).to(DEV)
        f2 = torch.randn(shape, generator=g).to(DEV)
        rc_s, lv_s, rm_s = _run(lib, _lib, True, f1, f2, mode_id)
        assert rc_s == _lib.OK, _lib.last_error()
        rc_w, lv_w, rm_w = _run(lib, _lib, False, f1, f2, mode_id)
        assert rc_w == _lib.OK, _lib.last_error()
        for a, b in zip(lv_s, lv_w):
            assert not torch.isnan(a).any()
            assert (a - b).abs().max().item() <= ATOL[mode]
        assert (rm_s - rm_w).abs().max().item() <= 1e-4
        if shape == shapes[0]:
            # and against the CPU oracle (reference arithmetic, raft_core/corr.py:13-27, 61-69), not only the sibling
            from oracle import woft_oracle as wo
            ref = wo.OracleCorrBlock(f1.cpu(), f2.cpu())
            for a, want in zip(lv_s, ref.corr_pyramid):
                assert (a.cpu() - want).abs().max().item() <= ATOL[mode]


def test_static_scratch_entry_point_at_the_frame_size_it_is_sized_for():
    """One 1280x720 frame pair (90x160 feature maps, 256 channels) in fp16: the largest problem the header promises
    to fit.  Checked against the caller-workspace entry point everywhere and against fp64 dot products on a sample."""
    from woft_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(5)
    f1 = torch.randn((1, 256, 90, 160), generator=g).to(DEV)
    f2 = torch.randn((1, 256, 90, 160), generator=g).to(DEV)
    for mode in ("fp16", "fp16x3"):
        mode_id = _lib.CORR_MODES[mode]
        assert lib.woft_corr_workspace_bytes(1, 256, 90, 160, mode_id) <= (32 << 20)
        rc_s, lv_s, rm_s = _run(lib, _lib, True, f1, f2, mode_id)
        assert rc_s == _lib.OK, _lib.last_error()
        rc_w, lv_w, rm_w = _run(lib, _lib, False, f1, f2, mode_id)
        assert rc_w == _lib.OK, _lib.last_error()
        for a, b in zip(lv_s, lv_w):
            assert torch.equal(a, b)   # same kernels, same inputs, deterministic summation order
        q = torch.randint(0, 14400, (256,), generator=g)
        t = torch.randint(0, 14400, (256,), generator=g)
        a64 = f1.double().view(256, 14400)[:, q.to(DEV)]
        b64 = f2.double().view(256, 14400)[:, t.to(DEV)]
        want = (a64 * b64).sum(0) / 16.0
        got = lv_s[0].view(14400, 14400)[q.to(DEV), t.to(DEV)].double()
        assert (got - want).abs().max().item() <= (6e-3 if mode == "fp16" else 1e-4)
        del lv_s, lv_w


def test_static_scratch_entry_point_refuses_oversized_problems():
    from woft_b200 import _lib
    lib = _lib.load()
    # 3xTF32 at 1280x720 needs 59 MB of scratch, more than the built-in 32 MiB: refused before any GPU work
    B, C, H, W = 1, 256, 90, 160
    assert lib.woft_corr_workspace_bytes(B, C, H, W, _lib.CORR_TF32X3) > (32 << 20)
    dummy = torch.zeros(16, device=DEV)
    arr, keep = _lib.ptr_array([dummy] * 4)
    rc = lib.woft_corr_pyramid_static(_lib.ptr(dummy), _lib.ptr(dummy), B, C, H, W, 4, arr, None,
                                      _lib.CORR_TF32X3, ctypes.c_void_p(0))
    assert rc == _lib.ERR_WORKSPACE
    assert "32" in _lib.last_error() or "scratch" in _lib.last_error()
    # ... while the single-pass modes at that size fit
    assert lib.woft_corr_workspace_bytes(B, C, H, W, _lib.CORR_TF32) <= (32 << 20)
