"""Parity of the CUDA path (through the C ABI) with the CPU oracle and the reference-generated golden vectors.

Tolerances (absolute, on correlation values of unit scale; stated per arithmetic mode of the GEMM):
  fp32 / tf32x3 / fp16x3   1e-4   same precision class as the reference's fp32 matmul (summation order differs)
  tf32, fp16      6e-3   operands rounded once to 10-bit mantissa, 256-term dot product / 16
  bf16            5e-2   operands rounded to 7-bit mantissa
Flow: mean end-point error <= 0.01 px against the reference network output in the fp32-class modes (north star),
<= 0.05 px stated for tf32 / bf16.  Homography: corner error <= 0.05 px on a 1280x720 frame.
"""
import functools

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import woft_oracle as wo

pytestmark = pytest.mark.gpu

# "fp16s": fp16 operands AND fp16 storage of the pyramid -- the stored value adds a relative rounding of 2^-11 (|corr| of the
# fixtures stays below 8, so 4e-3 absolute) to the 6e-3 of the fp16 operands
CORR_ATOL = {"fp32": 1e-4, "tf32x3": 1e-4, "fp16x3": 1e-4, "tf32": 6e-3, "fp16": 6e-3, "fp16s": 1e-2, "bf16": 5e-2}
DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _exact_convs():
    # keep cuDNN/cuBLAS out of TF32 so the fp32 tolerance is not spent on library noise
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


# ----------------------------------------------------------------------------------------------------
# K3 lookup
# ----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("flags", [0, 1, 8, 9])
def test_lookup_on_golden_pyramid(golden, flags):
    """K3 alone: feed the reference's own pyramid, compare with the reference's lookup."""
    from woft_b200.corr import CorrBlock
    g = golden("corr_small")
    blk = CorrBlock.__new__(CorrBlock)
    blk.num_levels, blk.radius, blk.shape = 4, 4, (1, 16, 16, 24)
    blk.corr_pyramid = [cu(g[f"level{l}"]) for l in range(4)]
    for tag in ("grid", "jitter", "oob"):
        out = blk(cu(g[f"coords_{tag}"]), flags=flags)
        assert out.shape == (1, 324, 16, 24)
        assert out.is_contiguous(memory_format=torch.channels_last if flags & 8 else torch.contiguous_format)
        np.testing.assert_allclose(out.cpu().numpy(), g[f"lookup_{tag}"], atol=2e-5, err_msg=tag)


@pytest.mark.parametrize("name", ["corr_small", "corr_odd", "corr_c256"])
@pytest.mark.parametrize("mode", ["fp32", "tf32x3", "fp16x3", "tf32", "fp16", "fp16s", "bf16"])
def test_corr_block_matches_golden(golden, name, mode):
    from woft_b200.corr import CorrBlock
    g = golden(name)
    blk = CorrBlock(cu(g["fmap1"]), cu(g["fmap2"]), num_levels=4, radius=4, mode=mode)
    blk.check()
    atol = CORR_ATOL[mode]
    for l, lvl in enumerate(blk.corr_pyramid):
        assert tuple(lvl.shape) == tuple(g[f"level{l}_shape"])
        if f"level{l}" in g:
            np.testing.assert_allclose(lvl.float().cpu().numpy(), g[f"level{l}"], atol=atol, err_msg=f"level {l}")
        assert lvl.double().sum().item() == pytest.approx(float(g[f"level{l}_sum"]), abs=atol * lvl.numel() ** 0.5 * 4)
    np.testing.assert_allclose(blk.mean_response().cpu().numpy(), g["mean_response"], atol=1e-4)
    for tag in ("grid", "jitter", "oob"):
        out = blk(cu(g[f"coords_{tag}"]))
        assert not torch.isnan(out).any()
        np.testing.assert_allclose(out.cpu().numpy(), g[f"lookup_{tag}"], atol=atol, err_msg=tag)


@pytest.mark.parametrize("shape", [(1, 256, 48, 64), (2, 64, 45, 80), (1, 40, 20, 36), (1, 256, 33, 70)])
@pytest.mark.parametrize("mode", ["fp32", "tf32x3", "fp16x3", "tf32", "fp16", "fp16s", "bf16"])
def test_corr_pyramid_vs_oracle_ragged_shapes(shape, mode):
    """Partial tiles in every direction, level widths that do and do not allow TMA stores, K padding (C=40)."""
    from woft_b200.corr import CorrBlock
    B, C, H, W = shape
    g = torch.Generator().manual_seed(H * 1000 + W)
    f1, f2 = torch.randn(shape, generator=g), torch.randn(shape, generator=g)
    ref = wo.corr_pyramid(f1, f2, 4)
    blk = CorrBlock(f1.to(DEV), f2.to(DEV), mode=mode)
    blk.check()
    for l in range(4):
        assert blk.corr_pyramid[l].shape == ref[l].shape
        np.testing.assert_allclose(blk.corr_pyramid[l].float().cpu().numpy(), ref[l].numpy(), atol=CORR_ATOL[mode],
                                   err_msg=f"{shape} {mode} level {l}")


@pytest.mark.parametrize("mode", ["fp16", "tf32x3", "bf16"])
def test_corr_block_takes_channels_last_feature_maps(mode):
    """Channels-last feature maps (WOFT_CORR_FMAPS_NHWC) give the volume the contiguous ones give, bit for bit (the
    packed operands are the same numbers); the mean response is summed in another order, hence a tolerance."""
    from woft_b200.corr import CorrBlock
    g = torch.Generator().manual_seed(21)
    for (B, C, H, W) in ((1, 256, 24, 40), (2, 40, 17, 23)):
        f1 = torch.randn((B, C, H, W), generator=g).to(DEV)
        f2 = torch.randn((B, C, H, W), generator=g).to(DEV)
        a = CorrBlock(f1, f2, mode=mode)
        b = CorrBlock(f1.contiguous(memory_format=torch.channels_last), f2.contiguous(memory_format=torch.channels_last),
                      mode=mode)
        b.check()
        for la, lb in zip(a.corr_pyramid, b.corr_pyramid):
            assert torch.equal(la, lb)
        assert torch.allclose(a.mean_response(), b.mean_response(), atol=1e-5)
    with pytest.raises(AssertionError):   # the CUDA-core mode reads (B,C,H,W) only; the Python class converts, the ABI refuses
        from woft_b200 import _lib
        lib = _lib.load()
        lv, keep = _lib.ptr_array(a.corr_pyramid)
        ws = torch.empty(1 << 20, dtype=torch.uint8, device=DEV)
        _lib.check(lib.woft_corr_pyramid(_lib.ptr(f1), _lib.ptr(f2), B, C, H, W, 4, lv, None, _lib.ptr(ws), 1 << 20,
                                         _lib.CORR_FP32_SIMT | _lib.CORR_FMAPS_NHWC, _lib.current_stream(torch.device(DEV))),
                   "woft_corr_pyramid")


@pytest.mark.parametrize("shape", [(1, 256, 90, 160), (2, 64, 45, 80), (1, 40, 20, 36), (1, 256, 33, 70)])
def test_fp16_storage_is_the_fp16_mode_rounded_once(shape):
    """WOFT_CORR_STORE_F16: every level is the fp32 value of mode "fp16" rounded once to half precision (the pooled means
    are formed from the unrounded accumulators), and the lookup interpolates those stored values in fp32 -- so it equals,
    bit for bit, the fp32-level lookup fed with the same numbers.  Shapes with and without TMA-storable level rows."""
    from woft_b200.corr import CorrBlock
    B, C, H, W = shape
    g = torch.Generator().manual_seed(H * 31 + W)
    f1, f2 = torch.randn(shape, generator=g).to(DEV), torch.randn(shape, generator=g).to(DEV)
    a = CorrBlock(f1, f2, mode="fp16")
    b = CorrBlock(f1, f2, mode="fp16s")
    b.check()
    for la, lb in zip(a.corr_pyramid, b.corr_pyramid):
        assert lb.dtype == torch.float16 and la.shape == lb.shape
        assert torch.equal(la.half(), lb)
    ys, xs = torch.meshgrid(torch.arange(H, device=DEV), torch.arange(W, device=DEV), indexing="ij")
    coords = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    coords = coords + (torch.rand(coords.shape, generator=g).to(DEV) * 2 - 1) * 9
    a.corr_pyramid = [l.float() for l in b.corr_pyramid]   # the same numbers as fp32 levels
    for flags in (0, 1, 8, 24):   # 24: fp16 output, channels padded to 328 (the padding is never written)
        assert torch.equal(a(coords, flags=flags)[:, :324], b(coords, flags=flags)[:, :324]), flags
    # 16-byte pieces (default where the level width allows) and 8-byte pieces (flag 2) fetch the same values
    assert torch.equal(b(coords, flags=8), b(coords, flags=8 | 2))


@pytest.mark.parametrize("mode", ["fp16", "fp16s"])
def test_corr_block_takes_fp16_feature_maps(mode):
    """WOFT_CORR_FMAPS_F16: channels-last fp16 maps are the GEMM's operands as they are (no packing pass) -- the volume
    is the one the same numbers give as fp32 maps, bit for bit; the mean response is formed from the same values."""
    from woft_b200 import _lib
    from woft_b200.corr import CorrBlock
    g = torch.Generator().manual_seed(23)
    cl = torch.channels_last
    for (B, C, H, W) in ((1, 256, 24, 40), (2, 64, 17, 23)):
        f1 = torch.randn((B, C, H, W), generator=g).to(DEV).half().contiguous(memory_format=cl)
        f2 = torch.randn((B, C, H, W), generator=g).to(DEV).half().contiguous(memory_format=cl)
        a = CorrBlock(f1.float(), f2.float(), mode=mode)
        b = CorrBlock(f1, f2, mode=mode)
        b.check()
        for la, lb in zip(a.corr_pyramid, b.corr_pyramid):
            assert torch.equal(la, lb)
        assert torch.allclose(a.mean_response(), b.mean_response(), atol=1e-6)
    # other modes and channel counts convert instead (the Python class), the ABI refuses
    c = CorrBlock(f1, f2, mode="fp16x3")
    assert torch.allclose(c.corr_pyramid[0], a.corr_pyramid[0].float(), atol=4e-3)   # fp16s: the stored value is rounded
    lib = _lib.load()
    lv, keep = _lib.ptr_array(a.corr_pyramid)
    ws = torch.empty(1 << 22, dtype=torch.uint8, device=DEV)
    with pytest.raises(AssertionError):
        _lib.check(lib.woft_corr_pyramid(_lib.ptr(f1), _lib.ptr(f2), B, C, H, W, 4, lv, None, _lib.ptr(ws), 1 << 22,
                                         _lib.CORR_FP16X3 | _lib.CORR_FMAPS_F16, _lib.current_stream(torch.device(DEV))),
                   "woft_corr_pyramid")


def test_corr_block_rejects_what_the_reference_turns_into_nan():
    from woft_b200.corr import CorrBlock
    f = torch.randn(1, 16, 8, 12, device=DEV)  # coarsest level would be 1x1 -> reference returns NaN
    with pytest.raises(AssertionError, match="too small"):
        CorrBlock(f, f)
    f = torch.randn(1, 16, 16, 16, device=DEV)
    blk = CorrBlock(f, f, mode="fp32")
    with pytest.raises(AssertionError):
        blk(torch.zeros(1, 2, 8, 8, device=DEV))


@pytest.mark.parametrize("mode", ["tf32", "tf32x3", "fp16x3", "fp16", "fp16s", "bf16"])
def test_full_size_properties(mode):
    """BASELINE config 2 size (90x160x256): properties that need no CPU volume."""
    from woft_b200.corr import CorrBlock
    torch.manual_seed(0)
    B, C, H, W = 1, 256, 90, 160
    f1 = torch.randn(B, C, H, W, device=DEV)
    f2 = torch.randn(B, C, H, W, device=DEV)
    blk = CorrBlock(f1, f2, mode=mode)
    blk.check()
    lv = blk.corr_pyramid
    assert [tuple(l.shape[2:]) for l in lv] == [(90, 160), (45, 80), (22, 40), (11, 20)]
    # (a) sampled entries against fp64 dot products
    idx = torch.randint(0, H * W, (2, 20000), device=DEV)
    a = f1.view(C, H * W)[:, idx[0]].double()
    b = f2.view(C, H * W)[:, idx[1]].double()
    want = (a * b).sum(0) / 16.0
    got = lv[0].view(H * W, H * W)[idx[0], idx[1]].double()
    assert (got - want).abs().max().item() < CORR_ATOL[mode]
    # (b) every level is the 2x2 mean of the one below (avg_pool2d semantics, floor on 45 -> 22)
    stored = 2e-6 if mode != "fp16s" else 4e-3   # fp16 storage: each level is rounded on its own
    for l in range(3):
        assert torch.allclose(lv[l + 1].float(), F.avg_pool2d(lv[l].float(), 2, stride=2), atol=stored)
    # (c) the mean response is the mean of the level-0 plane
    assert torch.allclose(blk.mean_response().view(-1), lv[0].float().mean(dim=(1, 2, 3)), atol=1e-4)
    # (d) lookup at integer coordinates returns volume entries; out-of-map taps are exactly zero
    ys, xs = torch.meshgrid(torch.arange(H, device=DEV), torch.arange(W, device=DEV), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None]
    out = blk(grid)
    q = 37 * W + 91
    assert out[0, (2 + 4) * 9 + (-3 + 4), 37, 91].item() == pytest.approx(lv[0][q, 0, 37 - 3, 91 + 2].item(), abs=1e-6)
    assert out[0, 0, 0, 0].item() == 0.0  # tap (-4, -4) of the corner query
    # (e) lookup of a block of queries against grid_sample on the same device pyramid
    coords = grid + (torch.rand_like(grid) * 2 - 1) * 8
    out = blk(coords)
    sel = torch.arange(3000, 3600, device=DEV)
    ref = wo.corr_lookup([l[sel].float().cpu() for l in lv], coords.view(2, -1)[:, sel].cpu().view(1, 2, 1, -1))
    got = out.view(324, -1)[:, sel].cpu().view(1, 324, 1, -1)
    assert (got - ref).abs().max().item() < 2e-5 * max(1.0, ref.abs().max().item())


# ----------------------------------------------------------------------------------------------------
# A'' alt_cuda_corr
# ----------------------------------------------------------------------------------------------------
def test_alt_corr_forward_vs_oracle():
    from woft_b200.corr import AlternateCorrBlock, CorrBlock, alt_cuda_corr
    g = torch.Generator().manual_seed(3)
    B, C, H, W = 2, 64, 16, 24
    f1 = torch.randn((B, H, W, C), generator=g)
    f2 = torch.randn((B, 8, 12, C), generator=g)
    coords = torch.rand((B, 1, H, W, 2), generator=g) * torch.tensor([14.0, 10.0]) - 1.0
    want = wo.alt_corr_forward(f1, f2, coords, 4)
    got, = alt_cuda_corr.forward(f1.to(DEV), f2.to(DEV), coords.to(DEV), 4)
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), atol=2e-4)
    with pytest.raises(RuntimeError):
        alt_cuda_corr.forward(f1, f2.to(DEV), coords.to(DEV), 4)
    # the on-the-fly block agrees with the materialised one (pooling is linear)
    n1 = torch.randn((1, 64, 16, 24), generator=g).to(DEV)
    n2 = torch.randn((1, 64, 16, 24), generator=g).to(DEV)
    c = (torch.rand((1, 2, 16, 24), generator=g) * 12 + 2).to(DEV)
    a = AlternateCorrBlock(n1, n2)(c)
    m = CorrBlock(n1, n2, mode="fp32")(c)
    np.testing.assert_allclose(a.cpu().numpy(), m.cpu().numpy(), atol=2e-4)


def test_alt_corr_forward_vs_reference_extension():
    """Row a8 pinned by the reference itself: `alt_cuda_corr` compiled from the reference's own two source files
    (oracle/build_ref.py -> oracle/_ref/alt_cuda_corr.so, built in the container, shipped with the tree) runs here as
    the oracle of woft_alt_corr_forward and of the pure-torch restatement woft_oracle.alt_corr_forward, on batched,
    multi-coordinate and out-of-bounds inputs (alt_cuda_corr/correlation.cpp:23-33, correlation_kernel.cu:18-119)."""
    from oracle import build_ref
    from woft_b200.corr import alt_cuda_corr
    ref = build_ref.load()
    if ref is None:
        pytest.skip("oracle/_ref/alt_cuda_corr.so was not built (python -m oracle.build_ref in the build container): "
                    "PARITY OF alt_corr_forward IS NOT PINNED BY THE REFERENCE IN THIS RUN")
    g = torch.Generator().manual_seed(7)
    for (B, C, H1, W1, H2, W2, N, r, spread) in ((2, 64, 16, 24, 8, 12, 1, 4, 1.0), (1, 256, 12, 20, 12, 20, 2, 4, 1.0),
                                                 (1, 32, 9, 13, 5, 7, 1, 3, 1.0), (1, 96, 8, 8, 8, 8, 1, 4, 3.0)):
        f1 = torch.randn((B, H1, W1, C), generator=g)
        f2 = torch.randn((B, H2, W2, C), generator=g)
        # `spread` 3: a third of the windows start outside the map
        coords = (torch.rand((B, N, H1, W1, 2), generator=g) * spread - (spread - 1) / 2) * torch.tensor([W2 - 1.0, H2 - 1.0])
        f1d, f2d, cd = f1.to(DEV), f2.to(DEV), coords.to(DEV)
        want, = ref.forward(f1d, f2d, cd, r)
        torch.cuda.synchronize()
        got, = alt_cuda_corr.forward(f1d, f2d, cd, r)
        assert got.shape == want.shape
        scale = C ** 0.5
        np.testing.assert_allclose(got.cpu().numpy() / scale, want.cpu().numpy() / scale, atol=2e-5)
        np.testing.assert_allclose(wo.alt_corr_forward(f1, f2, coords, r).numpy() / scale, want.cpu().numpy() / scale,
                                   atol=2e-5)


# ----------------------------------------------------------------------------------------------------
# K4 homography
# ----------------------------------------------------------------------------------------------------
def _h_cases(g):
    for tag in ("exact4", "plain100", "tracker500", "batch3", "ragged37"):
        p1, p2 = g[tag + "_p1"], g[tag + "_p2"]
        w = g[tag + "_w"] if tag + "_w" in g else None
        yield tag, p1, p2, w


def _corner_err(Ha, Hb):
    return wo.corner_error(torch.as_tensor(Ha), torch.as_tensor(Hb), 1280, 720)


def test_homography_matches_reference(golden):
    from woft_b200 import least_squares_H as lsq
    g = golden("homography")
    huber2 = functools.partial(lsq.IRLSq_Huber, k=2)
    fns = {"qr": (lsq.find_homography_nonhomogeneous_QR, "qr_f64"),
           "pinv": (lsq.find_homography_nonhomogeneous, "qr_f64"),
           "dlt": (lsq.find_homography_dlt, "dlt_f64"),
           "irls_l1": (lsq.find_homography_IRLSq_QR, "irls_l1_f64"),
           "irls_huber2": (functools.partial(lsq.find_homography_IRLSq_QR, reweighting_fn=huber2), None)}
    for tag, p1, p2, w in _h_cases(g):
        wd = cu(w) if w is not None else None
        for key, (fn, f64key) in fns.items():
            H = fn(cu(p1), cu(p2), wd)
            assert H.shape == (p1.shape[0], 3, 3) and H.dtype == torch.float32
            H = H.cpu()
            assert torch.allclose(H[:, 2, 2], torch.ones(p1.shape[0]), atol=1e-6)
            # against the reference's fp64 run (truth), and against its fp32 run within that run's own error
            if f64key is not None:
                e64 = _corner_err(H, g[f"{tag}_{f64key}"])
                ref_own = _corner_err(g[f"{tag}_{key}"], g[f"{tag}_{f64key}"])
                assert e64 <= 0.05, f"{tag}/{key}: {e64} px from the fp64 reference"
                assert _corner_err(H, g[f"{tag}_{key}"]) <= 0.05 + ref_own, f"{tag}/{key}"
            else:
                assert _corner_err(H, g[f"{tag}_{key}"]) <= 0.05 or tag == "exact4", f"{tag}/{key}"
        Hq = cu(g[f"{tag}_qr"])
        err = lsq.torch_proj_errors(Hq, cu(p1).permute(0, 2, 1).contiguous(), cu(p2).permute(0, 2, 1).contiguous())
        np.testing.assert_allclose(err.cpu().numpy(), g[f"{tag}_proj_err"], rtol=1e-4, atol=1e-3)


def test_homography_error_behaviour():
    from woft_b200 import least_squares_H as lsq
    p = torch.rand(1, 3, 2, device=DEV)
    with pytest.raises(AssertionError):
        lsq.find_homography_nonhomogeneous_QR(p, p)
    with pytest.raises(AssertionError):
        lsq.find_homography_dlt(torch.rand(1, 5, 2, device=DEV), torch.rand(1, 6, 2, device=DEV))
    with pytest.raises(AssertionError):
        lsq.find_homography_nonhomogeneous_QR(torch.rand(1, 5, 2, device=DEV), torch.rand(1, 5, 2, device=DEV),
                                              torch.rand(1, 4, device=DEV))


def test_homography_stress_config3_properties():
    """BASELINE config 3: B=1024 flow fields of 160x90 correspondences with weights; properties + a CPU subset."""
    from oracle.gen_golden import make_h_case
    from woft_b200 import least_squares_H as lsq
    B, N = 1024, 14400
    p1, p2, w, Ht = make_h_case(B=B, N=N, seed=0, noise=0.5, outlier_frac=0.3, with_weights=True)
    d1, d2, dw = p1.to(DEV), p2.to(DEV), w.to(DEV)
    H = lsq.find_homography_nonhomogeneous_QR(d1, d2, dw)
    assert torch.isfinite(H).all()
    ref = wo.find_homography_nonhomogeneous_QR(p1[:6].double(), p2[:6].double(), w[:6].double())
    assert _corner_err(H[:6].cpu(), ref) <= 0.05
    # order of the correspondences does not matter
    perm = torch.randperm(N, device=DEV)
    Hp = lsq.find_homography_nonhomogeneous_QR(d1[:32, perm].contiguous(), d2[:32, perm].contiguous(),
                                               dw[:32, perm].contiguous())
    assert _corner_err(H[:32].cpu(), Hp.cpu()) <= 1e-3
    # a common scale on the weights does not matter; Huber(k=2) reweighting never triggers on such data
    Hs = lsq.find_homography_nonhomogeneous_QR(d1[:32], d2[:32], dw[:32] * 3.0)
    assert _corner_err(H[:32].cpu(), Hs.cpu()) <= 1e-3
    Hh = lsq.find_homography_IRLSq_QR(d1[:32], d2[:32], dw[:32],
                                      reweighting_fn=functools.partial(lsq.IRLSq_Huber, k=2))
    assert _corner_err(H[:32].cpu(), Hh.cpu()) <= 1e-3
    # IRLS-L1 and DLT against the CPU oracle in fp64 on a few problems
    Hl = lsq.find_homography_IRLSq_QR(d1[:4], d2[:4], dw[:4])
    refl = wo.find_homography_IRLSq_QR(p1[:4].double(), p2[:4].double(), w[:4].double())
    assert _corner_err(Hl.cpu(), refl) <= 0.05
    Hd = lsq.find_homography_dlt(d1[:4], d2[:4], dw[:4])
    refd = wo.find_homography_dlt(p1[:4].double(), p2[:4].double(), w[:4].double())
    assert _corner_err(Hd.cpu(), refd) <= 0.05


# ----------------------------------------------------------------------------------------------------
# network + flow wrapper + tracker
# ----------------------------------------------------------------------------------------------------
def _net_inputs(g):
    t1 = torch.from_numpy(g["img1_bgr"][:, :, ::-1].copy()).permute(2, 0, 1).float()[None].to(DEV)
    t2 = torch.from_numpy(g["img2_bgr"][:, :, ::-1].copy()).permute(2, 0, 1).float()[None].to(DEV)
    return t1, t2


@pytest.mark.parametrize("mode,epe_tol", [("fp32", 0.01), ("tf32x3", 0.01), ("fp16x3", 0.01), ("tf32", 0.05), ("fp16", 0.05), ("bf16", 0.05)])
def test_flow_pass_matches_reference_network(golden, mode, epe_tol):
    from woft_b200.corr import CorrBlock
    from woft_b200.raft_net import WeightedRAFT, seeded_state_dict
    g = golden("net_small")
    net = WeightedRAFT().eval()
    net.load_state_dict(seeded_state_dict(net, 0))
    net = net.to(DEV)
    t1, t2 = _net_inputs(g)
    with torch.no_grad():
        flow_low, flow_up, vol, w_low, w_up = net(t1, t2, iters=12, corr_block=functools.partial(CorrBlock, mode=mode))
    epe = (flow_up.cpu() - torch.from_numpy(g["it12_flow_up"])).norm(dim=1).mean().item()
    assert epe <= epe_tol, f"{mode}: mean EPE {epe} px"
    assert (w_up.cpu() - torch.from_numpy(g["it12_w_up"])).abs().max().item() <= (2e-3 if epe_tol == 0.01 else 2e-2)
    assert vol.shape == (320, 1, 16, 20)


@pytest.mark.parametrize("graph", [False, True])
def test_flow_wrapper_matches_reference_wrapper(golden, graph):
    from types import SimpleNamespace
    from woft_b200.optical_flow import RAFTWrapper
    from woft_b200.corr import CorrBlock
    g, gw = golden("net_small"), golden("wrapper_small")
    conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False, alternate_corr=False),
                           model=None, seeded_weights=0, iters=12, padding_mode="nopad")
    wrapper = RAFTWrapper(conf, device=DEV, use_cuda_graph=graph, corr_block=functools.partial(CorrBlock, mode="tf32x3"))
    for _ in range(2):  # second call replays the graph
        src, dst, w = wrapper.compute_flow(g["img1_bgr"], g["img2_bgr"], mode="TC", do_sigmoid=True)
    assert src.dtype == torch.int64 and src.shape == (2, 128 * 160) and w.shape == (1, 128 * 160)
    np.testing.assert_array_equal(src.cpu().numpy(), gw["tc_src"])
    assert (dst.cpu() - torch.from_numpy(gw["tc_dst"])).norm(dim=0).mean().item() <= 0.01
    np.testing.assert_allclose(w.cpu().numpy(), gw["tc_w"], atol=1e-3)
    assert wrapper.last_flow_shape == {"batch": 1, "delta": 2, "H": 128, "W": 160}
    flow, w2 = wrapper.compute_flow(g["img1_bgr"], g["img2_bgr"], mode="flow", numpy_out=True)
    assert flow.shape == (2, 128, 160) and w2.shape == (1, 128, 160)
    assert np.abs(w2 - gw["flow_w"]).max() <= 5e-3


def test_flow_wrapper_mixed_precision(golden):
    """class_params.mixed_precision=True (optical_flow/raft.py:35): fp16 autocast where the reference has it.  No shipped
    config sets it and the CPU oracle cannot autocast to fp16, so the check is against the fp32 golden flow with a
    tolerance stated for half precision: mean EPE 0.1 px, weights 0.05."""
    from types import SimpleNamespace
    from woft_b200.optical_flow import RAFTWrapper
    g, gw = golden("net_small"), golden("wrapper_small")
    conf = SimpleNamespace(raft_type="weighted", model=None, seeded_weights=0, iters=12, padding_mode="nopad",
                           class_params=SimpleNamespace(small=False, alternate_corr=False, mixed_precision=True))
    wrapper = RAFTWrapper(conf, device=DEV)
    assert wrapper.mixed_precision and not wrapper.fused
    for _ in range(2):
        src, dst, w = wrapper.compute_flow(g["img1_bgr"], g["img2_bgr"], mode="TC", do_sigmoid=True)
    assert dst.dtype == torch.float32 and w.dtype == torch.float32
    assert (dst.cpu() - torch.from_numpy(gw["tc_dst"])).norm(dim=0).mean().item() <= 0.1
    assert np.abs(w.cpu().numpy() - gw["tc_w"]).max() <= 0.05


def test_tracker_runs_and_masks_like_the_reference():
    from types import SimpleNamespace
    import cv2
    from oracle.gen_golden import make_frames
    from woft_b200 import tracker
    from woft_b200.optical_flow import RAFTWrapper
    H, W = 256, 320
    flow_conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False), model=None,
                                seeded_weights=0, iters=12, padding_mode="nopad")
    trk = tracker.PlanarTracker(tracker.default_config(flow_conf), flower=RAFTWrapper(flow_conf, device=DEV))
    img0, img1 = make_frames(H, W, 3)
    mask = np.zeros((H, W), np.uint8)
    mask[64:192, 80:240] = 255
    trk.init(img0, mask)
    for frame in (img1, img0, img1):
        Hc, meta = trk.track(frame)
        assert Hc.shape == (3, 3) and np.all(np.isfinite(Hc)) and Hc[2, 2] == pytest.approx(1.0)
    assert 3 <= trk.flow_passes <= 6
    # the device-side correspondence mask against the oracle's restatement of _mask_coords
    g = torch.Generator().manual_seed(1)
    n = H * W
    from woft_b200.misc import torch_get_featuremap_coords
    src = torch_get_featuremap_coords((H, W), device="cpu")
    cur = src.float() + torch.randn((2, n), generator=g) * 30
    wts = torch.rand((1, n), generator=g)
    pw = torch.from_numpy(cv2.warpPerspective(np.ones((H, W)), np.array([[1, 0.02, 8], [0, 1, -5], [0, 0, 1.0]]), (W, H)) > 0)
    want = wo.mask_coords(torch.from_numpy(mask > 0), src, cur, wts, pw)
    got = trk._mask_coords(src.to(DEV), cur.to(DEV), wts.to(DEV), pw.to(DEV))
    for a, b in zip(got, want):
        assert torch.equal(a.cpu(), b)


# ----------------------------------------------------------------------------------------------------
# fused element-wise kernels and the fused inference schedule (started)
# ----------------------------------------------------------------------------------------------------
def test_fused_gru_and_instance_norm_kernels():
    from woft_b200 import fast_net
    g = torch.Generator().manual_seed(5)
    B, C, H, W = 2, 128, 22, 36
    zr = torch.randn((B, 2 * C, H, W), generator=g).to(DEV)
    h = torch.randn((B, C, H, W), generator=g).to(DEV)
    bias = torch.randn((2 * C,), generator=g).to(DEV)
    z, rh = fast_net.gru_gates(zr, bias, h)
    zb = zr + bias.view(1, -1, 1, 1)
    assert torch.allclose(z, torch.sigmoid(zb[:, :C]), atol=1e-6)
    assert torch.allclose(rh, torch.sigmoid(zb[:, C:]) * h, atol=1e-6)
    q = torch.randn((B, C, H, W), generator=g).to(DEV)
    bq = torch.randn((C,), generator=g).to(DEV)
    hn = fast_net.gru_update(q, bq, z, h)
    assert torch.allclose(hn, (1 - z) * h + z * torch.tanh(q + bq.view(1, -1, 1, 1)), atol=2e-6)
    for C in (16, 96):
        x = (torch.randn((3, C, 45, 80), generator=g) * 3 + 7).to(DEV)
        for relu in (True, False):
            want = F.instance_norm(x, eps=1e-5)
            want = F.relu(want) if relu else want
            assert torch.allclose(fast_net.instance_norm_relu(x, relu=relu), want, atol=2e-5)
            got = fast_net.instance_norm_relu_nhwc(x.contiguous(memory_format=torch.channels_last), relu=relu)
            assert got.is_contiguous(memory_format=torch.channels_last) and torch.allclose(got, want, atol=2e-5)
        # tail of a residual block in one pass: relu(res + relu(norm(x)))  (extractor.py:45-56)
        res = torch.randn((3, C, 45, 80), generator=g).to(DEV).contiguous(memory_format=torch.channels_last)
        got = fast_net.instance_norm_relu_nhwc(x.contiguous(memory_format=torch.channels_last), residual=res)
        assert torch.allclose(got, F.relu(res + F.relu(F.instance_norm(x, eps=1e-5))), atol=2e-5)
        for a in (x, x.contiguous(memory_format=torch.channels_last)):
            b = torch.randn_like(a)
            got = fast_net.add_relu(a, b)
            assert got.stride() == a.stride() and torch.equal(got, F.relu(a + b))
    with pytest.raises(AssertionError):
        fast_net.add_relu(x, x.contiguous(memory_format=torch.channels_last))


def test_fused_schedule_matches_layer_by_layer(golden):
    from woft_b200.corr import CorrBlock
    from woft_b200.fast_net import FastWeightedRAFT
    from woft_b200.raft_net import WeightedRAFT, seeded_state_dict
    g = golden("net_small")
    net = WeightedRAFT().eval()
    net.load_state_dict(seeded_state_dict(net, 0))
    net = net.to(DEV).requires_grad_(False)
    t1, t2 = _net_inputs(g)
    blk = functools.partial(CorrBlock, mode="tf32x3")
    with torch.no_grad():
        slow = net(t1, t2, iters=12, corr_block=blk)
    fast = FastWeightedRAFT(net)(t1, t2, iters=12, corr_block=blk)
    epe = (fast[1] - slow[1]).norm(dim=1).mean().item()
    assert epe <= 2e-3, f"fused vs layer-by-layer schedule: mean EPE {epe} px"
    assert (fast[4] - slow[4]).abs().max().item() <= 2e-3
    ref_epe = (fast[1].cpu() - torch.from_numpy(g["it12_flow_up"])).norm(dim=1).mean().item()
    assert ref_epe <= 0.01, f"fused schedule vs reference network: mean EPE {ref_epe} px"
    # the channel-first variant of the schedule (kept for callers whose correlation operator is not this library's)
    nchw = FastWeightedRAFT(net, nhwc_update=False)(t1, t2, iters=12, corr_block=blk)
    assert (nchw[1] - slow[1]).norm(dim=1).mean().item() <= 2e-3 and (nchw[4] - slow[4]).abs().max().item() <= 2e-3


def test_tensor_core_convolution_schedule_matches_layer_by_layer(golden):
    """The schedule bench.py times -- update block and weight head on woft_conv_tc (fp16 operands, fp32 accumulation, GRU
    arithmetic in the epilogues), encoders on cuDNN with TF32 allowed -- against the layer-by-layer fp32 network and the
    reference's golden output.  Tolerance stated for this arithmetic class: mean EPE 0.02 px, weight logits 0.02."""
    from woft_b200 import conv_tc
    from woft_b200.corr import CorrBlock
    from woft_b200.fast_net import FastWeightedRAFT
    from woft_b200.raft_net import WeightedRAFT, seeded_state_dict
    g = golden("net_small")
    net = WeightedRAFT().eval()
    net.load_state_dict(seeded_state_dict(net, 0))
    net = net.to(DEV).requires_grad_(False)
    t1, t2 = _net_inputs(g)
    blk = functools.partial(CorrBlock, mode="fp16")
    with torch.no_grad():
        slow = net(t1, t2, iters=12, corr_block=functools.partial(CorrBlock, mode="tf32x3"))
    fast_net_ = FastWeightedRAFT(net, tf32=True)
    assert fast_net_.own_convs and fast_net_.own_head
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    try:
        fast = fast_net_(t1, t2, iters=12, corr_block=blk)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    conv_tc.check(DEV)
    epe = (fast[1] - slow[1]).norm(dim=1).mean().item()
    werr = (fast[4] - slow[4]).abs().max().item()
    ref_epe = (fast[1].cpu() - torch.from_numpy(g["it12_flow_up"])).norm(dim=1).mean().item()
    print(f"\ntensor-core convolution schedule at 128x160: mean EPE {epe:.5f} px vs layer-by-layer, {ref_epe:.5f} px vs the "
          f"reference golden, weight-logit max error {werr:.5f}")
    assert epe <= 0.02 and ref_epe <= 0.02, f"mean EPE {epe} / {ref_epe} px"
    assert werr <= 0.02
    # a batch of two frame pairs (four images through the feature encoder: the three-pass instance norm path): each item
    # must reproduce the single-pair result
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    try:
        both = fast_net_(torch.cat([t1, t2]), torch.cat([t2, t1]), iters=12, corr_block=blk)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    conv_tc.check(DEV)
    assert both[1].shape[0] == 2
    assert (both[1][:1] - fast[1]).norm(dim=1).mean().item() <= 5e-3
    assert (both[4][:1] - fast[4]).abs().max().item() <= 5e-3
    # the flow head's gather inside the next lookup (woft_corr_lookup_gather; an option, measured slower) and as its own
    # launch after every iteration are the same arithmetic: every output bit for bit; likewise one stream instead of two
    from woft_b200 import fast_net as fn_mod
    saved = (fn_mod._FUSE_GATHER, fn_mod._TWO_STREAMS)
    torch.backends.cudnn.allow_tf32 = True
    try:
        fn_mod._FUSE_GATHER, fn_mod._TWO_STREAMS = True, False
        plain = fast_net_(t1, t2, iters=12, corr_block=blk)
    finally:
        fn_mod._FUSE_GATHER, fn_mod._TWO_STREAMS = saved
        torch.backends.cudnn.allow_tf32 = old
    for a, b in zip(fast, plain):
        assert torch.equal(a, b)


def test_lookup_with_the_flow_head_gather_in_the_same_launch():
    """woft_corr_lookup_gather against woft_flow_head_gather followed by woft_corr_lookup: coordinates, every flow copy
    and the lookup bit for bit, for both pyramid storage types and a ragged map."""
    from woft_b200 import _lib
    from woft_b200.corr import CorrBlock
    g = torch.Generator().manual_seed(77)
    for (B, H, W), mode in (((1, 24, 40), "fp16"), ((2, 17, 23), "fp16s")):
        f1, f2 = torch.randn((B, 64, H, W), generator=g).to(DEV), torch.randn((B, 64, H, W), generator=g).to(DEV)
        blk = CorrBlock(f1, f2, mode=mode)
        ys, xs = torch.meshgrid(torch.arange(H, device=DEV), torch.arange(W, device=DEV), indexing="ij")
        coords0 = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1).contiguous()
        start = coords0 + torch.randn((B, 2, H, W), generator=g).to(DEV) * 3
        taps = (torch.randn((B * H * W, 18), generator=g) * 0.3).to(DEV)
        bias = torch.tensor([0.05, -0.02], device=DEV)
        stride, off = 392, 382
        outs = []
        for fused in (False, True):
            coords = start.clone()
            flow, flow_cl = torch.zeros_like(coords), torch.zeros((B, H, W, 2), device=DEV)
            ha = torch.zeros((B * H * W, stride), device=DEV, dtype=torch.float16)
            hb = torch.zeros_like(ha)
            gd = dict(taps=taps, bias=bias, coords0=coords0, flow_nchw=flow, flow_nhwc=flow_cl, f16_a=ha, f16_b=hb,
                      f16_stride=stride, f16_offset=off)
            if fused:
                out = blk(coords, flags=24, gather=gd)
            else:
                rc = _lib.load().woft_flow_head_gather(_lib.ptr(taps), _lib.ptr(bias), B, H, W, None, _lib.ptr(coords),
                                                       _lib.ptr(coords0), _lib.ptr(flow), _lib.ptr(flow_cl), _lib.ptr(ha),
                                                       _lib.ptr(hb), stride, off, _lib.current_stream(torch.device(DEV)))
                assert rc == 0
                out = blk(coords, flags=24)
            outs.append((coords, flow, flow_cl, ha, hb, out[:, :324]))
        for a, b in zip(*outs):
            assert torch.equal(a, b)
        assert not torch.equal(outs[0][0], start)


def test_u8_input_path_is_the_float_path(golden):
    """woft_images_to_input against the torch expressions it replaces (bit for bit), and the fused schedule fed with
    uint8 frames against the same schedule fed the reference's way."""
    from woft_b200 import _lib
    from woft_b200.fast_net import FastWeightedRAFT
    from woft_b200.raft_net import WeightedRAFT, seeded_state_dict
    g = golden("net_small")
    a = torch.from_numpy(g["img1_bgr"]).to(DEV)
    b = torch.from_numpy(g["img2_bgr"]).to(DEV)
    H, W, _ = a.shape
    pair = torch.empty((2, 3, H, W), device=DEV).contiguous(memory_format=torch.channels_last)
    rc = _lib.load().woft_images_to_input(_lib.ptr(a), _lib.ptr(b), H, W, _lib.ptr(pair),
                                          _lib.current_stream(torch.device(DEV)))
    assert rc == 0
    f = [t.flip(-1).permute(2, 0, 1).float()[None] for t in (a, b)]
    want = torch.cat([2 * (t / 255.0) - 1.0 for t in f], dim=0)
    assert torch.equal(pair, want)
    net = WeightedRAFT().eval()
    net.load_state_dict(seeded_state_dict(net, 0))
    fast = FastWeightedRAFT(net.to(DEV).requires_grad_(False))
    out_f = fast(f[0], f[1], iters=4)
    out_u = fast.run_u8(a, b, iters=4)
    assert torch.equal(out_f[1], out_u[1]) and torch.equal(out_f[4], out_u[4])


def test_device_prewarp_matches_opencv():
    """woft_warp_perspective_u8 against cv2.warpPerspective (YAOF_tracker_single_control.py:89-95): byte work, so the
    frame and the validity mask must be bit-identical (OpenCV's fixed-point remap restated integer for integer)."""
    import cv2
    from oracle.gen_golden import make_frames
    from woft_b200.tracker import prewarp_on_device
    mats = (np.eye(3), np.array([[1.02, 0.015, 12.5], [-0.01, 0.97, -7.25], [2e-5, -1e-5, 1.0]]),
            np.array([[0.8, 0.2, 40.0], [-0.15, 1.1, 30.0], [1e-4, 2e-4, 1.0]]),
            np.array([[1.3, 0.4, -200.0], [0.3, 0.7, 100.0], [3e-4, -2e-4, 1.0]]))
    for (h, w) in ((360, 640), (720, 1280), (45, 50), (9, 200)):
        img, _ = make_frames(max(h, 16), max(w, 16), 9)
        img = np.ascontiguousarray(img[:h, :w])
        for M in mats:
            want = cv2.warpPerspective(img, M, (w, h), flags=cv2.INTER_LINEAR)
            want_mask = cv2.warpPerspective(np.ones((h, w)), M, (w, h), flags=cv2.INTER_LINEAR) > 0
            got, mask = prewarp_on_device(torch.from_numpy(img).to(DEV), M)
            np.testing.assert_array_equal(got.cpu().numpy(), want, err_msg=f"{h}x{w}")
            np.testing.assert_array_equal(mask.cpu().numpy(), want_mask, err_msg=f"mask {h}x{w}")


@pytest.mark.parametrize("pmode", ["RAFT", "crop"])
def test_flow_wrapper_padders_match_reference_wrapper(golden, pmode):
    """Frame sizes that are not multiples of 8 through the padders (optical_flow/raft.py:122-150, 221-271;
    raft_core/utils/utils.py:7-26), against the reference wrapper's output on the same 133x171 pair."""
    from types import SimpleNamespace
    from woft_b200.optical_flow import RAFTWrapper
    from woft_b200.corr import CorrBlock
    gw = golden("wrapper_small")
    conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False, alternate_corr=False),
                           model=None, seeded_weights=0, iters=12, padding_mode=pmode)
    wrapper = RAFTWrapper(conf, device=DEV, corr_block=functools.partial(CorrBlock, mode="tf32x3"))
    for _ in range(2):  # second call replays the graph
        flow, w = wrapper.compute_flow(gw["odd_img1_bgr"], gw["odd_img2_bgr"], mode="flow", do_sigmoid=False)
    want_flow, want_w = torch.from_numpy(gw[f"pad_{pmode}_flow"]), torch.from_numpy(gw[f"pad_{pmode}_w"])
    assert flow.shape == want_flow.shape and w.shape == want_w.shape
    epe = (flow.cpu() - want_flow).norm(dim=0).mean().item()
    assert epe <= 0.01, f"padding_mode {pmode}: mean EPE {epe} px"
    assert (w.cpu() - want_w).abs().max().item() <= 5e-3


def test_michal_padder_matches_reference_class(golden):
    """MichalPadder (raft.py:250-271) on the device against the reference class (the reference wrapper itself cannot
    run it with raft_type 'weighted': it unpads a mask only 'weighted_masked' produces, raft.py:150)."""
    from woft_b200.optical_flow import MichalPadder
    gw = golden("wrapper_small")
    mp = MichalPadder((1, 3, 125, 163))
    pa, pb = mp.pad(cu(gw["michal_in_a"]), cu(gw["michal_in_b"]))
    np.testing.assert_allclose(pa.cpu().numpy(), gw["michal_pad_a"], atol=2e-4)
    np.testing.assert_allclose(pb.cpu().numpy(), gw["michal_pad_b"], atol=2e-4)
    np.testing.assert_allclose(mp.unpad(cu(gw["michal_flow_in"])).cpu().numpy(), gw["michal_flow_out"], atol=1e-5)


def test_nhwc_gru_and_concat_kernels():
    from woft_b200 import _lib, fast_net
    lib = _lib.load()
    g = torch.Generator().manual_seed(6)
    B, C, H, W = 2, 128, 18, 28
    cl = torch.channels_last
    a = torch.randn((B, 192, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    b = torch.randn((B, 126, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    c = torch.randn((B, 2, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    d1 = torch.zeros((B, 384, H, W), device=DEV).contiguous(memory_format=cl)
    d2 = torch.zeros((B, 400, H, W), device=DEV).contiguous(memory_format=cl)
    fast_net.concat_nhwc([a, b, c], d1, 64, d2, 80)
    want = torch.cat([a, b, c], dim=1)
    assert torch.equal(d1[:, 64:], want) and torch.equal(d2[:, 80:], want)
    assert d1[:, :64].abs().sum().item() == 0
    # channel prefixes of wider sources (the 126 motion features in front of two padding channels, then the flow)
    wide = torch.randn((B, 128, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    d3 = torch.full((B, 384, H, W), -1.0, device=DEV).contiguous(memory_format=cl)
    fast_net.concat_nhwc([wide, c], d3, 256, take=[126, 2])
    assert torch.equal(d3[:, 256:], torch.cat([wide[:, :126], c], dim=1)) and (d3[:, :256] == -1).all()
    # gates / update on channel slices of wider buffers, against the NCHW formulas
    buf1 = torch.randn((B, 3 * C, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    buf2 = torch.zeros_like(buf1)
    zr = torch.randn((B, 2 * C, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    bias = torch.randn((2 * C,), generator=g).to(DEV)
    z = torch.empty((B, C, H, W), device=DEV).contiguous(memory_format=cl)
    h = buf1[:, :C].clone()
    st = _lib.current_stream(torch.device(DEV))
    rc = lib.woft_gru_gates_nhwc(_lib.ptr(zr), _lib.ptr(bias), _lib.ptr(buf1), 3 * C, B * H * W, C, _lib.ptr(z),
                                 _lib.ptr(buf2), 3 * C, st)
    assert rc == 0
    zb = zr + bias.view(1, -1, 1, 1)
    assert torch.allclose(z, torch.sigmoid(zb[:, :C]), atol=1e-6)
    assert torch.allclose(buf2[:, :C], torch.sigmoid(zb[:, C:]) * h, atol=1e-6)
    q = torch.randn((B, C, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    bq = torch.randn((C,), generator=g).to(DEV)
    zp, rp = zr[:, :C].contiguous(memory_format=cl), zr[:, C:].contiguous(memory_format=cl)
    z2, buf3 = torch.empty_like(z), torch.zeros_like(buf1)
    rc = lib.woft_gru_gates_nhwc_split(_lib.ptr(zp), _lib.ptr(rp), C, _lib.ptr(bias[:C].contiguous()),
                                       _lib.ptr(bias[C:].contiguous()), _lib.ptr(buf1), 3 * C, B * H * W, C, _lib.ptr(z2),
                                       _lib.ptr(buf3), 3 * C, st)
    assert rc == 0 and torch.equal(z2, z) and torch.equal(buf3[:, :C], buf2[:, :C])
    hd = torch.empty((B, C, H, W), device=DEV).contiguous(memory_format=cl)
    tail = buf1[:, C:].clone()
    rc = lib.woft_gru_update_nhwc(_lib.ptr(q), _lib.ptr(bq), _lib.ptr(z), _lib.ptr(buf1), 3 * C, B * H * W, C,
                                  _lib.ptr(buf1), 3 * C, _lib.ptr(hd), st)
    assert rc == 0
    want = (1 - z) * h + z * torch.tanh(q + bq.view(1, -1, 1, 1))
    assert torch.allclose(hd, want, atol=2e-6) and torch.allclose(buf1[:, :C], want, atol=2e-6)
    assert torch.equal(buf1[:, C:], tail)


def test_context_split_kernel():
    """woft_context_split_nhwc: tanh / relu halves of the context encoder's output written into the GRU buffers."""
    from woft_b200 import _lib
    g = torch.Generator().manual_seed(16)
    B, C, H, W = 2, 128, 9, 13
    cl = torch.channels_last
    y = torch.randn((B, 2 * C, H, W), generator=g).to(DEV).contiguous(memory_format=cl)
    b = torch.randn((2 * C,), generator=g).to(DEV)
    h = torch.empty((B, C, H, W), device=DEV).contiguous(memory_format=cl)
    buf1 = torch.full((B, 3 * C, H, W), -3.0, device=DEV).contiguous(memory_format=cl)
    buf2 = torch.full((B, 3 * C, H, W), -3.0, device=DEV).contiguous(memory_format=cl)
    rc = _lib.load().woft_context_split_nhwc(_lib.ptr(y), _lib.ptr(b), B * H * W, C, _lib.ptr(h), _lib.ptr(buf1),
                                             _lib.ptr(buf2), 3 * C, _lib.current_stream(torch.device(DEV)))
    assert rc == 0
    yb = y + b.view(1, -1, 1, 1)
    assert torch.allclose(h, torch.tanh(yb[:, :C]), atol=2e-6) and torch.equal(buf1[:, :C], h)
    assert torch.equal(buf1[:, C:2 * C], F.relu(yb[:, C:])) and torch.equal(buf2[:, C:2 * C], buf1[:, C:2 * C])
    assert (buf1[:, 2 * C:] == -3).all() and (buf2[:, :C] == -3).all() and (buf2[:, 2 * C:] == -3).all()


def test_thin_convolution_kernels():
    from woft_b200 import fast_net
    g = torch.Generator().manual_seed(8)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        for (B, H, W) in ((1, 22, 36), (2, 16, 24), (1, 90, 160), (3, 9, 7)):
            flow = (torch.randn((B, 2, H, W), generator=g) * 3).to(DEV)
            w7 = (torch.randn((128, 2, 7, 7), generator=g) * 0.1).to(DEV)
            b7 = torch.randn((128,), generator=g).to(DEV)
            got = fast_net.flow_conv7x7_relu(flow, w7, b7)
            assert got.is_contiguous(memory_format=torch.channels_last)
            assert torch.allclose(got, F.relu(F.conv2d(flow, w7, b7, padding=3)), atol=2e-5)
            x = torch.randn((B, 256, H, W), generator=g).to(DEV).contiguous(memory_format=torch.channels_last)
            w3 = (torch.randn((2, 256, 3, 3), generator=g) * 0.05).to(DEV)
            b3 = torch.randn((2,), generator=g).to(DEV)
            got = fast_net.flow_out3x3(x, w3, b3)
            want = F.conv2d(x, w3, b3, padding=1)
            assert got.is_contiguous() and torch.allclose(got, want, atol=5e-5)
            # the two-launch form, alone and fused with the coordinate update of the refinement loop
            assert torch.allclose(fast_net.flow_head_step(x, w3, b3), want, atol=5e-5)
            c0 = torch.randn((B, 2, H, W), generator=g).to(DEV)
            c1 = c0 + flow
            coords, f1 = c1.clone(), torch.empty_like(c1)
            f2 = torch.empty_like(c1).contiguous(memory_format=torch.channels_last)
            delta = fast_net.flow_head_step(x, w3, b3, coords=coords, coords0=c0, flow_nchw=f1, flow_nhwc=f2,
                                            want_delta=True)
            assert torch.allclose(delta, want, atol=5e-5)
            assert torch.allclose(coords, c1 + want, atol=1e-4)
            assert torch.allclose(f1, c1 + want - c0, atol=1e-4) and torch.equal(f1, f2)
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_thin_kernels_stay_inside_their_outputs():
    """Outputs embedded in sentinel-filled buffers: nothing may be written before or after them (ragged shapes, where
    tiles, row segments and pixel groups are partial)."""
    from woft_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(12)
    st = _lib.current_stream(torch.device(DEV))
    PAD = 4096

    def guarded(n):
        big = torch.full((n + 2 * PAD,), 12345.0, device=DEV)
        return big, big[PAD:PAD + n]

    def intact(big, n):
        return bool((big[:PAD] == 12345.0).all() and (big[PAD + n:] == 12345.0).all())

    for (B, H, W) in ((1, 9, 7), (2, 17, 23), (1, 90, 160), (3, 16, 41)):
        flow = (torch.randn((B, 2, H, W), generator=g) * 3).to(DEV)
        w7 = (torch.randn((128, 2, 7, 7), generator=g) * 0.1).to(DEV)
        b7 = torch.randn((128,), generator=g).to(DEV)
        n = B * H * W * 128
        big, out = guarded(n)
        assert lib.woft_flow_conv7x7_relu(_lib.ptr(flow), _lib.ptr(w7), _lib.ptr(b7), B, H, W, _lib.ptr(out), st) == 0
        want = F.relu(F.conv2d(flow, w7, b7, padding=3)).permute(0, 2, 3, 1).reshape(-1)
        assert intact(big, n) and torch.allclose(out, want, atol=2e-5)

        x = torch.randn((B, H, W, 64), generator=g).to(DEV)          # channels-last storage, C = 64
        w3 = (torch.randn((2, 64, 3, 3), generator=g) * 0.05).to(DEV)
        b3 = torch.randn((2,), generator=g).to(DEV)
        nb = lib.woft_flow_head_scratch_bytes(B, H, W)
        bigs, scratch = guarded(nb // 4)
        bigd, delta = guarded(B * 2 * H * W)
        bigc, coords = guarded(B * 2 * H * W)
        bign, fn = guarded(B * 2 * H * W)
        bigh, fh = guarded(B * 2 * H * W)
        c0 = torch.randn((B, 2, H, W), generator=g).to(DEV)
        coords.copy_(c0.reshape(-1))
        rc = lib.woft_flow_head_step(_lib.ptr(x), _lib.ptr(w3), _lib.ptr(b3), B, H, W, 64, _lib.ptr(scratch), nb,
                                     _lib.ptr(delta), _lib.ptr(coords), _lib.ptr(c0), _lib.ptr(fn), _lib.ptr(fh), st)
        assert rc == 0
        want = F.conv2d(x.permute(0, 3, 1, 2), w3, b3, padding=1)
        assert all(intact(b_, n_) for b_, n_ in ((bigs, nb // 4), (bigd, B * 2 * H * W), (bigc, B * 2 * H * W),
                                                  (bign, B * 2 * H * W), (bigh, B * 2 * H * W)))
        assert torch.allclose(delta.view(B, 2, H, W), want, atol=5e-5)
        assert torch.allclose(coords.view(B, 2, H, W), c0 + want, atol=1e-4)
        assert torch.allclose(fh.view(B, H, W, 2).permute(0, 3, 1, 2), fn.view(B, 2, H, W))
    with pytest.raises(AssertionError):   # channel count the product kernel cannot split
        from woft_b200 import fast_net
        fast_net.flow_head_step(torch.randn((1, 24, 8, 8), device=DEV), torch.randn((2, 24, 3, 3), device=DEV),
                                torch.randn((2,), device=DEV))


def test_stem_convolution_kernel():
    """woft_stem_conv7x7 (TF32 tensor cores) against the fp32 convolution: the error of 147 products with operands
    rounded to 11 bits, 1e-3 at these magnitudes; ragged sizes exercise the partial tiles and the zero padding."""
    from woft_b200 import fast_net
    g = torch.Generator().manual_seed(14)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        for (B, H, W) in ((1, 64, 96), (2, 45, 83), (3, 720, 1280), (1, 17, 9)):
            x = (torch.rand((B, 3, H, W), generator=g) * 2 - 1).to(DEV).contiguous(memory_format=torch.channels_last)
            w = (torch.randn((64, 3, 7, 7), generator=g) * 0.05).to(DEV)
            b = torch.randn((64,), generator=g).to(DEV)
            want = F.conv2d(x, w, None, stride=2, padding=3)
            got = fast_net.stem_conv7x7(x, w)
            assert got.shape == want.shape and got.is_contiguous(memory_format=torch.channels_last)
            assert (got - want).abs().max().item() <= 2e-3
            got = fast_net.stem_conv7x7(x, w, b, relu=True)
            assert (got - F.relu(want + b.view(1, -1, 1, 1))).abs().max().item() <= 2e-3
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_weight_head_first_layer_kernel():
    """woft_weight_head_conv1 against the reference's formulation: scramble the lookup channels into (level, hp, wp),
    append the broadcast mean response, conv3x3(5 -> 128) + ReLU in fp32 (weighted_raft.py:267-272, 347-384).
    Tolerance: TF32 operands on 45 products of O(1) values."""
    from woft_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(15)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        for n in (1, 37, 1000):   # 81 * n rows: partial last step, steps that straddle two images
            window = torch.randn((n, 81 * 4), generator=g).to(DEV)
            mr = torch.randn((n,), generator=g).to(DEV)
            w = (torch.randn((128, 5, 3, 3), generator=g) * 0.2).to(DEV)
            b = torch.randn((128,), generator=g).to(DEV)
            out = torch.full((n * 81 * 128 + 256,), 777.0, device=DEV)
            rc = lib.woft_weight_head_conv1(_lib.ptr(window), _lib.ptr(mr), _lib.ptr(w), _lib.ptr(b), n, 9, 4, 128,
                                            _lib.ptr(out), _lib.current_stream(torch.device(DEV)))
            assert rc == 0 and (out[n * 81 * 128:] == 777.0).all()
            x = torch.cat([window.view(n, 9, 9, 4).permute(0, 3, 1, 2), mr.view(n, 1, 1, 1).expand(n, 1, 9, 9)], dim=1)
            want = F.relu(F.conv2d(x, w, b, padding=1)).permute(0, 2, 3, 1).reshape(-1)
            assert (out[:n * 81 * 128] - want).abs().max().item() <= 1e-2
            assert (out[:n * 81 * 128] - want).abs().mean().item() <= 1e-3
        assert lib.woft_weight_head_conv1(_lib.ptr(window), _lib.ptr(mr), _lib.ptr(w), _lib.ptr(b), n, 7, 4, 128,
                                          _lib.ptr(out), _lib.current_stream(torch.device(DEV))) == _lib.ERR_INVALID_ARGUMENT
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_weight_head_first_layer_stacked_fp16():
    """woft_weight_head_conv1_stacked_f16 (tcgen05: fp16 operands, fp32 accumulation): the first layer written as fp16 into
    the stacked layout of the tensor-core layers (ten rows of nine pixels per image, the tenth zero) against the same
    convolution in fp64 on the fp16-rounded operands; nothing outside the images' rows touched except the zero rows.
    Image counts whose 128-pixel tiles straddle two and three images and end in a partial tile."""
    from woft_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(16)
    for n in (1, 2, 37, 1000):
        window = torch.randn((n, 81 * 4), generator=g).to(DEV)
        mr = torch.randn((n,), generator=g).to(DEV)
        w = (torch.randn((128, 5, 3, 3), generator=g) * 0.2).to(DEV)
        b = torch.randn((128,), generator=g).to(DEV)
        out = torch.full((n * 90 * 128 + 256,), 7.0, device=DEV, dtype=torch.float16)
        assert lib.woft_weight_head_conv1_stacked_f16(_lib.ptr(window), _lib.ptr(mr), _lib.ptr(w), _lib.ptr(b), n, 9, 4, 128,
                                                      _lib.ptr(out), _lib.current_stream(torch.device(DEV))) == 0
        torch.cuda.synchronize()
        assert (out[n * 90 * 128:] == 7.0).all()
        st = out[:n * 90 * 128].view(n, 90, 128)
        assert (st[:, 81:] == 0).all()
        x = torch.cat([window.view(n, 9, 9, 4).permute(0, 3, 1, 2), mr.view(n, 1, 1, 1).expand(n, 1, 9, 9)], dim=1)
        want = F.relu(F.conv2d(x.half().double(), w.half().double(), b.double(), padding=1)).permute(0, 2, 3, 1).reshape(n, 81, 128)
        scale = want.abs().max().item()
        assert (st[:, :81].double() - want).abs().max().item() <= 2e-3 * scale
        assert (st[:, :81].double() - want).abs().mean().item() <= 2e-4 * scale


def test_convex_upsample_kernel():
    from woft_b200.fast_net import convex_upsample_both
    from woft_b200.raft_net import convex_upsample
    g = torch.Generator().manual_seed(4)
    B, H, W = 2, 12, 20
    flow = (torch.randn((B, 2, H, W), generator=g) * 4).to(DEV)
    wl = torch.randn((B, 1, H, W), generator=g).to(DEV)
    mask = torch.randn((B, 576, H, W), generator=g).to(DEV)
    want_f, want_w = convex_upsample(flow, mask), convex_upsample(wl, mask) / 8
    for m in (mask, mask.contiguous(memory_format=torch.channels_last)):
        got_f, got_w = convex_upsample_both(flow, wl, m)
        assert torch.allclose(got_f, want_f, atol=2e-5) and torch.allclose(got_w, want_w, atol=2e-6)
        got_f, got_w = convex_upsample_both(flow, wl, m, sigmoid=True)
        assert torch.allclose(got_w, torch.sigmoid(want_w), atol=2e-6)
    only_f, none_w = convex_upsample_both(flow, None, mask)
    assert none_w is None and torch.allclose(only_f, want_f, atol=2e-5)


def test_full_size_properties_1080p():
    """BASELINE config 5 size (1920x1080 -> 135x240 feature maps): 4.2 GB volume, odd level sizes (135 -> 67 -> 33 -> 16,
    240 -> 120 -> 60 -> 30: level 3 rows are not 16-byte multiples, so that level takes the direct-store path)."""
    from woft_b200.corr import CorrBlock
    torch.manual_seed(1)
    B, C, H, W = 1, 256, 135, 240
    f1 = torch.randn(B, C, H, W, device=DEV)
    f2 = torch.randn(B, C, H, W, device=DEV)
    blk = CorrBlock(f1, f2, mode="fp16")
    blk.check()
    lv = blk.corr_pyramid
    assert [tuple(l.shape[2:]) for l in lv] == [(135, 240), (67, 120), (33, 60), (16, 30)]
    idx = torch.randint(0, H * W, (2, 20000), device=DEV)
    want = (f1.view(C, H * W)[:, idx[0]].double() * f2.view(C, H * W)[:, idx[1]].double()).sum(0) / 16.0
    got = lv[0].view(H * W, H * W)[idx[0], idx[1]].double()
    assert (got - want).abs().max().item() < CORR_ATOL["fp16"]
    for l in range(3):
        sel = slice(0, 4000)  # a slab of queries is enough; pooling is per query
        assert torch.allclose(lv[l + 1][sel], F.avg_pool2d(lv[l][sel], 2, stride=2), atol=2e-6)
        sel = slice(H * W - 3000, H * W)
        assert torch.allclose(lv[l + 1][sel], F.avg_pool2d(lv[l][sel], 2, stride=2), atol=2e-6)
    ys, xs = torch.meshgrid(torch.arange(H, device=DEV), torch.arange(W, device=DEV), indexing="ij")
    coords = torch.stack([xs, ys]).float()[None] + (torch.rand((1, 2, H, W), device=DEV) * 2 - 1) * 8
    out = blk(coords)
    out_cl = blk(coords, flags=8)
    assert torch.allclose(out, out_cl.contiguous(), atol=2e-5)  # separable lerp vs 4-tap weights: same values to rounding
    sel = torch.arange(20000, 20400, device=DEV)
    ref = wo.corr_lookup([l[sel].float().cpu() for l in lv], coords.view(2, -1)[:, sel].cpu().view(1, 2, 1, -1))
    got = out.view(324, -1)[:, sel].cpu().view(1, 324, 1, -1)
    assert (got - ref).abs().max().item() < 2e-5 * max(1.0, ref.abs().max().item())
    del blk, lv, out, out_cl
    torch.cuda.empty_cache()


def test_tracked_frame_matches_cpu_pipeline():
    """The whole tracked-frame path (flow wrapper -> mask -> Sobol pick -> weighted fit -> inlier test) on the GPU against
    the oracle's CPU pipeline on the same frames and weights: homography corner error and the success flag."""
    from types import SimpleNamespace
    from oracle import cpu_pipeline
    from oracle.gen_golden import make_frames
    from woft_b200 import tracker
    from woft_b200.corr import CorrBlock
    from woft_b200.optical_flow import RAFTWrapper
    H, W = 256, 320
    img0, img1 = make_frames(H, W, 21)
    mask = np.zeros((H, W), np.uint8)
    mask[48:208, 64:256] = 255
    net = cpu_pipeline.make_net(0)
    torch.set_num_threads(8)
    H_cpu, ok_cpu, n_cpu = cpu_pipeline.track_frame(net, img0, img1, mask, np.eye(3), 12)
    flow_conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False), model=None,
                                seeded_weights=0, iters=12, padding_mode="nopad")
    for device_prewarp in (False, True):
        flower = RAFTWrapper(flow_conf, device=DEV, corr_block=functools.partial(CorrBlock, mode="tf32x3"))
        trk = tracker.PlanarTracker(tracker.default_config(flow_conf, device_prewarp=device_prewarp), flower=flower)
        trk.init(img0, mask)
        H_gpu, meta = trk.track(img1)
        assert meta.global_H_success == ok_cpu
        err = wo.corner_error(torch.from_numpy(H_gpu), torch.from_numpy(H_cpu), W, H)
        assert err <= 0.05, f"tracked-frame homography differs from the CPU pipeline by {err} px at the corners"


@pytest.fixture(scope="module")
def cpu_frame_720p():
    """One 1280x720 pair of bench.py's synthetic sequence through the CPU oracle pipeline (about 5 s of host time)."""
    import bench
    from oracle import cpu_pipeline
    base, frames, mask = bench.make_sequence(1, 0)
    net = cpu_pipeline.make_net(0)
    torch.set_num_threads(max(1, min(16, torch.get_num_threads())))
    t1 = torch.from_numpy(base[:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    t2 = torch.from_numpy(frames[0][:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    _, flow_up, _, w_up = wo.flow_pass(net, t1, t2, iters=12)
    H_cpu, ok_cpu, _ = cpu_pipeline.track_frame(net, base, frames[0], mask, np.eye(3), 12)
    return base, frames[0], mask, flow_up, w_up, H_cpu, ok_cpu


@pytest.mark.parametrize("corr_mode,tf32,epe_tol,corner_tol", [("fp16s", True, 0.02, 0.05), ("fp16", True, 0.02, 0.05),
                                                                ("fp16x3", False, 0.01, 0.05)])
def test_benched_configuration_parity_720p(cpu_frame_720p, corr_mode, tf32, epe_tol, corner_tol):
    """What bench.py times, checked: the same constructor arguments (CUDA graph, fused channels-last schedule, correlation
    mode, TF32 convolutions allowed / forbidden) on a 1280x720 pair of bench.make_sequence, against the CPU oracle.
    ("fp16s", TF32 on) is the headline configuration with its stated tolerance -- fp16 operands everywhere and the
    correlation pyramid stored as fp16; ("fp16", ...) the same with the fp32 pyramid; ("fp16x3", TF32 off) the fp32-class one
    under the north star's budgets (0.01 px mean EPE, 0.05 px corner error)."""
    from types import SimpleNamespace
    from woft_b200 import tracker
    from woft_b200.corr import CorrBlock
    from woft_b200.optical_flow import RAFTWrapper
    base, frame, mask, flow_cpu, w_cpu, H_cpu, ok_cpu = cpu_frame_720p
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = tf32
    try:
        flow_conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False, alternate_corr=False),
                                    model=None, seeded_weights=0, iters=12, padding_mode="nopad")
        flower = RAFTWrapper(flow_conf, device=DEV, use_cuda_graph=True,
                             corr_block=functools.partial(CorrBlock, mode=corr_mode), tf32=tf32)
        trk = tracker.PlanarTracker(tracker.default_config(flow_conf), flower=flower)
        trk.init(base, mask)
        for _ in range(2):   # the second call replays the captured graph, as every timed step does
            flow, w = flower.compute_flow(base, frame, mode="flow", do_sigmoid=False)
        H_gpu, meta = trk.track(frame)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    epe = (flow.cpu().view(2, 720, 1280) - flow_cpu[0]).norm(dim=0).mean().item()
    w_err = (w.cpu().view(720, 1280) - w_cpu[0, 0]).abs().max().item()
    corner = wo.corner_error(torch.from_numpy(H_gpu), torch.from_numpy(H_cpu), 1280, 720)
    print(f"\nbenched configuration {corr_mode}, TF32 convolutions {tf32}: mean EPE {epe:.5f} px, "
          f"weight-logit max error {w_err:.4f}, homography corner error {corner:.5f} px")
    assert meta.global_H_success == ok_cpu
    assert epe <= epe_tol, f"mean EPE {epe} px"
    assert corner <= corner_tol, f"corner error {corner} px"


def test_lost_branch_matches_cpu_pipeline():
    """A frame whose global fit fails the inlier test (YAOF_tracker_single_control.py:167-204): the tracker must chain a
    local homography from the previous frame.  The failure is forced through the config's own hook
    (redet_success_fn), the local homography is compared with the oracle's restatement of that branch."""
    from types import SimpleNamespace
    from oracle import cpu_pipeline
    from oracle.gen_golden import make_frames
    from woft_b200 import tracker
    from woft_b200.corr import CorrBlock
    from woft_b200.optical_flow import RAFTWrapper
    H, W = 256, 320
    img0, img1 = make_frames(H, W, 33)
    mask = np.zeros((H, W), np.uint8)
    mask[48:208, 64:256] = 255
    net = cpu_pipeline.make_net(0)
    torch.set_num_threads(8)
    flow_conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False), model=None,
                                seeded_weights=0, iters=12, padding_mode="nopad")
    for device_prewarp in (False, True):
        conf = tracker.default_config(flow_conf, device_prewarp=device_prewarp)
        conf.fused_select = False                      # the fused path has the default success test built in
        conf.redet_success_fn = lambda *a, **k: torch.zeros((), dtype=torch.bool, device=DEV)
        flower = RAFTWrapper(flow_conf, device=DEV, corr_block=functools.partial(CorrBlock, mode="tf32x3"))
        trk = tracker.PlanarTracker(conf, flower=flower)
        trk.init(img0, mask)
        H_gpu, meta = trk.track(img1)
        assert meta.lost and meta.N_lost == 1 and not meta.global_H_success and trk.flow_passes == 2
        H_cpu, n_cpu = cpu_pipeline.track_frame_local(net, img0, img1, mask, np.eye(3), 12)
        err = wo.corner_error(torch.from_numpy(meta.H_local_cur2init), torch.from_numpy(H_cpu), W, H)
        assert err <= 0.05, f"local-branch homography differs from the CPU pipeline by {err} px at the corners"
        assert np.array_equal(H_gpu, meta.H_local_cur2init)
        # second lost frame chains onto the first one's result and the pre-warp keeps using the last good homography
        H2, meta2 = trk.track(img0)
        assert meta2.N_lost == 2 and np.all(np.isfinite(H2))
        H_cpu2, _ = cpu_pipeline.track_frame_local(net, img1, img0, mask, H_gpu, 12)
        assert wo.corner_error(torch.from_numpy(H2), torch.from_numpy(H_cpu2), W, H) <= 0.1
    # the same branch with everything after the flow on the device (woft_select_correspondences_local): the global
    # stage's inlier fraction is zeroed on its way back to force the branch
    conf = tracker.default_config(flow_conf, device_prewarp=True)
    assert conf.fused_select
    flower = RAFTWrapper(flow_conf, device=DEV, corr_block=functools.partial(CorrBlock, mode="tf32x3"))
    trk = tracker.PlanarTracker(conf, flower=flower)
    trk.init(img0, mask)
    fused = trk.estimate_global_fused

    def fail_global(*a):
        res = fused(*a).clone()
        res[9] = 0.0
        return res

    trk.estimate_global_fused = fail_global
    H_gpu, meta = trk.track(img1)
    assert meta.lost and trk.flow_passes == 2
    H_cpu, _ = cpu_pipeline.track_frame_local(net, img0, img1, mask, np.eye(3), 12)
    assert wo.corner_error(torch.from_numpy(H_gpu), torch.from_numpy(H_cpu), W, H) <= 0.05
    H2, meta2 = trk.track(img0)
    H_cpu2, _ = cpu_pipeline.track_frame_local(net, img1, img0, mask, H_gpu, 12)
    assert meta2.N_lost == 2 and wo.corner_error(torch.from_numpy(H2), torch.from_numpy(H_cpu2), W, H) <= 0.1


def test_tracker_1080p_with_a_lost_frame():
    """BASELINE config 5 in small: 1920x1080 frames (a 32 400-query volume, 5.6 GB pyramid) through the whole tracker with
    one frame forced into the lost branch.  The weights are seeded random ones, so there is no known answer at this size
    (parity with the CPU pipeline is checked at 256x320): every homography must be finite and normalised and the pass
    counts must match the branches taken."""
    from types import SimpleNamespace
    from oracle.gen_golden import make_frames
    from woft_b200 import tracker
    from woft_b200.optical_flow import RAFTWrapper
    H, W = 1080, 1920
    img0, img1 = make_frames(H, W, 5)
    mask = np.zeros((H, W), np.uint8)
    mask[300:800, 500:1400] = 255
    flow_conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False), model=None,
                                seeded_weights=0, iters=12, padding_mode="nopad")
    trk = tracker.PlanarTracker(tracker.default_config(flow_conf), flower=RAFTWrapper(flow_conf, device=DEV))
    trk.init(img0, mask)
    fused = trk.estimate_global_fused
    lose = {"now": False}

    def maybe_fail(*a):
        res = fused(*a)
        if lose["now"]:
            res = res.clone()
            res[9] = 0.0
        return res

    trk.estimate_global_fused = maybe_fail
    H_first, meta = trk.track(img0)
    assert np.all(np.isfinite(H_first)) and trk.flow_passes >= 1
    passes = trk.flow_passes
    lose["now"] = True
    H_lost, meta = trk.track(img1)
    assert meta.lost and trk.flow_passes == passes + 2 and np.all(np.isfinite(H_lost))
    lose["now"] = False
    H_next, meta = trk.track(img1)
    assert np.all(np.isfinite(H_next)) and H_next[2, 2] == pytest.approx(1.0)
    del trk
    torch.cuda.empty_cache()


def test_fused_select_fit_check_matches_reference_pieces():
    """woft_select_correspondences + woft_homography_fit_n + woft_inlier_fraction against the oracle's restatement of
    _mask_coords, the Sobol subsampler, find_homography_nonhomogeneous_QR and redet_success_fn."""
    import cv2
    from woft_b200 import tracker
    from woft_b200.misc import torch_get_featuremap_coords
    g = torch.Generator().manual_seed(12)
    for (H, W, box) in ((256, 320, (48, 208, 64, 256)), (64, 96, (20, 38, 30, 52))):  # the second keeps < 500 points
        flow = torch.randn((2, H, W), generator=g) * 6
        flow[:, : H // 3] += 40  # push a band out of the frame
        logits = torch.randn((1, H, W), generator=g)
        tmask = torch.zeros((H, W), dtype=torch.bool)
        tmask[box[0]:box[1], box[2]:box[3]] = True
        pw = torch.from_numpy(cv2.warpPerspective(np.ones((H, W)), np.array([[1, 0.03, 9], [0.01, 1, -6], [0, 0, 1.0]]),
                                                  (W, H)) > 0)
        src = torch_get_featuremap_coords((H, W), device="cpu")
        dst = src + flow.reshape(2, -1)
        wts = torch.sigmoid(logits).reshape(1, -1)
        t_c, c_c, w_c, _ = wo.mask_coords(tmask, src, dst, wts, pw)
        keep = wo.sobol_subsample_mask(t_c.shape[1])
        if keep is not None:
            t_c, c_c, w_c = t_c[:, keep], c_c[:, keep], w_c[:, keep]
        res, (pts_t, pts_c, w, n) = tracker.select_fit_check(flow.to(DEV), logits.to(DEV), tmask.to(DEV), pw.to(DEV))
        m = int(n.item())
        assert m == t_c.shape[1]
        assert torch.equal(pts_t[:m].cpu(), t_c.float().T) and torch.equal(pts_c[:m].cpu(), c_c.T)
        assert torch.allclose(w[:m].cpu(), w_c[0], atol=1e-6)
        res = res.cpu()
        assert int(res[10].item()) == m
        Href = wo.find_homography_nonhomogeneous_QR(c_c.T[None].double(), t_c.T[None].double(), w_c.double())
        assert wo.corner_error(res[:9].view(1, 3, 3), Href, W, H) <= 0.05
        errs = wo.torch_proj_errors(res[:9].view(1, 3, 3), c_c[None], t_c.float()[None])
        assert res[9].item() == pytest.approx((errs <= 5).float().mean().item(), abs=1e-6)


@pytest.mark.parametrize("shape", [(1, 32, 90, 160), (2, 16, 48, 64), (1, 8, 36, 52)])
def test_tma_lookup_is_the_load_store_lookup(shape):
    """lookup_tma.cu (flag 64: footprints fetched as TMA boxes, zero padding by out-of-bounds fill) against lookup.cu's
    kernel on jittered and far out-of-range coordinates: same arithmetic, so the channels-last outputs are equal bit for
    bit, fp32 and fp16; and both agree with the oracle."""
    from woft_b200 import _lib
    from woft_b200.corr import CorrBlock
    B, C, H, W = shape
    g = torch.Generator().manual_seed(H)
    f1, f2 = torch.randn(shape, generator=g), torch.randn(shape, generator=g)
    blk = CorrBlock(f1.to(DEV), f2.to(DEV), mode="fp32")
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    coords = grid + (torch.rand(grid.shape, generator=g) * 2 - 1) * 9
    far = torch.rand((B, 1, H, W), generator=g) < 0.1
    coords = torch.where(far, coords + (torch.rand(grid.shape, generator=g) - 0.5) * 8 * max(H, W), coords).to(DEV)
    for flags in (8, 24):
        a = blk(coords, flags=flags | 64)
        b = blk(coords, flags=flags)
        with torch.cuda.device(DEV):
            assert _lib.load().woft_corr_lookup_check(_lib.current_stream(torch.device(DEV))) == _lib.OK, _lib.last_error()
        assert torch.equal(a[:, :324], b[:, :324]), f"flags {flags}"
    want = wo.OracleCorrBlock(f1, f2)(coords.cpu())
    np.testing.assert_allclose(blk(coords, flags=8).cpu().numpy(), want.numpy(), atol=1e-4)
