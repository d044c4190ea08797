"""woft_conv_tc (tcgen05 implicit-GEMM convolution with fused epilogues) against plain PyTorch expressions of the layers
it replaces (raft_core/update.py:6-14, 33-60, 79-97, 122-125; raft_core/weighted_raft.py:318-339, 347-384).

Two references per case:
  exact   the same operand rounding as the kernel (activations truncated to TF32 as the tensor core reads them, filters
          rounded to nearest TF32 by pack_conv) convolved in fp64: only the fp32 accumulation order differs -> 2e-4
          relative to the output scale; this is the layout / indexing / padding check
  fp32    torch's fp32 convolution (TF32 off) of the unrounded operands: TF32-class tolerance 4e-3 of the output scale
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _exact_convs():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def trunc_tf32(t):
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


def make(B, cin, cout, H, W, kh, kw, seed, ctot=None, off=0):
    g = torch.Generator().manual_seed(seed)
    ctot = ctot or cin
    x = cl(torch.randn((B, ctot, H, W), generator=g)).to(DEV)
    w = (torch.randn((cout, cin, kh, kw), generator=g) / (cin * kh * kw) ** 0.5).to(DEV)
    b = (torch.randn((cout,), generator=g) * 0.1).to(DEV)
    return x, w, b


def refs(x, w, b, pad):
    from woft_b200.conv_tc import round_to_tf32
    exact = F.conv2d(trunc_tf32(x).double(), round_to_tf32(w).double(), b.double(), padding=pad).float()
    plain = F.conv2d(x, w, b, padding=pad)
    return exact, plain


def check(got, exact, plain, what):
    scale = plain.abs().max().item()
    e1 = (got - exact).abs().max().item() / scale
    e2 = (got - plain).abs().max().item() / scale
    assert e1 <= 2e-4, f"{what}: differs from the same-rounding reference by {e1:.2e} of the output scale"
    assert e2 <= 4e-3, f"{what}: differs from the fp32 convolution by {e2:.2e} of the output scale"


@pytest.mark.parametrize("shape", [
    # B, cin, cout, H, W, kh, kw, pad
    (1, 256, 192, 90, 160, 3, 3, (1, 1)),    # convc2 at 1280x720
    (1, 324, 256, 90, 160, 1, 1, (0, 0)),    # convc1: Cin not a multiple of 32 (zero-filled by TMA)
    (1, 128, 64, 90, 160, 3, 3, (1, 1)),     # convf2
    (2, 64, 126, 13, 21, 3, 3, (1, 1)),      # ragged pixel grid, batch 2, 126 of 128 output channels stored
    (1, 256, 576, 24, 40, 1, 1, (0, 0)),     # mask head 1x1: three output-channel tiles
    (1, 96, 32, 17, 9, 5, 1, (2, 0)),        # vertical 5x1
    (1, 96, 32, 9, 33, 1, 5, (0, 2)),        # horizontal 1x5
])
def test_conv_bias_relu_matches_torch(shape):
    from woft_b200 import _lib, conv_tc
    B, cin, cout, H, W, kh, kw, pad = shape
    x, w, b = make(B, cin, cout, H, W, kh, kw, seed=cin + cout)
    pc = conv_tc.PackedConv(w, b, pad)
    out = cl(torch.full((B, cout + 8, H, W), 7.0, device=DEV))   # written at channel offset 4 of a wider buffer
    out2 = cl(torch.full((B, cout, H, W), 7.0, device=DEV))
    conv_tc.run(pc, x, relu=True, out=out, out_off=4, out2=out2)
    conv_tc.check(DEV)
    exact, plain = refs(x, w, b, pad)
    check(out[:, 4:4 + cout], exact.relu(), plain.relu(), f"{shape}")
    assert torch.equal(out[:, 4:4 + cout], out2)
    assert (out[:, :4] == 7.0).all() and (out[:, 4 + cout:] == 7.0).all(), "wrote outside its channel slice"


def test_conv_reads_a_channel_slice_and_scales():
    from woft_b200 import conv_tc
    x, w, b = make(1, 128, 64, 30, 44, 3, 3, seed=3, ctot=384)
    pc = conv_tc.PackedConv(w, b, 1)
    out = cl(torch.empty((1, 64, 30, 44), device=DEV))
    conv_tc.run(pc, x, x_off=128, relu=False, scale=0.25, out=out)
    exact, plain = refs(x[:, 128:256], w, b, 1)
    check(out, exact * 0.25, plain * 0.25, "slice")


@pytest.mark.parametrize("kh,kw", [(1, 5), (5, 1)])
@pytest.mark.parametrize("H,W", [(90, 160), (11, 27)])
def test_gru_half_step_matches_torch(kh, kw, H, W):
    """SepConvGRU half step (update.py:45-60) as two launches: stacked convz|convr with the gate epilogue, convq with the
    blend epilogue, on the persistent [h | x] / [r*h | x] buffers of the fused schedule."""
    from woft_b200 import _lib, conv_tc
    C, X = 128, 256
    g = torch.Generator().manual_seed(kh * 10 + W)
    pad = (kh // 2, kw // 2)
    buf1 = cl(torch.randn((1, C + X, H, W), generator=g)).to(DEV)
    buf1[:, :C] = torch.tanh(buf1[:, :C])
    buf2 = buf1.clone(memory_format=torch.channels_last)
    wz, wr, wq = ((torch.randn((C, C + X, kh, kw), generator=g) / ((C + X) * 5) ** 0.5).to(DEV) for _ in range(3))
    bz, br, bq = ((torch.randn((C,), generator=g) * 0.1).to(DEV) for _ in range(3))
    h0 = buf1[:, :C].clone()
    # reference, fp32
    hx = buf1.clone()
    zt = torch.sigmoid(F.conv2d(hx, wz, bz, padding=pad))
    rt = torch.sigmoid(F.conv2d(hx, wr, br, padding=pad))
    qt = torch.tanh(F.conv2d(torch.cat([rt * h0, hx[:, C:]], 1), wq, bq, padding=pad))
    h_ref = (1 - zt) * h0 + zt * qt
    pzr = conv_tc.PackedConv(torch.cat([wz, wr], 0), torch.cat([bz, br], 0), pad)
    pq = conv_tc.PackedConv(wq, bq, pad)
    z = torch.empty((H * W, C), device=DEV)
    conv_tc.run(pzr, buf1, epilogue=_lib.CONV_GRU_ZR, h=buf1, z=z, out=buf2)
    conv_tc.run(pq, buf2, epilogue=_lib.CONV_GRU_Q, h=buf1, z=z, out=buf1)
    conv_tc.check(DEV)
    assert (z.view(1, H, W, C).permute(0, 3, 1, 2) - zt).abs().max().item() <= 2e-3
    assert (buf2[:, :C] - rt * h0).abs().max().item() <= 2e-3
    assert torch.equal(buf2[:, C:], hx[:, C:]) and torch.equal(buf1[:, C:], hx[:, C:])
    assert (buf1[:, :C] - h_ref).abs().max().item() <= 3e-3


def test_fp16_operands_and_output():
    from woft_b200 import conv_tc
    x, w, b = make(1, 128, 128, 40, 56, 3, 3, seed=9)
    xh = cl(x.half())
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    out = cl(torch.empty((1, 128, 40, 56), device=DEV, dtype=torch.float16))
    conv_tc.run(pc, xh, relu=True, out=out)
    exact = F.conv2d(xh.double(), w.half().double(), b.double(), padding=1).relu()
    scale = exact.abs().max().item()
    assert (out.double() - exact).abs().max().item() / scale <= 1.5e-3   # fp16 rounding of the stored result


@pytest.mark.parametrize("f16,halo", [(False, False), (True, False), (True, True), (False, True)])
def test_weight_head_layers_on_stacked_images(f16, halo):
    """Layers 2-3 of the weight head + 1x1 + patch mean (weighted_raft.py:318-339, 380-384) on 9x9 images stacked along H
    with one zero row between them: conv+ReLU writes zeros into the separator rows, the last layer reduces each image."""
    from woft_b200 import _lib, conv_tc
    n, P, C = 53, 9, 128
    g = torch.Generator().manual_seed(1)
    imgs = torch.randn((n, C, P, P), generator=g).relu().to(DEV)
    w2, w3 = ((torch.randn((C, C, 3, 3), generator=g) / (C * 9) ** 0.5).to(DEV) for _ in range(2))
    b2, b3 = ((torch.randn((C,), generator=g) * 0.1).to(DEV) for _ in range(2))
    w1 = (torch.randn((C,), generator=g) / C ** 0.5).to(DEV)
    b1 = 0.3
    dt = torch.float16 if f16 else torch.float32
    if f16:
        imgs = imgs.half().float()
    want2 = F.conv2d(imgs, w2.half().float() if f16 else w2, b2, padding=1).relu()
    x3 = want2.half().float() if f16 else want2
    want3 = F.conv2d(x3, w3.half().float() if f16 else w3, b3, padding=1).relu()
    want = (want3 * w1.view(1, C, 1, 1)).sum(1).mean(dim=(1, 2)) + b1
    # stacked layout: (1, C, n*10, 9) channels-last, row 9 of every image zero
    stacked = torch.zeros((n, P + 1, P, C), device=DEV, dtype=dt)
    stacked[:, :P] = imgs.permute(0, 2, 3, 1).to(dt)
    x = stacked.view(1, n * (P + 1), P, C).permute(0, 3, 1, 2)
    mid = torch.full((1, n * (P + 1), P, C), 5.0, device=DEV, dtype=dt).permute(0, 3, 1, 2)
    p2, p3 = conv_tc.PackedConv(w2, b2, 1, f16=f16), conv_tc.PackedConv(w3, b3, 1, f16=f16)
    # halo: all nine taps read one 11 x 13 activation box (11 output rows of 9 pixels per tile) at their own row offsets
    kw = dict(box=(9, 11), halo=True) if halo else dict(box=(9, 14))
    conv_tc.run(p2, x, relu=True, out=mid, row_period=P + 1, row_valid=P, **kw)
    logits = torch.full((n,), 9.0, device=DEV)
    conv_tc.run(p3, mid, epilogue=_lib.CONV_HEAD_MEAN, w1x1=w1, head_out=logits, head_bias=b1, row_period=P + 1,
                row_valid=P, **kw)
    conv_tc.check(DEV)
    got2 = mid.permute(0, 2, 3, 1).reshape(n, P + 1, P, C)
    assert (got2[:, P] == 0).all(), "separator rows must be zero"
    tol = 4e-3 if f16 else 3e-3
    s2 = want2.abs().max().item()
    assert (got2[:, :P].permute(0, 3, 1, 2).float() - want2).abs().max().item() / s2 <= tol
    assert (logits - want).abs().max().item() <= tol * max(1.0, want.abs().max().item())
    # deterministic: a second run gives the same bits
    logits2 = torch.empty_like(logits)
    conv_tc.run(p3, mid, epilogue=_lib.CONV_HEAD_MEAN, w1x1=w1, head_out=logits2, head_bias=b1, row_period=P + 1,
                row_valid=P, **kw)
    assert torch.equal(logits, logits2)


def test_conv_tc_argument_errors():
    from woft_b200 import conv_tc
    x, w, b = make(1, 32, 32, 8, 8, 3, 3, seed=0)
    pc = conv_tc.PackedConv(w, b, 1)
    with pytest.raises(AssertionError):
        conv_tc.run(pc, x.contiguous(), out=x)            # not channels-last
    with pytest.raises(AssertionError):
        conv_tc.run(pc, x, out=None)                       # no output


@pytest.mark.parametrize("shape", [
    # B, cin, cout, H, W, k, pad
    (2, 64, 96, 72, 128, 3, 1),     # first convolution of the encoders' second stage (stride 2), two images
    (1, 96, 128, 45, 80, 3, 1),     # third stage, odd input height
    (2, 64, 96, 72, 128, 1, 0),     # the 1x1 stride-2 projection of the shortcut
    (1, 64, 64, 37, 51, 3, 1),      # ragged, odd x odd
])
def test_stride_two_convolution_fp16(shape):
    """Stride-2 convolutions of the encoders (raft_core/extractor.py:16-24, 36-37): TMA boxes with an element stride of 2."""
    from woft_b200 import conv_tc
    B, cin, cout, H, W, k, pad = shape
    x, w, b = make(B, cin, cout, H, W, k, k, seed=H + cout)
    xh = cl(x.half())
    pc = conv_tc.PackedConv(w, b, pad, f16=True)
    Ho, Wo = (H + 2 * pad - k) // 2 + 1, (W + 2 * pad - k) // 2 + 1
    out = cl(torch.full((B, cout, Ho, Wo), 7.0, device=DEV, dtype=torch.float16))
    conv_tc.run(pc, xh, relu=True, out=out, stride=2)
    conv_tc.check(DEV)
    want = F.conv2d(xh.double(), w.half().double(), b.double(), stride=2, padding=pad).relu()
    assert tuple(want.shape) == tuple(out.shape)
    scale = want.abs().max().item()
    assert (out.double() - want).abs().max().item() / scale <= 1.5e-3


def test_residual_tail_in_the_epilogue():
    """relu(x + relu(conv(y) + b)): the tail of a batch-norm residual block (extractor.py:45-56, norm folded) in one launch."""
    from woft_b200 import conv_tc
    x, w, b = make(2, 64, 64, 45, 80, 3, 3, seed=4)
    xh = cl(x.half())
    res = cl(torch.randn((2, 64, 45, 80), generator=torch.Generator().manual_seed(5)).half().to(DEV))
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    out = cl(torch.empty((2, 64, 45, 80), device=DEV, dtype=torch.float16))
    conv_tc.run(pc, xh, relu=True, out=out, residual=res)
    want = (res.double() + F.conv2d(xh.double(), w.half().double(), b.double(), padding=1).relu()).relu()
    scale = want.abs().max().item()
    assert (out.double() - want).abs().max().item() / scale <= 1.5e-3


@pytest.mark.parametrize("shape", [(2, 64, 64, 90, 160, 1), (2, 64, 96, 90, 160, 2), (1, 96, 96, 23, 37, 1)])
def test_convolution_gathers_instance_norm_statistics(shape):
    """conv -> InstanceNorm -> ReLU (-> residual add) of the feature encoder (extractor.py:30-32, 45-56) with the norm's
    statistics gathered in the convolution's epilogue, against torch on the convolution's own fp16 output."""
    from woft_b200 import fast_net
    from woft_b200 import conv_tc
    B, cin, cout, H, W, stride = shape
    x, w, _ = make(B, cin, cout, H, W, 3, 3, seed=cout + stride)
    xh = cl((x * 1.7 + 0.4).half())
    pc = conv_tc.PackedConv(w, None, 1, f16=True)
    y = fast_net._conv_f16(pc, xh, stride=stride)                    # the plain convolution, fp16
    got = fast_net.conv_instance_norm_f16(pc, xh, stride=stride)
    want = F.relu(F.instance_norm(y.float(), eps=1e-5))
    assert (got.float() - want).abs().max().item() <= 4e-3          # fp16 rounding of a unit-variance result
    res = cl(torch.randn(tuple(y.shape), generator=torch.Generator().manual_seed(3)).half().to(DEV))
    got_r = fast_net.conv_instance_norm_f16(pc, xh, stride=stride, residual=res)
    assert (got_r.float() - F.relu(res.float() + want)).abs().max().item() <= 6e-3
    conv_tc.check(DEV)


@pytest.mark.parametrize("H,W", [(90, 160), (13, 21)])
def test_flow_head_epilogue_and_gather(H, W):
    """FlowHead (update.py:6-14): conv1 + ReLU + the tap products of conv2 in the convolution's epilogue, then the gather
    that also advances the coordinates and copies the flow (fp32 and fp16), against the two torch convolutions."""
    from woft_b200 import _lib, conv_tc
    g = torch.Generator().manual_seed(H)
    hbuf = cl((torch.randn((1, 384, H, W), generator=g)).half().to(DEV))
    w1 = (torch.randn((256, 128, 3, 3), generator=g) / (128 * 9) ** 0.5).to(DEV)
    b1 = (torch.randn((256,), generator=g) * 0.1).to(DEV)
    w2 = (torch.randn((2, 256, 3, 3), generator=g) / (256 * 9) ** 0.5).to(DEV)
    b2 = (torch.randn((2,), generator=g) * 0.1).to(DEV)
    want = F.conv2d(F.conv2d(hbuf[:, :128].float(), w1.half().float(), b1, padding=1).relu(), w2, b2, padding=1)
    wt = w2.reshape(2, 256, 9).permute(1, 2, 0)
    fh2 = torch.zeros((256, 24), device=DEV)
    fh2[:, 0:10] = wt[:, 0:5].reshape(-1, 10)
    fh2[:, 12:20] = wt[:, 5:9].reshape(-1, 8)
    pc = conv_tc.PackedConv(w1, b1, 1, f16=True)
    taps = torch.full((H * W, 18), float("nan"), device=DEV)
    conv_tc.run(pc, hbuf, epilogue=_lib.CONV_FLOW_TAPS, w1x1=fh2, head_out=taps)
    conv_tc.check(DEV)
    coords0 = torch.rand((1, 2, H, W), generator=g).to(DEV) * 50
    coords = coords0 + 1.5
    c_before = coords.clone()
    flow_nchw = torch.empty((1, 2, H, W), device=DEV)
    flow_nhwc = torch.empty((H * W, 2), device=DEV)
    fa = cl(torch.zeros((1, 384, H, W), device=DEV, dtype=torch.float16))
    fb = cl(torch.zeros((1, 384, H, W), device=DEV, dtype=torch.float16))
    with torch.cuda.device(DEV):
        rc = _lib.load().woft_flow_head_gather(_lib.ptr(taps), _lib.ptr(b2), 1, H, W, None, _lib.ptr(coords), _lib.ptr(coords0),
                                               _lib.ptr(flow_nchw), _lib.ptr(flow_nhwc), _lib.ptr(fa), _lib.ptr(fb), 384, 382,
                                               _lib.current_stream(torch.device(DEV)))
    assert rc == _lib.OK, _lib.last_error()
    assert not torch.isnan(taps).any()
    delta = coords - c_before
    assert (delta - want).abs().max().item() <= 3e-3 * max(1.0, want.abs().max().item())
    assert torch.allclose(flow_nchw, coords - coords0)
    assert torch.allclose(flow_nhwc.view(H, W, 2).permute(2, 0, 1)[None], flow_nchw)
    assert torch.equal(fa[:, 382:384], flow_nchw.half()) and torch.equal(fb[:, 382:384], fa[:, 382:384])
    assert (fa[:, :382] == 0).all()


@pytest.mark.parametrize("shape", [(2, 64, 64, 45, 80, 3, 3), (1, 128, 256, 90, 160, 3, 3), (1, 384, 128, 90, 160, 5, 1),
                                   (1, 96, 96, 23, 37, 3, 3)])
@pytest.mark.parametrize("f16", [True, False])
def test_row_halo_mode_is_the_plain_convolution(shape, f16):
    """halo=2: one activation box per channel block and filter column, the column's taps reading it ky * box_w rows further
    in (8-row aligned operand starts).  The products are summed in a different order (channel block, column, row instead
    of tap, channel block), so the two modes agree to fp32 summation noise, and both with torch."""
    from woft_b200 import conv_tc
    B, cin, cout, H, W, kh, kw = shape
    x, w, b = make(B, cin, cout, H, W, kh, kw, seed=cin + kh)
    pad = (kh // 2, kw // 2)
    if f16:
        x = cl(x.half())
    pc = conv_tc.PackedConv(w, b, pad, f16=f16)
    outs = []
    for kwargs in (dict(), dict(halo=2, box=(32, 4))):
        out = cl(torch.full((B, cout, H, W), 7.0, device=DEV, dtype=x.dtype))
        conv_tc.run(pc, x, relu=True, out=out, **kwargs)
        outs.append(out)
    conv_tc.check(DEV)
    want = F.conv2d(x.double(), (w.half() if f16 else conv_tc.round_to_tf32(w)).double(), b.double(), padding=pad).relu()
    if not f16:
        want = F.conv2d(trunc_tf32(x).double(), conv_tc.round_to_tf32(w).double(), b.double(), padding=pad).relu()
    scale = want.abs().max().item()
    tol = 1.5e-3 if f16 else 2e-4          # fp16: rounding of the stored result
    assert (outs[1].double() - want).abs().max().item() / scale <= tol
    assert (outs[0].double() - outs[1].double()).abs().max().item() / scale <= tol


@pytest.mark.parametrize("shape", [(2, 64, 64, 45, 80, 1), (1, 96, 96, 23, 37, 1), (2, 64, 96, 90, 160, 2), (1, 128, 128, 33, 47, 1)])
def test_two_tiles_per_k_step_is_the_same_convolution(shape):
    """M = 256 per CTA (two pixel tiles share every filter tile; debug_flags 2048 forces the mode on a small problem,
    1024 forbids it): every output pixel sees the same MMAs in the same order, so results are bit-identical -- with an odd
    tile count (the second tile of the last pair lies outside the batch), stride 2, the residual tail and the
    instance-norm sums (whose per-CTA rows are dealt differently, hence compared after the combination)."""
    from woft_b200 import conv_tc
    B, cin, cout, H, W, stride = shape
    x, w, b = make(B, cin, cout, H, W, 3, 3, seed=cin + H)
    xh = cl(x.half())
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    res = cl(torch.randn((B, cout, Ho, Wo), generator=torch.Generator().manual_seed(9)).half().to(DEV))
    outs, sums = [], []
    for flag in (1024, 2048):
        out = cl(torch.full((B, cout, Ho, Wo), 3.0, device=DEV, dtype=torch.float16))
        stats = torch.zeros((148, B, 2, cout), device=DEV)
        conv_tc.run(pc, xh, relu=True, out=out, stride=stride, residual=res, stats=stats, debug_flags=flag)
        outs.append(out)
        sums.append(stats.double().sum(0))
    conv_tc.check(DEV)
    assert torch.equal(outs[0], outs[1])
    assert torch.allclose(sums[0], sums[1], rtol=1e-6, atol=1e-3)
    want = (res.double() + F.conv2d(xh.double(), w.half().double(), b.double(), padding=1, stride=stride).relu()).relu()
    assert (outs[1].double() - want).abs().max().item() / want.abs().max().item() <= 1.5e-3


def test_two_tiles_per_k_step_weight_head():
    """The stacked weight-head layers (9 x 14 boxes, HEAD_MEAN epilogue) in the two-tile mode against the one-tile mode."""
    from woft_b200 import _lib, conv_tc
    n, P, C = 61, 9, 128
    g = torch.Generator().manual_seed(2)
    stacked = torch.zeros((n, P + 1, P, C), dtype=torch.float16)
    stacked[:, :P] = torch.randn((n, P, P, C), generator=g).relu().half()
    x = stacked.to(DEV).view(1, n * (P + 1), P, C).permute(0, 3, 1, 2)
    w = (torch.randn((C, C, 3, 3), generator=g) / (C * 9) ** 0.5).to(DEV)
    b = (torch.randn((C,), generator=g) * 0.1).to(DEV)
    w1 = (torch.randn((C,), generator=g) / C ** 0.5).to(DEV)
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    mids, logits = [], []
    for flag in (1024, 2048):
        mid = torch.full((1, n * (P + 1), P, C), 5.0, device=DEV, dtype=torch.float16).permute(0, 3, 1, 2)
        conv_tc.run(pc, x, relu=True, out=mid, row_period=P + 1, row_valid=P, box=(9, 14), debug_flags=flag)
        lg = torch.zeros((n,), device=DEV)
        conv_tc.run(pc, x, epilogue=_lib.CONV_HEAD_MEAN, w1x1=w1, head_out=lg, head_bias=0.2, row_period=P + 1,
                    row_valid=P, box=(9, 14), debug_flags=flag)
        mids.append(mid)
        logits.append(lg)
    conv_tc.check(DEV)
    assert torch.equal(mids[0], mids[1])
    assert torch.equal(logits[0], logits[1])
