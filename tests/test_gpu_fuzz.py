"""Randomised shapes for the hand-written kernels of the fused schedule, each against the torch expression it
replaces.  hypothesis draws the extents (odd sizes, single rows and columns, sizes that leave tiles, row segments and
pixel groups partial); the data are seeded normal deviates.  Sixty cases per kernel keep the file to seconds."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

pytestmark = pytest.mark.gpu
DEV = "cuda"
CL = torch.channels_last
FUZZ = settings(max_examples=60, deadline=None, derandomize=True,
                suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(DEV)


@pytest.fixture(autouse=True)
def _fp32_convs():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32 = old


@FUZZ
@given(b=st.integers(1, 3), h=st.integers(1, 40), w=st.integers(1, 70), seed=st.integers(0, 10**6))
def test_flow_conv7x7(b, h, w, seed):
    from woft_b200 import fast_net
    flow, w7, b7 = _rand((b, 2, h, w), seed, 3.0), _rand((128, 2, 7, 7), seed + 1, 0.1), _rand((128,), seed + 2)
    got = fast_net.flow_conv7x7_relu(flow, w7, b7)
    assert torch.allclose(got, F.relu(F.conv2d(flow, w7, b7, padding=3)), atol=3e-5)


@FUZZ
@given(b=st.integers(1, 3), h=st.integers(1, 30), w=st.integers(1, 50), c=st.sampled_from([32, 64, 256, 320]),
       seed=st.integers(0, 10**6))
def test_flow_head_step(b, h, w, c, seed):
    from woft_b200 import fast_net
    x = _rand((b, c, h, w), seed).contiguous(memory_format=CL)
    w3, b3 = _rand((2, c, 3, 3), seed + 1, 0.05), _rand((2,), seed + 2)
    want = F.conv2d(x, w3, b3, padding=1)
    c0 = _rand((b, 2, h, w), seed + 3, 5.0)
    coords = c0 + _rand((b, 2, h, w), seed + 4)
    before = coords.clone()
    f1, f2 = torch.empty_like(c0), torch.empty_like(c0).contiguous(memory_format=CL)
    delta = fast_net.flow_head_step(x, w3, b3, coords=coords, coords0=c0, flow_nchw=f1, flow_nhwc=f2, want_delta=True)
    assert torch.allclose(delta, want, atol=1e-4)
    assert torch.allclose(coords, before + want, atol=2e-4) and torch.equal(f1, coords - c0) and torch.equal(f1, f2)


@FUZZ
@given(b=st.integers(1, 3), h=st.integers(1, 70), w=st.integers(1, 90), bias=st.booleans(), seed=st.integers(0, 10**6))
def test_stem_conv(b, h, w, bias, seed):
    from woft_b200 import fast_net
    x = (torch.rand((b, 3, h, w), generator=torch.Generator().manual_seed(seed)) * 2 - 1).to(DEV).contiguous(memory_format=CL)
    wt, bs = _rand((64, 3, 7, 7), seed + 1, 0.05), _rand((64,), seed + 2)
    want = F.conv2d(x, wt, bs if bias else None, stride=2, padding=3)
    got = fast_net.stem_conv7x7(x, wt, bs if bias else None, relu=bias)
    assert got.shape == want.shape
    assert (got - (F.relu(want) if bias else want)).abs().max().item() <= 2e-3   # TF32 operands, 147 products


@FUZZ
@given(n=st.integers(1, 300), seed=st.integers(0, 10**6))
def test_weight_head_first_layer(n, seed):
    from woft_b200 import _lib
    window, mr = _rand((n, 324), seed), _rand((n,), seed + 1)
    w, b = _rand((128, 5, 3, 3), seed + 2, 0.2), _rand((128,), seed + 3)
    out = torch.empty((n * 81 * 128,), device=DEV)
    rc = _lib.load().woft_weight_head_conv1(_lib.ptr(window), _lib.ptr(mr), _lib.ptr(w), _lib.ptr(b), n, 9, 4, 128,
                                            _lib.ptr(out), _lib.current_stream(torch.device(DEV)))
    assert rc == 0
    x = torch.cat([window.view(n, 9, 9, 4).permute(0, 3, 1, 2), mr.view(n, 1, 1, 1).expand(n, 1, 9, 9)], dim=1)
    want = F.relu(F.conv2d(x, w, b, padding=1)).permute(0, 2, 3, 1).reshape(-1)
    assert (out - want).abs().max().item() <= 1e-2


@FUZZ
@given(b=st.integers(1, 3), h=st.integers(2, 40), w=st.integers(2, 60), c=st.sampled_from([4, 16, 64, 96, 128]),
       relu=st.booleans(), res=st.booleans(), seed=st.integers(0, 10**6))
def test_instance_norm_nhwc(b, h, w, c, relu, res, seed):
    from woft_b200 import fast_net
    x = (_rand((b, c, h, w), seed, 3.0) + 5.0).contiguous(memory_format=CL)
    want = F.instance_norm(x, eps=1e-5)
    if res:
        r = _rand((b, c, h, w), seed + 1).contiguous(memory_format=CL)
        got, want = fast_net.instance_norm_relu_nhwc(x, residual=r), F.relu(r + F.relu(want))
    else:
        got, want = fast_net.instance_norm_relu_nhwc(x, relu=relu), (F.relu(want) if relu else want)
    assert torch.allclose(got, want, atol=5e-5)


@FUZZ
@given(b=st.integers(1, 2), h=st.integers(1, 20), w=st.integers(1, 30),
       chans=st.lists(st.sampled_from([2, 3, 4, 6, 64, 126, 128]), min_size=1, max_size=4),
       off=st.sampled_from([0, 2, 4, 128]), two=st.booleans(), seed=st.integers(0, 10**6))
def test_concat_nhwc(b, h, w, chans, off, two, seed):
    from woft_b200 import fast_net
    srcs = [_rand((b, c + 2, h, w), seed + i).contiguous(memory_format=CL) for i, c in enumerate(chans)]
    total = sum(chans)
    d1 = torch.full((b, off + total + 3, h, w), -7.0, device=DEV).contiguous(memory_format=CL)
    d2 = torch.full((b, off + total + 8, h, w), -7.0, device=DEV).contiguous(memory_format=CL) if two else None
    fast_net.concat_nhwc(srcs, d1, off, d2, off + 4 if two else 0, take=chans)   # leading `c` of c + 2 channels each
    want = torch.cat([s[:, :c] for s, c in zip(srcs, chans)], dim=1)
    assert torch.equal(d1[:, off:off + total], want) and (d1[:, :off] == -7).all() and (d1[:, off + total:] == -7).all()
    if two:
        assert torch.equal(d2[:, off + 4:off + 4 + total], want) and (d2[:, :off + 4] == -7).all()


@FUZZ
@given(b=st.integers(1, 2), h=st.integers(16, 40), w=st.integers(16, 56), nhwc=st.booleans(), seed=st.integers(0, 10**6))
def test_lookup_against_oracle(b, h, w, nhwc, seed):
    from oracle import woft_oracle as wo
    from woft_b200.corr import CorrBlock
    f1, f2 = _rand((b, 8, h, w), seed), _rand((b, 8, h, w), seed + 1)
    blk = CorrBlock(f1, f2, mode="fp32")
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    coords = torch.stack([xs, ys]).float()[None].repeat(b, 1, 1, 1) + _rand((b, 2, h, w), seed + 2, 6.0).cpu()
    got = blk(coords.to(DEV), flags=8 if nhwc else 0)
    levels = [lv.cpu() for lv in blk.corr_pyramid]
    want = wo.corr_lookup(levels, coords, radius=4)
    assert torch.allclose(got.cpu(), want, atol=2e-5)


@FUZZ
@given(b=st.integers(1, 2), h=st.integers(1, 20), w=st.integers(1, 30), cin=st.sampled_from([4, 16]),
       cout=st.sampled_from([4, 8, 96, 576]), scale=st.sampled_from([1.0, 0.25]), nhwc=st.booleans(),
       seed=st.integers(0, 10**6))
def test_conv_bias(b, h, w, cin, cout, scale, nhwc, seed):
    from woft_b200 import fast_net
    x, wt, bs = _rand((b, cin, h, w), seed), _rand((cout, cin, 1, 1), seed + 1, 0.3), _rand((cout,), seed + 2)
    if nhwc:
        x, wt = x.contiguous(memory_format=CL), wt.contiguous(memory_format=CL)
    got = fast_net.conv_bias_nhwc(x, wt, bs, scale=scale)
    assert torch.allclose(got, scale * F.conv2d(x, wt, bs), atol=1e-5)
