"""Inference schedule of the flow network for the GPU: same arithmetic as raft_net.WeightedRAFT, fewer passes.

`raft_net.WeightedRAFT` is the layer-by-layer mirror of the reference (and what the CPU oracle runs).  On the GPU
one frame of that schedule is ~1 500 launches of which two thirds are bias adds, ReLUs, layout conversions,
concatenations and the three-pass instance norm around the cuDNN convolutions (profiles/r1_step_breakdown.txt).
`FastWeightedRAFT` runs the same network from the same parameters with those stages fused:

  feature encoder   conv -> InstanceNorm -> ReLU   : bias-free cuDNN conv (a per-channel constant cancels in the
                    norm) + woft_instance_norm_relu (one kernel, two reads, one write)
  context encoder   conv -> BatchNorm(eval) -> ReLU: BN folded into the conv weights, cuDNN fused conv+bias+ReLU
  motion encoder, flow head, mask head, weight head: cuDNN fused conv+bias+ReLU; the weight head runs channels-last
                    end to end (its 597 MB activations are never transposed)
  SepConvGRU        convz and convr as ONE bias-free conv with stacked weights + woft_gru_gates / woft_gru_update
                    (update.py:45-60: 11 element-wise kernels per half step become 2)

Results agree with the layer-by-layer schedule to fp32 rounding (tests/test_gpu_parity.py).
"""
import os

import torch
import torch.nn.functional as F

from . import _lib, conv_tc
from .corr import CorrBlock as _CorrBlock
from .raft_net import pixel_grid


def _conv_relu(x, w, b, stride=(1, 1), padding=(0, 0)):
    return torch.cudnn_convolution_relu(x, w, b, stride, padding, (1, 1), 1)


def instance_norm_relu(x, relu=True, eps=1e-5):
    """(N, C, H, W) fp32 contiguous -> same shape; woft_instance_norm_relu."""
    x = x.contiguous()
    N, C, H, W = x.shape
    y = torch.empty_like(x)
    with torch.cuda.device(x.device):
        rc = _lib.load().woft_instance_norm_relu(_lib.ptr(x), N * C, H * W, eps, int(relu), _lib.ptr(y),
                                                 _lib.current_stream(x.device))
    _lib.check(rc, "woft_instance_norm_relu")
    return y


def instance_norm_relu_nhwc(x, relu=True, eps=1e-5, residual=None):
    """channels-last (N, C, H, W) tensor -> same; woft_instance_norm_relu_nhwc.  With `residual` (same shape,
    channels-last) the result is relu(residual + relu(norm(x))): woft_instance_norm_relu_add_nhwc."""
    x = _cl(x)
    N, C, H, W = x.shape
    y = torch.empty_like(x)
    lib = _lib.load()
    nbytes = lib.woft_instance_norm_nhwc_scratch_bytes(N, C)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        if residual is None:
            what = "woft_instance_norm_relu_nhwc"
            rc = lib.woft_instance_norm_relu_nhwc(_lib.ptr(x), N, H * W, C, eps, int(relu), _lib.ptr(y),
                                                  _lib.ptr(scratch), nbytes, _lib.current_stream(x.device))
        else:
            what = "woft_instance_norm_relu_add_nhwc"
            if not relu or residual.shape != x.shape:
                raise AssertionError("the fused residual tail needs relu=True and a residual of the same shape")
            residual = _cl(residual)
            rc = lib.woft_instance_norm_relu_add_nhwc(_lib.ptr(x), N, H * W, C, eps, _lib.ptr(residual), _lib.ptr(y),
                                                      _lib.ptr(scratch), nbytes, _lib.current_stream(x.device))
    _lib.check(rc, what)
    return y


def conv_bias_nhwc(x, weight, bias, scale=1.0, **kw):
    """scale * (conv2d(x, weight) + bias) for channels-last operands, the bias through woft_bias_act_nhwc (in place on
    the convolution's output) instead of ATen's broadcast add; anything else goes the ATen way."""
    y = F.conv2d(x, weight, None, **kw)
    if bias is None:
        return y if scale == 1.0 else y * scale
    B, C, H, W = y.shape
    if C % 4 or not y.is_contiguous(memory_format=torch.channels_last):
        y = y + bias.view(1, -1, 1, 1)
        return y if scale == 1.0 else y * scale
    with torch.cuda.device(y.device):
        rc = _lib.load().woft_bias_act_nhwc(_lib.ptr(y), _lib.ptr(bias), B * H * W, C, float(scale), 0,
                                            _lib.current_stream(y.device))
    _lib.check(rc, "woft_bias_act_nhwc")
    return y


def _dense(t):
    return t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last))


def add_relu(a, b):
    """relu(a + b) in one pass for two tensors of the same shape and memory layout; woft_add_relu."""
    if a.shape != b.shape or a.stride() != b.stride() or not _dense(a):
        raise AssertionError("add_relu needs two dense tensors of the same shape and strides")
    y = torch.empty_like(a)
    if y.stride() != a.stride():
        raise AssertionError("add_relu: unexpected output layout")
    with torch.cuda.device(a.device):
        rc = _lib.load().woft_add_relu(_lib.ptr(a), _lib.ptr(b), a.numel(), _lib.ptr(y), _lib.current_stream(a.device))
    _lib.check(rc, "woft_add_relu")
    return y


def gru_gates(zr, bias_zr, h):
    B, C2, H, W = zr.shape
    C = C2 // 2
    z, rh = torch.empty_like(h), torch.empty_like(h)
    with torch.cuda.device(h.device):
        rc = _lib.load().woft_gru_gates(_lib.ptr(zr), _lib.ptr(bias_zr), _lib.ptr(h), B, C, H * W, _lib.ptr(z), _lib.ptr(rh),
                                        _lib.current_stream(h.device))
    _lib.check(rc, "woft_gru_gates")
    return z, rh


def gru_update(q, bias_q, z, h):
    B, C, H, W = h.shape
    out = torch.empty_like(h)
    with torch.cuda.device(h.device):
        rc = _lib.load().woft_gru_update(_lib.ptr(q), _lib.ptr(bias_q), _lib.ptr(z), _lib.ptr(h), B, C, H * W, _lib.ptr(out),
                                         _lib.current_stream(h.device))
    _lib.check(rc, "woft_gru_update")
    return out


def _cl(t):
    return t if t.is_contiguous(memory_format=torch.channels_last) else t.contiguous(memory_format=torch.channels_last)


def _empty_cl(b, c, h, w, like):
    return torch.empty((b, c, h, w), device=like.device, dtype=torch.float32, memory_format=torch.channels_last)


def concat_nhwc(srcs, dst, dst_offset, dst2=None, dst2_offset=0, take=None):
    """Channels-last tensors `srcs` (dense) written side by side into channel slice [dst_offset, ...) of `dst` (and of
    `dst2`); woft_concat_nhwc.  `take[i]` (optional) copies only the first take[i] channels of source i
    (woft_concat_nhwc_strided)."""
    import ctypes
    B, _, H, W = srcs[0].shape
    ptrs, keep = _lib.ptr_array(srcs)
    strides = (ctypes.c_int * len(srcs))(*[t.shape[1] for t in srcs])
    lib = _lib.load()
    args = (B * H * W, _lib.ptr(dst), dst.shape[1], dst_offset, _lib.ptr(dst2),
            dst2.shape[1] if dst2 is not None else 0, dst2_offset, _lib.current_stream(dst.device))
    with torch.cuda.device(dst.device):
        if take is None:
            rc = lib.woft_concat_nhwc(ptrs, strides, len(srcs), *args)
        else:
            chans = (ctypes.c_int * len(srcs))(*take)
            rc = lib.woft_concat_nhwc_strided(ptrs, chans, strides, len(srcs), *args)
    del keep
    _lib.check(rc, "woft_concat_nhwc")


def flow_conv7x7_relu(flow, weight, bias):
    """(B,2,H,W) NCHW -> relu(conv7x7) as a channels-last (B,128,H,W) tensor; woft_flow_conv7x7_relu."""
    flow = flow.contiguous()
    B, _, H, W = flow.shape
    out = _empty_cl(B, weight.shape[0], H, W, flow)
    with torch.cuda.device(flow.device):
        rc = _lib.load().woft_flow_conv7x7_relu(_lib.ptr(flow), _lib.ptr(weight), _lib.ptr(bias), B, H, W, _lib.ptr(out),
                                                _lib.current_stream(flow.device))
    _lib.check(rc, "woft_flow_conv7x7_relu")
    return out


def flow_conv7x7_relu_f16(flow, weight, bias, out=None):
    """The same with an fp16 output tensor (`out` reuses one); woft_flow_conv7x7_relu_f16."""
    flow = flow.contiguous()
    B, _, H, W = flow.shape
    if out is None:
        out = torch.empty((B, weight.shape[0], H, W), device=flow.device, dtype=torch.float16,
                          memory_format=torch.channels_last)
    with torch.cuda.device(flow.device):
        rc = _lib.load().woft_flow_conv7x7_relu_f16(_lib.ptr(flow), _lib.ptr(weight), _lib.ptr(bias), B, H, W, _lib.ptr(out),
                                                    _lib.current_stream(flow.device))
    _lib.check(rc, "woft_flow_conv7x7_relu_f16")
    return out


def cast_nhwc_f16(src, channels, dst, dst_off, dst2=None, dst2_off=0):
    """The first `channels` channels of the channels-last fp32 tensor `src`, rounded to fp16, into channel slice
    [dst_off, ...) of the channels-last fp16 tensor `dst` (and of `dst2`); woft_cast_nhwc_f16."""
    B, _, H, W = src.shape
    with torch.cuda.device(src.device):
        rc = _lib.load().woft_cast_nhwc_f16(_lib.ptr(src), src.shape[1], channels, B * H * W, _lib.ptr(dst), dst.shape[1],
                                            dst_off, _lib.ptr(dst2), dst2.shape[1] if dst2 is not None else 0, dst2_off,
                                            _lib.current_stream(src.device))
    _lib.check(rc, "woft_cast_nhwc_f16")


def flow_out3x3(x_cl, weight, bias):
    """channels-last (B,C,H,W) -> conv3x3 to 2 channels, NCHW (B,2,H,W); woft_flow_out3x3."""
    x_cl = _cl(x_cl)
    B, C, H, W = x_cl.shape
    out = torch.empty((B, 2, H, W), device=x_cl.device, dtype=torch.float32)
    with torch.cuda.device(x_cl.device):
        rc = _lib.load().woft_flow_out3x3(_lib.ptr(x_cl), _lib.ptr(weight), _lib.ptr(bias), B, H, W, C, _lib.ptr(out),
                                          _lib.current_stream(x_cl.device))
    _lib.check(rc, "woft_flow_out3x3")
    return out


def stem_conv7x7(x_cl, weight, bias=None, relu=False):
    """channels-last (B,3,H,W) -> 7x7 stride-2 pad-3 convolution to 64 channels, channels-last; woft_stem_conv7x7
    (TF32 tensor-core implicit GEMM).  `weight` is the (64,3,7,7) tensor in its plain contiguous layout."""
    x_cl = _cl(x_cl)
    B, C, H, W = x_cl.shape
    if C != 3 or tuple(weight.shape) != (64, 3, 7, 7) or not weight.is_contiguous():
        raise AssertionError("stem_conv7x7 takes a 3-channel input and a contiguous (64,3,7,7) weight")
    out = _empty_cl(B, 64, (H - 1) // 2 + 1, (W - 1) // 2 + 1, x_cl)
    with torch.cuda.device(x_cl.device):
        rc = _lib.load().woft_stem_conv7x7(_lib.ptr(x_cl), _lib.ptr(weight), _lib.ptr(bias), B, H, W, int(relu),
                                           _lib.ptr(out), _lib.current_stream(x_cl.device))
    _lib.check(rc, "woft_stem_conv7x7")
    return out


def stem_conv7x7_f16(x_cl, weight, bias=None, relu=False):
    """woft_stem_conv7x7_f16: the same first layer with an fp16 output tensor."""
    x_cl = _cl(x_cl)
    B, C, H, W = x_cl.shape
    out = torch.empty((B, 64, (H - 1) // 2 + 1, (W - 1) // 2 + 1), device=x_cl.device, dtype=torch.float16,
                      memory_format=torch.channels_last)
    with torch.cuda.device(x_cl.device):
        rc = _lib.load().woft_stem_conv7x7_f16(_lib.ptr(x_cl), _lib.ptr(weight), _lib.ptr(bias), B, H, W, int(relu),
                                               _lib.ptr(out), _lib.current_stream(x_cl.device))
    _lib.check(rc, "woft_stem_conv7x7_f16")
    return out


def instance_norm_relu_f16(x, relu=True, eps=1e-5, residual=None):
    """Channels-last fp16 tensor -> same; woft_instance_norm_relu_nhwc_f16 (with `residual`: relu(residual + relu(norm)))."""
    N, C, H, W = x.shape
    y = torch.empty_like(x)
    lib = _lib.load()
    nbytes = lib.woft_instance_norm_nhwc_scratch_bytes(N, C)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.woft_instance_norm_relu_nhwc_f16(_lib.ptr(x), N, H * W, C, eps, int(relu), _lib.ptr(residual), _lib.ptr(y),
                                                  _lib.ptr(scratch), nbytes, _lib.current_stream(x.device))
    _lib.check(rc, "woft_instance_norm_relu_nhwc_f16")
    return y


def _conv_f16(pc, x, stride=1, relu=False, residual=None, out_dtype=torch.float16, stats=None):
    """One woft_conv_tc launch on a channels-last fp16 tensor; returns the new channels-last tensor."""
    B, _, H, W = x.shape
    Ho = (H + 2 * pc.pad_h - pc.kh) // stride + 1
    Wo = (W + 2 * pc.pad_w - pc.kw) // stride + 1
    out = torch.empty((B, pc.cout, Ho, Wo), device=x.device, dtype=out_dtype, memory_format=torch.channels_last)
    # stride-1 filters with more than one row: fetch the activations once per filter column instead of once per tap
    tile = dict(halo=2, box=(32, 4)) if (_ROW_HALO and stride == 1 and pc.kh > 1) else {}
    conv_tc.run(pc, x, relu=relu, out=out, stride=stride, residual=residual, stats=stats, **tile)
    return out


# flow-head gather inside the next correlation lookup (woft_corr_lookup_gather): bit-identical, 12 launches fewer per frame,
# and measured SLOWER (4.82 vs 4.64 ms per frame): the flow branch of the next iteration then starts behind the lookup
# instead of beside it, and its 7x7 kernel finds the SMs taken by convc1's one-CTA-per-SM launch.  Off by default.
_FUSE_GATHER = os.environ.get("WOFT_FUSE_GATHER", "0") == "1"
_F16_FMAPS = os.environ.get("WOFT_F16_FMAPS", "1") == "1"   # fp16 feature maps straight into the fp16 correlation modes


def _corr_default_mode():
    from . import corr
    return corr.DEFAULT_MODE


_CTX_STREAM = os.environ.get("WOFT_CTX_STREAM", "1") == "1"   # context encoder beside the feature encoder
_TWO_STREAMS = os.environ.get("WOFT_TWO_STREAMS", "1") == "1"   # independent branches of the schedule on a second stream
_ROW_HALO = os.environ.get("WOFT_ROW_HALO", "0") == "1"   # measured: no gain on the encoders (205.9 vs 208.7 frames/s)
_STAT_ROWS = 148   # one row of partial sums per CTA of a persistent woft_conv_tc launch (at most one CTA per SM)


def conv_instance_norm_f16(pc, x, stride=1, relu=True, residual=None, eps=1e-5):
    """conv -> InstanceNorm -> ReLU (extractor.py:30-32, 45-56) as convolution + combine + apply: the convolution's epilogue
    gathers the per-channel sums of what it stores, so the norm's own statistics pass over the tensor disappears."""
    B = x.shape[0]
    if B > 2:   # the epilogue keeps sums for two images (one frame pair); larger batches take the three-pass norm
        return instance_norm_relu_f16(_conv_f16(pc, x, stride=stride), relu=relu, eps=eps, residual=residual)
    stats = torch.empty((_STAT_ROWS, B, 2, pc.cout), device=x.device, dtype=torch.float32)
    y = _conv_f16(pc, x, stride=stride, stats=stats)
    N, C, H, W = y.shape
    out = y if C % 8 == 0 else torch.empty_like(y)   # in place: the 16-byte apply kernel reads and writes the same vector
    scratch = torch.empty((N * 2 * C,), device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        rc = _lib.load().woft_instance_norm_from_sums_f16(_lib.ptr(y), _lib.ptr(stats), _STAT_ROWS, N, H * W, C, eps, int(relu),
                                                          _lib.ptr(residual), _lib.ptr(out), _lib.ptr(scratch),
                                                          scratch.numel() * 4, _lib.current_stream(x.device))
    _lib.check(rc, "woft_instance_norm_from_sums_f16")
    return out


def flow_head_step(x_cl, weight, bias, coords=None, coords0=None, flow_nchw=None, flow_nhwc=None, want_delta=False):
    """FlowHead.conv2 on a channels-last (B,C,H,W) tensor; woft_flow_head_step.  With `coords` the refinement step is
    fused: coords += delta in place, flow_nchw / flow_nhwc (preallocated, optional) = coords - coords0.  Returns the
    delta (B,2,H,W) when `want_delta` or when there is nothing else to return."""
    x_cl = _cl(x_cl)
    B, C, H, W = x_cl.shape
    lib = _lib.load()
    dev = x_cl.device
    nbytes = lib.woft_flow_head_scratch_bytes(B, H, W)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    delta = torch.empty((B, 2, H, W), device=dev, dtype=torch.float32) if (want_delta or coords is None) else None
    with torch.cuda.device(dev):
        rc = lib.woft_flow_head_step(_lib.ptr(x_cl), _lib.ptr(weight), _lib.ptr(bias), B, H, W, C, _lib.ptr(scratch), nbytes,
                                     _lib.ptr(delta), _lib.ptr(coords), _lib.ptr(coords0), _lib.ptr(flow_nchw),
                                     _lib.ptr(flow_nhwc), _lib.current_stream(dev))
    _lib.check(rc, "woft_flow_head_step")
    return delta


def convex_upsample_both(flow_low, weights_low, mask, sigmoid=False):
    """Flow (B,2,H,W) and weight logits (B,1,H,W) or None -> (flow_up, weights_up) at 8x; woft_convex_upsample.
    `mask` (B,576,H,W) may be NCHW or channels-last."""
    flow_low = flow_low.contiguous()
    B, _, H, W = flow_low.shape
    nhwc = mask.is_contiguous(memory_format=torch.channels_last) and not mask.is_contiguous()
    if not nhwc:
        mask = mask.contiguous()
    flow_up = torch.empty((B, 2, 8 * H, 8 * W), device=flow_low.device, dtype=torch.float32)
    w_up = None
    if weights_low is not None:
        weights_low = weights_low.contiguous()
        w_up = torch.empty((B, 1, 8 * H, 8 * W), device=flow_low.device, dtype=torch.float32)
    with torch.cuda.device(flow_low.device):
        rc = _lib.load().woft_convex_upsample(_lib.ptr(flow_low), _lib.ptr(weights_low), _lib.ptr(mask), int(nhwc), B, H, W,
                                              int(sigmoid), _lib.ptr(flow_up), _lib.ptr(w_up),
                                              _lib.current_stream(flow_low.device))
    _lib.check(rc, "woft_convex_upsample")
    return flow_up, w_up


class FastWeightedRAFT:
    """Callable with the signature of WeightedRAFT.forward; built from a WeightedRAFT that already sits on the GPU."""

    def __init__(self, net, nhwc_update=True, tf32=None):
        self.net = net
        # may convolution operands be rounded to 10-bit mantissas (TF32 / fp16 with fp32 accumulation)?  None follows
        # torch.backends.cudnn.allow_tf32; False keeps every convolution in fp32 on cuDNN (the reference's class)
        self.tf32 = bool(torch.backends.cudnn.allow_tf32 if tf32 is None else tf32)
        self.with_weights = net.with_weights
        self._side = {}   # device index -> second stream (see _update_own)
        dev = next(net.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("FastWeightedRAFT needs the network on a CUDA device")
        with torch.no_grad():
            self.cnet = self._fold_encoder(net.cnet)
            g = net.update_block.gru
            self.gru = {}
            for tag in ("1", "2"):
                cz, cr, cq = (getattr(g, f"conv{k}{tag}") for k in "zrq")
                self.gru[tag] = dict(w_zr=torch.cat([cz.weight, cr.weight], 0).contiguous(),
                                     b_zr=torch.cat([cz.bias, cr.bias], 0).contiguous(),
                                     w_q=cq.weight.contiguous(), b_q=cq.bias.contiguous(), padding=cz.padding)
            if self.with_weights:
                self.head = [(m.weight.contiguous(memory_format=torch.channels_last), m.bias, m.padding)
                             for m in net.weight_head.net if isinstance(m, torch.nn.Conv2d)]
            # channels-last copies of the update block's weights: with NHWC activations cuDNN's sm_100 kernels run
            # without the NCHW<->NHWC conversion kernels that otherwise wrap every convolution
            self.nhwc = nhwc_update
            self.own_convs = self.own_head = self.own_encoders = False
            # both encoders' 3-channel first layer through woft_stem_conv7x7 instead of cuDNN's sm_80 fallback
            self.own_stem = os.environ.get("WOFT_OWN_STEM", "1") == "1"
            self.own_head_conv1 = os.environ.get("WOFT_OWN_HEAD_CONV1", "1") == "1"
            if nhwc_update:
                fn = net.fnet
                self.fw = {"stem": _cl(fn.conv1.weight), "out": (_cl(fn.conv2.weight), fn.conv2.bias), "blocks": []}
                for layer in (fn.layer1, fn.layer2, fn.layer3):
                    for blk in layer:
                        self.fw["blocks"].append({"c1": _cl(blk.conv1.weight), "c2": _cl(blk.conv2.weight),
                                                  "stride": blk.conv1.stride,
                                                  "down": _cl(blk.downsample[0].weight) if blk.downsample is not None else None})
                c = self.cnet  # folded context-encoder weights, channels-last as well
                c["stem_plain"] = (c["stem"][0].contiguous(), c["stem"][1].contiguous())  # for woft_stem_conv7x7
                c["stem"] = (_cl(c["stem"][0]), c["stem"][1])
                c["out"] = (_cl(c["out"][0]), c["out"][1])
                for d in c["blocks"]:
                    d["c1"] = (_cl(d["c1"][0]), d["c1"][1])
                    d["c2"] = (_cl(d["c2"][0]), d["c2"][1])
                    if d["down"] is not None:
                        d["down"] = (_cl(d["down"][0]), d["down"][1])
                ub, e = net.update_block, net.update_block.encoder
                self.ubw = {name: (_cl(m.weight), m.bias) for name, m in (
                    ("convc1", e.convc1), ("convc2", e.convc2), ("convf1", e.convf1), ("convf2", e.convf2),
                    ("conv", e.conv), ("fh1", ub.flow_head.conv1), ("fh2", ub.flow_head.conv2), ("m0", ub.mask[0]),
                    ("m2", ub.mask[2]))}
                # BasicMotionEncoder.conv has 126 output channels (update.py:87), which cuDNN's sm_100 kernels do not
                # take; two all-zero filters make it 128, and the flow that the reference concatenates behind the 126
                # channels (update.py:97) is written over those two
                wc, bc = e.conv.weight, e.conv.bias
                pad = (-wc.shape[0]) % 8
                self.conv_pad = pad
                self.ubw["conv"] = (_cl(torch.cat([wc, wc.new_zeros((pad,) + tuple(wc.shape[1:]))], 0)),
                                    torch.cat([bc, bc.new_zeros(pad)], 0).contiguous())
                # the update block's and the weight head's wide convolutions on woft_conv_tc (tcgen05 implicit GEMM with
                # the element-wise stage of each layer in its epilogue) instead of cuDNN: TF32 operands, so only when
                # the caller allows TF32-class convolutions
                self.own_convs = (self.tf32 and os.environ.get("WOFT_OWN_CONVS", "1") == "1"
                                  and net.hidden_dim == net.context_dim == 128 and e.conv.weight.shape[0] == 126)
                if self.own_convs:
                    # fp16 operands (the 10-bit mantissa of TF32 at half the bytes; every tensor that feeds a convolution
                    # is bounded: ReLU / tanh / sigmoid outputs, correlation values, flow in pixels), fp32 accumulation;
                    # the recurrent state h and the gate z stay fp32
                    def P(w, b, pad):
                        return conv_tc.PackedConv(w, b, pad, f16=True)
                    self.tc = {name: P(m.weight, m.bias, m.padding) for name, m in (
                        ("convc1", e.convc1), ("convc2", e.convc2), ("convf2", e.convf2), ("conv", e.conv),
                        ("fh1", ub.flow_head.conv1), ("m0", ub.mask[0]), ("m2", ub.mask[2]))}
                    # FlowHead.conv2 (2, 256, 3, 3) as (256, 24): [taps 0-4 x (dx, dy) | 2 pad | taps 5-8 x (dx, dy) | 4 pad],
                    # the operand of the flow-head epilogue of woft_conv_tc
                    w2 = ub.flow_head.conv2.weight.detach().float()
                    wt = w2.reshape(2, w2.shape[1], 9).permute(1, 2, 0)            # (c, tap, o)
                    self.fh2 = wt.new_zeros((w2.shape[1], 24))
                    self.fh2[:, 0:10] = wt[:, 0:5].reshape(-1, 10)
                    self.fh2[:, 12:20] = wt[:, 5:9].reshape(-1, 8)
                    self.fh2 = self.fh2.contiguous()
                    self.fh2_bias = ub.flow_head.conv2.bias.detach().float().contiguous()
                    for tag in ("1", "2"):
                        gt = self.gru[tag]
                        self.tc["zr" + tag] = P(gt["w_zr"], gt["b_zr"], gt["padding"])
                        self.tc["q" + tag] = P(gt["w_q"], gt["b_q"], gt["padding"])
                    # both encoders' residual stages on the same kernel (stride-2 convolutions through TMA element
                    # strides, fp16 activations between the layers, the batch-norm encoder's block tails in the epilogue)
                    self.own_encoders = os.environ.get("WOFT_OWN_ENCODERS", "1") == "1" and self.own_stem
                    if self.own_encoders:
                        fn = net.fnet
                        self.enc_f = {"blocks": [], "out": P(fn.conv2.weight, fn.conv2.bias, 0)}
                        for layer in (fn.layer1, fn.layer2, fn.layer3):
                            for blk in layer:
                                self.enc_f["blocks"].append({
                                    "c1": P(blk.conv1.weight, None, 1), "c2": P(blk.conv2.weight, None, 1),
                                    "stride": blk.conv1.stride[0],
                                    "down": P(blk.downsample[0].weight, None, 0) if blk.downsample is not None else None})
                        folded = self._fold_encoder(net.cnet)   # plain (w, b) pairs, batch norm folded in
                        self.enc_c = {"stem": (folded["stem"][0].contiguous(), folded["stem"][1].contiguous()),
                                      "blocks": [], "out": (P(folded["out"][0], None, 0), folded["out"][1])}
                        for d in folded["blocks"]:
                            self.enc_c["blocks"].append({
                                "c1": P(d["c1"][0], d["c1"][1], 1), "c2": P(d["c2"][0], d["c2"][1], 1),
                                "stride": d["stride"][0],
                                "down": P(d["down"][0], d["down"][1], 0) if d["down"] is not None else None})
                    self.own_head = False
                    if self.with_weights and os.environ.get("WOFT_OWN_HEAD", "1") == "1":
                        convs = [m for m in net.weight_head.net if isinstance(m, torch.nn.Conv2d)]
                        if (len(convs) == 4 and tuple(convs[0].weight.shape) == (128, 5, 3, 3)
                                and all(tuple(c.weight.shape) == (128, 128, 3, 3) for c in convs[1:3])
                                and tuple(convs[3].weight.shape) == (1, 128, 1, 1)):
                            # layers 2 and 3 read and write fp16 activations (10-bit mantissa like TF32, half the
                            # 2.6 GB the three layers would otherwise move), fp32 accumulation
                            self.tc["head2"] = P(convs[1].weight, convs[1].bias, 1)
                            self.tc["head3"] = P(convs[2].weight, convs[2].bias, 1)
                            self.head_w1x1 = convs[3].weight.detach().reshape(-1).float().contiguous()
                            self.head_b1x1 = float(convs[3].bias.detach().item())
                            self.own_head = True
                            self.head_halo = os.environ.get("WOFT_HEAD_HALO", "0") == "1"
                for tag in ("1", "2"):
                    gt = self.gru[tag]
                    gt["w_zr"], gt["w_q"] = _cl(gt["w_zr"]), _cl(gt["w_q"])
                    # cuDNN 9 has no sm_100 kernel for the stacked horizontal 1x5 384->256 shape and falls back to an
                    # sm_80 one (83 us + a layout conversion, against 14 us for each 128-filter half): run that
                    # pair as two convolutions (scripts/conv_times.py, profiles/r1_conv_times.txt)
                    gt["split"] = tuple(gt["w_zr"].shape[2:]) == (1, 5)
                    if gt["split"]:
                        cz, cr = (getattr(g, f"conv{k}{tag}") for k in "zr")
                        gt["w_z"], gt["w_r"] = _cl(cz.weight), _cl(cr.weight)
                        gt["b_z"], gt["b_r"] = cz.bias.contiguous(), cr.bias.contiguous()

    # ---- context encoder: fold eval-mode BatchNorm into the convolutions ---------------------------------
    @staticmethod
    def _fold(conv, bn):
        if not isinstance(bn, torch.nn.BatchNorm2d):
            return conv.weight, conv.bias
        s = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        return (conv.weight * s.view(-1, 1, 1, 1)).contiguous(), ((conv.bias - bn.running_mean) * s + bn.bias).contiguous()

    def _fold_encoder(self, enc):
        out = {"stem": self._fold(enc.conv1, enc.norm1), "blocks": [], "out": (enc.conv2.weight, enc.conv2.bias)}
        for layer in (enc.layer1, enc.layer2, enc.layer3):
            for blk in layer:
                d = {"c1": self._fold(blk.conv1, blk.norm1), "c2": self._fold(blk.conv2, blk.norm2),
                     "stride": blk.conv1.stride, "down": None}
                if blk.downsample is not None:
                    d["down"] = self._fold(blk.downsample[0], blk.downsample[1])
                out["blocks"].append(d)
        return out

    def _context(self, x, raw_tail=False):
        c = self.cnet
        if self.nhwc:
            x = _cl(x)
        if self.nhwc and self.own_stem:
            x = stem_conv7x7(x, *c["stem_plain"], relu=True)
        else:
            x = _conv_relu(x, *c["stem"], stride=(2, 2), padding=(3, 3))
        for d in c["blocks"]:
            y = _conv_relu(x, *d["c1"], stride=d["stride"], padding=(1, 1))
            y = _conv_relu(y, *d["c2"], padding=(1, 1))
            if d["down"] is not None:
                x = conv_bias_nhwc(x, d["down"][0], d["down"][1], stride=d["stride"])
            x = add_relu(x, y) if x.stride() == y.stride() and _dense(x) else F.relu(x + y)
        if raw_tail:   # the caller adds the bias while splitting into hidden state and context (woft_context_split_nhwc)
            return F.conv2d(x, c["out"][0], None), c["out"][1]
        return conv_bias_nhwc(x, *c["out"])

    # ---- feature encoder: bias-free conv + fused instance norm / ReLU --------------------------------------
    def _features(self, x):
        if self.nhwc:
            return self._features_nhwc(x)
        enc = self.net.fnet
        x = instance_norm_relu(F.conv2d(x, enc.conv1.weight, None, stride=2, padding=3))
        for layer in (enc.layer1, enc.layer2, enc.layer3):
            for blk in layer:
                y = instance_norm_relu(F.conv2d(x, blk.conv1.weight, None, stride=blk.conv1.stride, padding=1))
                y = instance_norm_relu(F.conv2d(y, blk.conv2.weight, None, padding=1))
                if blk.downsample is not None:
                    x = instance_norm_relu(F.conv2d(x, blk.downsample[0].weight, None, stride=blk.downsample[0].stride),
                                           relu=False)
                x = F.relu(x + y)
        return F.conv2d(x, enc.conv2.weight, enc.conv2.bias)

    def _features_tc(self, x, f16_out=False):
        """Feature encoder (extractor.py:118-192, instance norm) on woft_conv_tc: fp16 activations, bias-free convolutions
        (a per-channel constant cancels in the norm), the residual tails fused into the norm pass."""
        x = instance_norm_relu_f16(stem_conv7x7_f16(x, self.net.fnet.conv1.weight))
        for d in self.enc_f["blocks"]:
            y = conv_instance_norm_f16(d["c1"], x, stride=d["stride"])
            if d["down"] is not None:
                x = conv_instance_norm_f16(d["down"], x, stride=d["stride"], relu=False)
            x = conv_instance_norm_f16(d["c2"], y, residual=x)              # relu(x + relu(norm(conv2(y))))
        return _conv_f16(self.enc_f["out"], x, out_dtype=torch.float16 if f16_out else torch.float32)

    def _context_tc(self, x):
        """Context encoder (batch norm folded into the filters) on woft_conv_tc; returns the last convolution WITHOUT its
        bias (fp32) and the bias, for woft_context_split_nhwc."""
        c = self.enc_c
        x = stem_conv7x7_f16(_cl(x), *c["stem"], relu=True)
        for d in c["blocks"]:
            y = _conv_f16(d["c1"], x, stride=d["stride"], relu=True)
            if d["down"] is not None:
                x = _conv_f16(d["down"], x, stride=d["stride"])
            x = _conv_f16(d["c2"], y, relu=True, residual=x)                 # relu(x + relu(conv2(y) + b))
        return _conv_f16(c["out"][0], x, out_dtype=torch.float32), c["out"][1]

    def _features_nhwc(self, x):
        """Same encoder with channels-last activations and weights (no cuDNN layout conversions)."""
        if self.own_encoders:
            return self._features_tc(x)
        w = self.fw
        x = instance_norm_relu_nhwc(stem_conv7x7(x, self.net.fnet.conv1.weight) if self.own_stem else
                                    F.conv2d(_cl(x), w["stem"], None, stride=2, padding=3))
        for d in w["blocks"]:
            y = instance_norm_relu_nhwc(F.conv2d(x, d["c1"], None, stride=d["stride"], padding=1))
            if d["down"] is not None:
                x = instance_norm_relu_nhwc(F.conv2d(x, d["down"], None, stride=d["stride"]), relu=False)
            x = instance_norm_relu_nhwc(F.conv2d(y, d["c2"], None, padding=1), residual=x)  # relu(x + relu(norm(..)))
        return conv_bias_nhwc(x, w["out"][0], w["out"][1])

    # ---- update block ---------------------------------------------------------------------------------------
    def _gru_half(self, h, x, tag):
        g = self.gru[tag]
        zr = F.conv2d(torch.cat([h, x], dim=1), g["w_zr"], None, padding=g["padding"])
        z, rh = gru_gates(zr, g["b_zr"], h)
        q = F.conv2d(torch.cat([rh, x], dim=1), g["w_q"], None, padding=g["padding"])
        return gru_update(q, g["b_q"], z, h)

    def _update(self, net, inp, corr, flow, need_mask):
        ub = self.net.update_block
        e = ub.encoder
        cor = _conv_relu(corr, e.convc1.weight, e.convc1.bias)
        cor = _conv_relu(cor, e.convc2.weight, e.convc2.bias, padding=(1, 1))
        flo = _conv_relu(flow, e.convf1.weight, e.convf1.bias, padding=(3, 3))
        flo = _conv_relu(flo, e.convf2.weight, e.convf2.bias, padding=(1, 1))
        out = _conv_relu(torch.cat([cor, flo], dim=1), e.conv.weight, e.conv.bias, padding=(1, 1))
        x = torch.cat([inp, out, flow], dim=1)
        net = self._gru_half(self._gru_half(net, x, "1"), x, "2")
        fh = ub.flow_head
        dflow = F.conv2d(_conv_relu(net, fh.conv1.weight, fh.conv1.bias, padding=(1, 1)), fh.conv2.weight, fh.conv2.bias,
                         padding=1)
        mask = None
        if need_mask:
            m = _conv_relu(net, ub.mask[0].weight, ub.mask[0].bias, padding=(1, 1))
            mask = 0.25 * F.conv2d(m, ub.mask[2].weight, ub.mask[2].bias)
        return net, mask, dflow

    # ---- update block with channels-last activations and no torch.cat -----------------------------------------------
    def _nhwc_state(self, ctx_out, ctx_bias, coords0, coords1):
        """State of the channels-last update block from the context encoder's output (B, 2C, H, W), bias not yet added:
        buf1 = [h | x], buf2 = [r*h | x] with x = [inp | motion | flow]; inp is the same every iteration."""
        ctx_out = _cl(ctx_out)
        B, C2, H, W = ctx_out.shape
        C = C2 // 2
        flow = (coords1 - coords0).contiguous()
        st = {"ci": C, "h": _empty_cl(B, C, H, W, ctx_out), "z": _empty_cl(B, C, H, W, ctx_out),
              "buf1": _empty_cl(B, 3 * C, H, W, ctx_out), "buf2": _empty_cl(B, 3 * C, H, W, ctx_out), "C": C,
              "pixels": B * H * W, "coords0": coords0.contiguous(), "coords1": coords1.contiguous().clone(),
              "flow": flow, "flow_cl": _cl(flow).clone()}   # flow_cl: logical (B,2,H,W), stored (B,H,W,2)
        if self.own_convs:
            def half_cl(c):
                return torch.empty((B, c, H, W), device=ctx_out.device, dtype=torch.float16, memory_format=torch.channels_last)
            st["cor1"], st["cf"], st["fh"] = half_cl(256), half_cl(256), _empty_cl(B, 256, H, W, ctx_out)
            # fp16 copies of the two GRU operand buffers [h | inp | motion | flow] and [r*h | inp | motion | flow]
            st["buf1h"], st["buf2h"] = half_cl(3 * C), half_cl(3 * C)
        with torch.cuda.device(ctx_out.device):
            rc = _lib.load().woft_context_split_nhwc(_lib.ptr(ctx_out), _lib.ptr(ctx_bias), B * H * W, C, _lib.ptr(st["h"]),
                                                     _lib.ptr(st["buf1"]), _lib.ptr(st["buf2"]), 3 * C,
                                                     _lib.current_stream(ctx_out.device))
        _lib.check(rc, "woft_context_split_nhwc")
        if self.own_convs:
            cast_nhwc_f16(st["buf1"], 2 * C, st["buf1h"], 0, st["buf2h"], 0)   # [h | inp] into both operand buffers
        return st

    def _gru_half_nhwc(self, st, tag):
        g, C, lib = self.gru[tag], st["C"], _lib.load()
        buf1, buf2, dev = st["buf1"], st["buf2"], st["h"].device
        if g["split"]:
            zp = _cl(F.conv2d(buf1, g["w_z"], None, padding=g["padding"]))
            rp = _cl(F.conv2d(buf1, g["w_r"], None, padding=g["padding"]))
            with torch.cuda.device(dev):
                rc = lib.woft_gru_gates_nhwc_split(_lib.ptr(zp), _lib.ptr(rp), C, _lib.ptr(g["b_z"]), _lib.ptr(g["b_r"]),
                                                   _lib.ptr(buf1), buf1.shape[1], st["pixels"], C, _lib.ptr(st["z"]),
                                                   _lib.ptr(buf2), buf2.shape[1], _lib.current_stream(dev))
            _lib.check(rc, "woft_gru_gates_nhwc_split")
        else:
            zr = _cl(F.conv2d(buf1, g["w_zr"], None, padding=g["padding"]))
            with torch.cuda.device(dev):
                rc = lib.woft_gru_gates_nhwc(_lib.ptr(zr), _lib.ptr(g["b_zr"]), _lib.ptr(buf1), buf1.shape[1],
                                             st["pixels"], C, _lib.ptr(st["z"]), _lib.ptr(buf2), buf2.shape[1],
                                             _lib.current_stream(dev))
            _lib.check(rc, "woft_gru_gates_nhwc")
        q = _cl(F.conv2d(buf2, g["w_q"], None, padding=g["padding"]))
        with torch.cuda.device(dev):
            rc = lib.woft_gru_update_nhwc(_lib.ptr(q), _lib.ptr(g["b_q"]), _lib.ptr(st["z"]), _lib.ptr(buf1), buf1.shape[1],
                                          st["pixels"], C, _lib.ptr(buf1), buf1.shape[1], _lib.ptr(st["h"]),
                                          _lib.current_stream(dev))
        _lib.check(rc, "woft_gru_update_nhwc")

    def _update_nhwc(self, st, corr, need_mask):
        """One refinement iteration on the state of _nhwc_state: st["coords1"] is advanced in place and st["flow"] /
        st["flow_cl"] (coords1 - coords0, NCHW and (pixels, 2)) are rewritten by the flow head's last kernel."""
        w = self.ubw
        e = self.net.update_block.encoder
        flo = flow_conv7x7_relu(st["flow"], e.convf1.weight, e.convf1.bias)   # NCHW flow in, channels-last out
        cor = _conv_relu(corr, *w["convc1"])
        cor = _conv_relu(cor, *w["convc2"], padding=(1, 1))
        flo = _conv_relu(flo, *w["convf2"], padding=(1, 1))
        B, _, H, W = cor.shape
        cf = _empty_cl(B, cor.shape[1] + flo.shape[1], H, W, cor)
        concat_nhwc([_cl(cor), _cl(flo)], cf, 0)
        out = _cl(_conv_relu(cf, *w["conv"], padding=(1, 1)))      # 126 motion channels + `conv_pad` zero channels
        C, ci = st["C"], st["ci"]
        # x = [inp | motion | flow]: inp sits in both GRU buffers since _nhwc_state; the motion features go in without
        # the padding channels of their convolution, the flow behind them (update.py:97)
        concat_nhwc([out, st["flow_cl"]], st["buf1"], C + ci, st["buf2"], C + ci,
                    take=[out.shape[1] - self.conv_pad, st["flow_cl"].shape[1]])
        self._gru_half_nhwc(st, "1")
        self._gru_half_nhwc(st, "2")
        h = st["h"]
        fh = self.net.update_block.flow_head
        flow_head_step(_conv_relu(h, *w["fh1"], padding=(1, 1)), fh.conv2.weight, fh.conv2.bias, coords=st["coords1"],
                       coords0=st["coords0"], flow_nchw=st["flow"], flow_nhwc=st["flow_cl"])
        mask = None
        if need_mask:
            mask = conv_bias_nhwc(_conv_relu(h, *w["m0"], padding=(1, 1)), *w["m2"], scale=0.25)
        return mask

    def _side_stream(self, device):
        """The second stream of the two-branch sections (one per device, created on first use)."""
        key = torch.device(device).index
        if key not in self._side:
            self._side[key] = torch.cuda.Stream(device=device)
        return self._side[key]

    def _update_own(self, st, lookup):
        """The iteration on woft_conv_tc: every wide convolution is one launch that also applies its bias / ReLU / GRU gate
        arithmetic and writes where the next layer reads (no concatenation, no separate element-wise kernels).  `lookup()`
        returns the fp16 channels-last correlation lookup (flags 8 | 16).

        The motion encoder's two branches (update.py:88-95) are independent until `conv`: the flow branch (7x7 + 3x3) runs
        on a second stream beside the correlation branch (lookup + 1x1 + 3x3), which leaves 33 SMs and every prologue and
        epilogue bubble of its one-tile-per-CTA launches free.  Under graph capture the fork and join become graph edges."""
        tc, C, ci = self.tc, st["C"], st["ci"]
        buf1, buf2, cf, h, z = st["buf1h"], st["buf2h"], st["cf"], st["h"], st["z"]
        e = self.net.update_block.encoder
        main = torch.cuda.current_stream(h.device)
        side = self._side_stream(h.device) if _TWO_STREAMS else main
        corr = None
        if st.get("pending_gather"):
            # the previous iteration left its flow-head taps ungathered: this lookup sums them, advances the coordinates
            # and writes the new flow first (woft_corr_lookup_gather), so the flow branch starts behind it
            corr = lookup(self._gather_args(st, True))
            st["pending_gather"] = False
        if side is not main:
            side.wait_stream(main)
        with torch.cuda.stream(side):
            flo = flow_conv7x7_relu_f16(st["flow"], e.convf1.weight, e.convf1.bias, out=st.get("flo"))
            st["flo"] = flo
            conv_tc.run(tc["convf2"], flo, relu=True, out=cf, out_off=tc["convc2"].cout)
        if corr is None:
            corr = lookup(None)
        conv_tc.run(tc["convc1"], corr, relu=True, out=st["cor1"])
        conv_tc.run(tc["convc2"], st["cor1"], relu=True, out=cf)                        # cf = [cor (192) | flo (64)]
        if side is not main:
            main.wait_stream(side)
        # x = [inp | motion (126) | flow (2)] behind h resp. r*h in both GRU buffers (update.py:96-97)
        conv_tc.run(tc["conv"], cf, relu=True, out=buf1, out_off=C + ci, out2=buf2, out2_off=C + ci)
        if "taps" not in st:   # first iteration: the flow head's last kernel writes the flow into both buffers afterwards
            cast_nhwc_f16(st["flow_cl"], 2, buf1, C + ci + tc["conv"].cout, buf2, C + ci + tc["conv"].cout)
            st["taps"] = torch.empty((st["pixels"], 18), device=h.device, dtype=torch.float32)
        for tag in ("1", "2"):
            conv_tc.run(tc["zr" + tag], buf1, epilogue=_lib.CONV_GRU_ZR, h=h, z=z, out=buf2)
            conv_tc.run(tc["q" + tag], buf2, epilogue=_lib.CONV_GRU_Q, h=h, z=z, out=h, out2=buf1)
        # flow head: conv1 + ReLU + the 18 tap products of conv2 in one launch (reads the h slice of buf1), then the gather
        # that sums the shifted products, advances the coordinates and writes the flow where the next iteration reads it
        conv_tc.run(tc["fh1"], buf1, epilogue=_lib.CONV_FLOW_TAPS, w1x1=self.fh2, head_out=st["taps"])
        if _FUSE_GATHER:
            st["pending_gather"] = True   # the next lookup does it (or _flush_gather after the last iteration)
        else:
            self._flush_gather(st, True)

    def _gather_args(self, st, write_operands):
        """Arguments of the flow head's gather for CorrBlock.__call__(gather=...); `write_operands`: also write the fp16
        flow into the two GRU operand buffers (not needed once the iterations are over)."""
        buf1, buf2 = st["buf1h"], st["buf2h"]
        off = st["C"] + st["ci"] + self.tc["conv"].cout
        g = dict(taps=st["taps"], bias=self.fh2_bias, coords0=st["coords0"], flow_nchw=st["flow"], flow_nhwc=st["flow_cl"])
        if write_operands:
            g.update(f16_a=buf1, f16_b=buf2, f16_stride=buf1.shape[1], f16_offset=off)
        return g

    def _flush_gather(self, st, write_operands=False):
        """The gather as its own launch (woft_flow_head_gather)."""
        buf1, buf2, h = st["buf1h"], st["buf2h"], st["h"]
        B, _, H, W = buf1.shape
        off = st["C"] + st["ci"] + self.tc["conv"].cout
        with torch.cuda.device(h.device):
            rc = _lib.load().woft_flow_head_gather(_lib.ptr(st["taps"]), _lib.ptr(self.fh2_bias), B, H, W, None,
                                                   _lib.ptr(st["coords1"]), _lib.ptr(st["coords0"]), _lib.ptr(st["flow"]),
                                                   _lib.ptr(st["flow_cl"]), _lib.ptr(buf1 if write_operands else None),
                                                   _lib.ptr(buf2 if write_operands else None), buf1.shape[1], off,
                                                   _lib.current_stream(h.device))
        _lib.check(rc, "woft_flow_head_gather")
        st["pending_gather"] = False

    def _mask_own(self, st):
        """The upsampling-mask head (update.py:122-125, 135) on the final hidden state."""
        tc, buf1 = self.tc, st["buf1h"]
        B, _, H, W = buf1.shape
        m = torch.empty((B, tc["m0"].cout, H, W), device=buf1.device, dtype=torch.float16, memory_format=torch.channels_last)
        conv_tc.run(tc["m0"], buf1, relu=True, out=m)
        mask = _empty_cl(B, tc["m2"].cout, H, W, st["h"])
        conv_tc.run(tc["m2"], m, relu=False, scale=0.25, out=mask)
        return mask

    def _weights_own(self, window, mean_response):
        """The weight head on woft_weight_head_conv1_stacked_f16 + two woft_conv_tc launches: layer 1 writes the 14 400
        9x9x128 images stacked along H with a zero row between them (fp16), layer 2 is a plain 3x3 convolution over that
        stack, layer 3 ends in the 1x1 convolution and the patch mean (weighted_raft.py:318-339, 380-384)."""
        B, ch, H1, W1 = window.shape
        n, P = B * H1 * W1, 9
        first = self.net.weight_head.net[0]
        dev = window.device
        x1 = torch.empty((1, n * (P + 1), P, 128), device=dev, dtype=torch.float16).permute(0, 3, 1, 2)
        x2 = torch.empty((1, n * (P + 1), P, 128), device=dev, dtype=torch.float16).permute(0, 3, 1, 2)
        mr = mean_response.contiguous()
        with torch.cuda.device(dev):
            rc = _lib.load().woft_weight_head_conv1_stacked_f16(_lib.ptr(window), _lib.ptr(mr), _lib.ptr(first.weight),
                                                                _lib.ptr(first.bias), n, P, 4, 128, _lib.ptr(x1),
                                                                _lib.current_stream(dev))
        _lib.check(rc, "woft_weight_head_conv1_stacked_f16")
        # default: tiles of 14 stacked rows (126 of the 128 GEMM rows), one shifted box per tap.  The halo mode (WOFT_HEAD_HALO=1:
        # tiles of 11 rows, one 11 x 13 activation box per tile and channel block read by all nine taps at their own row
        # offsets) fetches a ninth of the activation bytes yet measured slower, 0.64 ms against 0.50 ms per layer: MMAs whose
        # operand starts off an 8-row boundary of the swizzled tile run about three times longer
        tile = dict(box=(P, 11), halo=True) if self.head_halo else dict(box=(P, 14))
        conv_tc.run(self.tc["head2"], x1, relu=True, out=x2, row_period=P + 1, row_valid=P, **tile)
        logits = torch.empty((n,), device=dev, dtype=torch.float32)
        conv_tc.run(self.tc["head3"], x2, epilogue=_lib.CONV_HEAD_MEAN, w1x1=self.head_w1x1, head_out=logits,
                    head_bias=self.head_b1x1, row_period=P + 1, row_valid=P, **tile)
        return logits.view(B, H1, W1).unsqueeze(1)

    # ---- weight head, channels-last -----------------------------------------------------------------------------
    def _weights(self, window, mean_response):
        B, ch, H1, W1 = window.shape
        L = self.head[0][0].shape[1] - 1
        P = int(round((ch // L) ** 0.5))
        n = B * H1 * W1
        # NHWC buffer (n, P, P, L+1): CNN input [lvl', hp, wp] is lookup channel hp*P*L + wp*L + lvl'
        # (the reference's reinterpretation of the lookup channels, weighted_raft.py:267-272)
        nhwc_window = window.is_contiguous(memory_format=torch.channels_last) and not window.is_contiguous()
        first = self.net.weight_head.net[0]
        rest = self.head[:-1]
        if (self.own_head_conv1 and nhwc_window and (P, L) == (9, 4) and tuple(first.weight.shape) == (128, 5, 3, 3)
                and first.weight.is_contiguous()):
            # first layer and its input assembly in one kernel (woft_weight_head_conv1)
            x = torch.empty((n, 128, P, P), device=window.device, dtype=torch.float32, memory_format=torch.channels_last)
            mr = mean_response.contiguous()
            with torch.cuda.device(window.device):
                rc = _lib.load().woft_weight_head_conv1(_lib.ptr(window), _lib.ptr(mr), _lib.ptr(first.weight),
                                                        _lib.ptr(first.bias), n, P, L, 128, _lib.ptr(x),
                                                        _lib.current_stream(window.device))
            _lib.check(rc, "woft_weight_head_conv1")
            rest = rest[1:]
        else:
            buf = torch.empty((n, P, P, L + 1), device=window.device, dtype=window.dtype)
            if nhwc_window:
                buf[..., :L] = window.permute(0, 2, 3, 1).reshape(n, P, P, L)  # NHWC lookup: already (n, hp, wp, lvl')
            else:
                buf[..., :L] = window.view(B, P, P, L, H1, W1).permute(0, 4, 5, 1, 2, 3).reshape(n, P, P, L)
            buf[..., L] = mean_response.reshape(n, 1, 1)
            x = buf.permute(0, 3, 1, 2)  # logical NCHW, channels-last in memory
        for w, b, pad in rest:
            x = _conv_relu(x, w, b, padding=pad)
        w, b, _ = self.head[-1]
        # mean over the patch of a 1x1 convolution = 1x1 convolution of the patch mean
        pooled = x.mean(dim=(2, 3))
        logit = pooled @ w.reshape(1, -1).T + b
        return logit.view(B, H1, W1).unsqueeze(1)

    @torch.no_grad()
    def __call__(self, image1, image2, iters=12, flow_init=None, test_mode=True, corr_block=None):
        image1 = 2 * (image1 / 255.0) - 1.0
        image2 = 2 * (image2 / 255.0) - 1.0
        return self._run(torch.cat([image1, image2], dim=0), iters, flow_init, corr_block)

    @torch.no_grad()
    def run_u8(self, img1_bgr, img2_bgr, iters=12, flow_init=None, test_mode=True, corr_block=None):
        """The same pass from two (H,W,3) uint8 BGR device tensors: one kernel (woft_images_to_input) forms the scaled
        RGB pair in channels-last memory instead of flips, casts, three scalings per image, a cat and layout copies."""
        _lib.require_cuda(img1_bgr, img2_bgr)
        if img1_bgr.dtype != torch.uint8 or img1_bgr.shape != img2_bgr.shape or img1_bgr.dim() != 3 \
                or img1_bgr.shape[2] != 3:
            raise AssertionError("run_u8 takes two (H,W,3) uint8 tensors of equal shape")
        H, W, _ = img1_bgr.shape
        img1_bgr, img2_bgr = img1_bgr.contiguous(), img2_bgr.contiguous()
        pair = torch.empty((2, 3, H, W), device=img1_bgr.device, dtype=torch.float32, memory_format=torch.channels_last)
        with torch.cuda.device(pair.device):
            rc = _lib.load().woft_images_to_input(_lib.ptr(img1_bgr), _lib.ptr(img2_bgr), H, W, _lib.ptr(pair),
                                                  _lib.current_stream(pair.device))
        _lib.check(rc, "woft_images_to_input")
        return self._run(pair, iters, flow_init, corr_block)

    def _run(self, images, iters, flow_init, corr_block):
        """`images`: (2B,3,H,W), first the B source frames then the B target frames, scaled to [-1, 1]."""
        net_ = self.net
        if corr_block is None:
            from .corr import CorrBlock as corr_block
        image1 = images[:images.shape[0] // 2]
        # the context encoder needs nothing from the feature encoder or the correlation: on the second stream its
        # tensor-core convolutions fill the feature encoder's normalisation passes and vice versa (WOFT_CTX_STREAM)
        ctx_early, ctx_join = None, None
        if _CTX_STREAM and _TWO_STREAMS and self.nhwc and self.own_encoders:
            main, side = torch.cuda.current_stream(images.device), self._side_stream(images.device)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                ctx_early = self._context_tc(image1)
            ctx_join = (main, side)
        own_corr = corr_block is _CorrBlock or getattr(corr_block, "func", None) is _CorrBlock
        nhwc_maps = self.nhwc and os.environ.get("WOFT_NHWC_FMAPS", "1") == "1" and own_corr
        # single-pass fp16 correlation: the encoder's last convolution writes the fp16 maps the GEMM multiplies (no packing
        # pass, WOFT_CORR_FMAPS_F16)
        corr_mode = (getattr(corr_block, "keywords", None) or {}).get("mode") or _corr_default_mode()
        f16_maps = (nhwc_maps and self.own_encoders and _F16_FMAPS and corr_mode in ("fp16", "fp16s")
                    and net_.fnet.conv2.out_channels % 64 == 0)
        fmap1, fmap2 = (self._features_tc(images, f16_out=True) if f16_maps else self._features(images)).chunk(2, dim=0)
        if not nhwc_maps:
            fmap1, fmap2 = fmap1.contiguous(), fmap2.contiguous()   # this library's CorrBlock takes channels-last maps as is
        corr_fn = corr_block(fmap1, fmap2, num_levels=net_.corr_levels, radius=net_.corr_radius)
        nhwc = self.nhwc and isinstance(corr_fn, _CorrBlock) and net_.hidden_dim == net_.context_dim
        B, _, H, W = image1.shape
        coords0 = pixel_grid(B, H // 8, W // 8, image1.device)
        coords1 = coords0.clone()
        if flow_init is not None:
            coords1 = coords1 + flow_init
        up_mask, mask_join = None, None
        lookup_flags = 8 if nhwc else 0
        if ctx_join:
            ctx_join[0].wait_stream(ctx_join[1])
            for t in ctx_early:
                if torch.is_tensor(t):
                    t.record_stream(ctx_join[0])
        if nhwc:
            ctx = ctx_early if ctx_early is not None else (
                self._context_tc(image1) if self.own_encoders else self._context(image1, raw_tail=True))
            state = self._nhwc_state(*ctx, coords0, coords1)
        else:
            net, inp = torch.split(self._context(image1), [net_.hidden_dim, net_.context_dim], dim=1)
            net, inp = torch.tanh(net), torch.relu(inp)
            state = None
        for it in range(iters):
            last = it == iters - 1
            if nhwc and self.own_convs:
                coords1 = state["coords1"]  # advanced in place by the flow head's last kernel
                self._update_own(state, lambda g: corr_fn(coords1, flags=lookup_flags | 16, gather=g))
                if last and _TWO_STREAMS:   # the mask head runs beside the last lookup and the weight head
                    main, side = torch.cuda.current_stream(image1.device), self._side_stream(image1.device)
                    side.wait_stream(main)
                    with torch.cuda.stream(side):
                        up_mask = self._mask_own(state)
                    up_mask.record_stream(main)
                    mask_join = (main, side)
                elif last:
                    up_mask = self._mask_own(state)
            elif nhwc:
                coords1 = state["coords1"]
                corr = corr_fn(coords1, flags=lookup_flags)
                up_mask = self._update_nhwc(state, corr, need_mask=last)
            else:
                corr = corr_fn(coords1)
                net, up_mask, dflow = self._update(net, inp, corr, coords1 - coords0, need_mask=last)
                coords1 = coords1 + dflow
        pending = bool(state is not None and state.get("pending_gather"))
        if pending and not self.with_weights:
            self._flush_gather(state)
        if not self.with_weights:
            flow_low = coords1 - coords0
            if mask_join:
                mask_join[0].wait_stream(mask_join[1])
            return flow_low, convex_upsample_both(flow_low, None, up_mask)[0]
        if pending:   # the weight head's window lookup gathers the last iteration's taps first
            window = corr_fn(coords1, flags=lookup_flags, gather=self._gather_args(state, False))
            state["pending_gather"] = False
        else:
            window = corr_fn(coords1, flags=lookup_flags) if nhwc else corr_fn(coords1)
        flow_low = coords1 - coords0
        if nhwc and self.own_head and tuple(window.shape[1:2]) == (324,):
            weights = self._weights_own(window, corr_fn.mean_response())
        else:
            weights = self._weights(window, corr_fn.mean_response())
        if mask_join:
            mask_join[0].wait_stream(mask_join[1])
        flow_up, weights_up = convex_upsample_both(flow_low, weights, up_mask)  # one kernel, softmax computed once
        return flow_low, flow_up, corr_fn.corr_pyramid[0], weights, weights_up
