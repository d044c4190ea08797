"""Host side of woft_conv_tc (csrc/conv_tc.cu): the update block's and the weight head's convolutions as tcgen05
implicit GEMMs with fused epilogues.

The reference runs these layers as nn.Conv2d + separate element-wise kernels
(pytracking/external/RAFT/raft_core/update.py:6-14, 33-60, 79-97, 122-125; raft_core/weighted_raft.py:318-339).
Here a layer's filter bank is repacked once (`pack_conv`) to the [tap][Cout][Cin] K-major form the kernel's TMA loads
take, and `run` launches one convolution on a channels-last tensor (or on a channel slice of one) with the bias,
activation, GRU gate arithmetic or patch mean in its epilogue.  No fallback: the call raises if the library refuses.
"""
import ctypes
import os

import torch

from . import _lib


#: two-CTA mode (tcgen05 cta_group::2) unless a call says otherwise; WOFT_CONV_PAIR=0 switches it off
PAIR_DEFAULT = os.environ.get("WOFT_CONV_PAIR", "0") == "1"
#: programmatic dependent launch of the convolution kernels (WOFT_CONV_PDL=0 switches it off)
PDL_DEFAULT = os.environ.get("WOFT_CONV_PDL", "1") == "1"


def round_to_tf32(t):
    """fp32 -> nearest TF32 (10-bit mantissa, ties away from zero: cvt.rna.tf32.f32), still stored as fp32."""
    i = t.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)


class PackedConv:
    """Filters of one stride-1 convolution in the kernel's layout."""

    def __init__(self, weight, bias, padding, f16=False):
        cout, cin, kh, kw = weight.shape
        k_elems = 64 if f16 else 32
        self.cin, self.cout, self.kh, self.kw = cin, cout, kh, kw
        self.pad_h, self.pad_w = (padding, padding) if isinstance(padding, int) else tuple(padding)
        self.cin_pad = (cin + k_elems - 1) // k_elems * k_elems
        self.cout_pad = (cout + 31) // 32 * 32
        self.f16 = f16
        w = weight.detach().float().permute(2, 3, 0, 1).reshape(kh * kw, cout, cin)
        packed = w.new_zeros((kh * kw, self.cout_pad, self.cin_pad))
        packed[:, :cout, :cin] = w
        self.weight = packed.half().contiguous() if f16 else round_to_tf32(packed)
        self.bias = bias.detach().float().contiguous() if bias is not None else None


def _slice(t, ch_off, elem=None):
    """(pointer to channel `ch_off` of pixel 0, pixel stride in elements) of a channels-last (B,C,H,W) tensor, or of a
    plain (pixels, C) matrix."""
    if t is None:
        return ctypes.c_void_p(0), 0
    if t.dim() == 4:
        if not t.is_contiguous(memory_format=torch.channels_last):
            raise AssertionError("conv_tc operands must be channels-last")
        stride = t.shape[1]
    else:
        if not t.is_contiguous():
            raise AssertionError("conv_tc matrix operands must be contiguous")
        stride = t.shape[-1]
    return ctypes.c_void_p(t.data_ptr() + ch_off * t.element_size()), stride


def run(pc, x, x_off=0, epilogue=_lib.CONV_BIAS_ACT, relu=False, scale=1.0, out=None, out_off=0, out2=None, out2_off=0,
        h=None, h_off=0, z=None, w1x1=None, head_out=None, head_bias=0.0, row_period=0, row_valid=0, box=(0, 0),
        dims=None, trace=None, debug_flags=0, halo=False, pair=None, stride=1, residual=None, stats=None):
    """One convolution.  `x` (B,C,H,W) channels-last fp32 or fp16; channels [x_off, x_off + pc.cin) are the input.
    `dims=(B, H, W)` overrides the pixel grid (the weight head stacks its 9x9 images along H)."""
    _lib.require_cuda(x)
    if (x.dtype == torch.float16) != pc.f16:
        raise AssertionError("conv_tc: input type and packed filter type differ")
    B, H, W = dims if dims is not None else (x.shape[0], x.shape[2], x.shape[3])
    a = _lib.ConvTcArgs()
    a.in_, a.in_stride = _slice(x, x_off)
    a.in_f16 = int(pc.f16)
    a.batch, a.height, a.width, a.cin, a.cin_pad = B, H, W, pc.cin, pc.cin_pad
    a.weight_packed = _lib.ptr(pc.weight)
    a.cout, a.cout_pad, a.kh, a.kw, a.pad_h, a.pad_w = pc.cout, pc.cout_pad, pc.kh, pc.kw, pc.pad_h, pc.pad_w
    a.box_w, a.box_h = box
    a.epilogue, a.relu, a.scale = epilogue, int(relu), float(scale)
    a.bias = _lib.ptr(pc.bias)
    a.out, a.out_stride = _slice(out, out_off)
    a.out_f16 = int(out is not None and out.dtype == torch.float16)
    a.out2, a.out2_stride = _slice(out2, out2_off)
    a.out2_f16 = int(out2 is not None and out2.dtype == torch.float16)
    a.h, a.h_stride = _slice(h, h_off)
    a.z, a.z_stride = _slice(z, 0)
    a.w1x1 = _lib.ptr(w1x1)
    a.head_out = _lib.ptr(head_out)
    a.head_bias = float(head_bias)
    a.row_period, a.row_valid = row_period, row_valid
    a.debug_flags = debug_flags
    a.pair = int(PAIR_DEFAULT if pair is None else pair)   # two-CTA clusters sharing the filter tile (cta_group::2)
    a.pdl = int(PDL_DEFAULT)
    a.stride = stride
    a.residual, a.residual_stride = _slice(residual, 0)
    a.stats = _lib.ptr(stats)          # (rows, batch, 2, cout) fp32 partial sums for an instance norm that follows
    a.stats_ctas = stats.shape[0] if stats is not None else 0
    a.halo = int(halo)   # 1: one activation box per tile and channel block for all taps; 2: one per filter column (rows only)
    a.debug_trace = _lib.ptr(trace)   # debug: int64 tensor of 8 * 148 per-CTA time stamps
    with torch.cuda.device(x.device):
        rc = _lib.load().woft_conv_tc(ctypes.byref(a), _lib.current_stream(x.device))
    _lib.check(rc, "woft_conv_tc")


def check(device):
    """Synchronise and raise if a convolution kernel aborted (its pipeline waits are bounded)."""
    with torch.cuda.device(device):
        rc = _lib.load().woft_conv_tc_check(_lib.current_stream(device))
    _lib.check(rc, "woft_conv_tc_check")
