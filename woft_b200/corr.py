"""Seam A': the correlation operator inside the flow network.

Drop-in for `raft_core.corr.CorrBlock` / `AlternateCorrBlock` and for the `alt_cuda_corr` extension module
(pytracking/external/RAFT/raft_core/corr.py:12-100, alt_cuda_corr/correlation.cpp:23-54): same constructor,
same call, same `corr_pyramid` attribute, same output layout.  The arithmetic runs in libwoft_b200.so:

  CorrBlock.__init__   -> woft_corr_pyramid   (K1 GEMM + K2 pooled pyramid, one pass, every level written once)
  CorrBlock.__call__   -> woft_corr_lookup    (K3, one launch for all levels)
  alt_cuda_corr.forward-> woft_alt_corr_forward

To use it under the reference's own network:  `raft_core.weighted_raft.CorrBlock = woft_b200.corr.CorrBlock`.
"""
import os

import torch
import torch.nn.functional as F

from . import _lib

#: arithmetic of the all-pairs GEMM, all with fp32 accumulation:
#:   "fp16x3" tensor cores, error-compensated fp16 hi/lo split (3 MMAs), fp32-class result -- default: the reference
#:            multiplies in fp32 (corr.py:67), and the split keeps full accuracy whatever the feature magnitudes are
#:   "fp16"   tensor cores, operands rounded once to fp16 (10-bit mantissa like tf32, half the operand bytes, inputs
#:            saturated at +-65504): opt-in, 0.05 px tolerance class
#:   "fp16s"  the same arithmetic, and the pyramid itself is stored as fp16 (every level rounded once from fp32): half
#:            the bytes K1 writes and K3 gathers; `corr_pyramid` then holds fp16 tensors
#:   "tf32"   tensor cores, operands rounded to tf32        "bf16"  tensor cores, 7-bit mantissa
#:   "tf32x3" tensor cores, the same with tf32 hi/lo planes (twice the operand bytes)
#:   "fp32"   CUDA cores, the reference's own precision class
DEFAULT_MODE = os.environ.get("WOFT_CORR_MODE", "fp16x3")


def _as_f32c(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class CorrBlock:
    """corr.py:12-69.  `corr_pyramid[l]` is a real (B*H*W, 1, H_l, W_l) tensor, as the weight head expects: fp32, or
    fp16 in the storage mode "fp16s"."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4, mode=None):
        _lib.require_cuda(fmap1, fmap2)
        if fmap1.shape != fmap2.shape or fmap1.dim() != 4:
            raise AssertionError(f"fmaps must be two (B,C,H,W) tensors of equal shape, got {tuple(fmap1.shape)} "
                                 f"and {tuple(fmap2.shape)}")
        lib = _lib.load()
        self.num_levels, self.radius = num_levels, radius
        self.mode = mode or DEFAULT_MODE
        if self.mode not in _lib.CORR_MODES:
            raise ValueError(f"unknown correlation mode '{self.mode}', expected one of {sorted(_lib.CORR_MODES)}")
        # channels-last feature maps (what the channels-last encoder produces) are packed as they are by the
        # tensor-core modes; anything else is brought to contiguous (B,C,H,W)
        cl = torch.channels_last
        is_cl = fmap1.shape[1] > 1 and all(t.is_contiguous(memory_format=cl) and not t.is_contiguous() for t in (fmap1, fmap2))
        nhwc = self.mode != "fp32" and fmap1.dtype == fmap2.dtype == torch.float32 and is_cl
        # channels-last fp16 maps with C % 64 == 0 ARE the K-major fp16 operands of the single-pass fp16 modes: no packing
        f16_maps = (self.mode in ("fp16", "fp16s") and fmap1.dtype == fmap2.dtype == torch.float16 and is_cl
                    and fmap1.shape[1] % 64 == 0)
        if not (nhwc or f16_maps):
            fmap1, fmap2 = _as_f32c(fmap1), _as_f32c(fmap2)
        B, C, H, W = fmap1.shape
        self.shape = (B, C, H, W)
        dev = fmap1.device
        sizes = [(H >> l, W >> l) for l in range(num_levels)]
        if min(sizes[-1]) < 2:
            # the reference returns NaN for a level of width/height 1 (utils.py:63-64); refuse instead
            raise AssertionError(f"feature map {H}x{W} is too small for {num_levels} levels: "
                                 f"coarsest level would be {sizes[-1][0]}x{sizes[-1][1]}, need at least 2x2")
        with torch.cuda.device(dev):
            self.store_f16 = bool(_lib.CORR_MODES[self.mode] & _lib.CORR_STORE_F16)
            level_dtype = torch.float16 if self.store_f16 else torch.float32
            self.corr_pyramid = [torch.empty((B * H * W, 1, h, w), device=dev, dtype=level_dtype) for h, w in sizes]
            self._row_mean = torch.empty((B, H, W), device=dev, dtype=torch.float32)
            mode_id = (_lib.CORR_MODES[self.mode] | (_lib.CORR_FMAPS_NHWC if nhwc else 0)
                       | (_lib.CORR_FMAPS_F16 if f16_maps else 0))
            nbytes = lib.woft_corr_workspace_bytes(B, C, H, W, mode_id)
            self._workspace = torch.empty((nbytes,), device=dev, dtype=torch.uint8)
            levels, keep = _lib.ptr_array(self.corr_pyramid)
            rc = lib.woft_corr_pyramid(_lib.ptr(fmap1), _lib.ptr(fmap2), B, C, H, W, num_levels, levels,
                                       _lib.ptr(self._row_mean), _lib.ptr(self._workspace), nbytes, mode_id,
                                       _lib.current_stream(dev))
            del keep
        _lib.check(rc, "woft_corr_pyramid")

    def __call__(self, coords, flags=0, gather=None):
        """corr.py:29-59: coords (B,2,H1,W1) -> (B, L*(2r+1)^2, H1, W1) fp32 contiguous.

        `gather` (fused schedule only): dict(taps, bias, coords0, flow_nchw, flow_nhwc, f16_a, f16_b, f16_stride, f16_offset)
        -- the flow head's 9-tap gather and coordinate update run in the same launch first (woft_corr_lookup_gather):
        `coords` is then advanced IN PLACE and must be the contiguous fp32 tensor itself."""
        _lib.require_cuda(coords)
        B, C, H, W = self.shape
        if gather is not None and not (coords.dtype == torch.float32 and coords.is_contiguous()):
            raise AssertionError("the fused gather advances `coords` in place: contiguous fp32 needed")
        coords = _as_f32c(coords)
        if coords.shape[0] != B or coords.shape[1] != 2 or tuple(coords.shape[2:]) != (H, W):
            raise AssertionError(f"coords must be ({B},2,{H},{W}), got {tuple(coords.shape)}")
        P = 2 * self.radius + 1
        dev = coords.device
        # flags bit 3: channels-last output for an NHWC consumer (same logical shape and values); bit 4 (with bit 3):
        # fp16, channels padded to a multiple of 8 (the padding is never written) -- the operand of woft_conv_tc
        fmt = torch.channels_last if flags & 8 else torch.contiguous_format
        ch = self.num_levels * P * P
        if flags & 16:
            out = torch.empty((B, (ch + 7) // 8 * 8, H, W), device=dev, dtype=torch.float16, memory_format=fmt)
        else:
            out = torch.empty((B, ch, H, W), device=dev, dtype=torch.float32, memory_format=fmt)
        with torch.cuda.device(dev):
            levels, keep = _lib.ptr_array(self.corr_pyramid)
            flags |= 32 if self.corr_pyramid[0].dtype == torch.float16 else 0
            if gather is None:
                rc = _lib.load().woft_corr_lookup(levels, self.num_levels, _lib.ptr(coords), B, H, W, H, W, self.radius,
                                                  _lib.ptr(out), flags, _lib.current_stream(dev))
            else:
                g = gather
                rc = _lib.load().woft_corr_lookup_gather(
                    levels, self.num_levels, _lib.ptr(coords), B, H, W, H, W, self.radius, _lib.ptr(out), flags,
                    _lib.ptr(g["taps"]), _lib.ptr(g["bias"]), _lib.ptr(g["coords0"]), _lib.ptr(g.get("flow_nchw")),
                    _lib.ptr(g.get("flow_nhwc")), _lib.ptr(g.get("f16_a")), _lib.ptr(g.get("f16_b")),
                    g.get("f16_stride", 0), g.get("f16_offset", 0), _lib.current_stream(dev))
            del keep
        _lib.check(rc, "woft_corr_lookup")
        return out

    def mean_response(self):
        """(B, H1, W1) mean of level 0 over (H2, W2): what WeightHead.forward reduces from the whole volume
        (weighted_raft.py:358-361), produced without re-reading it."""
        return self._row_mean

    def check(self):
        """Synchronise and raise if the correlation kernel aborted (its pipeline waits are bounded)."""
        dev = self._workspace.device
        with torch.cuda.device(dev):
            rc = _lib.load().woft_corr_check(_lib.ptr(self._workspace), _lib.current_stream(dev))
        _lib.check(rc, "woft_corr_check")

    @staticmethod
    def corr(fmap1, fmap2):
        """corr.py:61-69: the level-0 volume alone, shaped (B,H,W,1,H,W)."""
        B, C, H, W = fmap1.shape
        return CorrBlock(fmap1, fmap2, num_levels=1).corr_pyramid[0].view(B, H, W, 1, H, W)


def alt_corr_forward(fmap1, fmap2, coords, radius):
    """`alt_cuda_corr.forward` (correlation.cpp:23-33): channels-last fmaps, coords (B,N,H1,W1,2) ->
    [corr (B,N,(2r+1)^2,H1,W1)].  Inputs must be CUDA and contiguous, as the reference op demands (:19-21)."""
    for name, t in (("fmap1", fmap1), ("fmap2", fmap2), ("coords", coords)):
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be a CUDA tensor")
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    N = coords.shape[1]
    P = 2 * radius + 1
    dev = fmap1.device
    out = torch.empty((B, N, P * P, H1, W1), device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        rc = _lib.load().woft_alt_corr_forward(_lib.ptr(fmap1), _lib.ptr(fmap2), _lib.ptr(coords), B, H1, W1, H2, W2, C,
                                               N, radius, _lib.ptr(out), _lib.current_stream(dev))
    _lib.check(rc, "woft_alt_corr_forward")
    return [out]


class _AltCudaCorrModule:
    """Stands where the reference does `import alt_cuda_corr` (corr.py:5-9); forward only, backward is training."""

    forward = staticmethod(alt_corr_forward)

    @staticmethod
    def backward(*args, **kwargs):
        raise NotImplementedError("alt_cuda_corr.backward is the training path, out of scope")


alt_cuda_corr = _AltCudaCorrModule()


class AlternateCorrBlock:
    """corr.py:72-100: on-the-fly correlation against average-pooled fmap2, no materialised volume."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels, self.radius = num_levels, radius
        self.pyramid = [(fmap1, fmap2)]
        for _ in range(num_levels):
            fmap1 = F.avg_pool2d(fmap1, 2, stride=2)
            fmap2 = F.avg_pool2d(fmap2, 2, stride=2)
            self.pyramid.append((fmap1, fmap2))

    def __call__(self, coords):
        coords = coords.permute(0, 2, 3, 1)
        B, H, W, _ = coords.shape
        dim = self.pyramid[0][0].shape[1]
        f1 = self.pyramid[0][0].permute(0, 2, 3, 1).contiguous()
        outs = []
        for i in range(self.num_levels):
            f2 = self.pyramid[i][1].permute(0, 2, 3, 1).contiguous()
            c = (coords / 2 ** i).reshape(B, 1, H, W, 2).contiguous()
            corr, = alt_corr_forward(f1, f2, c, self.radius)
            outs.append(corr.squeeze(1))
        corr = torch.stack(outs, dim=1).reshape(B, -1, H, W)
        return corr / torch.sqrt(torch.tensor(dim).float())
