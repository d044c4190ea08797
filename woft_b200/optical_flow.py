"""Seam A: the flow wrapper the tracker calls.

Drop-in for `pytracking.optical_flow.raft.RAFTWrapper` (pytracking/optical_flow/raft.py:29-218): constructed as
`RAFTWrapper(flow_config)`, reads the same config keys, and `compute_flow` takes the same two (H,W,3) uint8 BGR
frames and returns the same tensors in `mode='TC'` / `mode='flow'`.  What differs is where the work happens:

  - frames are uploaded as uint8 from pinned staging buffers (2 x 2.8 MB at 1280x720) and converted
    BGR->RGB / HWC->CHW / fp32 on the device; the reference converts on the host and uploads fp32 (2 x 11 MB,
    raft.py:113-120);
  - the network is woft_b200.raft_net.WeightedRAFT, whose correlation stages are the CUDA kernels of
    libwoft_b200.so (seam A');
  - with `use_cuda_graph` (default on) the whole pass for a given frame shape is captured once into a CUDA graph
    and replayed, instead of ~900 eager launches per pass.

Pre-computed flow caches (`src_img_identifier`, raft.py:96-111) are not read: the tracker disables them on its
hot path (YAOF_tracker_single_control.py:101,183) and the flow is simply computed.
"""
import logging
from pathlib import Path

import numpy as np
import torch
import torch.nn.functional as F

from .misc import torch_get_featuremap_coords
from .raft_net import WeightedRAFT, seeded_state_dict

logger = logging.getLogger(__name__)


def _cfg(config, name, default=None):
    """Read a key from the reference's Config (missing keys are falsy Config objects, utils/config.py:9-17), a
    SimpleNamespace or a dict."""
    if isinstance(config, dict):
        val = config.get(name, default)
    else:
        val = getattr(config, name, default)
    if val is None or type(val).__name__ == "Config" and not val:
        return default
    return val


class NoPadder:
    """raft.py:221-232: only accepts sizes that are multiples of 8."""

    def __init__(self, shape):
        H, W = shape[-2:]
        assert H % 8 == 0
        assert W % 8 == 0

    def pad(self, src, dst):
        return src, dst

    def unpad(self, x):
        return x


class CropPadder:
    """raft.py:235-247: crop to a multiple of 8 from the right and bottom; nothing is restored."""

    def __init__(self, shape):
        H, W = shape[-2:]
        self.h, self.w = (H // 8) * 8, (W // 8) * 8

    def pad(self, src, dst):
        return src[:, :, :self.h, :self.w], dst[:, :, :self.h, :self.w]

    def unpad(self, x):
        return x


class MichalPadder:
    """raft.py:250-271: bilinear resize up to a multiple of 8; flow is resized back and rescaled."""

    def __init__(self, shape):
        self.h0, self.w0 = shape[-2:]
        self.h1 = int(np.ceil(float(self.h0) / 8) * 8)
        self.w1 = int(np.ceil(float(self.w0) / 8) * 8)

    def pad(self, src, dst):
        size = (self.h1, self.w1)
        return F.interpolate(src, size=size, mode="bilinear"), F.interpolate(dst, size=size, mode="bilinear")

    def unpad(self, flow):
        assert tuple(flow.shape[2:]) == (self.h1, self.w1)
        f = F.interpolate(flow, size=(self.h0, self.w0), mode="bilinear")
        return torch.cat([f[:, 0:1] * self.w0 / self.w1, f[:, 1:2] * self.h0 / self.h1], dim=1)


class InputPadder:
    """raft_core/utils/utils.py:7-26 ('sintel' centring): replicate-pad to a multiple of 8."""

    def __init__(self, shape):
        self.ht, self.wd = shape[-2:]
        ph = (((self.ht // 8) + 1) * 8 - self.ht) % 8
        pw = (((self.wd // 8) + 1) * 8 - self.wd) % 8
        self._pad = [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2]

    def pad(self, *inputs):
        return [F.pad(x, self._pad, mode="replicate") for x in inputs]

    def unpad(self, x):
        if x is None:
            return None
        ht, wd = x.shape[-2:]
        return x[..., self._pad[2]:ht - self._pad[3], self._pad[0]:wd - self._pad[1]]


_PADDERS = {"RAFT": InputPadder, "Michal": MichalPadder, "nopad": NoPadder, "crop": CropPadder}


def sigmoid(xs):
    """raft.py:274-280."""
    if xs is None:
        return None
    if isinstance(xs, torch.Tensor):
        return torch.sigmoid(xs)
    return 1 / (1 + np.exp(-xs))


class RAFTWrapper:
    def __init__(self, config, device="cuda", use_cuda_graph=True, corr_block=None, channels_last=False,
                 cudnn_benchmark=True, fused=True, tf32=None):
        """`tf32`: may the convolutions round their operands to TF32 (the hand-written tensor-core convolution kernels
        of the fused schedule do)?  None follows torch.backends.cudnn.allow_tf32 at construction; False keeps every
        convolution in fp32 (the reference's precision class)."""
        self.C = config
        self.channels_last = channels_last
        self.fused = fused
        if cudnn_benchmark:
            torch.backends.cudnn.benchmark = True  # fixed shapes per sequence: let cuDNN pick its fastest kernels
        params = _cfg(config, "class_params")
        raft_type = _cfg(config, "raft_type", "weighted")
        if raft_type not in ("orig", "weighted"):
            if raft_type == "weighted_masked":
                raise NotImplementedError("raft_type 'weighted_masked' (mask head) is not used by any shipped WOFT "
                                          "tracker config and is not implemented")
            raise ValueError(f"Unknown RAFT type {raft_type}")
        if _cfg(params, "small", False):
            raise NotImplementedError("the small RAFT variant is not used by any WOFT config and is not implemented")
        if _cfg(params, "alternate_corr", False):
            # the reference itself cannot run WeightedRAFT with alternate_corr (weighted_raft.py:265 needs corr_pyramid)
            raise NotImplementedError("alternate_corr=True is incompatible with the weighted network")
        self.raft_type = raft_type
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("RAFTWrapper runs on a CUDA device only (no CPU path)")
        structure = _cfg(params, "weight_head_structure", [(128, 3), (128, 3), (128, 3)])
        net = WeightedRAFT(structure, with_weights=(raft_type == "weighted"))
        net.load_state_dict(self._load_weights(net), strict=not _cfg(config, "non_strict_loading", False))
        self.net = net.to(self.device).requires_grad_(False).eval()
        if channels_last:
            # cuDNN's sm_100 convolution kernels are NHWC: keep weights and activations channels-last so it does not
            # wrap every convolution in a pair of layout-conversion kernels
            self.net = self.net.to(memory_format=torch.channels_last)
        # class_params.mixed_precision (raft.py:35): the reference's fp16 autocast regions, honoured on the layer-by-layer
        # mirror; the fused schedule computes in fp32 / TF32 only
        self.mixed_precision = bool(_cfg(params, "mixed_precision", False))
        self.net.mixed_precision = self.mixed_precision
        if self.mixed_precision:
            fused = self.fused = False
        # fused inference schedule (fast_net.py) unless the layer-by-layer mirror is asked for
        if fused and not channels_last:
            from .fast_net import FastWeightedRAFT
            self.tf32 = bool(torch.backends.cudnn.allow_tf32 if tf32 is None else tf32)
            self._run = FastWeightedRAFT(self.net, tf32=self.tf32)
            self._u8_input = True
        else:
            self._run = self.net
            self._u8_input = False
        self.iters = int(_cfg(config, "iters", 12))
        self.use_cuda_graph = use_cuda_graph
        self.corr_block = corr_block
        self._graphs = {}
        self._pinned = {}
        self.last_flow_shape = None

    # ---- weights ---------------------------------------------------------------------------------
    def _load_weights(self, net):
        model = _cfg(self.C, "model")
        seed = _cfg(self.C, "seeded_weights")
        if model is None or (seed is not None and not Path(str(model)).exists()):
            if seed is None:
                raise FileNotFoundError("flow config has no `model` checkpoint (the WOFT weights are not shipped with "
                                        "the reference: .MISSING_LARGE_BLOBS); set `seeded_weights = <int>` to run "
                                        "with deterministic random-init weights of the same architecture")
            logger.info("using seeded random-init weights (seed %s)", seed)
            return seeded_state_dict(net, int(seed))
        logger.info(f"Loading weights from: {model}")
        state = torch.load(model, map_location="cpu")
        # the reference wraps the net in DataParallel and prefixes keys with `module.` (raft.py:45-58): undo both
        state = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state.items()}
        if _cfg(self.C, "backbone_model"):
            state = {k: v for k, v in state.items() if not any(s in k for s in ("fnet", "cnet", "update_block"))}
        return state

    # ---- reference API -----------------------------------------------------------------------------
    def postprocess_weights(self, flat_weights, fn):
        """raft.py:73-79."""
        s = self.last_flow_shape
        w = fn(flat_weights.view(s["batch"], 1, s["H"], s["W"]))
        return w.reshape(s["batch"], s["H"] * s["W"])

    def upload_u8(self, img, slot):
        """(H,W,3) uint8 host array -> device uint8 tensor through a pinned staging buffer (asynchronous copy).
        Device tensors pass through."""
        if isinstance(img, torch.Tensor):
            return img.to(self.device, non_blocking=True)
        key = (slot, img.shape)
        entry = self._pinned.get(key)
        if entry is None:
            entry = (torch.empty(img.shape, dtype=torch.uint8).pin_memory(), torch.cuda.Event())
            self._pinned[key] = entry
        buf, done = entry
        done.synchronize()  # the previous upload from this buffer must have left the host
        try:
            buf.copy_(torch.from_numpy(img))   # multi-threaded: 0.02 ms for a 720p frame against 0.29 ms for numpy's copy
        except (TypeError, ValueError):        # negative strides or a dtype torch cannot wrap
            buf.numpy()[...] = img
        t = buf.to(self.device, non_blocking=True)
        done.record(torch.cuda.current_stream(self.device))
        return t

    def _upload(self, img, slot):
        """uint8 BGR (H,W,3) -> float RGB (1,3,H,W) on the device (raft.py:113-120, conversion moved to the device)."""
        return self.upload_u8(img, slot).flip(-1).permute(2, 0, 1).float()[None]

    def _forward(self, src, dst, u8=False):
        """`u8`: src / dst are (H,W,3) uint8 BGR device tensors and the fused schedule converts them itself."""
        kwargs = dict(iters=self.iters, test_mode=True, corr_block=self.corr_block)
        run = self._run.run_u8 if u8 else self._run
        if not self.use_cuda_graph:
            return run(src, dst, **kwargs)
        key = (u8,) + tuple(src.shape)
        entry = self._graphs.get(key)
        if entry is None:
            s_in, d_in = torch.empty_like(src), torch.empty_like(dst)
            s_in.copy_(src)
            d_in.copy_(dst)
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):  # warm-up outside capture: cuDNN autotune, workspace allocation
                for _ in range(2):
                    run(s_in, d_in, **kwargs)
            torch.cuda.current_stream(self.device).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = run(s_in, d_in, **kwargs)
            entry = (graph, s_in, d_in, out)
            self._graphs[key] = entry
        graph, s_in, d_in, out = entry
        s_in.copy_(src)
        d_in.copy_(dst)
        graph.replay()
        # the graph's outputs are static buffers that the next replay overwrites: hand out copies
        return tuple(o.clone() if i in (1, 4) else o for i, o in enumerate(out))

    @torch.no_grad()
    def compute_flow(self, src_img, dst_img, mode="TC", vis=False, src_img_identifier=None, numpy_out=False,
                     do_sigmoid=False):
        """raft.py:81-218.  src_img, dst_img: (H, W, 3) uint8 BGR; mode 'flow' or 'TC'."""
        assert mode in ["flow", "TC"]
        assert src_img.shape == dst_img.shape
        with torch.cuda.device(self.device):
            padding_mode = _cfg(self.C, "padding_mode", "nopad")
            if padding_mode not in _PADDERS:
                raise ValueError(f"invalid padding_mode '{padding_mode}'")
            h_img, w_img = src_img.shape[:2]
            if self._u8_input and padding_mode == "nopad" and h_img % 8 == 0 and w_img % 8 == 0:
                # nothing to pad: the uint8 frames go to the fused schedule as they are (one conversion kernel)
                padder = _PADDERS[padding_mode]((1, 3, h_img, w_img))
                out = self._forward(self.upload_u8(src_img, 0), self.upload_u8(dst_img, 1), u8=True)
            else:
                src = self._upload(src_img, 0)
                dst = self._upload(dst_img, 1)
                padder = _PADDERS[padding_mode](dst.shape)
                src, dst = padder.pad(src, dst)
                fmt = torch.channels_last if self.channels_last else torch.contiguous_format
                out = self._forward(src.contiguous(memory_format=fmt), dst.contiguous(memory_format=fmt))
            if self.raft_type == "orig":
                flow_up, weights_up = out[1], None
            else:
                flow_up, weights_up = out[1], out[4]
            flow = padder.unpad(flow_up)
            weights_up = padder.unpad(weights_up) if weights_up is not None else None
            post = _cfg(self.C, "weights_postprocessing_fn")
            if post:
                weights_up = post(weights_up)
            if do_sigmoid:
                weights_up = sigmoid(weights_up)

            B, _, H, W = flow.shape
            if mode == "flow":
                flow = flow.reshape(B * 2, H, W)
                weights_up = weights_up.reshape(B, H, W) if weights_up is not None else None
                if numpy_out:
                    flow = flow.cpu().numpy()
                    weights_up = weights_up.cpu().numpy() if weights_up is not None else None
                return flow, weights_up

            flow_flat = flow.reshape(B * 2, H * W)
            self.last_flow_shape = {"batch": B, "delta": 2, "H": H, "W": W}
            src_coords = torch_get_featuremap_coords((H, W), device=flow_flat.device)
            dst_coords = src_coords + flow_flat
            weights_flat = weights_up.reshape(B, H * W) if weights_up is not None else None
            if numpy_out:
                src_coords, dst_coords = src_coords.cpu().numpy(), dst_coords.cpu().numpy()
                weights_flat = weights_flat.cpu().numpy() if weights_flat is not None else None
            return src_coords, dst_coords, weights_flat
