"""The flow network around the correlation operator: encoders, GRU update block, weight head.

These stages stay ordinary PyTorch/cuDNN layers; the three correlation
stages between them (volume, pyramid, lookup) are the hand-written CUDA kernels behind `corr_block`.
Parameter names and shapes follow the reference so its checkpoints load unchanged:

  WeightedRAFT      pytracking/external/RAFT/raft_core/weighted_raft.py:29-315
  encoders          pytracking/external/RAFT/raft_core/extractor.py:6-56, 118-192
  update block      pytracking/external/RAFT/raft_core/update.py:6-14, 33-60, 79-97, 114-136
  weight head       weighted_raft.py:318-384

Inference only (`test_mode=True` semantics).  Two things differ from the reference's schedule, neither
changes a result: the upsampling-mask head runs on the last GRU iteration only (the reference evaluates
it every iteration and discards all but the last, weighted_raft.py:228-255), and the weight head's mean
response comes from `corr_block.mean_response()` when the operator offers one instead of re-reading the
whole level-0 volume (weighted_raft.py:358-361).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


def _norm(kind, channels):
    if kind == "batch":
        return nn.BatchNorm2d(channels)
    if kind == "instance":
        return nn.InstanceNorm2d(channels)
    if kind == "none":
        return nn.Sequential()
    raise ValueError(f"unsupported norm '{kind}'")


class ResidualBlock(nn.Module):
    """extractor.py:6-56: two 3x3 convs with norm+ReLU, 1x1 strided projection when stride != 1."""

    def __init__(self, c_in, c_out, norm_fn, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(c_in, c_out, 3, padding=1, stride=stride)
        self.conv2 = nn.Conv2d(c_out, c_out, 3, padding=1)
        self.norm1 = _norm(norm_fn, c_out)
        self.norm2 = _norm(norm_fn, c_out)
        self.downsample = None
        if stride != 1:
            self.norm3 = _norm(norm_fn, c_out)
            self.downsample = nn.Sequential(nn.Conv2d(c_in, c_out, 1, stride=stride), self.norm3)

    def forward(self, x):
        y = F.relu(self.norm1(self.conv1(x)))
        y = F.relu(self.norm2(self.conv2(y)))
        if self.downsample is not None:
            x = self.downsample(x)
        return F.relu(x + y)


class BasicEncoder(nn.Module):
    """extractor.py:118-192: 7x7/2 stem, residual stages 64/96/128 at strides 1/2/2, 1x1 output conv."""

    def __init__(self, output_dim, norm_fn):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3)
        self.norm1 = _norm(norm_fn, 64)
        stages, c_in = [], 64
        for c_out, stride in ((64, 1), (96, 2), (128, 2)):
            stages.append(nn.Sequential(ResidualBlock(c_in, c_out, norm_fn, stride),
                                        ResidualBlock(c_out, c_out, norm_fn, 1)))
            c_in = c_out
        self.layer1, self.layer2, self.layer3 = stages
        self.conv2 = nn.Conv2d(128, output_dim, 1)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = F.relu(self.norm1(self.conv1(x)))
        x = self.layer3(self.layer2(self.layer1(x)))
        return self.conv2(x)


class FlowHead(nn.Module):
    """update.py:6-14."""

    def __init__(self, c_in, c_hidden):
        super().__init__()
        self.conv1 = nn.Conv2d(c_in, c_hidden, 3, padding=1)
        self.conv2 = nn.Conv2d(c_hidden, 2, 3, padding=1)

    def forward(self, x):
        return self.conv2(F.relu(self.conv1(x)))


class SepConvGRU(nn.Module):
    """update.py:33-60: a 1x5 GRU step followed by a 5x1 GRU step."""

    def __init__(self, hidden, c_in):
        super().__init__()
        for tag, k, p in (("1", (1, 5), (0, 2)), ("2", (5, 1), (2, 0))):
            for gate in "zrq":
                setattr(self, f"conv{gate}{tag}", nn.Conv2d(hidden + c_in, hidden, k, padding=p))

    def _step(self, h, x, tag):
        hx = torch.cat([h, x], dim=1)
        z = torch.sigmoid(getattr(self, "convz" + tag)(hx))
        r = torch.sigmoid(getattr(self, "convr" + tag)(hx))
        q = torch.tanh(getattr(self, "convq" + tag)(torch.cat([r * h, x], dim=1)))
        return (1 - z) * h + z * q

    def forward(self, h, x):
        return self._step(self._step(h, x, "1"), x, "2")


class BasicMotionEncoder(nn.Module):
    """update.py:79-97: (lookup 324ch, flow 2ch) -> 126ch, flow appended -> 128ch."""

    def __init__(self, corr_levels, corr_radius):
        super().__init__()
        planes = corr_levels * (2 * corr_radius + 1) ** 2
        self.convc1 = nn.Conv2d(planes, 256, 1)
        self.convc2 = nn.Conv2d(256, 192, 3, padding=1)
        self.convf1 = nn.Conv2d(2, 128, 7, padding=3)
        self.convf2 = nn.Conv2d(128, 64, 3, padding=1)
        self.conv = nn.Conv2d(64 + 192, 128 - 2, 3, padding=1)

    def forward(self, flow, corr):
        cor = F.relu(self.convc2(F.relu(self.convc1(corr))))
        flo = F.relu(self.convf2(F.relu(self.convf1(flow))))
        out = F.relu(self.conv(torch.cat([cor, flo], dim=1)))
        return torch.cat([out, flow], dim=1)


class BasicUpdateBlock(nn.Module):
    """update.py:114-136."""

    def __init__(self, corr_levels, corr_radius, hidden=128):
        super().__init__()
        self.encoder = BasicMotionEncoder(corr_levels, corr_radius)
        self.gru = SepConvGRU(hidden, 128 + hidden)
        self.flow_head = FlowHead(hidden, 256)
        self.mask = nn.Sequential(nn.Conv2d(128, 256, 3, padding=1), nn.ReLU(inplace=True),
                                  nn.Conv2d(256, 64 * 9, 1))

    def forward(self, net, inp, corr, flow, need_mask=True):
        x = torch.cat([inp, self.encoder(flow, corr)], dim=1)
        net = self.gru(net, x)
        mask = 0.25 * self.mask(net) if need_mask else None
        return net, mask, self.flow_head(net)


class WeightHead(nn.Module):
    """weighted_raft.py:318-384.  Per query pixel a (levels+1)-channel 9x9 image -> CNN -> mean -> logit."""

    def __init__(self, structure, n_levels=4):
        super().__init__()
        layers, c = [], n_levels + 1
        for item in structure:
            c_out, k = item if isinstance(item, (list, tuple)) else (item, 3)
            assert k % 2 == 1
            layers += [nn.Conv2d(c, c_out, k, padding=k // 2), nn.ReLU()]
            c = c_out
        layers.append(nn.Conv2d(c, 1, 1))
        self.net = nn.Sequential(*layers)

    def forward(self, window, mean_response):
        """window: (B, P*P*L, H1, W1) lookup output; mean_response: (B, H1, W1).

        The reference re-reads the lookup channels as (H_patch, W_patch, N_levels)
        (weighted_raft.py:267-272) although they are laid out (level, x, y): CNN input [lvl', hp, wp] is
        lookup channel hp*P*L + wp*L + lvl'.  Trained weights bake this in, so it is reproduced as is."""
        B, ch, H1, W1 = window.shape
        L = self.net[0].in_channels - 1
        P = int(round((ch // L) ** 0.5))
        x = window.view(B, P, P, L, H1, W1).permute(0, 4, 5, 3, 1, 2).reshape(B * H1 * W1, L, P, P)
        m = mean_response.reshape(B * H1 * W1, 1, 1, 1).expand(-1, 1, P, P)
        dense = self.net(torch.cat((x, m), dim=1))
        return dense.mean(dim=(2, 3)).view(B, H1, W1).unsqueeze(1)


def convex_upsample(field, mask):
    """weighted_raft.py:92-103: (N,C,H,W) -> (N,C,8H,8W); each fine pixel is a softmax-weighted sum of
    the 3x3 coarse neighbourhood of 8*field."""
    N, C, H, W = field.shape
    mask = torch.softmax(mask.view(N, 1, 9, 8, 8, H, W), dim=2)
    taps = F.unfold(8 * field, [3, 3], padding=1).view(N, C, 9, 1, 1, H, W)
    up = torch.sum(mask * taps, dim=2)
    return up.permute(0, 1, 4, 2, 5, 3).reshape(N, C, 8 * H, 8 * W)


def pixel_grid(batch, ht, wd, device):
    """raft_core/utils/utils.py:76-79: (batch, 2, ht, wd) with channel 0 = x, channel 1 = y."""
    ys, xs = torch.meshgrid(torch.arange(ht, device=device), torch.arange(wd, device=device), indexing="ij")
    return torch.stack([xs, ys], dim=0).float()[None].repeat(batch, 1, 1, 1)


class WeightedRAFT(nn.Module):
    """weighted_raft.py:29-315 (large model).  With `with_weights=False` this is the stock RAFT of
    raft_core/raft.py:169-262 (`raft_type='orig'`): same trunk, no weight head."""

    corr_levels = 4
    corr_radius = 4
    hidden_dim = 128
    context_dim = 128

    def __init__(self, weight_head_structure=((128, 3), (128, 3), (128, 3)), with_weights=True):
        super().__init__()
        self.fnet = BasicEncoder(256, "instance")
        self.cnet = BasicEncoder(self.hidden_dim + self.context_dim, "batch")
        self.update_block = BasicUpdateBlock(self.corr_levels, self.corr_radius, self.hidden_dim)
        self.with_weights = with_weights
        # class_params.mixed_precision (optical_flow/raft.py:35): fp16 autocast around the two encoders and the update
        # block exactly where the reference has it (weighted_raft.py:204, 215, 233); the correlation, the coordinate
        # arithmetic, the upsampling and the weight head stay fp32.  False in every shipped config.
        self.mixed_precision = False
        if with_weights:
            self.weight_head = WeightHead(weight_head_structure, self.corr_levels)

    def _autocast(self, device):
        return torch.autocast(device_type=device.type, dtype=torch.float16,
                              enabled=self.mixed_precision and device.type == "cuda")

    def encode(self, image1, image2):
        """weighted_raft.py:194-219: scale to [-1,1], feature maps of both frames, context of frame 1."""
        image1 = 2 * (image1 / 255.0) - 1.0   # memory format (NCHW or channels-last) is kept as given
        image2 = 2 * (image2 / 255.0) - 1.0
        with self._autocast(image1.device):
            fmaps = self.fnet(torch.cat([image1, image2], dim=0))
        fmap1, fmap2 = fmaps.float().chunk(2, dim=0)
        with self._autocast(image1.device):
            net, inp = torch.split(self.cnet(image1), [self.hidden_dim, self.context_dim], dim=1)
            net, inp = torch.tanh(net), torch.relu(inp)
        return fmap1.contiguous(), fmap2.contiguous(), net, inp

    def forward(self, image1, image2, iters=12, flow_init=None, test_mode=True, corr_block=None):
        if not test_mode:
            raise NotImplementedError("inference only: the training path of the reference is out of scope")
        if corr_block is None:
            from .corr import CorrBlock as corr_block
        fmap1, fmap2, net, inp = self.encode(image1, image2)
        corr_fn = corr_block(fmap1, fmap2, num_levels=self.corr_levels, radius=self.corr_radius)

        B, _, H, W = image1.shape
        coords0 = pixel_grid(B, H // 8, W // 8, image1.device)
        coords1 = coords0.clone()
        if flow_init is not None:
            coords1 = coords1 + flow_init

        up_mask = None
        for it in range(iters):
            corr = corr_fn(coords1)
            with self._autocast(image1.device):
                net, up_mask, dflow = self.update_block(net, inp, corr, coords1 - coords0,
                                                        need_mask=(it == iters - 1))
            coords1 = coords1 + dflow   # fp32 + fp16 promotes to fp32, as in the reference
        if up_mask is not None:
            up_mask = up_mask.float()
        flow_low = coords1 - coords0
        flow_up = convex_upsample(flow_low, up_mask)
        if not self.with_weights:
            return flow_low, flow_up

        # 13th lookup at the final coordinates feeds the weight head (weighted_raft.py:265-279)
        window = corr_fn(coords1)
        if hasattr(corr_fn, "mean_response"):
            mean_response = corr_fn.mean_response()
        else:
            vol = corr_fn.corr_pyramid[0]
            mean_response = vol.mean(dim=(1, 2, 3)).view(B, H // 8, W // 8)
        weights = self.weight_head(window, mean_response)
        weights_up = convex_upsample(weights, up_mask) / 8   # undo the x8 of flow upsampling (:285-288)
        return flow_low, flow_up, corr_fn.corr_pyramid[0], weights, weights_up


def seeded_state_dict(model, seed=0):
    """Deterministic stand-in for the absent WOFT checkpoints (.MISSING_LARGE_BLOBS:1-3).

    Values depend only on (seed, parameter name, shape), not on construction order or torch's global RNG,
    so the reference model and this one can be given identical weights.  Distributions follow what the
    reference's constructors would draw: encoder convs ~ N(0, 2/fan_out) (extractor.py:143-145), every other
    conv and all biases ~ U(+-1/sqrt(fan_in)) (torch's Conv2d default); BatchNorm layers get weight ~ U(0.8, 1.2),
    bias and running_mean ~ N(0, 0.1), running_var ~ U(0.5, 1.5) as a trained checkpoint would hold some."""
    import hashlib

    ref_sd = model.state_dict()
    out = {}
    for name, ref in ref_sd.items():
        digest = hashlib.sha256(f"{seed}:{name}".encode()).digest()
        g = torch.Generator().manual_seed(int.from_bytes(digest[:8], "little") % (2 ** 63))
        is_norm = "norm" in name or "downsample.1" in name
        if name.endswith("num_batches_tracked"):
            t = torch.zeros_like(ref)
        elif name.endswith("running_mean"):
            t = torch.randn(ref.shape, generator=g) * 0.1
        elif name.endswith("running_var"):
            t = 0.5 + torch.rand(ref.shape, generator=g)
        elif is_norm:
            # eval-mode BatchNorm of the context encoder (extractor.py:118-192): non-trivial affine parameters and
            # statistics, so that folding them into the convolutions is exercised by every parity test
            t = 0.8 + 0.4 * torch.rand(ref.shape, generator=g) if name.endswith("weight") else \
                torch.randn(ref.shape, generator=g) * 0.1
        elif ref.dim() == 4 and name.startswith(("fnet.", "cnet.")):
            fan_out = ref.shape[0] * ref.shape[2] * ref.shape[3]
            t = torch.randn(ref.shape, generator=g) * (2.0 / fan_out) ** 0.5
        else:
            w = ref if ref.dim() == 4 else ref_sd[name[: -len("bias")] + "weight"]
            bound = 1.0 / (w.shape[1] * w.shape[2] * w.shape[3]) ** 0.5
            t = (torch.rand(ref.shape, generator=g) * 2 - 1) * bound
        out[name] = t.to(ref.dtype)
    return out
