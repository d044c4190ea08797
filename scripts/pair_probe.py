"""First contact with the two-CTA mode of woft_conv_tc: same convolution with and without pair=True."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import conv_tc  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
for (cin, cout, kh, kw, H, W, f16) in ((64, 64, 1, 1, 8, 32, True), (128, 256, 3, 3, 90, 160, True), (384, 256, 1, 5, 90, 160, True),
                                      (96, 128, 3, 3, 13, 21, False)):
    x = torch.randn((1, cin, H, W), generator=g, device=dev)
    x = (x.half() if f16 else x).contiguous(memory_format=torch.channels_last)
    w = torch.randn((cout, cin, kh, kw), generator=g, device=dev) * 0.05
    b = torch.randn((cout,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, (kh // 2, kw // 2), f16=f16)
    outs = []
    for pair in (False, True):
        out = torch.zeros((1, cout, H, W), device=dev).contiguous(memory_format=torch.channels_last)
        print('launch pair =', pair, flush=True)
        conv_tc.run(pc, x, relu=True, out=out, pair=pair)
        torch.cuda.synchronize()
        conv_tc.check(dev)
        outs.append(out)
    print(f"cin {cin} cout {cout} {kh}x{kw} {H}x{W} f16={f16}: max |pair - single| = {(outs[0] - outs[1]).abs().max().item():.3g}, "
          f"scale {outs[0].abs().max().item():.3g}", flush=True)
