"""Experiment: does a tcgen05 shared-memory descriptor whose start address lies one 128-byte row inside a 128B-swizzled
tile (written by TMA at a 1 KiB aligned address) address the rows behind it, and does it need the base-offset field?"""
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import conv_tc  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn((1, 64, 1, 128), generator=g, device=dev).half().contiguous(memory_format=torch.channels_last)
w = torch.randn((32, 64, 1, 1), generator=g, device=dev) * 0.1
pc = conv_tc.PackedConv(w, None, 0, f16=True)
want = F.conv2d(x.float(), w.half().float())
for flags in (0, 1, 2):
    out = torch.zeros((1, 32, 1, 128), device=dev).contiguous(memory_format=torch.channels_last)
    conv_tc.run(pc, x, out=out, box=(128, 1), debug_flags=flags)
    torch.cuda.synchronize()
    err = (out - want)[..., :127].abs().max().item()
    print(f"debug_flags {flags}: max error on pixels 0..126 = {err:.4g}")
conv_tc.check(dev)
