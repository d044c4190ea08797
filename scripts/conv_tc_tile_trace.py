"""SM-clock timeline of CTA 0's three roles over its first tiles of a persistent woft_conv_tc launch (debug_flags bit 16)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import conv_tc  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


for name, cin, cout, H, W, B in (("layer1 64->64 360x640 x2", 64, 64, 360, 640, 2), ):
    x = cl(torch.randn((B, cin, H, W), generator=g, device=dev).half())
    w = torch.randn((cout, cin, 3, 3), generator=g, device=dev) * 0.05
    b = torch.randn((cout,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    out = cl(torch.empty((B, pc.cout_pad, H, W), device=dev, dtype=torch.float16))
    for dbg in (0, 11):
        tr = torch.zeros(8 * 148, dtype=torch.int64, device=dev)
        for _ in range(3):
            conv_tc.run(pc, x, relu=True, out=out, trace=tr, debug_flags=16 | dbg)
        torch.cuda.synchronize()
        conv_tc.check(dev)
        t = tr.cpu().numpy().reshape(148, 8)[:18].astype("float64")
        t0 = t[0, 0]
        print(f"== {name}, debug_flags {dbg}: cycles since the producer's first tile began")
        print("   tile | producer start  end | MMA start (accumulator free)  end (all issued) | epilogue: accumulator ready  after barrier  done")
        for i in range(7):
            r = t[i] - t0
            print(f"   {i:4d} | {r[0]:8.0f} {r[1]:8.0f} | {r[2]:8.0f} {r[3]:8.0f} | {r[4]:8.0f} {r[5]:8.0f} {r[6]:8.0f}")
