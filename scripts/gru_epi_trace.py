"""SM-clock stamps inside the GRU blend epilogue (debug_flags 16): CTA 0, first epilogue warp, per chunk."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import _lib, conv_tc  # noqa: E402

dev = "cuda:0"
H, W = 90, 160
g = torch.Generator(device=dev).manual_seed(0)


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


x = cl(torch.randn((1, 384, H, W), generator=g, device=dev).half())
w = torch.randn((128, 384, 1, 5), generator=g, device=dev) * 0.05
b = torch.randn((128,), generator=g, device=dev)
pc = conv_tc.PackedConv(w, b, (0, 2), f16=True)
out = cl(torch.empty((1, 384, H, W), device=dev, dtype=torch.float16))
hbuf = cl(torch.randn((1, 128, H, W), generator=g, device=dev))
z = cl(torch.rand((1, 128, H, W), generator=g, device=dev))
for dbg in (16, 16 | 32):
    tr = torch.zeros(8 * 148, dtype=torch.int64, device=dev)
    for _ in range(3):
        conv_tc.run(pc, x, epilogue=_lib.CONV_GRU_Q, h=hbuf, z=z, out=hbuf, out2=out, trace=tr, debug_flags=dbg)
    torch.cuda.synchronize()
    conv_tc.check(dev)
    t = tr.cpu().numpy().astype("float64")
    t0 = t[4]   # epilogue leader: accumulator ready
    print(f"debug_flags {dbg}: cycles since the accumulator was ready (tile stamps: ready {t[4]-t0:.0f}, barrier {t[5]-t0:.0f}, done {t[6]-t0:.0f})")
    print("   chunk | before tmem ld | after | transposed | operands landed | math done | stores issued")
    for ci in range(2):
        r = t[64 + 8 * ci: 64 + 8 * ci + 6] - t0
        print(f"   {ci:5d} | {r[0]:8.0f} {r[1]:8.0f} {r[2]:8.0f} {r[3]:8.0f} {r[4]:8.0f} {r[5]:8.0f}")
