"""Phase timeline of woft_conv_tc (debug): per-CTA %globaltimer stamps for the update block's shapes at 1280x720, fp16."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import _lib, conv_tc  # noqa: E402

import os

dev = "cuda:0"
DBG = int(os.environ.get("CONV_DBG", "0"))
H, W = 90, 160
g = torch.Generator(device=dev).manual_seed(0)


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


def report(name, tr, n_cta):
    t = tr.cpu().numpy().reshape(148, 8)[:n_cta].astype("float64")
    t0 = t[:, 0].min()
    rel = (t - t0) / 1e3
    names = ["entry", "setup done", "first operands", "last MMA issued", "accumulator ready", "epilogue done", "exit"]
    print(f"== {name}: {n_cta} CTAs, kernel span {rel[:, 6].max():.1f} us")
    for i, nm in enumerate(names):
        print(f"   {nm:20s} mean {rel[:, i].mean():7.2f}  min {rel[:, i].min():7.2f}  max {rel[:, i].max():7.2f} us")


for name, cin, cout, kh, kw, epi in (("convc1 1x1 324->256", 328, 256, 1, 1, 0), ("convc2 3x3 256->192", 256, 192, 3, 3, 0),
                                     ("fh1 3x3 128->256", 128, 256, 3, 3, 0), ("gru zr 1x5 384->256", 384, 256, 1, 5, 1),
                                     ("gru q 1x5 384->128", 384, 128, 1, 5, 2)):
    pad = (kh // 2, kw // 2)
    x = cl(torch.randn((1, cin, H, W), generator=g, device=dev).half())
    real_cin = 324 if cin == 328 else cin
    w = torch.randn((cout, real_cin, kh, kw), generator=g, device=dev) * 0.05
    b = torch.randn((cout,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, pad, f16=True)
    out = cl(torch.empty((1, 384, H, W), device=dev, dtype=torch.float16))
    hbuf = cl(torch.randn((1, 128, H, W), generator=g, device=dev))
    z = cl(torch.rand((1, 128, H, W), generator=g, device=dev))
    tr = torch.zeros(8 * 148, dtype=torch.int64, device=dev)
    kw_ = dict(trace=tr, debug_flags=DBG)
    if epi == 0:
        kw_.update(relu=True, out=out)
    elif epi == 1:
        kw_.update(epilogue=_lib.CONV_GRU_ZR, h=hbuf, z=z, out=out)
    else:
        kw_.update(epilogue=_lib.CONV_GRU_Q, h=hbuf, z=z, out=hbuf, out2=out)
    for _ in range(3):
        conv_tc.run(pc, x, **kw_)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    conv_tc.run(pc, x, **kw_)
    ev[1].record()
    torch.cuda.synchronize()
    report(f"{name} (event {ev[0].elapsed_time(ev[1]) * 1e3:.1f} us)", tr, 115)
conv_tc.check(dev)
