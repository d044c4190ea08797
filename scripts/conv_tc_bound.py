"""Which resource bounds a persistent woft_conv_tc launch: the same launch with the activation loads, the filter loads, the
output stores left out (debug_flags bits 1, 2, 8 -- results are meaningless, only the time is read)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import conv_tc  # noqa: E402

dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=10, reps=4):
    for _ in range(3):
        fn()
    best = 1e9
    for _ in range(reps):
        flush.zero_()
        torch.cuda._sleep(1_000_000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record()
        b.synchronize()
        best = min(best, a.elapsed_time(b) * 1e3 / n)
    return best


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


NAMES = {0: "everything", 1: "no A loads", 2: "no B loads", 3: "no loads", 8: "no stores", 11: "no loads, no stores"}
g = torch.Generator(device=dev).manual_seed(0)
for name, cin, cout, H, W, stride, B in (("layer1 64->64 360x640 x2", 64, 64, 360, 640, 1, 2), ("layer2 96->96 180x320 x2", 96, 96, 180, 320, 1, 2),
                                         ("layer3 128->128 90x160 x2", 128, 128, 90, 160, 1, 2), ("mask 128->256 90x160", 128, 256, 90, 160, 1, 1)):
    x = cl(torch.randn((B, cin, H, W), generator=g, device=dev).half())
    w = torch.randn((cout, cin, 3, 3), generator=g, device=dev) * 0.05
    b = torch.randn((cout,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    out = cl(torch.empty((B, pc.cout_pad, H // stride, W // stride), device=dev, dtype=torch.float16))
    print(f"== {name}")
    for dbg in sorted(NAMES):
        t = timeit(lambda: conv_tc.run(pc, x, relu=True, out=out, stride=stride, debug_flags=dbg))
        conv_tc.check(dev)
        print(f"   {NAMES[dbg]:30s} {t:7.1f} us")
    t = timeit(lambda: conv_tc.run(pc, x, relu=True, out=out, stride=stride, pair=True))
    conv_tc.check(dev)
    print(f"   {'everything, CTA pairs':30s} {t:7.1f} us")

n, P, C = 14400, 9, 128
x = torch.randn((1, n * (P + 1), P, C), generator=g, device=dev).half().permute(0, 3, 1, 2)
mid = torch.empty((1, n * (P + 1), P, C), device=dev, dtype=torch.float16).permute(0, 3, 1, 2)
w = torch.randn((C, C, 3, 3), generator=g, device=dev) * 0.03
b = torch.randn((C,), generator=g, device=dev)
pc = conv_tc.PackedConv(w, b, 1, f16=True)
print("== weight head layer 2 (14400 x 9x9x128)")
for dbg in sorted(NAMES):
    t = timeit(lambda: conv_tc.run(pc, x, relu=True, out=mid, row_period=P + 1, row_valid=P, box=(9, 14), debug_flags=dbg), n=3)
    conv_tc.check(dev)
    print(f"   {NAMES[dbg]:30s} {t:7.1f} us")
t = timeit(lambda: conv_tc.run(pc, x, relu=True, out=mid, row_period=P + 1, row_valid=P, box=(9, 14), pair=True), n=3)
conv_tc.check(dev)
print(f"   {'everything, CTA pairs':30s} {t:7.1f} us")
