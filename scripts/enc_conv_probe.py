"""Device time of woft_conv_tc on the encoder's 3x3 shapes at 1280x720 (two frames per launch, fp16 operands): plain, halo
and row-halo modes, with and without the instance-norm partial sums, over a few pixel-tile boxes."""
import sys
from pathlib import Path

import os

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import conv_tc  # noqa: E402

dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=10, reps=4):
    """(warm, cold): back-to-back launches (inputs mostly L2 resident), and single launches after an L2 flush."""
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
    cold = []
    for _ in range(7):
        flush.zero_()
        torch.cuda._sleep(200_000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        cold.append(a.elapsed_time(b) * 1e3)
    return best, sorted(cold)[3]


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


g = torch.Generator(device=dev).manual_seed(0)
print("WOFT_CONV_PREFETCH =", os.environ.get("WOFT_CONV_PREFETCH", "(default)"))
for name, cin, cout, H, W, stride in (("layer1 64->64 360x640", 64, 64, 360, 640, 1), ("layer2a 64->96 s2", 64, 96, 360, 640, 2),
                                      ("layer2 96->96 180x320", 96, 96, 180, 320, 1), ("layer3a 96->128 s2", 96, 128, 180, 320, 2),
                                      ("layer3 128->128 90x160", 128, 128, 90, 160, 1)):
    x = cl(torch.randn((2, cin, H, W), generator=g, device=dev).half())
    w = torch.randn((cout, cin, 3, 3), generator=g, device=dev) * 0.05
    b = torch.randn((cout,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, 1, f16=True)
    Ho, Wo = H // stride, W // stride
    out = cl(torch.empty((2, pc.cout_pad, Ho, Wo), device=dev, dtype=torch.float16))
    stats = torch.zeros((148, 2, 2, pc.cout_pad), device=dev)
    gf = 2.0 * 2 * Ho * Wo * cin * cout * 9 / 1e9
    mb = (x.numel() + out.numel()) * 2 / 1e6
    print(f"== {name}: {gf:.1f} GFLOP, {mb:.0f} MB in+out")
    for box in ((0, 0), (16, 8), (64, 2)):
        if box[0] > Wo:
            continue
        for halo in (0, 1, 2):
            if stride != 1 and halo:
                continue
            for st in (None, stats):
                try:
                    t, tcold = timeit(lambda: conv_tc.run(pc, x, relu=True, out=out, box=box, halo=halo, stride=stride, stats=st))
                    conv_tc.check(dev)
                except Exception as e:  # noqa: BLE001
                    print(f"   box {box} halo {halo} stats {st is not None}: {str(e)[:80]}")
                    continue
                print(f"   box {str(box):9s} halo {halo} stats {int(st is not None)}: {t:7.1f} us warm  {tcold:7.1f} us cold  {gf / t * 1e3:6.1f} TFLOP/s  {mb / t:5.2f} TB/s")

# weight head layer 2: 14 400 stacked 9x9x128 images
n, P, C = 14400, 9, 128
x = torch.randn((1, n * (P + 1), P, C), generator=g, device=dev).half().permute(0, 3, 1, 2)
mid = torch.empty((1, n * (P + 1), P, C), device=dev, dtype=torch.float16).permute(0, 3, 1, 2)
w = torch.randn((C, C, 3, 3), generator=g, device=dev) * 0.03
b = torch.randn((C,), generator=g, device=dev)
pc = conv_tc.PackedConv(w, b, 1, f16=True)
t, tcold = timeit(lambda: conv_tc.run(pc, x, relu=True, out=mid, row_period=P + 1, row_valid=P, box=(9, 14)), n=3)
conv_tc.check(dev)
gf = 2.0 * n * 81 * C * C * 9 / 1e9
print(f"== weight head layer 2: {gf:.0f} GFLOP  {t:7.1f} us warm  {tcold:7.1f} us cold  {gf / t * 1e3:6.1f} TFLOP/s")
