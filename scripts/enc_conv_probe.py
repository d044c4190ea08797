"""Device time of woft_conv_tc on the encoder's 3x3 shapes at 1280x720 (two frames per launch, fp16 operands): plain, halo
and row-halo modes, with and without the instance-norm partial sums, over a few pixel-tile boxes."""
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


g = torch.Generator(device=dev).manual_seed(0)
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
    for box in ((0, 0), (32, 4), (16, 8), (64, 2), (128, 1)):
        if box[0] > Wo:
            continue
        for halo in (0, 1, 2):
            if stride != 1 and halo:
                continue
            for st in (None, stats):
                try:
                    t = timeit(lambda: conv_tc.run(pc, x, relu=True, out=out, box=box, halo=halo, stride=stride, stats=st))
                    conv_tc.check(dev)
                except Exception as e:  # noqa: BLE001
                    print(f"   box {box} halo {halo} stats {st is not None}: {str(e)[:80]}")
                    continue
                print(f"   box {str(box):9s} halo {halo} stats {int(st is not None)}: {t:7.1f} us  {gf / t / 1e3:6.1f} TFLOP/s  {mb / t / 1e3:5.2f} TB/s")
