"""Per-convolution device time of the update block at 1280x720 (channels-last, cuDNN benchmark mode), with the kernel
cuDNN picked for each.  Shows which library calls dominate one GRU iteration.

    python scripts/conv_times.py > gpurun_out/conv_times.txt
"""
import sys
from pathlib import Path

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda", 0)
H, W = 90, 160


def cl(t):
    return t.contiguous(memory_format=torch.channels_last)


def time_conv(name, cin, cout, k, pad, relu=True, n=1, h=H, w=W, stride=(1, 1), dtype=torch.float32, bias=False,
              nchw=False):
    lay = (lambda t: t.contiguous()) if nchw else cl
    x = lay(torch.randn(n, cin, h, w, device=dev).to(dtype))
    wt = lay((torch.randn(cout, cin, *k, device=dev) * 0.05).to(dtype))
    b = torch.randn(cout, device=dev).to(dtype)

    def run():
        if relu:
            return torch.cudnn_convolution_relu(x, wt, b, stride, pad, (1, 1), 1)
        return torch.nn.functional.conv2d(x, wt, b if bias else None, stride=stride, padding=pad)

    for _ in range(5):
        y = run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        run()
        torch.cuda.synchronize()
    kern = sorted(((e.device_time_total, e.key) for e in prof.key_averages() if e.device_time_total > 0), reverse=True)
    ho, wo = y.shape[2], y.shape[3]
    gf = 2.0 * n * ho * wo * cout * cin * k[0] * k[1] / 1e9
    names = " + ".join(f"{k_[:60]}({t:.0f}us)" for t, k_ in kern[:3])
    print(f"{name:28s} {cin:4d}->{cout:4d} {k[0]}x{k[1]} n={n:<5d} {us:8.1f} us {gf:7.1f} GF {gf / us * 1e3 / 1e3:7.1f} TF/s  {names}")


HALF = torch.float16
print("--- variants of the slow 1x5 zr conv")
time_conv("gru zr 1x5 bias", 384, 256, (1, 5), (0, 2), relu=False, bias=True)
time_conv("gru zr 1x5 relu-form", 384, 256, (1, 5), (0, 2), relu=True)
time_conv("gru zr 1x5 nchw", 384, 256, (1, 5), (0, 2), relu=False, nchw=True)
time_conv("gru zr 1x5 cout 384", 384, 384, (1, 5), (0, 2), relu=False)
time_conv("gru zr 1x5 cout 320", 384, 320, (1, 5), (0, 2), relu=False)
time_conv("gru z 1x5 cout 128", 384, 128, (1, 5), (0, 2), relu=False)
time_conv("gru zr 1x7", 384, 256, (1, 7), (0, 3), relu=False)
time_conv("gru zr 3x5", 384, 256, (3, 5), (1, 2), relu=False)
print("--- fp16 operands (fp32 accumulate inside cuDNN)")
time_conv("h convc1 1x1", 324, 256, (1, 1), (0, 0), dtype=HALF)
time_conv("h convc1 1x1 Cin 328", 328, 256, (1, 1), (0, 0), dtype=HALF)
time_conv("h convc2 3x3", 256, 192, (3, 3), (1, 1), dtype=HALF)
time_conv("h convf2 3x3", 128, 64, (3, 3), (1, 1), dtype=HALF)
time_conv("h conv 3x3", 256, 128, (3, 3), (1, 1), dtype=HALF)
time_conv("h gru zr 1x5", 384, 256, (1, 5), (0, 2), relu=False, dtype=HALF)
time_conv("h gru q 1x5", 384, 128, (1, 5), (0, 2), relu=False, dtype=HALF)
time_conv("h gru zr 5x1", 384, 256, (5, 1), (2, 0), relu=False, dtype=HALF)
time_conv("h gru q 5x1", 384, 128, (5, 1), (2, 0), relu=False, dtype=HALF)
time_conv("h flow head 3x3", 128, 256, (3, 3), (1, 1), dtype=HALF)
time_conv("h weight head c1", 5, 128, (3, 3), (1, 1), n=14400, h=9, w=9, dtype=HALF)
time_conv("h weight head c1 cin8", 8, 128, (3, 3), (1, 1), n=14400, h=9, w=9, dtype=HALF)
time_conv("h weight head c2", 128, 128, (3, 3), (1, 1), n=14400, h=9, w=9, dtype=HALF)
time_conv("h fnet stem 7x7 s2", 3, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2), dtype=HALF)
time_conv("h fnet stem cin4", 4, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2), dtype=HALF)
time_conv("h fnet stem cin8", 8, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2), dtype=HALF)
time_conv("h fnet l1 3x3", 64, 64, (3, 3), (1, 1), relu=False, n=2, h=360, w=640, dtype=HALF)
time_conv("h fnet l2 3x3", 96, 96, (3, 3), (1, 1), relu=False, n=2, h=180, w=320, dtype=HALF)
time_conv("h fnet l3 3x3", 128, 128, (3, 3), (1, 1), relu=False, n=2, h=90, w=160, dtype=HALF)
print("--- fp32 stem with padded input channels")
time_conv("fnet stem cin4", 4, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2))
time_conv("fnet stem cin8", 8, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2))
time_conv("fnet stem nchw", 3, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2), nchw=True)
time_conv("weight head c1 cin8", 8, 128, (3, 3), (1, 1), n=14400, h=9, w=9)
print("--- the schedule as it runs today (tf32)")
time_conv("convc1 1x1", 324, 256, (1, 1), (0, 0))
time_conv("convc1 1x1 Cin 328", 328, 256, (1, 1), (0, 0))
time_conv("convc1 1x1 Cin 352", 352, 256, (1, 1), (0, 0))
time_conv("convc2 3x3", 256, 192, (3, 3), (1, 1))
time_conv("convc2 3x3 Cout 256", 256, 256, (3, 3), (1, 1))
time_conv("convf2 3x3", 128, 64, (3, 3), (1, 1))
time_conv("conv 3x3 (126->128)", 256, 128, (3, 3), (1, 1))
time_conv("gru zr 1x5", 384, 256, (1, 5), (0, 2), relu=False)
time_conv("gru q 1x5", 384, 128, (1, 5), (0, 2), relu=False)
time_conv("gru zr 5x1", 384, 256, (5, 1), (2, 0), relu=False)
time_conv("gru q 5x1", 384, 128, (5, 1), (2, 0), relu=False)
time_conv("flow head 3x3", 128, 256, (3, 3), (1, 1))
time_conv("mask head 3x3", 128, 256, (3, 3), (1, 1))
time_conv("mask head 1x1", 256, 576, (1, 1), (0, 0), relu=False)
time_conv("weight head c1", 5, 128, (3, 3), (1, 1), n=14400, h=9, w=9)
time_conv("weight head c2", 128, 128, (3, 3), (1, 1), n=14400, h=9, w=9)
# encoders (fnet batch 2, cnet batch 1)
time_conv("fnet stem 7x7 s2", 3, 64, (7, 7), (3, 3), relu=False, n=2, h=720, w=1280, stride=(2, 2))
time_conv("fnet l1 3x3", 64, 64, (3, 3), (1, 1), relu=False, n=2, h=360, w=640)
time_conv("fnet l2a 3x3 s2", 64, 96, (3, 3), (1, 1), relu=False, n=2, h=360, w=640, stride=(2, 2))
time_conv("fnet l2 3x3", 96, 96, (3, 3), (1, 1), relu=False, n=2, h=180, w=320)
time_conv("fnet l3a 3x3 s2", 96, 128, (3, 3), (1, 1), relu=False, n=2, h=180, w=320, stride=(2, 2))
time_conv("fnet l3 3x3", 128, 128, (3, 3), (1, 1), relu=False, n=2, h=90, w=160)
time_conv("fnet out 1x1", 128, 256, (1, 1), (0, 0), relu=False, n=2, h=90, w=160)
time_conv("cnet stem 7x7 s2", 3, 64, (7, 7), (3, 3), n=1, h=720, w=1280, stride=(2, 2))
time_conv("cnet l1 3x3", 64, 64, (3, 3), (1, 1), n=1, h=360, w=640)
time_conv("cnet l2 3x3", 96, 96, (3, 3), (1, 1), n=1, h=180, w=320)
time_conv("cnet l3 3x3", 128, 128, (3, 3), (1, 1), n=1, h=90, w=160)
