"""Device time of woft_conv_tc against cuDNN (TF32 allowed, channels-last) for the update block's and the weight head's
convolution shapes at 1280x720.  CUDA events around bursts of 20 launches, L2 flushed before each burst."""
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import _lib, conv_tc  # noqa: E402

dev = "cuda:0"
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=20, reps=5):
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


H, W = 90, 160
g = torch.Generator(device=dev).manual_seed(0)
print(f"{'layer':34s} {'GFLOP':>7s} {'conv_tc us':>10s} {'TFLOP/s':>8s} {'cuDNN us':>9s}")
for name, cin, cout, kh, kw in (("convc1 1x1 324->256", 324, 256, 1, 1), ("convc2 3x3 256->192", 256, 192, 3, 3),
                                ("convf2 3x3 128->64", 128, 64, 3, 3), ("conv 3x3 256->126", 256, 126, 3, 3),
                                ("gru zr 1x5 384->256", 384, 256, 1, 5), ("gru q 1x5 384->128", 384, 128, 1, 5),
                                ("gru zr 5x1 384->256", 384, 256, 5, 1), ("gru q 5x1 384->128", 384, 128, 5, 1),
                                ("flow head 3x3 128->256", 128, 256, 3, 3), ("mask 3x3 128->256", 128, 256, 3, 3),
                                ("mask 1x1 256->576", 256, 576, 1, 1)):
    pad = (kh // 2, kw // 2)
    x = cl(torch.randn((1, cin, H, W), generator=g, device=dev))
    w = torch.randn((cout, cin, kh, kw), generator=g, device=dev) * 0.05
    b = torch.randn((cout,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, pad)
    out = cl(torch.empty((1, pc.cout_pad, H, W), device=dev))
    t_own = timeit(lambda: conv_tc.run(pc, x, relu=True, out=out))
    wc = cl(w)
    t_ref = timeit(lambda: torch.cudnn_convolution_relu(x, wc, b, (1, 1), pad, (1, 1), 1)) if cout % 8 == 0 else float("nan")
    gf = 2.0 * H * W * cin * cout * kh * kw / 1e9
    print(f"{name:34s} {gf:7.2f} {t_own:10.1f} {gf / t_own * 1e3:8.1f} {t_ref:9.1f}")
conv_tc.check(dev)

# weight head layers 2/3: 14 400 images of 9x9x128, stacked with a separator row
n, P, C = 14400, 9, 128
for f16 in (False, True):
    dt = torch.float16 if f16 else torch.float32
    x = torch.randn((1, n * (P + 1), P, C), generator=g, device=dev).to(dt).permute(0, 3, 1, 2)
    mid = torch.empty((1, n * (P + 1), P, C), device=dev, dtype=dt).permute(0, 3, 1, 2)
    w = torch.randn((C, C, 3, 3), generator=g, device=dev) * 0.03
    b = torch.randn((C,), generator=g, device=dev)
    w1 = torch.randn((C,), generator=g, device=dev)
    pc = conv_tc.PackedConv(w, b, 1, f16=f16)
    logits = torch.empty((n,), device=dev)
    t2 = timeit(lambda: conv_tc.run(pc, x, relu=True, out=mid, row_period=P + 1, row_valid=P, box=(9, 14)), n=3)
    t3 = timeit(lambda: conv_tc.run(pc, x, epilogue=_lib.CONV_HEAD_MEAN, w1x1=w1, head_out=logits, head_bias=0.1,
                                    row_period=P + 1, row_valid=P, box=(9, 14)), n=3)
    gf = 2.0 * n * 81 * C * C * 9 / 1e9
    print(f"weight head layer 2 ({'fp16' if f16 else 'tf32'}) {gf:7.1f} GFLOP {t2:8.1f} us {gf / t2 * 1e3:6.1f} TFLOP/s;  "
          f"layer 3 + 1x1 + mean {t3:8.1f} us")
    if not f16:
        xi = cl(torch.randn((n, C, P, P), generator=g, device=dev))
        wc = cl(w)
        t_ref = timeit(lambda: torch.cudnn_convolution_relu(xi, wc, b, (1, 1), (1, 1), (1, 1), 1), n=3)
        print(f"   cuDNN TF32 on (n,128,9,9) channels-last: {t_ref:8.1f} us")
conv_tc.check(dev)
