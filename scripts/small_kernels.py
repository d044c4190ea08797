"""Launch the update block's small custom kernels on the 1280x720 shapes (for event timing or ncu).

    python scripts/small_kernels.py            # CUDA-event times
    ncu --set full --import-source on --clock-control none -o gpurun_out/small python scripts/small_kernels.py --once
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import fast_net  # noqa: E402

dev = torch.device("cuda", 0)
once = "--once" in sys.argv
g = torch.Generator().manual_seed(0)
B, H, W = 1, 90, 160
cl = torch.channels_last
flow = (torch.randn((B, 2, H, W), generator=g) * 3).to(dev)
w7, b7 = (torch.randn((128, 2, 7, 7), generator=g) * 0.1).to(dev), torch.randn((128,), generator=g).to(dev)
x256 = torch.randn((B, 256, H, W), generator=g).to(dev).contiguous(memory_format=cl)
w3, b3 = (torch.randn((2, 256, 3, 3), generator=g) * 0.05).to(dev), torch.randn((2,), generator=g).to(dev)
c0 = torch.randn((B, 2, H, W), generator=g).to(dev)
coords = c0.clone()
f1, f2 = torch.empty_like(c0), torch.empty_like(c0).contiguous(memory_format=cl)
cor = torch.randn((B, 192, H, W), generator=g).to(dev).contiguous(memory_format=cl)
flo = torch.randn((B, 64, H, W), generator=g).to(dev).contiguous(memory_format=cl)
cf = torch.empty((B, 256, H, W), device=dev).contiguous(memory_format=cl)
out = torch.randn((B, 128, H, W), generator=g).to(dev).contiguous(memory_format=cl)
buf1 = torch.zeros((B, 384, H, W), device=dev).contiguous(memory_format=cl)
buf2 = torch.zeros_like(buf1)
big = torch.empty(64 << 20, device=dev)  # 256 MiB, flushes L2 between timed launches

img2 = (torch.rand((2, 3, 720, 1280), generator=g) * 2 - 1).to(dev).contiguous(memory_format=cl)
wst, bst = (torch.randn((64, 3, 7, 7), generator=g) * 0.05).to(dev), torch.randn((64,), generator=g).to(dev)

from woft_b200 import _lib  # noqa: E402
nq = 14400
win = torch.randn((nq, 324), generator=g).to(dev)
mresp = torch.randn((nq,), generator=g).to(dev)
wh_w, wh_b = (torch.randn((128, 5, 3, 3), generator=g) * 0.2).to(dev), torch.randn((128,), generator=g).to(dev)
wh_out = torch.empty((nq, 128, 9, 9), device=dev).contiguous(memory_format=cl)


def head_conv1():
    rc = _lib.load().woft_weight_head_conv1(_lib.ptr(win), _lib.ptr(mresp), _lib.ptr(wh_w), _lib.ptr(wh_b), nq, 9, 4, 128,
                                            _lib.ptr(wh_out), _lib.current_stream(dev))
    assert rc == 0


cases = {
    "weight_head_conv1 (14400 windows)": head_conv1,
    "stem_conv7x7 pair (fnet)": lambda: fast_net.stem_conv7x7(img2, wst),
    "stem_conv7x7 one frame + bias + relu": lambda: fast_net.stem_conv7x7(img2[:1], wst, bst, relu=True),
    "flow_conv7x7_relu": lambda: fast_net.flow_conv7x7_relu(flow, w7, b7),
    "flow_out3x3 (one launch)": lambda: fast_net.flow_out3x3(x256, w3, b3),
    "flow_head_step (taps + gather)": lambda: fast_net.flow_head_step(x256, w3, b3, coords=coords, coords0=c0,
                                                                      flow_nchw=f1, flow_nhwc=f2),
    "concat [cor|flo]": lambda: fast_net.concat_nhwc([cor, flo], cf, 0),
    "concat [out126|flow] x2": lambda: fast_net.concat_nhwc([out, f2], buf1, 256, buf2, 256, take=[126, 2]),
}
for name, fn in cases.items():
    fn()
    torch.cuda.synchronize()
    if once:
        continue
    ts = []
    for warm in (False, True):
        best = []
        for _ in range(10):
            if not warm:
                big.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(200000)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best.append(e0.elapsed_time(e1) * 1e3)
        ts.append(min(best))
    print(f"{name:34s} cold L2 {ts[0]:7.1f} us   warm {ts[1]:7.1f} us")
