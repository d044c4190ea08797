"""Launch each hand-written kernel a few times at BASELINE sizes, for ncu (-k regex:...) and quick timing.

    python scripts/profile_kernels.py [--time]
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import least_squares_H as lsq  # noqa: E402
from woft_b200.corr import CorrBlock  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
B, C, H, W = 1, 256, 90, 160
f1 = torch.randn((B, C, H, W), generator=g, device=dev)
f2 = torch.randn((B, C, H, W), generator=g, device=dev)
ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
coords = torch.stack([xs, ys]).float()[None] + torch.randn((1, 2, H, W), generator=g, device=dev) * 2.0
scratch = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(name, fn, n=5):
    ts = []
    for _ in range(n + 2):
        scratch.zero_()
        torch.cuda._sleep(400_000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = ts[2:]
    print(f"{name:34s} {min(ts):9.1f} us min {sum(ts)/len(ts):9.1f} us avg", flush=True)


blocks = {}
for mode in ("tf32", "tf32x3", "bf16", "fp16", "fp16x3"):
    def build(mode=mode):
        blocks[mode] = CorrBlock(f1, f2, mode=mode)
    timed(f"corr_pyramid[{mode}]", build, 3)
    blocks[mode].check()
blk = blocks["tf32"]
timed("corr_lookup flags=0 (float4 rows)", lambda: blk(coords, flags=0), 5)
timed("corr_lookup flags=1 (scalar rows)", lambda: blk(coords, flags=1), 5)
timed("corr_lookup flags=8 (NHWC output)", lambda: blk(coords, flags=8), 5)
# 13 lookups back to back as in one flow pass (L2 warm after the first): amortises the ~5 us event/launch overhead
cs = [coords + 0.3 * i for i in range(13)]
timed("13 x corr_lookup (one flow pass)", lambda: [blk(c) for c in cs], 5)
if "--l2gran" in sys.argv:
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    for gran in (32, 64, 128):
        rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran))  # cudaLimitMaxL2FetchGranularity
        timed(f"corr_lookup, L2 fetch granularity {gran} (rc {rc})", lambda: blk(coords), 5)
        timed(f"corr_pyramid[tf32], granularity {gran}", lambda: CorrBlock(f1, f2, mode="tf32"), 2)
f1b = torch.randn((4, C, H, W), generator=g, device=dev)
del blocks["tf32x3"], blocks["bf16"], blocks["fp16"], blocks["fp16x3"]
Bh, Nh = 1024, 14400
p1 = torch.rand((Bh, Nh, 2), generator=g, device=dev) * torch.tensor([1280.0, 720.0], device=dev)
p2 = p1 + torch.randn((Bh, Nh, 2), generator=g, device=dev)
w = torch.rand((Bh, Nh), generator=g, device=dev)
timed("homography QR  B=1024 N=14400", lambda: lsq.find_homography_nonhomogeneous_QR(p1, p2, w), 3)
timed("homography IRLS-L1 B=1024 N=14400", lambda: lsq.find_homography_IRLSq_QR(p1, p2, w), 2)
timed("homography DLT B=1024 N=14400", lambda: lsq.find_homography_dlt(p1, p2, w), 2)
timed("homography QR  B=1 N=500", lambda: lsq.find_homography_nonhomogeneous_QR(p1[:1, :500], p2[:1, :500], w[:1, :500]), 5)
torch.cuda.synchronize()
