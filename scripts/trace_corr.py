"""Per-role timeline of corr_tc_kernel (debug): woft_corr_set_trace_buffer hands the library a zeroed device buffer into
which the first 4 CTAs write globaltimer stamps per tile:
[producer tile begin, producer tile end, MMA begin, MMA issue end, epilogue begin, epilogue end]."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from woft_b200 import _lib  # noqa: E402
from woft_b200.corr import CorrBlock  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
f1 = torch.randn((1, 256, 90, 160), generator=g, device=dev)
f2 = torch.randn((1, 256, 90, 160), generator=g, device=dev)
for mode in sys.argv[1:] or ["bf16", "tf32", "tf32x3"]:
    CorrBlock(f1, f2, mode=mode)
    torch.cuda.synchronize()
    buf = torch.zeros(4 * 64 * 6, dtype=torch.int64, device=dev)
    _lib.load().woft_corr_set_trace_buffer(_lib.ptr(buf))
    CorrBlock(f1, f2, mode=mode)
    torch.cuda.synchronize()
    _lib.load().woft_corr_set_trace_buffer(None)
    st_all = buf.cpu().numpy().reshape(4, 64, 6)
    rows = np.array([[c, t, *st_all[c, t]] for c in range(4) for t in range(64) if st_all[c, t].any()], dtype=np.int64)
    rows = rows[rows[:, 0] == 0]
    t0 = rows[0, 2]
    st = (rows[:, 2:] - t0) / 1e3  # us
    n = len(st)
    print(f"== {mode}: CTA 0, {n} tiles, total {st[-1,5]:.1f} us  ({st[-1,5]/n:.2f} us/tile)")
    prod = st[:, 1] - st[:, 0]
    mma = st[:, 3] - st[:, 2]
    epi = st[:, 5] - st[:, 4]
    epi_wait = st[1:, 4] - st[:-1, 5]           # epilogue idle between tiles (waiting for the accumulator)
    mma_wait = st[1:, 2] - st[:-1, 3]           # MMA issuer idle between tiles (waiting for a free accumulator)
    for name, a in (("producer issue span", prod), ("MMA issue span", mma), ("epilogue span", epi),
                    ("epilogue idle before tile", epi_wait), ("MMA idle before tile", mma_wait)):
        a = a[2:-2] if len(a) > 8 else a
        print(f"   {name:28s} mean {a.mean():7.2f} us  p10 {np.percentile(a,10):7.2f}  p90 {np.percentile(a,90):7.2f}")
    print("   first tiles [prod0 prod1 mma0 mma1 epi0 epi1] us:")
    for r in st[:5]:
        print("   ", " ".join(f"{v:8.2f}" for v in r))
