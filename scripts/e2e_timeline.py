"""Host-side timeline of PlanarTracker.track at 1280x720: where the end-to-end frame time goes beyond the device step.

    python scripts/e2e_timeline.py
"""
import sys
import time
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from woft_b200 import tracker as wtracker  # noqa: E402
from woft_b200.optical_flow import RAFTWrapper  # noqa: E402

dev = torch.device("cuda", 0)
conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False, alternate_corr=False), model=None,
                       seeded_weights=0, iters=12, padding_mode="nopad")
flower = RAFTWrapper(conf, device=dev)
trk = wtracker.PlanarTracker(wtracker.default_config(conf), flower=flower)
base, frames, mask = bench.make_sequence(24, 0)
trk.init(base, mask)
for f in frames[:4]:
    trk.track(f)
torch.cuda.synchronize()

marks = []
orig_upload, orig_fused = flower.upload_u8, trk.estimate_global_fused


def upload(img, slot):
    t0 = time.perf_counter()
    out = orig_upload(img, slot)
    marks.append(("upload (pinned staging copy + H2D enqueue)", time.perf_counter() - t0))
    return out


def fused(*a):
    t0 = time.perf_counter()
    out = orig_fused(*a)
    marks.append(("enqueue flow graph + select + fit", time.perf_counter() - t0))
    return out


flower.upload_u8, trk.estimate_global_fused = upload, fused
tot = []
for f in frames[4:]:
    t0 = time.perf_counter()
    trk.track(f)
    tot.append(time.perf_counter() - t0)
agg = {}
for k, v in marks:
    agg.setdefault(k, []).append(v)
print(f"track(): {1e3 * np.median(tot):.3f} ms median over {len(tot)} frames")
for k, v in agg.items():
    print(f"  {k:50s} {1e3 * np.median(v):.3f} ms")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
tmpl, pw = trk.template_dev, flower.upload_u8(frames[5], 1)
pwm = torch.ones(pw.shape[:2], dtype=torch.bool, device=dev)
torch.cuda.synchronize()
e0.record()
for _ in range(10):
    orig_fused(tmpl, pw, pwm)
e1.record()
torch.cuda.synchronize()
print(f"  device step (estimate_global_fused, events)          {e0.elapsed_time(e1) / 10:.3f} ms")
