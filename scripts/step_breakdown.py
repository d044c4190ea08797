"""Where one tracked frame spends its device time: kernel table of an eager (no CUDA graph) step via torch.profiler.

    python scripts/step_breakdown.py [--mode tf32] > gpurun_out/step_breakdown.txt
"""
import functools
import sys
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from woft_b200 import tracker as wtracker  # noqa: E402
from woft_b200.corr import CorrBlock  # noqa: E402
from woft_b200.optical_flow import RAFTWrapper  # noqa: E402

mode = sys.argv[sys.argv.index("--mode") + 1] if "--mode" in sys.argv else "tf32"
dev = torch.device("cuda", 0)
conf = SimpleNamespace(raft_type="weighted", class_params=SimpleNamespace(small=False, alternate_corr=False), model=None,
                       seeded_weights=0, iters=12, padding_mode="nopad")
flower = RAFTWrapper(conf, device=dev, use_cuda_graph=False, corr_block=functools.partial(CorrBlock, mode=mode))
trk = wtracker.PlanarTracker(wtracker.default_config(conf), flower=flower)
base, frames, mask = bench.make_sequence(4, 0)
trk.init(base, mask)
for f in frames[:3]:
    trk.track(f)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    trk.track(frames[3])
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total if hasattr(e, "device_time_total") else e.cuda_time_total, e.count)
        for e in prof.key_averages() if (getattr(e, "device_time_total", 0) or getattr(e, "cuda_time_total", 0)) > 0
        and e.device_type.name == "CUDA"]
rows.sort(key=lambda r: -r[1])
total = sum(r[1] for r in rows)
print(f"device time of one tracked frame (eager, corr mode {mode}): {total/1e3:.2f} ms over {sum(r[2] for r in rows)} kernels")
for name, t, n in rows[:45]:
    print(f"{t/1e3:9.3f} ms {100*t/total:5.1f}% x{n:<4d} {name[:150]}")
