"""One eager tracked frame between cudaProfilerStart/Stop, for `ncu --profile-from-start off -k regex:...`.

    ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:'flow_|concat' \
        -o gpurun_out/frame python scripts/ncu_frame.py
"""
import functools
import sys
from pathlib import Path
from types import SimpleNamespace

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from woft_b200 import tracker as wtracker  # noqa: E402
from woft_b200.corr import CorrBlock  # noqa: E402
from woft_b200.optical_flow import RAFTWrapper  # noqa: E402

mode = sys.argv[sys.argv.index("--mode") + 1] if "--mode" in sys.argv else "fp16"
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
torch.cuda.profiler.start()
trk.track(frames[3])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
