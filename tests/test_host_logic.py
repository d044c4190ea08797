"""Host-side mirrors of the reference interface that need no GPU: config reading, padders, sharding."""
import os
import sys
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def test_cfg_reads_namespace_dict_and_falsy_config():
    from woft_b200.optical_flow import _cfg

    class Config:  # the reference's attribute bag: unknown keys are falsy Config objects (utils/config.py:5-17)
        def __getattr__(self, name):
            return Config()

        def __bool__(self):
            return False

    c = Config()
    c.iters = 12
    c.seeded_weights = 0
    assert _cfg(c, "iters") == 12 and _cfg(c, "seeded_weights") == 0
    assert _cfg(c, "padding_mode", "nopad") == "nopad"
    assert _cfg({"a": 1}, "a") == 1 and _cfg(SimpleNamespace(a=2), "b", 7) == 7


def test_padders_follow_reference_shapes():
    from woft_b200.optical_flow import CropPadder, InputPadder, MichalPadder, NoPadder
    x = torch.arange(2 * 3 * 37 * 50, dtype=torch.float32).view(2, 3, 37, 50)
    with pytest.raises(AssertionError):
        NoPadder(x.shape)
    NoPadder((1, 3, 720, 1280))
    a, b = CropPadder(x.shape).pad(x, x)
    assert a.shape[-2:] == (32, 48)
    p = InputPadder(x.shape)
    a, b = p.pad(x, x)
    assert a.shape[-2:] == (40, 56) and torch.equal(p.unpad(a), x)
    m = MichalPadder(x.shape)
    a, _ = m.pad(x, x)
    assert a.shape[-2:] == (40, 56) and m.unpad(a[:, :2]).shape[-2:] == (37, 50)


def test_subsampler_keeps_raster_order_and_count():
    from woft_b200 import tracker
    n = 5000
    a = torch.arange(2 * n).view(2, n).float()
    w = torch.rand(1, n)
    ca, cb, cw = tracker.subsampler(a, a, w)
    assert ca.shape == (2, 500) and cw.shape == (1, 500)
    assert torch.all(ca[0, 1:] > ca[0, :-1])
    ca2, _, _ = tracker.subsampler(a[:, :100], a[:, :100], w[:, :100])
    assert ca2.shape == (2, 100)


def test_sequence_sharding_is_a_partition():
    from woft_b200.sharding import shard_sequences
    for world in (1, 2, 4, 8):
        got = sorted(s for r in range(world) for s in shard_sequences(11, r, world))
        assert got == list(range(11))
    assert shard_sequences(8, 3, 8) == [3]
    assert shard_sequences(8, 1, 2) == [1, 3, 5, 7]


def _gather_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, str(ROOT))
    import torch.distributed as dist
    from woft_b200.sharding import gather_results, max_over_ranks, shard_sequences
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = {s: np.full((3, 3, 3), float(s)) for s in shard_sequences(5, rank, world)}
    elapsed = max_over_ranks(0.5 + rank)
    merged = gather_results(mine)
    if rank == 0:
        assert sorted(merged) == [0, 1, 2, 3, 4]
        assert all(float(merged[s][0, 0, 0]) == float(s) for s in merged)
        assert elapsed == pytest.approx(0.5 + world - 1)
        (Path(out_dir) / "ok").write_text("ok")
    dist.destroy_process_group()


def test_host_side_result_gather_two_ranks(tmp_path):
    """N>1 path: independent sequences per rank, no data-path collective, a host-side gather of the (T,3,3)
    homographies and a max-over-ranks of the elapsed time (gloo, world_size 2, CPU)."""
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_gather_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").exists()


def test_padders_match_reference_values(golden):
    """Padder classes on CPU tensors against outputs of the reference classes (tests/golden/wrapper_small.npz)."""
    import torch
    from woft_b200.optical_flow import CropPadder, InputPadder, MichalPadder
    gw = golden("wrapper_small")
    a, b = torch.from_numpy(gw["michal_in_a"]), torch.from_numpy(gw["michal_in_b"])
    mp = MichalPadder(a.shape)
    pa, pb = mp.pad(a, b)
    np.testing.assert_array_equal(pa.numpy(), gw["michal_pad_a"])
    np.testing.assert_array_equal(pb.numpy(), gw["michal_pad_b"])
    np.testing.assert_array_equal(mp.unpad(torch.from_numpy(gw["michal_flow_in"])).numpy(), gw["michal_flow_out"])
    ip = InputPadder(a.shape)
    xa, xb = ip.pad(a, b)
    assert tuple(xa.shape[-2:]) == (128, 168) and torch.equal(ip.unpad(xa), a) and torch.equal(ip.unpad(xb), b)
    assert torch.equal(xa[..., 0, :], xa[..., 1, :])   # replicate padding: 1 row on top, 2 at the bottom (sintel centring)
    ca, cb = CropPadder(a.shape).pad(a, b)
    assert tuple(ca.shape[-2:]) == (120, 160) and torch.equal(ca, a[..., :120, :160])


def test_opencv_order_inverse_is_bit_identical_to_cv_invert():
    import cv2
    from woft_b200.tracker import invert3x3_opencv
    rng = np.random.RandomState(0)
    for _ in range(2000):
        M = np.eye(3) + rng.randn(3, 3) * np.array([[0.2, 0.2, 50], [0.2, 0.2, 50], [1e-4, 1e-4, 0.05]])
        np.testing.assert_array_equal(invert3x3_opencv(M), cv2.invert(M)[1])
