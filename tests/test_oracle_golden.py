"""The CPU oracle against the golden vectors that oracle/gen_golden.py made by running the reference itself."""
import functools

import numpy as np
import pytest
import torch

from oracle import kornia_restated as kr
from oracle import woft_oracle as wo


@pytest.mark.parametrize("name", ["corr_small", "corr_odd", "corr_c256"])
def test_pyramid_and_lookup_match_reference(golden, name):
    g = golden(name)
    blk = wo.OracleCorrBlock(torch.from_numpy(g["fmap1"]), torch.from_numpy(g["fmap2"]))
    for l, lvl in enumerate(blk.corr_pyramid):
        assert tuple(lvl.shape) == tuple(g[f"level{l}_shape"])
        assert lvl.double().sum().item() == pytest.approx(float(g[f"level{l}_sum"]), abs=1e-9)
        if f"level{l}" in g:
            np.testing.assert_array_equal(lvl.numpy(), g[f"level{l}"])
    for tag in ("grid", "jitter", "oob"):
        out = blk(torch.from_numpy(g[f"coords_{tag}"]))
        assert not torch.isnan(out).any()
        np.testing.assert_array_equal(out.numpy(), g[f"lookup_{tag}"])


def test_pyramid_floor_semantics(golden):
    g = golden("corr_odd")  # 18 x 21 -> 9 x 10 -> 4 x 5 -> 2 x 2
    assert [tuple(g[f"level{l}_shape"][2:]) for l in range(4)] == [(18, 21), (9, 10), (4, 5), (2, 2)]


def test_lookup_channel_order_x_major(golden):
    """channel = level*81 + (xoff+4)*9 + (yoff+4): at integer coords the tap (xoff, yoff) is the volume entry."""
    g = golden("corr_small")
    lv0 = torch.from_numpy(g["level0"])  # (HW, 1, 16, 24)
    out = torch.from_numpy(g["lookup_grid"])  # coords = pixel grid
    H, W = 16, 24
    y, x = 7, 11
    q = y * W + x
    for xo, yo in ((-4, 0), (3, -2), (0, 4)):
        ch = (xo + 4) * 9 + (yo + 4)
        assert out[0, ch, y, x].item() == pytest.approx(lv0[q, 0, y + yo, x + xo].item(), abs=1e-5)


def _cases(g):
    for tag in ("exact4", "plain100", "tracker500", "batch3", "ragged37"):
        p1, p2 = torch.from_numpy(g[tag + "_p1"]), torch.from_numpy(g[tag + "_p2"])
        w = torch.from_numpy(g[tag + "_w"]) if tag + "_w" in g else None
        yield tag, p1, p2, w


def test_homography_estimators_match_reference(golden):
    g = golden("homography")
    huber2 = functools.partial(wo.IRLSq_Huber, k=2)
    fns = {"qr": wo.find_homography_nonhomogeneous_QR, "pinv": wo.find_homography_nonhomogeneous,
           "dlt": wo.find_homography_dlt, "irls_l1": wo.find_homography_IRLSq_QR,
           "irls_huber2": functools.partial(wo.find_homography_IRLSq_QR, reweighting_fn=huber2)}
    for tag, p1, p2, w in _cases(g):
        for key, fn in fns.items():
            np.testing.assert_array_equal(fn(p1, p2, w).numpy(), g[f"{tag}_{key}"], err_msg=f"{tag}/{key}")
        Hq = torch.from_numpy(g[f"{tag}_qr"])
        err = wo.torch_proj_errors(Hq, p1.permute(0, 2, 1), p2.permute(0, 2, 1))
        np.testing.assert_array_equal(err.numpy(), g[f"{tag}_proj_err"])


def test_homography_shape_errors():
    p = torch.zeros(1, 3, 2)
    with pytest.raises(AssertionError):
        wo.find_homography_nonhomogeneous_QR(p, p)
    with pytest.raises(AssertionError):
        wo.find_homography_nonhomogeneous_QR(torch.zeros(1, 5, 2), torch.zeros(1, 6, 2))


def test_hartley_normalisation_regression(golden):
    g = golden("homography")
    pn, T = kr.normalize_points(torch.from_numpy(g["norm_in"]))
    np.testing.assert_array_equal(pn.numpy(), g["norm_pts"])
    np.testing.assert_array_equal(T.numpy(), g["norm_T"])
    mean_dist, target = wo.hartley_scale_sanity(torch.from_numpy(g["norm_in"]))
    assert torch.allclose(mean_dist, torch.full_like(mean_dist, target), atol=1e-5)


def test_flow_net_matches_reference(golden):
    """woft_b200's torch mirror of WeightedRAFT, given the same seeded weights and the oracle correlation
    operator, reproduces the reference network's outputs (generated with the reference's own code)."""
    from woft_b200.raft_net import WeightedRAFT, seeded_state_dict
    g = golden("net_small")
    net = WeightedRAFT().eval()
    net.load_state_dict(seeded_state_dict(net, int(g["state_seed"])))
    t1 = torch.from_numpy(g["img1_bgr"][:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    t2 = torch.from_numpy(g["img2_bgr"][:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    torch.set_num_threads(8)
    for iters in (1, 12):
        flow_low, flow_up, w_low, w_up = wo.flow_pass(net, t1, t2, iters=iters)
        np.testing.assert_allclose(flow_low.numpy(), g[f"it{iters}_flow_low"], atol=2e-5)
        np.testing.assert_allclose(flow_up.numpy(), g[f"it{iters}_flow_up"], atol=2e-4)
        np.testing.assert_allclose(w_low.numpy(), g[f"it{iters}_w_low"], atol=2e-5)
        np.testing.assert_allclose(w_up.numpy(), g[f"it{iters}_w_up"], atol=2e-5)


def test_tracker_pieces_match_reference(golden):
    from woft_b200 import tracker
    g = golden("tracker_pieces")
    for n in (300, 777, 20000, 400000):
        keep = tracker.sobol_keep_indices(n)
        ref = g[f"sobol_keep_{n}"]
        if keep is None:
            np.testing.assert_array_equal(ref, np.arange(n))
        else:
            np.testing.assert_array_equal(keep, ref)
        m = wo.sobol_subsample_mask(n)
        np.testing.assert_array_equal(np.arange(n) if m is None else np.nonzero(m)[0], ref)
    np.testing.assert_allclose(tracker.compose_H(g["compose_a"], g["compose_b"]), g["compose_ab"], rtol=1e-12)
    ok = wo.redet_success(torch.from_numpy(g["redet_H"]), torch.from_numpy(g["redet_p2"][0].T.copy()),
                          torch.from_numpy(g["redet_p1"][0].T.copy()))
    assert ok == bool(g["redet_ok"])


def test_wrapper_golden_is_consistent(golden):
    """TC mode = pixel grid + flow (optical_flow/raft.py:185-218); checks the host-side grid helper."""
    from woft_b200.misc import torch_get_featuremap_coords
    g = golden("wrapper_small")
    b, d, H, W = g["last_flow_shape"]
    src = torch_get_featuremap_coords((int(H), int(W)), device="cpu")
    np.testing.assert_array_equal(src.numpy(), g["tc_src"])
    np.testing.assert_allclose(g["tc_dst"], g["tc_src"] + g["flow_flow"].reshape(2, -1), atol=1e-4)
    np.testing.assert_allclose(g["tc_w"].reshape(H, W), 1 / (1 + np.exp(-g["flow_w"][0])), atol=1e-6)


def test_kornia_restatement_against_the_package_when_present():
    """Row a17: oracle/kornia_restated.py against the real kornia wherever one is installed.  The image this
    repository is built and judged in ships no kornia wheel (pinned 0.5.11 by the reference, README.org:28,33), so
    here the test SKIPS and the three functions stay 'parity unpinned' -- the skip message says so in every run."""
    kornia = pytest.importorskip("kornia", reason="kornia is not installed: oracle/kornia_restated.py (Hartley "
                                                  "normalisation under every K4 golden vector) REMAINS UNPINNED")
    from oracle import kornia_restated as kr
    g = torch.Generator().manual_seed(0)
    for B, N in ((1, 4), (2, 50), (3, 500)):
        pts = torch.rand((B, N, 2), generator=g) * torch.tensor([1280.0, 720.0])
        want_p, want_T = kornia.geometry.epipolar.normalize_points(pts)
        got_p, got_T = kr.normalize_points(pts)
        np.testing.assert_allclose(got_T.numpy(), want_T.numpy(), rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(got_p.numpy(), want_p.numpy(), rtol=1e-5, atol=1e-6)
        ph = torch.randn((B, N, 3), generator=g)
        ph[0, 0, 2] = 0.0   # the |z| <= eps branch
        np.testing.assert_allclose(kr.convert_points_from_homogeneous(ph).numpy(),
                                   kornia.geometry.conversions.convert_points_from_homogeneous(ph).numpy(), rtol=1e-6)
        np.testing.assert_array_equal(kr.convert_points_to_homogeneous(pts).numpy(),
                                      kornia.geometry.conversions.convert_points_to_homogeneous(pts).numpy())
