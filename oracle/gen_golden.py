"""Generate tests/golden/*.npz by running the UNMODIFIED reference on seeded inputs.

TEST INFRASTRUCTURE ONLY; runs in the build container only (needs /root/reference):

    python -m oracle.gen_golden            # rewrites tests/golden/

The reference ships no tests, golden vectors or fixtures of its own, so these outputs of
the reference's own code are what pins the oracle (tests/test_oracle_golden.py) and, through it, the CUDA
path.  Everything is fp32 on CPU, 8 threads, torch 2.11.
"""
import functools
import sys
import types
import warnings
from pathlib import Path

import numpy as np
import torch

from . import ref_import

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def _np(t):
    return t.detach().cpu().numpy()


def make_coords(B, H, W, seed, jitter, oob_frac=0.0):
    """Pixel grid + U(-jitter, jitter); a fraction of queries is thrown far outside the map."""
    g = torch.Generator().manual_seed(seed)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    c = torch.stack([xs, ys], dim=0).float()[None].repeat(B, 1, 1, 1)
    c = c + (torch.rand(c.shape, generator=g) * 2 - 1) * jitter
    if oob_frac > 0:
        far = torch.rand((B, 1, H, W), generator=g) < oob_frac
        c = torch.where(far, c + (torch.rand(c.shape, generator=g) - 0.5) * 6 * max(H, W), c)
    return c


def gen_corr(ref, name, B, C, H, W, seed, keep_levels):
    g = torch.Generator().manual_seed(seed)
    f1 = torch.randn((B, C, H, W), generator=g)
    f2 = torch.randn((B, C, H, W), generator=g)
    blk = ref.corr.CorrBlock(f1, f2, num_levels=4, radius=4)
    out = {"fmap1": _np(f1), "fmap2": _np(f2)}
    for l, lvl in enumerate(blk.corr_pyramid):
        out[f"level{l}_shape"] = np.array(lvl.shape)
        out[f"level{l}_sum"] = np.array(lvl.double().sum().item())
        out[f"level{l}_abs_sum"] = np.array(lvl.double().abs().sum().item())
        if l in keep_levels:
            out[f"level{l}"] = _np(lvl)
    out["mean_response"] = _np(blk.corr_pyramid[0].mean(dim=(1, 2, 3)).view(B, H, W))
    for tag, (jit, oob) in {"grid": (0.0, 0.0), "jitter": (8.0, 0.0), "oob": (8.0, 0.1)}.items():
        c = make_coords(B, H, W, seed + 17, jit, oob)
        out[f"coords_{tag}"] = _np(c)
        out[f"lookup_{tag}"] = _np(blk(c))
    np.savez_compressed(OUT / f"{name}.npz", **out)
    print(name, {k: v.shape for k, v in out.items() if v.ndim > 1})


def make_h_case(B, N, seed, noise, outlier_frac, with_weights):
    g = torch.Generator().manual_seed(seed)
    p1 = torch.rand((B, N, 2), generator=g) * torch.tensor([1280.0, 720.0])
    H = torch.eye(3)[None].repeat(B, 1, 1)
    H[:, :2, :2] += (torch.rand((B, 2, 2), generator=g) - 0.5) * 0.2
    H[:, :2, 2] += (torch.rand((B, 2), generator=g) - 0.5) * 60
    H[:, 2, :2] += (torch.rand((B, 2), generator=g) - 0.5) * 2e-4
    ph = torch.cat([p1, torch.ones(B, N, 1)], dim=-1) @ H.transpose(1, 2)
    p2 = ph[..., :2] / ph[..., 2:]
    p2 = p2 + torch.randn((B, N, 2), generator=g) * noise
    w = torch.rand((B, N), generator=g) if with_weights else None
    if outlier_frac > 0:
        bad = torch.rand((B, N), generator=g) < outlier_frac
        p2 = p2 + bad[..., None] * torch.randn((B, N, 2), generator=g) * 30
        if w is not None:
            w = torch.where(bad, w * 0.05, w)
    return p1.contiguous(), p2.contiguous(), w, H


def gen_homography(ref):
    lsq = ref.lsq
    irls = ref_import.irls_without_cuda_check(lsq)
    huber2 = functools.partial(lsq.IRLSq_Huber, k=2)
    cases = {
        "exact4": dict(B=1, N=4, seed=1, noise=0.0, outlier_frac=0.0, with_weights=False),
        "plain100": dict(B=1, N=100, seed=2, noise=0.5, outlier_frac=0.0, with_weights=False),
        "tracker500": dict(B=1, N=500, seed=3, noise=0.5, outlier_frac=0.1, with_weights=True),
        "batch3": dict(B=3, N=2000, seed=4, noise=0.5, outlier_frac=0.3, with_weights=True),
        "ragged37": dict(B=2, N=37, seed=5, noise=1.0, outlier_frac=0.2, with_weights=True),
    }
    out = {}
    for tag, kw in cases.items():
        p1, p2, w, Ht = make_h_case(**kw)
        out[f"{tag}_p1"], out[f"{tag}_p2"], out[f"{tag}_Htrue"] = _np(p1), _np(p2), _np(Ht)
        if w is not None:
            out[f"{tag}_w"] = _np(w)
        out[f"{tag}_qr"] = _np(lsq.find_homography_nonhomogeneous_QR(p1, p2, w))
        out[f"{tag}_pinv"] = _np(lsq.find_homography_nonhomogeneous(p1, p2, w))
        out[f"{tag}_dlt"] = _np(lsq.find_homography_dlt(p1, p2, w))
        out[f"{tag}_irls_l1"] = _np(irls(p1, p2, w, reweighting_fn=lsq.IRLSq_L1, n_iter=5))
        out[f"{tag}_irls_huber2"] = _np(irls(p1, p2, w, reweighting_fn=huber2, n_iter=5))
        wd = None if w is None else w.double()
        out[f"{tag}_qr_f64"] = _np(lsq.find_homography_nonhomogeneous_QR(p1.double(), p2.double(), wd))
        out[f"{tag}_dlt_f64"] = _np(lsq.find_homography_dlt(p1.double(), p2.double(), wd))
        out[f"{tag}_irls_l1_f64"] = _np(irls(p1.double(), p2.double(), wd, reweighting_fn=lsq.IRLSq_L1, n_iter=5))
        # reprojection error helper (least_squares_H.py:474-489) on the QR result
        Hq = torch.from_numpy(out[f"{tag}_qr"])
        out[f"{tag}_proj_err"] = _np(lsq.torch_proj_errors(Hq, p1.permute(0, 2, 1), p2.permute(0, 2, 1)))
    # Hartley normalisation at the (stubbed, parity-unpinned) kornia boundary, recorded for regression
    p1, _, _, _ = make_h_case(B=2, N=50, seed=9, noise=0, outlier_frac=0, with_weights=False)
    import kornia
    pn, T = kornia.geometry.epipolar.normalize_points(p1)
    out["norm_in"], out["norm_pts"], out["norm_T"] = _np(p1), _np(pn), _np(T)
    np.savez_compressed(OUT / "homography.npz", **out)
    print("homography", len(out), "arrays")


def ref_args():
    return types.SimpleNamespace(small=False, mixed_precision=False, alternate_corr=False,
                                 weight_head_structure=[(128, 3), (128, 3), (128, 3)],
                                 mask_estimation=False, model=None)


def make_frames(H, W, seed):
    """Smooth random texture and a shifted + slightly sheared copy (BGR uint8), so flow is non-degenerate."""
    import cv2
    rng = np.random.RandomState(seed)
    base = rng.randint(0, 256, (H // 4, W // 4, 3)).astype(np.uint8)
    img1 = cv2.resize(base, (W, H), interpolation=cv2.INTER_CUBIC)
    M = np.array([[1.01, 0.01, 2.5], [-0.008, 0.995, -1.5], [0, 0, 1.0]])
    img2 = cv2.warpPerspective(img1, M, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
    return img1, img2


def gen_net(ref):
    from woft_b200.raft_net import seeded_state_dict
    model = ref.weighted_raft.WeightedRAFT(ref_args()).eval()
    model.load_state_dict(seeded_state_dict(model, 0))
    H, W = 128, 160
    img1, img2 = make_frames(H, W, 5)
    t1 = torch.from_numpy(img1[:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    t2 = torch.from_numpy(img2[:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    out = {"img1_bgr": img1, "img2_bgr": img2, "state_seed": np.array(0)}
    with torch.no_grad():
        for iters in (1, 12):
            flow_low, flow_up, vol, w_low, w_up = model(t1, t2, iters=iters, test_mode=True)
            out[f"it{iters}_flow_low"], out[f"it{iters}_flow_up"] = _np(flow_low), _np(flow_up)
            out[f"it{iters}_w_low"], out[f"it{iters}_w_up"] = _np(w_low), _np(w_up)
    np.savez_compressed(OUT / "net_small.npz", **out)
    print("net_small", {k: v.shape for k, v in out.items()})

    # the flow-wrapper seam: reference RAFTWrapper.compute_flow in TC and flow mode (optical_flow/raft.py:81-218)
    ckpt = Path("/tmp/woft_seeded_ckpt.pth")
    torch.save(seeded_state_dict(model, 0), ckpt)
    conf = ref.config.Config()
    conf.of_class = ref.flow.RAFTWrapper
    conf.raft_type = "weighted"
    conf.class_params = ref.config.Config()
    conf.class_params.small = False
    conf.class_params.mixed_precision = False
    conf.class_params.alternate_corr = False
    conf.class_params.weight_head_structure = [(128, 3), (128, 3), (128, 3)]
    conf.model = ckpt
    conf.add_module_to_statedict = True
    conf.non_strict_loading = False
    conf.iters = 12
    conf.padding_mode = "nopad"
    wout = {}
    with ref_import.cpu_redirect(), torch.no_grad():
        wrapper = ref.flow.RAFTWrapper(conf)
        src, dst, w = wrapper.compute_flow(img1, img2, mode="TC", do_sigmoid=True)
        wout["tc_src"], wout["tc_dst"], wout["tc_w"] = _np(src), _np(dst), _np(w)
        flow, w2 = wrapper.compute_flow(img1, img2, mode="flow", do_sigmoid=False)
        wout["flow_flow"], wout["flow_w"] = _np(flow), _np(w2)
        wout["last_flow_shape"] = np.array([wrapper.last_flow_shape[k] for k in ("batch", "delta", "H", "W")])
        # the padders (optical_flow/raft.py:221-271, raft_core/utils/utils.py:7-26) on a frame size that is not a
        # multiple of 8: replicate padding, bilinear resize and crop, each through the reference wrapper
        oimg1, oimg2 = make_frames(133, 171, 6)   # RAFT pads to 136x176, crop cuts to 128x168 (a 15x20 feature map
        # would leave a 1x2 coarsest level, on which the reference's bilinear sampler divides by zero: all-NaN flow)
        wout["odd_img1_bgr"], wout["odd_img2_bgr"] = oimg1, oimg2
        for pmode in ("RAFT", "crop"):
            conf.padding_mode = pmode
            flow, w2 = wrapper.compute_flow(oimg1, oimg2, mode="flow", do_sigmoid=False)
            wout[f"pad_{pmode}_flow"], wout[f"pad_{pmode}_w"] = _np(flow), _np(w2)
        conf.padding_mode = "nopad"
        # 'Michal' cannot run through the reference wrapper with raft_type 'weighted': raft.py:150 unpads a mask that
        # only the 'weighted_masked' network produces (AttributeError on None).  Its padder class is recorded directly.
        mp = ref.flow.MichalPadder((1, 3, 125, 163))  # -> 128x168
        g = torch.Generator().manual_seed(21)
        a, b = torch.rand((1, 3, 125, 163), generator=g) * 255, torch.rand((1, 3, 125, 163), generator=g) * 255
        pa, pb = mp.pad(a, b)
        fl = torch.randn((1, 2, 128, 168), generator=g) * 3
        wout["michal_in_a"], wout["michal_in_b"], wout["michal_pad_a"], wout["michal_pad_b"] = _np(a), _np(b), _np(pa), _np(pb)
        wout["michal_flow_in"], wout["michal_flow_out"] = _np(fl), _np(mp.unpad(fl))
    np.savez_compressed(OUT / "wrapper_small.npz", **wout)
    print("wrapper_small", {k: v.shape for k, v in wout.items()})


def gen_tracker_pieces(ref):
    """Sobol subsampler, redetection test and compose_H of the default tracker config, on seeded inputs."""
    from pytracking.utils.config import load_config
    with ref_import.cpu_redirect():
        cfg = load_config(ref_import.REF_ROOT / "pytracking/configs/YAOFT_single_control_repRAFT_sub500_noreliableinl_wLSq.py")
    out = {}
    for n in (300, 777, 20000, 400000):
        a = torch.arange(2 * n).view(2, n).float()
        w = torch.ones(1, n)
        ca, cb, cw = cfg.subsampler_fn(a, a, w)
        out[f"sobol_keep_{n}"] = _np(ca[0]).astype(np.int64)
    p1, p2, w, Ht = make_h_case(B=1, N=500, seed=11, noise=0.5, outlier_frac=0.5, with_weights=True)
    Hq = ref.lsq.find_homography_nonhomogeneous_QR(p1, p2, w)
    ok = cfg.redet_success_fn(Hq, p2[0].T.contiguous(), p1[0].T.contiguous(), w)
    out["redet_p1"], out["redet_p2"], out["redet_H"], out["redet_ok"] = _np(p1), _np(p2), _np(Hq), np.array(bool(ok))
    Ha = np.array([[1.1, 0.02, 5], [0.01, 0.9, -3], [1e-5, 2e-5, 1.0]])
    Hb = np.array([[0.9, -0.02, 1], [0.03, 1.05, 2], [-2e-5, 1e-5, 0.97]])
    out["compose_a"], out["compose_b"], out["compose_ab"] = Ha, Hb, ref.geom.compose_H(Ha, Hb)
    np.savez_compressed(OUT / "tracker_pieces.npz", **out)
    print("tracker_pieces", {k: v.shape for k, v in out.items()})


def main():
    warnings.simplefilter("ignore")
    torch.set_num_threads(8)
    OUT.mkdir(parents=True, exist_ok=True)
    ref = ref_import.import_reference()
    which = set(sys.argv[1:]) or {"corr", "homography", "net", "tracker"}
    if "corr" in which:
        gen_corr(ref, "corr_small", B=1, C=16, H=16, W=24, seed=0, keep_levels=(0, 1, 2, 3))
        gen_corr(ref, "corr_odd", B=2, C=8, H=18, W=21, seed=1, keep_levels=(1, 2, 3))
        gen_corr(ref, "corr_c256", B=1, C=256, H=24, W=32, seed=2, keep_levels=(2, 3))
    if "homography" in which:
        gen_homography(ref)
    if "net" in which:
        gen_net(ref)
    if "tracker" in which:
        gen_tracker_pieces(ref)


if __name__ == "__main__":
    main()
