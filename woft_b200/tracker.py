"""The caller of both seams: WOFT's per-frame tracker state machine, kept as the reference has it.

Mirror of `YAOFTrackerSingleControl` (pytracking/tracker/YAOF_tracker_single_control.py:18-327) and of the
default config's plug-ins (pytracking/configs/YAOFT_single_control_repRAFT_sub500_noreliableinl_wLSq.py:14-53):
pre-warp by the last good homography, template->pre-warped flow + weights, mask, Sobol-subsample to 500,
weighted homography, inlier check, and on failure the frame-to-frame fallback chained onto the previous
homography.  It exists so the whole tracked-frame path can be run and timed where the reference is absent;
with the reference present, its own tracker runs unchanged on top of `RAFTWrapper` and `find_homography_*`
(INTEGRATION.md).

Kept on the host exactly as in the reference: cv2.warpPerspective pre-warp (:89-95), compose_H in float64
numpy (utils/geom_utils.py:365-373).  Moved to the device: the correspondence mask and the Sobol pick, which the
reference does with boolean indexing plus a GPU->CPU->GPU round trip of the mask (:287-312, config :31-53);
here the (deterministic) Sobol positions are drawn once per point count and applied as a device gather.
"""
from types import SimpleNamespace

import cv2
import numpy as np
import torch

from . import least_squares_H as lsq
from .optical_flow import RAFTWrapper, _cfg


def compose_H(*Hs):
    """geom_utils.py:365-373: product in reverse order, normalised by the bottom-right element."""
    for H in Hs:
        if H is None:
            return None
    out = np.eye(3)
    for H in Hs:
        out = np.dot(H, out)
    return out / out[2, 2]


_SOBOL = {}


def sobol_positions(to_draw=500):
    """First `to_draw` points of the unscrambled 1-D Sobol sequence (config :44-46), cached."""
    if to_draw not in _SOBOL:
        _SOBOL[to_draw] = torch.quasirandom.SobolEngine(dimension=1).draw(to_draw).numpy().flatten()
    return _SOBOL[to_draw]


def sobol_keep_indices(n_pts, to_draw=500):
    """Sorted unique indices round(n_pts * u_k) the reference's boolean mask selects (config :31-53).
    None when to_draw >= n_pts (everything is kept).  np.round is round-half-to-even, as in the reference."""
    if to_draw >= n_pts:
        return None
    idx = np.round(n_pts * sobol_positions(to_draw)).astype(np.int32)
    return np.unique(idx)


def subsampler(coords_a, coords_b, weights, to_draw=500):
    """config :31-53 on device tensors: (2,N), (2,N), (1,N) -> at most `to_draw` columns, raster order kept."""
    assert coords_a.shape == coords_b.shape
    n_pts = coords_a.shape[1]
    assert weights.shape == (1, n_pts)
    keep = sobol_keep_indices(n_pts, to_draw)
    if keep is None:
        return coords_a, coords_b, weights
    idx = torch.from_numpy(keep.astype(np.int64)).to(coords_a.device, non_blocking=True)
    return coords_a[:, idx], coords_b[:, idx], weights[:, idx]


REDET_THRESHOLD_PX, REDET_MIN_FRACTION = 5.0, 0.2   # config :14-21


def redet_success_fn(H, template_coords, cur_coords, weights=None, thr=REDET_THRESHOLD_PX, min_frac=REDET_MIN_FRACTION):
    """config :14-21: fraction of correspondences reprojecting within `thr` px must exceed `min_frac`."""
    errs = lsq.torch_proj_errors(H, cur_coords[None], template_coords[None])
    return torch.mean((errs <= thr).float()) > min_frac


def find_homography(pts_A, pts_B, weights=None):
    """config :24-28."""
    return lsq.find_homography_nonhomogeneous_QR(pts_A, pts_B, weights=weights)


_SOBOL_DEV = {}


def select_fit_check(flow_up, weight_logits, template_mask, pw_mask, to_draw=500, thr=REDET_THRESHOLD_PX,
                     min_frac=REDET_MIN_FRACTION,
                     prev_to_template=None):
    """The default config's post-flow stage as five launches and no host round trip: correspondence mask + Sobol pick
    (woft_select_correspondences), weighted fit (woft_homography_fit_n), inlier test (woft_inlier_fraction).
    flow_up (2,H,W), weight_logits (1,H,W) before the sigmoid, masks (H,W) bool/uint8 -- all on the device.
    Returns a device tensor of 12 floats: H row-major in [0:9], inlier fraction in [9], number of points in [10].
    With `prev_to_template` (3x3, host: the previous frame's homography to the template) this is the lost branch:
    the flow starts in the previous frame and its pixels are masked by the template mask seen through that matrix
    (woft_select_correspondences_local); H then maps current points onto previous-frame points."""
    from . import _lib
    lib = _lib.load()
    dev = flow_up.device
    _, H, W = flow_up.shape
    key = (dev, to_draw)
    if key not in _SOBOL_DEV:
        _SOBOL_DEV[key] = torch.from_numpy(np.sort(sobol_positions(to_draw)).astype(np.float32)).to(dev)
    flow_up = flow_up.contiguous()
    weight_logits = weight_logits.contiguous() if weight_logits is not None else None   # None: unit weights
    pts_t = torch.empty((to_draw, 2), device=dev, dtype=torch.float32)
    pts_c = torch.empty((to_draw, 2), device=dev, dtype=torch.float32)
    w = torch.empty((to_draw,), device=dev, dtype=torch.float32)
    n = torch.empty((1,), device=dev, dtype=torch.int32)
    res = torch.zeros((12,), device=dev, dtype=torch.float32)
    ok = torch.empty((1,), device=dev, dtype=torch.uint8)
    nbytes = lib.woft_select_scratch_bytes(H, W)
    scratch = torch.empty((nbytes,), device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        st = _lib.current_stream(dev)
        if prev_to_template is None:
            rc = lib.woft_select_correspondences(_lib.ptr(flow_up), _lib.ptr(weight_logits), _lib.ptr(template_mask),
                                                 _lib.ptr(pw_mask), H, W, _lib.ptr(_SOBOL_DEV[key]), to_draw, 1,
                                                 _lib.ptr(pts_t), _lib.ptr(pts_c), _lib.ptr(w), _lib.ptr(n),
                                                 _lib.ptr(scratch), nbytes, st)
        else:
            import ctypes
            m = np.ascontiguousarray(np.asarray(prev_to_template, dtype=np.float64).reshape(9))
            rc = lib.woft_select_correspondences_local(_lib.ptr(flow_up), _lib.ptr(weight_logits), _lib.ptr(template_mask),
                                                       m.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), H, W,
                                                       _lib.ptr(_SOBOL_DEV[key]), to_draw, 1, _lib.ptr(pts_t),
                                                       _lib.ptr(pts_c), _lib.ptr(w), _lib.ptr(n), _lib.ptr(scratch),
                                                       nbytes, st)
        _lib.check(rc, "woft_select_correspondences")
        # H maps the current (pre-warped) points onto the template points (tracker :152-155)
        rc = lib.woft_homography_fit_n(_lib.ptr(pts_c), _lib.ptr(pts_t), _lib.ptr(w), _lib.ptr(n), to_draw,
                                       _lib.H_NONHOMOGENEOUS, 0, 1.0, _lib.ptr(res), st)
        _lib.check(rc, "woft_homography_fit_n")
        rc = lib.woft_inlier_fraction(_lib.ptr(res), _lib.ptr(pts_c), _lib.ptr(pts_t), _lib.ptr(n), thr, min_frac,
                                      ctypes_offset(res, 9), _lib.ptr(ok), st)
        _lib.check(rc, "woft_inlier_fraction")
    return res, (pts_t, pts_c, w, n)


def ctypes_offset(t, index):
    import ctypes
    return ctypes.c_void_p(t.data_ptr() + index * t.element_size())


def default_config(flow_config, device_prewarp=True):
    """get_config() of the default tracker config (:56-71).  `device_prewarp` (not a reference key) moves the
    cv2.warpPerspective pre-warp and its validity mask onto the GPU (woft_warp_perspective_u8)."""
    return SimpleNamespace(flow_config=flow_config, H_estimator=find_homography, redet_success_fn=redet_success_fn,
                           pw_mask=True, no_prewarp_after_N=10, subsampler_fn=subsampler, no_local_H=False,
                           do_not_mask_TCs_by_prewarped=False, device_prewarp=device_prewarp,
                           fused_select=device_prewarp)


def invert3x3_opencv(H):
    """Inverse of a 3x3 matrix by the cofactor formula in the order of operations of cv::invert (the matrix
    cv2.warpPerspective inverts internally), so that the device pre-warp sees bit-identical coefficients."""
    S = np.asarray(H, dtype=np.float64)
    d = (S[0, 0] * (S[1, 1] * S[2, 2] - S[1, 2] * S[2, 1]) - S[0, 1] * (S[1, 0] * S[2, 2] - S[1, 2] * S[2, 0])
         + S[0, 2] * (S[1, 0] * S[2, 1] - S[1, 1] * S[2, 0]))
    if d == 0.0:
        raise np.linalg.LinAlgError("singular pre-warp homography")
    d = 1.0 / d
    t = np.empty((3, 3), dtype=np.float64)
    t[0, 0] = (S[1, 1] * S[2, 2] - S[1, 2] * S[2, 1]) * d
    t[0, 1] = (S[0, 2] * S[2, 1] - S[0, 1] * S[2, 2]) * d
    t[0, 2] = (S[0, 1] * S[1, 2] - S[0, 2] * S[1, 1]) * d
    t[1, 0] = (S[1, 2] * S[2, 0] - S[1, 0] * S[2, 2]) * d
    t[1, 1] = (S[0, 0] * S[2, 2] - S[0, 2] * S[2, 0]) * d
    t[1, 2] = (S[0, 2] * S[1, 0] - S[0, 0] * S[1, 2]) * d
    t[2, 0] = (S[1, 0] * S[2, 1] - S[1, 1] * S[2, 0]) * d
    t[2, 1] = (S[0, 1] * S[2, 0] - S[0, 0] * S[2, 1]) * d
    t[2, 2] = (S[0, 0] * S[1, 1] - S[0, 1] * S[1, 0]) * d
    return t


def prewarp_on_device(frame_u8, H, want_mask=True):
    """(H,W,3) uint8 device tensor warped by the 3x3 matrix `H` the way cv2.warpPerspective(frame, H, (W,H),
    flags=INTER_LINEAR) does, and the `> 0` mask of an image of ones warped alike (:89-95)."""
    from . import _lib
    import ctypes
    h, w, c = frame_u8.shape
    frame_u8 = frame_u8.contiguous()
    out = torch.empty_like(frame_u8)
    mask = torch.empty((h, w), dtype=torch.uint8, device=frame_u8.device) if want_mask else None
    inv = invert3x3_opencv(H)
    with torch.cuda.device(frame_u8.device):
        rc = _lib.load().woft_warp_perspective_u8(_lib.ptr(frame_u8), h, w, c,
                                                  inv.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), _lib.ptr(out),
                                                  _lib.ptr(mask), _lib.current_stream(frame_u8.device))
    _lib.check(rc, "woft_warp_perspective_u8")
    return out, (mask.bool() if want_mask else None)


class PlanarTracker:
    """YAOF_tracker_single_control.py:18-327 (no debug drawing, no fast-forward, no input downscaling)."""

    def __init__(self, config, flower=None):
        self.C = config
        self.flower = flower if flower is not None else RAFTWrapper(config.flow_config)
        self.device = self.flower.device
        self.flow_passes = 0

    def init(self, img, mask):
        """:26-47."""
        self.template_img = img
        self.template_dev = self.flower.upload_u8(img, slot=2).clone() if _cfg(self.C, "device_prewarp", False) else None
        self.prev_dev = self.template_dev
        self.template_mask = torch.from_numpy(mask > 0).to(self.device)
        self.np_template_mask = mask
        self.prev_H2init = np.eye(3)
        self.last_good_H2init = np.eye(3)
        self.prev_img = img
        self.lost = False
        self.N_lost = 0

    def track(self, input_img):
        """:57-285.  Returns (H_cur2init (3,3) float64 numpy, meta)."""
        meta = SimpleNamespace()
        n_max = _cfg(self.C, "no_prewarp_after_N")
        if n_max and self.N_lost > n_max:
            self.last_good_H2init = np.eye(3)
        prewarp_H = self.last_good_H2init
        h, w = input_img.shape[:2]
        on_device = _cfg(self.C, "device_prewarp", False)
        if on_device:
            frame_dev = self.flower.upload_u8(input_img, slot=1)
            prewarped, pw_mask = prewarp_on_device(frame_dev, prewarp_H)
            template = self.template_dev
        else:
            frame_dev = None
            prewarped = cv2.warpPerspective(input_img, prewarp_H, (w, h), flags=cv2.INTER_LINEAR)
            pw_mask = cv2.warpPerspective(np.ones((h, w)), prewarp_H, (w, h), flags=cv2.INTER_LINEAR)
            pw_mask = torch.from_numpy(pw_mask > 0).to(self.device, non_blocking=True)
            template = self.template_img

        fused = on_device and self._fused_select_applies()
        if fused:
            # mask + Sobol pick + fit + inlier test stay on the device; ONE 48-byte copy brings H and the flag back
            res = self.estimate_global_fused(template, prewarped, pw_mask).cpu().numpy()
            H_np = res[:9].reshape(3, 3).astype(np.float64)
            ok = bool(res[9] > REDET_MIN_FRACTION)
        else:
            H_pw2init, ok, t_coords, c_coords, weights = self.estimate_global(template, prewarped, pw_mask)
            H_np = H_pw2init.detach().cpu().numpy()[0].astype(np.float64)   # D2H of the 3x3 (reference :156)
        H_global = compose_H(prewarp_H, H_np)
        meta.H_global_cur2init = H_global.copy()

        if bool(ok):  # device sync on the flag, as in the reference (:167)
            H_cur2init = H_global
            self.lost, self.N_lost = False, 0
        else:
            self.lost = True
            self.N_lost += 1
            if _cfg(self.C, "no_local_H", False):
                H_cur2init = H_global
            elif fused:
                # local flow (T-1 -> T), mask, Sobol pick and fit on the device; one 48-byte copy back (:178-204)
                flow, w_logits = self.flower.compute_flow(self.prev_dev, frame_dev, mode="flow", do_sigmoid=False)
                self.flow_passes += 1
                res, _ = select_fit_check(flow, w_logits, self.template_mask, None, prev_to_template=self.prev_H2init)
                H_flow = res[:9].cpu().numpy().reshape(3, 3).astype(np.float64)
                H_local = compose_H(H_flow, self.prev_H2init) if np.all(np.isfinite(H_flow)) else None
                if H_local is None or not np.all(np.isfinite(H_local)):
                    H_local = self.prev_H2init   # fewer than 4 points or a degenerate fit: the reference's except branch
                meta.H_local_cur2init = H_local.copy()
                H_cur2init = H_local
            else:
                p_coords, c2_coords, w2 = self.flower.compute_flow(self.prev_dev if on_device else self.prev_img,
                                                                   frame_dev if on_device else input_img, mode="TC",
                                                                   do_sigmoid=True)
                self.flow_passes += 1
                p_coords, c2_coords, w2 = self._mask_coords_flow(p_coords, c2_coords, w2)
                if self.C.subsampler_fn:
                    p_coords, c2_coords, w2 = self.C.subsampler_fn(p_coords.float(), c2_coords, w2)
                try:
                    H_flow = self.C.H_estimator(c2_coords.T[None].contiguous(), p_coords.float().T[None].contiguous(), w2)
                    H_flow = H_flow.detach().cpu().numpy()[0].astype(np.float64)
                    H_local = compose_H(H_flow, self.prev_H2init)
                    if not np.all(np.isfinite(H_local)):
                        raise FloatingPointError("degenerate local homography")
                except Exception:
                    H_local = self.prev_H2init
                meta.H_local_cur2init = H_local.copy()
                H_cur2init = H_local

        if on_device:
            self.prev_dev = frame_dev
        else:
            self.prev_img = input_img.copy()
        self.prev_H2init = H_cur2init.copy()
        if not self.lost:
            self.last_good_H2init = H_cur2init.copy()
        meta.lost, meta.N_lost, meta.global_H_success = self.lost, self.N_lost, bool(ok)
        return H_cur2init, meta

    def _fused_select_applies(self):
        """The fused post-flow stage (woft_select_correspondences + fit + inlier test) IS the default config's three
        plug-ins; a config that swapped any of them (wIRLSq, plainLSq, cvransac, alwayswarp variants of the reference)
        goes through the plug-ins themselves."""
        C = self.C
        return bool(_cfg(C, "fused_select", False)) and getattr(C, "H_estimator", None) is find_homography \
            and getattr(C, "subsampler_fn", None) is subsampler and getattr(C, "redet_success_fn", None) is redet_success_fn

    def estimate_global_fused(self, template_img, prewarped_img, pw_mask):
        """estimate_global for the default config's plug-ins, without boolean indexing or host round trips."""
        flow, w_logits = self.flower.compute_flow(template_img, prewarped_img, mode="flow", do_sigmoid=False)
        self.flow_passes += 1
        mask = pw_mask if not _cfg(self.C, "do_not_mask_TCs_by_prewarped", False) else None
        res, _ = select_fit_check(flow, w_logits, self.template_mask, mask)
        return res

    def estimate_global(self, template_img, prewarped_img, pw_mask):
        """The device part of one tracked frame (:101-163): template -> pre-warped flow + weights, correspondence
        mask, Sobol pick, weighted homography, inlier test.  Images may be host uint8 arrays or device uint8
        tensors; nothing is copied back to the host here.  Returns (H (1,3,3), ok (0-dim bool tensor), and the
        correspondences that were fitted)."""
        t_coords, c_coords, weights = self.flower.compute_flow(template_img, prewarped_img, mode="TC", do_sigmoid=True)
        self.flow_passes += 1
        t_coords, c_coords, weights, _ = self._mask_coords(
            t_coords, c_coords, weights, pw_mask, do_pw_mask=not _cfg(self.C, "do_not_mask_TCs_by_prewarped", False))
        t_coords = t_coords.float()
        if self.C.subsampler_fn:
            t_coords, c_coords, weights = self.C.subsampler_fn(t_coords, c_coords, weights)
        H_pw2init = self.C.H_estimator(c_coords.T[None].contiguous(), t_coords.T[None].contiguous(), weights).float()
        ok = self.C.redet_success_fn(H_pw2init, t_coords, c_coords, weights)
        return H_pw2init, ok, t_coords, c_coords, weights

    def _mask_coords(self, template_coords, cur_coords, weights, pw_mask, do_pw_mask=True):
        """:287-312 without the in-place boolean scatter: keep = template mask & in bounds & pre-warp mask."""
        in_template = self.template_mask[template_coords[1], template_coords[0]]
        H, W = pw_mask.shape
        cur_int = cur_coords.round().long()
        inb = ~(torch.any(cur_coords < 0, dim=0) | (cur_int[0] >= W) | (cur_int[1] >= H))
        if do_pw_mask:
            # (cur >= 0 as floats can still round to -0 -> index 0; clamp only guards the masked-out columns)
            inb = inb & pw_mask[cur_int[1].clamp(0, H - 1), cur_int[0].clamp(0, W - 1)]
        keep = in_template & inb
        return template_coords[:, keep], cur_coords[:, keep], weights[:, keep], keep

    def _mask_coords_flow(self, prev_coords, cur_coords, weights):
        """:314-327."""
        h, w = self.template_mask.shape
        prev_mask = cv2.warpPerspective(self.np_template_mask, np.linalg.inv(self.prev_H2init), (w, h),
                                        flags=cv2.INTER_NEAREST) > 0
        prev_mask = torch.from_numpy(prev_mask).to(prev_coords.device)
        keep = prev_mask[prev_coords[1], prev_coords[0]]
        return prev_coords[:, keep], cur_coords[:, keep], weights[:, keep]
