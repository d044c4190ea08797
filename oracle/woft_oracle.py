"""torch-CPU restatement of WOFT's per-frame hot path.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Every function names the reference lines it follows (paths relative to /root/reference).  The arithmetic is
the reference's: fp32 torch ops in the same order (matmul then divide, avg_pool2d, grid_sample with
align_corners=True and zero padding, thin QR + triangular solve).  Pinned against the reference's own
code through tests/golden (made by oracle/gen_golden.py); the Hartley normalisation it calls lives in
kornia_restated.py and is parity-unpinned.
"""
import math

import torch
import torch.nn.functional as F

from . import kornia_restated as kr


# --------------------------------------------------------------------------------------------------
# K1 / K2: all-pairs correlation volume and its average-pooled pyramid
# --------------------------------------------------------------------------------------------------
def corr_volume(fmap1, fmap2):
    """pytracking/external/RAFT/raft_core/corr.py:61-69.  (B,C,H,W)x2 -> (B,H,W,1,H,W), scaled by 1/sqrt(C)."""
    B, C, H, W = fmap1.shape
    a = fmap1.reshape(B, C, H * W).transpose(1, 2)
    b = fmap2.reshape(B, C, H * W)
    vol = torch.matmul(a, b).view(B, H, W, 1, H, W)
    return vol / torch.sqrt(torch.tensor(C).float())


def corr_pyramid(fmap1, fmap2, num_levels=4):
    """corr.py:13-27.  Level l has shape (B*H*W, 1, H>>l, W>>l) with floor at every halving."""
    vol = corr_volume(fmap1, fmap2)
    B, H1, W1, one, H2, W2 = vol.shape
    level = vol.reshape(B * H1 * W1, one, H2, W2)
    levels = [level]
    for _ in range(num_levels - 1):
        level = F.avg_pool2d(level, 2, stride=2)
        levels.append(level)
    return levels


# --------------------------------------------------------------------------------------------------
# K3: radius-r bilinear lookup
# --------------------------------------------------------------------------------------------------
def sample_pixels(img, coords):
    """raft_core/utils/utils.py:59-73 (`bilinear_sampler`): pixel coords -> [-1,1] -> grid_sample
    (bilinear, zero padding, align_corners=True)."""
    H, W = img.shape[-2:]
    x, y = coords.split([1, 1], dim=-1)
    x = 2 * x / (W - 1) - 1
    y = 2 * y / (H - 1) - 1
    return F.grid_sample(img, torch.cat([x, y], dim=-1), align_corners=True)


def corr_lookup(levels, coords, radius=4):
    """corr.py:29-59.  coords (B,2,H1,W1) as (x,y) -> (B, L*(2r+1)^2, H1, W1) fp32 contiguous.

    Channel = level*(2r+1)^2 + (xoff+r)*(2r+1) + (yoff+r): the reference adds meshgrid(dy, dx) to (x, y)
    coordinates (corr.py:37-43), so the FIRST window index walks x.
    """
    r = radius
    coords = coords.permute(0, 2, 3, 1)
    B, H1, W1, _ = coords.shape
    outs = []
    for i, level in enumerate(levels):
        d = torch.linspace(-r, r, 2 * r + 1, device=coords.device)
        delta = torch.stack(torch.meshgrid(d, d, indexing="ij"), dim=-1)
        centre = coords.reshape(B * H1 * W1, 1, 1, 2) / 2 ** i
        window = sample_pixels(level, centre + delta.view(1, 2 * r + 1, 2 * r + 1, 2))
        outs.append(window.view(B, H1, W1, -1))
    return torch.cat(outs, dim=-1).permute(0, 3, 1, 2).contiguous().float()


class OracleCorrBlock:
    """Same constructor / call / `corr_pyramid` attribute as the reference CorrBlock (corr.py:12-69)."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels = num_levels
        self.radius = radius
        self.corr_pyramid = corr_pyramid(fmap1, fmap2, num_levels)

    def __call__(self, coords):
        return corr_lookup(self.corr_pyramid, coords, self.radius)


def alt_corr_forward(fmap1, fmap2, coords, radius):
    """What `alt_cuda_corr.forward` returns (alt_cuda_corr/correlation_kernel.cu:18-119, 260-286):
    fmap1 (B,H1,W1,C), fmap2 (B,H2,W2,C) channels-last, coords (B,N,H1,W1,2) -> (B,N,(2r+1)^2,H1,W1),
    UNSCALED dot products bilinearly interpolated at coords + integer offsets, zero outside fmap2,
    channel = (yoff+r) + (2r+1)*(xoff+r)  (correlation_kernel.cu:92-95)."""
    B, H1, W1, C = fmap1.shape
    H2, W2 = fmap2.shape[1:3]
    N = coords.shape[1]
    vol = torch.matmul(fmap1.reshape(B, H1 * W1, C), fmap2.reshape(B, H2 * W2, C).transpose(1, 2))
    vol = vol.reshape(B * H1 * W1, 1, H2, W2)
    r = radius
    d = torch.arange(-r, r + 1, dtype=coords.dtype)
    delta = torch.stack(torch.meshgrid(d, d, indexing="ij"), dim=-1).view(1, 2 * r + 1, 2 * r + 1, 2)
    out = []
    for n in range(N):
        centre = coords[:, n].reshape(B * H1 * W1, 1, 1, 2)
        # the CUDA kernel interpolates in pixel units directly; this goes through grid_sample's normalise
        # round trip, which agrees to ~1e-6 relative
        win = sample_pixels(vol, centre + delta)
        out.append(win.view(B, H1, W1, -1).permute(0, 3, 1, 2))
    return torch.stack(out, dim=1).contiguous()


# --------------------------------------------------------------------------------------------------
# K4: weighted homography estimators
# --------------------------------------------------------------------------------------------------
def _check_points(points1, points2):
    """least_squares_H.py:158-163 (same three checks in every estimator)."""
    if points1.shape != points2.shape:
        raise AssertionError(points1.shape)
    if not (len(points1.shape) >= 1 and points1.shape[-1] == 2):
        raise AssertionError(points1.shape)
    if points1.shape[1] < 4:
        raise AssertionError(points1.shape)


def _design(points1, points2, weights, columns):
    """Rows of the (non-)homogeneous DLT system in Hartley-normalised coordinates.

    least_squares_H.py:168-193 (8 columns + rhs) and :238-246 (9 columns).  Two consecutive rows per
    correspondence:   ax = [0,0,0,-x1,-y1,-1, y2x1, y2y1 | y2]   ay = [x1,y1,1,0,0,0,-x2x1,-x2y1 | -x2]
    """
    p1, T1 = kr.normalize_points(points1)
    p2, T2 = kr.normalize_points(points2)
    x1, y1 = torch.chunk(p1, dim=-1, chunks=2)
    x2, y2 = torch.chunk(p2, dim=-1, chunks=2)
    one, zero = torch.ones_like(x1), torch.zeros_like(x1)
    ax = [zero, zero, zero, -x1, -y1, -one, y2 * x1, y2 * y1]
    ay = [x1, y1, one, zero, zero, zero, -x2 * x1, -x2 * y1]
    if columns == 9:
        ax.append(y2)
        ay.append(-x2)
    ax, ay = torch.cat(ax, dim=-1), torch.cat(ay, dim=-1)
    B, N = x1.shape[:2]
    A = torch.stack((ax, ay), dim=2).reshape(B, 2 * N, columns)
    b = torch.stack((-y2, x2), dim=2).reshape(B, 2 * N, 1)
    w = None
    if weights is not None:
        w = weights.repeat_interleave(2, dim=1).unsqueeze(-1)
    return A, b, w, T1, T2


def _finish(solution8, T1, T2, eps=1e-8):
    """least_squares_H.py:204-209: append h33 = 1, de-normalise, divide by (H33 + eps)."""
    B = solution8.shape[0]
    h = torch.cat([solution8, torch.ones((B, 1, 1), dtype=solution8.dtype)], dim=1).reshape(B, 3, 3)
    H = T2.inverse() @ (h @ T1)
    return H / (H[..., -1:, -1:] + eps)


def _qr_solve(A, b):
    """least_squares_H.py:195-202: thin QR, then R x = Q^T b."""
    Q, R = torch.linalg.qr(A)
    return torch.linalg.solve_triangular(R, Q.transpose(1, 2) @ b, upper=True)


def find_homography_nonhomogeneous_QR(points1, points2, weights=None):
    """least_squares_H.py:142-210.  Rows are scaled by w (not sqrt(w)): the objective is sum (w_i r_i)^2."""
    _check_points(points1, points2)
    A, b, w, T1, T2 = _design(points1, points2, weights, 8)
    if w is not None:
        A, b = w * A, w * b
    return _finish(_qr_solve(A, b), T1, T2)


def find_homography_nonhomogeneous(points1, points2, weights=None):
    """least_squares_H.py:56-139.  Same system solved with pinv."""
    _check_points(points1, points2)
    A, b, w, T1, T2 = _design(points1, points2, weights, 8)
    if w is not None:
        A, b = w * A, w * b
    return _finish(torch.linalg.pinv(A) @ b, T1, T2)


def find_homography_dlt(points1, points2, weights=None):
    """least_squares_H.py:213-265.  9-column homogeneous system, Gram A^T (w A) (weight w, NOT w^2), SVD,
    last right-singular vector."""
    _check_points(points1, points2)
    A, _, w, T1, T2 = _design(points1, points2, weights, 9)
    if weights is None:
        G = A.transpose(-2, -1) @ A
    else:
        if not (len(weights.shape) == 2 and weights.shape == points1.shape[:2]):
            raise AssertionError(weights.shape)
        G = A.transpose(-2, -1) @ (w * A)
    _, _, V = torch.svd(G)
    H = V[..., -1].view(-1, 3, 3)
    H = T2.inverse() @ (H @ T1)
    return H / (H[..., -1:, -1:] + 1e-8)


def IRLSq_L1(residuals, eps=1e-8):
    """least_squares_H.py:268-269."""
    return 1 / (torch.abs(residuals) + eps)


def IRLSq_Huber(residuals, k=1, eps=1e-8):
    """least_squares_H.py:272-277.  1/|r| beyond k, exactly 1 inside (discontinuous at k; reference behaviour)."""
    a = torch.abs(residuals)
    w = 1 / (a + eps)
    w[a < k] = 1
    return w


def find_homography_IRLSq_QR(points1, points2, weights=None, reweighting_fn=IRLSq_L1, n_iter=5):
    """least_squares_H.py:280-346 without the `is_cuda` assertion (:292-293), which only gates the device.

    n_iter+1 solves; the residual is taken per ROW of the w-scaled system in normalised units, and the
    next solve scales rows by sqrt(fn(residual))."""
    _check_points(points1, points2)
    A, b, w, T1, T2 = _design(points1, points2, weights, 8)
    if w is not None:
        A, b = w * A, w * b
    rew = torch.ones_like(b)
    for _ in range(n_iter + 1):
        x = _qr_solve(rew * A, rew * b)
        rew = torch.sqrt(reweighting_fn(A @ x - b))
    return _finish(x, T1, T2)


def torch_proj_errors(H, pts_A, pts_B):
    """least_squares_H.py:474-489.  H (B,3,3), pts (B,2,N) -> L2 distance (B,N) between H*A and B."""
    ph = kr.convert_points_to_homogeneous(pts_A.permute(0, 2, 1)).permute(0, 2, 1)
    proj = kr.convert_points_from_homogeneous(torch.matmul(H, ph).permute(0, 2, 1)).permute(0, 2, 1)
    return torch.sqrt(torch.square(proj - pts_B).sum(dim=1))


def redet_success(H, template_coords, cur_coords, thr=5.0, min_frac=0.2):
    """pytracking/configs/YAOFT_single_control_repRAFT_sub500_noreliableinl_wLSq.py:14-21."""
    errs = torch_proj_errors(H, cur_coords[None], template_coords[None])
    return bool(torch.mean((errs <= thr).float()) > min_frac)


def sobol_subsample_mask(n_pts, to_draw=500):
    """configs/..._wLSq.py:31-53: boolean mask over n_pts points selecting round(n_pts * sobol_1d[k]) for the
    first `to_draw` points of the unscrambled 1-D Sobol sequence.  Returns None when to_draw >= n_pts
    (the reference then keeps everything)."""
    import numpy as np
    if to_draw >= n_pts:
        return None
    mask = np.zeros(n_pts) > 0
    u = torch.quasirandom.SobolEngine(dimension=1).draw(to_draw).cpu().numpy().flatten()
    mask[np.round(n_pts * u).astype(np.int32)] = True
    return mask


def mask_coords(template_mask, template_coords, cur_coords, weights, pw_mask):
    """pytracking/tracker/YAOF_tracker_single_control.py:287-312 with do_pw_mask=True, post_weights=None."""
    in_template = template_mask[template_coords[1, :], template_coords[0, :]]
    H, W = pw_mask.shape
    cur_int = cur_coords.round().long()
    oob = torch.logical_or(torch.any(cur_coords < 0, dim=0),
                           torch.logical_or(cur_int[0, :] >= W, cur_int[1, :] >= H))
    in_pw = torch.ones_like(in_template) > 0
    in_pw[oob] = False
    in_pw[in_pw.clone()] = pw_mask[cur_int[1, in_pw], cur_int[0, in_pw]]
    keep = torch.logical_and(in_template, in_pw)
    return template_coords[:, keep], cur_coords[:, keep], weights[:, keep], keep


# --------------------------------------------------------------------------------------------------
# whole flow pass (used as the timed CPU baseline and for end-to-end parity)
# --------------------------------------------------------------------------------------------------
def flow_pass(net, image1, image2, iters=12):
    """One WeightedRAFT pass with the oracle correlation operator.

    `net` is a torch module with the reference's WeightedRAFT layout (raft_core/weighted_raft.py:29-315);
    its conv stacks are ordinary torch layers, and the three correlation stages go through OracleCorrBlock.
    Returns (flow_low, flow_up, weights_low, weights_up) like weighted_raft.py:315 minus the volume."""
    with torch.no_grad():
        out = net(image1, image2, iters=iters, test_mode=True, corr_block=OracleCorrBlock)
    flow_low, flow_up, _, w_low, w_up = out
    return flow_low, flow_up, w_low, w_up


def corner_error(H_a, H_b, width, height):
    """Max distance (px) between the images of the four frame corners under two homographies."""
    c = torch.tensor([[0.0, 0.0, 1.0], [width, 0.0, 1.0], [width, height, 1.0], [0.0, height, 1.0]],
                     dtype=torch.float64).T
    pa = H_a.double() @ c
    pb = H_b.double() @ c
    pa = pa[..., :2, :] / pa[..., 2:, :]
    pb = pb[..., :2, :] / pb[..., 2:, :]
    return torch.sqrt(((pa - pb) ** 2).sum(dim=-2)).max().item()


def hartley_scale_sanity(points):
    """After normalisation the mean distance to the origin is sqrt(2): used by a property test."""
    pn, _ = kr.normalize_points(points)
    return pn.norm(dim=-1).mean(dim=-1), math.sqrt(2.0)
