"""Restatement of the kornia==0.5.11 functions the reference calls.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: kornia is pinned at 0.5.11 by the reference (README.org:28,33) but is neither
vendored under /root/reference nor installable here, so these three functions follow the
published 0.5.x algorithm and could not be diffed against the real package.

Call sites in the reference:
  normalize_points                  pytracking/utils/least_squares_H.py:82-83,168-169,238-239,296-297
  convert_points_to_homogeneous     pytracking/utils/least_squares_H.py:415,430,447,485
  convert_points_from_homogeneous   pytracking/utils/least_squares_H.py:416,432,434,459,486
"""
import torch


def convert_points_to_homogeneous(points):
    """(..., D) -> (..., D+1) by appending a constant 1."""
    return torch.nn.functional.pad(points, [0, 1], "constant", 1.0)


def convert_points_from_homogeneous(points, eps=1e-8):
    """(..., D+1) -> (..., D).  Divides by (z + eps) where |z| > eps, else leaves the point unscaled."""
    z = points[..., -1:]
    usable = torch.abs(z) > eps
    scale = torch.where(usable, 1.0 / (z + eps), torch.ones_like(z))
    return scale * points[..., :-1]


def transform_points(trans, points):
    """Apply (B, D+1, D+1) homogeneous transforms to (B, N, D) points."""
    ph = convert_points_to_homogeneous(points)
    ph = torch.matmul(ph, trans.permute(0, 2, 1))
    return convert_points_from_homogeneous(ph)


def normalize_points(points, eps=1e-8):
    """Hartley isotropic normalisation, unweighted, over all N points of each batch item.

    mu = mean(p);  s = sqrt(2) / (mean ||p - mu|| + eps);  T = [[s,0,-s*mu_x],[0,s,-s*mu_y],[0,0,1]]
    returns (T p, T) with shapes (B, N, 2), (B, 3, 3).
    """
    mu = torch.mean(points, dim=1, keepdim=True)
    dist = (points - mu).norm(dim=-1).mean(dim=-1)
    s = torch.sqrt(torch.tensor(2.0, dtype=points.dtype, device=points.device)) / (dist + eps)
    one, zero = torch.ones_like(s), torch.zeros_like(s)
    T = torch.stack([s, zero, -s * mu[..., 0, 0],
                     zero, s, -s * mu[..., 0, 1],
                     zero, zero, one], dim=-1).view(-1, 3, 3)
    return transform_points(T, points), T
