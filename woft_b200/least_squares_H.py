"""Seam B: the weighted homography estimators of pytracking/utils/least_squares_H.py.

Same names, arguments, return shapes and error behaviour as the reference; one kernel launch each
(woft_homography_fit, K4) instead of ~25 launches around a batched QR / pinv / SVD:

  find_homography_nonhomogeneous_QR   least_squares_H.py:142-210   (the tracker's default H_estimator)
  find_homography_nonhomogeneous      :56-139   (pinv of the same system: same least-squares solution)
  find_homography_dlt                 :213-265
  find_homography_IRLSq_QR            :280-346  with IRLSq_L1 (:268-269) / IRLSq_Huber (:272-277)
  torch_proj_errors                   :474-489

`points1`, `points2`: (B, N, 2); `weights`: (B, N) or None; returns (B, 3, 3) mapping points1 -> points2 with
H[2,2] ~ 1.  Shape errors raise AssertionError exactly where the reference does (:158-163).  Inputs must be
CUDA tensors (the reference's IRLS variant already insists on that, :292-293); there is no CPU path here.
"""
import functools

import torch

from . import _lib


def IRLSq_L1(residuals, eps=1e-8):
    """least_squares_H.py:268-269 (kept callable for configs that pass it as `reweighting_fn`)."""
    return 1 / (torch.abs(residuals) + eps)


def IRLSq_Huber(residuals, k=1, eps=1e-8):
    """least_squares_H.py:272-277."""
    a = torch.abs(residuals)
    w = 1 / (a + eps)
    w[a < k] = 1
    return w


def _check(points1, points2):
    if points1.shape != points2.shape:
        raise AssertionError(points1.shape)
    if not (len(points1.shape) >= 1 and points1.shape[-1] == 2):
        raise AssertionError(points1.shape)
    if points1.shape[1] < 4:
        raise AssertionError(points1.shape)


def _prep(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def _fit(points1, points2, weights, method, n_iter=0, huber_k=1.0):
    _check(points1, points2)
    if not points1.is_cuda:
        raise AssertionError("correspondences should be on GPU")
    _lib.require_cuda(points2, weights)
    B, N = points1.shape[:2]
    if weights is not None and tuple(weights.shape) != (B, N):
        raise AssertionError(weights.shape)
    p1, p2 = _prep(points1), _prep(points2)
    w = _prep(weights) if weights is not None else None
    dev = p1.device
    H = torch.empty((B, 3, 3), device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        rc = _lib.load().woft_homography_fit(_lib.ptr(p1), _lib.ptr(p2), _lib.ptr(w), B, N, method, n_iter,
                                             float(huber_k), _lib.ptr(H), _lib.current_stream(dev))
    _lib.check(rc, "woft_homography_fit")
    return H.to(points1.dtype) if points1.dtype in (torch.float64, torch.float16, torch.bfloat16) else H


def find_homography_nonhomogeneous_QR(points1, points2, weights=None):
    return _fit(points1, points2, weights, _lib.H_NONHOMOGENEOUS)


def find_homography_nonhomogeneous(points1, points2, weights=None):
    return _fit(points1, points2, weights, _lib.H_NONHOMOGENEOUS)


def find_homography_dlt(points1, points2, weights=None):
    return _fit(points1, points2, weights, _lib.H_DLT)


def _reweighting_to_method(fn):
    """Map the reference's `reweighting_fn` callables onto the kernel's modes."""
    base, kwargs = fn, {}
    while isinstance(base, functools.partial):
        kwargs = {**base.keywords, **kwargs}
        if base.args:
            raise NotImplementedError("positional partial arguments on a reweighting function are not supported")
        base = base.func
    name = getattr(base, "__name__", "")
    if name == "IRLSq_L1" and not kwargs:
        return _lib.H_IRLS_L1, 1.0
    if name == "IRLSq_Huber" and set(kwargs) <= {"k"}:
        return _lib.H_IRLS_HUBER, float(kwargs.get("k", 1))
    raise NotImplementedError(f"reweighting function {fn!r}: the fused kernel implements IRLSq_L1 and IRLSq_Huber(k) "
                              "with the reference's eps")


def find_homography_IRLSq_QR(points1, points2, weights=None, reweighting_fn=IRLSq_L1, n_iter=5):
    method, k = _reweighting_to_method(reweighting_fn)
    return _fit(points1, points2, weights, method, n_iter=n_iter, huber_k=k)


def torch_proj_errors(GT_H, pts_A, pts_B):
    """least_squares_H.py:474-489: H (B,3,3), pts (B,2,N) -> (B,N) distance between H*pts_A and pts_B."""
    _lib.require_cuda(GT_H, pts_A, pts_B)
    H, a, b = _prep(GT_H), _prep(pts_A), _prep(pts_B)
    B, two, N = a.shape
    if two != 2 or b.shape != a.shape or tuple(H.shape) != (B, 3, 3):
        raise AssertionError((tuple(H.shape), tuple(a.shape), tuple(b.shape)))
    err = torch.empty((B, N), device=a.device, dtype=torch.float32)
    with torch.cuda.device(a.device):
        rc = _lib.load().woft_proj_errors(_lib.ptr(H), _lib.ptr(a), _lib.ptr(b), B, N, _lib.ptr(err),
                                          _lib.current_stream(a.device))
    _lib.check(rc, "woft_proj_errors")
    return err
