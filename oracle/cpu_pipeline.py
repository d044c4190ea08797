"""One tracked frame on the host cores, the way the reference does it.  TEST INFRASTRUCTURE ONLY.

Used as the checker of the end-to-end path and as the thing timed by bench.py's `cpu_baseline` leg and its
`--impl reference` arm (the reference is Python and cannot travel to the GPU box, so this port is what runs
there).  It strings the oracle restatements together in the order of
pytracking/tracker/YAOF_tracker_single_control.py:57-170 (global branch):
  cv2 pre-warp (:89-95) -> RAFTWrapper.compute_flow 'TC' on CPU (optical_flow/raft.py:113-218) ->
  _mask_coords (:287-312) -> Sobol subsample (config :31-53) -> find_homography_nonhomogeneous_QR ->
  redet_success_fn (config :14-21) -> compose_H.
`track_frame_local` is the branch taken when that test fails (:178-204, 314-327): flow from the previous frame,
template mask warped into the previous frame, Sobol subsample, fit, composition with the previous homography.
"""
import cv2
import numpy as np
import torch

from . import woft_oracle as wo


def make_net(seed=0):
    """The flow network on CPU with the seeded stand-in weights (the WOFT checkpoints are not shipped)."""
    from woft_b200.raft_net import WeightedRAFT, seeded_state_dict
    net = WeightedRAFT().eval()
    net.load_state_dict(seeded_state_dict(net, seed))
    return net.requires_grad_(False)


def compute_flow_tc(net, src_bgr, dst_bgr, iters=12):
    """optical_flow/raft.py:113-218, mode='TC', do_sigmoid=True, padding 'nopad', on CPU in fp32."""
    H, W = src_bgr.shape[:2]
    assert H % 8 == 0 and W % 8 == 0
    src = torch.from_numpy(src_bgr[:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    dst = torch.from_numpy(dst_bgr[:, :, ::-1].copy()).permute(2, 0, 1).float()[None]
    _, flow_up, _, w_up = wo.flow_pass(net, src, dst, iters=iters)
    ys, xs = np.unravel_index(np.arange(H * W), (H, W))
    src_coords = torch.from_numpy(np.stack((xs, ys), axis=0))
    dst_coords = src_coords + flow_up.reshape(2, H * W)
    return src_coords, dst_coords, torch.sigmoid(w_up).reshape(1, H * W)


def track_frame(net, template_bgr, frame_bgr, template_mask, prewarp_H, iters=12):
    """Returns (H_cur2init (3,3) float64, success flag, number of fitted correspondences)."""
    h, w = frame_bgr.shape[:2]
    prewarped = cv2.warpPerspective(frame_bgr, prewarp_H, (w, h), flags=cv2.INTER_LINEAR)
    pw_mask = cv2.warpPerspective(np.ones((h, w)), prewarp_H, (w, h), flags=cv2.INTER_LINEAR)
    pw_mask = torch.from_numpy(pw_mask > 0)
    t_coords, c_coords, weights = compute_flow_tc(net, template_bgr, prewarped, iters)
    t_coords, c_coords, weights, _ = wo.mask_coords(torch.from_numpy(template_mask > 0), t_coords, c_coords, weights,
                                                    pw_mask)
    t_coords = t_coords.float()
    keep = wo.sobol_subsample_mask(t_coords.shape[1])
    if keep is not None:
        t_coords, c_coords, weights = t_coords[:, keep], c_coords[:, keep], weights[:, keep]
    H = wo.find_homography_nonhomogeneous_QR(c_coords.T[None].contiguous(), t_coords.T[None].contiguous(), weights)
    ok = wo.redet_success(H, t_coords, c_coords)
    H_global = np.dot(H[0].double().numpy(), prewarp_H)
    return H_global / H_global[2, 2], ok, t_coords.shape[1]


def track_frame_local(net, prev_bgr, frame_bgr, template_mask, prev_H2init, iters=12):
    """The lost branch of YAOF_tracker_single_control.py:178-204.  Returns (H_local_cur2init (3,3) float64, number of
    fitted correspondences)."""
    h, w = frame_bgr.shape[:2]
    p_coords, c_coords, weights = compute_flow_tc(net, prev_bgr, frame_bgr, iters)
    # _mask_coords_flow (:314-327): the template mask seen from the previous frame, nearest-neighbour warp
    prev_mask = cv2.warpPerspective(template_mask, np.linalg.inv(prev_H2init), (w, h), flags=cv2.INTER_NEAREST) > 0
    keep = torch.from_numpy(prev_mask)[p_coords[1], p_coords[0]]
    p_coords, c_coords, weights = p_coords[:, keep].float(), c_coords[:, keep], weights[:, keep]
    pick = wo.sobol_subsample_mask(p_coords.shape[1])
    if pick is not None:
        p_coords, c_coords, weights = p_coords[:, pick], c_coords[:, pick], weights[:, pick]
    H_flow = wo.find_homography_nonhomogeneous_QR(c_coords.T[None].contiguous(), p_coords.T[None].contiguous(), weights)
    # compose_H(H_flow, prev_H2init) multiplies in reverse order (geom_utils.py:365-373)
    H_local = np.dot(prev_H2init, H_flow[0].double().numpy())
    return H_local / H_local[2, 2], p_coords.shape[1]
