"""Pixel-grid helpers of pytracking/utils/misc.py used at the flow-wrapper seam."""
import torch


def unravel_indices(indices, shape, stack_dim=-1, np_order=False):
    """misc.py:71-96: flat indices -> coordinates; default order is (x, y), i.e. column then row."""
    coord = []
    for dim in reversed(shape):
        coord.append(indices % dim)
        indices = torch.div(indices, dim, rounding_mode="floor")
    if np_order:
        coord = coord[::-1]
    return torch.stack(coord, dim=stack_dim)


def torch_get_featuremap_coords(feature_map, device=None, keep_shape=False):
    """misc.py:45-68: (2, H*W) int64 [x; y] in raster order (or (2, H, W) with keep_shape)."""
    if type(feature_map) is tuple and len(feature_map) == 2:
        H, W = feature_map
        assert device is not None
    else:
        H, W = feature_map.shape[-2:]
        if device is None:
            device = feature_map.device
    xy = unravel_indices(torch.arange(H * W, device=device), (H, W), stack_dim=0)
    return xy.view(2, H, W) if keep_shape else xy
