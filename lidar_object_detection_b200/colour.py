"""Headless colour buffers.

The reference shows its results in Open3D / matplotlib windows; neither a window nor a
graphics stack exists on the GPU box ("unavailable on this box"), so the colour
arithmetic that feeds them is written to device buffers instead:

* ``point_colors`` / ``colored_points``: the per-point RGB the V4 / V5 scripts give their
  point clouds -- car colour per mask (BGR -> RGB, / 255), grey 0.5 for the points no mask
  claimed (V4_BBox_IoU_filtering.py:290-307; V5_ProjectingBBoxes.py:524-550);
* ``depth_overlay``: ``cm(depthMap / max)[..., :3]`` over the segmented image where
  ``depthMap > 0``, ``np.uint8(x * 255)``, RGB -> BGR (seg_with_pointcloud.py:173-180).

``jet_lut`` restates matplotlib's 'jet' lookup table (``LinearSegmentedColormap`` with
N = 256).  matplotlib is absent from this image, so that table is a restatement of a
third-party library from its published algorithm: **parity unpinned** for the LUT values
(everything around them is checked against numpy).
"""
import numpy as np
import torch

from . import lidarfuse as LF
from ._abi import check, ptr

# matplotlib._cm._jet_data
_JET = {'red': ((0.00, 0, 0), (0.35, 0, 0), (0.66, 1, 1), (0.89, 1, 1), (1.00, 0.5, 0.5)),
        'green': ((0.000, 0, 0), (0.125, 0, 0), (0.375, 1, 1), (0.640, 1, 1), (0.910, 0, 0), (1.000, 0, 0)),
        'blue': ((0.00, 0.5, 0.5), (0.11, 1, 1), (0.34, 1, 1), (0.65, 0, 0), (1.00, 0, 0))}


def _lookup_table(N, data):
    """matplotlib.colors._create_lookup_table for gamma = 1."""
    adata = np.array(data, dtype=np.float64)
    x, y0, y1 = adata[:, 0] * (N - 1), adata[:, 1], adata[:, 2]
    xind = (N - 1) * np.linspace(0, 1, N)
    ind = np.searchsorted(x, xind)[1:-1]
    distance = (xind[1:-1] - x[ind - 1]) / (x[ind] - x[ind - 1])
    lut = np.concatenate([[y1[0]], distance * (y0[ind] - y1[ind - 1]) + y1[ind - 1], [y0[-1]]])
    return np.clip(lut, 0.0, 1.0)


def jet_lut(N=256):
    """float64 [N,3]: rows of plt.get_cmap('jet')'s lookup table (RGB)."""
    return np.stack([_lookup_table(N, _JET[c]) for c in ('red', 'green', 'blue')], axis=1)


def point_colors(car_bits, colors_bgr, background=(0.5, 0.5, 0.5), device=None, policy='last', swap_bgr=True):
    """Membership bit fields (``lidarfuse.mask_membership``) -> (rgb f64 [n,3], assigned bool [n]).

    ``colors_bgr``: one (b, g, r) 0..255 tuple per mask, as the scripts keep them; a point
    gets ``[r, g, b] / 255`` of the last mask that contains it, or ``background`` (V4:290-307,
    V5:524-550, vis.py:175-205).  ``policy='first', swap_bgr=False`` is Same_color.py:112-127:
    the first matching mask wins and its colour is used in the order it is stored."""
    if policy not in ('last', 'first'):
        raise ValueError("policy must be 'last' or 'first'")
    ctx = LF.get_ctx(device)
    was_np = not isinstance(car_bits, torch.Tensor)
    bits = LF._dev(car_bits, ctx, torch.int64)
    n = bits.shape[0]
    cols = np.array([[c[2], c[1], c[0]] if swap_bgr else [c[0], c[1], c[2]] for c in colors_bgr],
                    dtype=np.float64).reshape(-1, 3) / 255.0
    d_cols = torch.from_numpy(np.ascontiguousarray(cols)).to(ctx.device)
    bg = np.ascontiguousarray(np.asarray(background, dtype=np.float64))
    rgb = torch.empty((n, 3), dtype=torch.float64, device=ctx.device)
    assigned = torch.empty(n, dtype=torch.uint8, device=ctx.device)
    check(ctx.lib.lf_point_colors(ctx.h, ptr(bits), n, ptr(d_cols), len(cols), bg.ctypes.data, ptr(rgb),
                                  ptr(assigned), 1 if policy == 'first' else 0, ctx.stream), ctx.h)
    assigned = assigned.to(torch.bool)
    return (rgb.cpu().numpy(), assigned.cpu().numpy()) if was_np else (rgb, assigned)


def colored_points(points, TrVeloToRect, masks, colors_bgr, camera, max_depth=30, device=None, policy='last',
                   swap_bgr=True):
    """(points_valid [Nv,3] f32, rgb [Nv,3] f64): the coloured cloud of V4:264-307 /
    V5:508-550 -- projection, the 30 m gate, mask membership and colours, all on the device.
    ``policy='first', swap_bgr=False``: the cloud of Same_color.py:106-132 (car points keep
    their place in the valid order here; the script lists them before the background)."""
    u, v, d = LF.project_points(points, TrVeloToRect, camera, device=device)
    valid = LF.filter_valid(u, v, d, camera, max_depth=max_depth, device=device)
    pts = np.asarray(points)[valid, :3]
    if len(masks) == 0:
        return pts, np.tile(np.asarray([0.5, 0.5, 0.5]), (len(pts), 1))
    bits = LF.mask_membership(u[valid], v[valid], masks, camera, device=device)
    rgb, _ = point_colors(bits, colors_bgr, device=device, policy=policy, swap_bgr=swap_bgr)
    return pts, rgb.cpu().numpy() if isinstance(rgb, torch.Tensor) else rgb


def depth_overlay(depth_map, image_rgb, lut=None, device=None):
    """uint8 [H,W,3] BGR: seg_with_pointcloud.py:176-180 for one car's depth map
    (f64 [H,W], 0 = no point) over the segmented RGB image (uint8 [H,W,3])."""
    ctx = LF.get_ctx(device)
    was_np = not isinstance(depth_map, torch.Tensor)
    dm = LF._dev(depth_map, ctx, torch.float64)
    img = LF._dev(image_rgb, ctx, torch.uint8)
    H, W = dm.shape
    if tuple(img.shape) != (H, W, 3):
        raise ValueError('image must be [H,W,3] uint8')
    lut = jet_lut() if lut is None else np.asarray(lut, dtype=np.float64).reshape(-1, 3)
    d_lut = torch.from_numpy(np.ascontiguousarray(lut)).to(ctx.device)
    dmax = dm.amax().reshape(1)                                    # np.max(depthMap), stays on the device
    out = torch.empty((H, W, 3), dtype=torch.uint8, device=ctx.device)
    check(ctx.lib.lf_depth_overlay(ctx.h, ptr(dm), ptr(dmax), ptr(img), ptr(d_lut), len(lut), H, W, ptr(out),
                                   ctx.stream), ctx.h)
    return out.cpu().numpy() if was_np else out


def depth_overlays(depth_maps, image_rgb, lut=None, device=None):
    """``depth_overlay`` for a stack of depth maps [n,H,W] in one launch: the loop over cars of
    seg_with_pointcloud.py:173-180.  ``image_rgb`` [H,W,3] is shared by all maps, or [n,H,W,3]
    gives every map its own image (a batch of frames).  Maps whose maximum is 0 (no point; the
    script skips them, :174-175) come back as the plain image.  -> uint8 [n,H,W,3] BGR."""
    ctx = LF.get_ctx(device)
    was_np = not isinstance(depth_maps, torch.Tensor)
    dm = LF._dev(depth_maps, ctx, torch.float64)
    img = LF._dev(image_rgb, ctx, torch.uint8)
    if dm.ndim != 3:
        raise ValueError('depth_maps must be [n,H,W]')
    n, H, W = dm.shape
    if tuple(img.shape) == (H, W, 3):
        stride = 0
    elif tuple(img.shape) == (n, H, W, 3):
        stride = H * W * 3
    else:
        raise ValueError('image must be [H,W,3] or [n,H,W,3] uint8')
    lut = jet_lut() if lut is None else np.asarray(lut, dtype=np.float64).reshape(-1, 3)
    d_lut = torch.from_numpy(np.ascontiguousarray(lut)).to(ctx.device)
    out = torch.empty((n, H, W, 3), dtype=torch.uint8, device=ctx.device)
    if n:
        dmax = dm.reshape(n, -1).amax(dim=1).contiguous()
        for a in range(0, n, 65535):
            b = min(n, a + 65535)
            check(ctx.lib.lf_depth_overlay_batch(ctx.h, ptr(dm[a:]), ptr(dmax[a:]), ptr(img[a:] if stride else img), stride,
                                                 ptr(d_lut), len(lut), b - a, H, W, ptr(out[a:]), ctx.stream), ctx.h)
    return out.cpu().numpy() if was_np else out
