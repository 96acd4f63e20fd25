"""Function-level surface of the reference's V1-V5 / cvs_erosion / seg_with_pointcloud
scripts, served by the sm_100a kernels behind include/lidarfuse.h.

Same names, argument order, defaults and return types as the reference
functions they replace (each docstring cites file:line).  numpy in -> numpy
out (the reference's dtypes); CUDA torch tensors in -> CUDA tensors out.
There is no CPU path: every call needs the CUDA library and a CUDA device.
"""
import numpy as np
import torch

from . import _abi
from ._abi import check, ptr

_contexts = {}


class _Ctx:
    """One lf_ctx per device (not thread-safe, like the C object)."""

    def __init__(self, device):
        import ctypes as C
        self.lib = _abi.load()
        if not torch.cuda.is_available():
            raise _abi.LidarFuseError('no CUDA device: lidarfuse has no CPU fallback')
        self.device = torch.device('cuda', device)
        h = C.c_void_p()
        check(self.lib.lf_create(C.byref(h), device))
        self.h = h
        self._cam_key = None

    def set_camera(self, TrVeloToRect, K, W, H):
        T = np.ascontiguousarray(np.asarray(TrVeloToRect, dtype=np.float64).reshape(4, 4))
        Kc = np.ascontiguousarray(np.asarray(K, dtype=np.float64).reshape(3, 4))
        key = (T.tobytes(), Kc.tobytes(), int(W), int(H))
        if key != self._cam_key:
            check(self.lib.lf_set_camera(self.h, T.ctypes.data, Kc.ctypes.data, int(W), int(H)), self.h)
            self._cam_key = key

    @property
    def stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream


def get_ctx(device=None):
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    if isinstance(device, torch.device):
        device = device.index if device.index is not None else torch.cuda.current_device()
    if device not in _contexts:
        _contexts[device] = _Ctx(device)
    return _contexts[device]


def _is_np(x):
    return not isinstance(x, torch.Tensor)


def _dev(x, ctx, dtype=None):
    """numpy / tensor -> contiguous tensor on the context's device."""
    if isinstance(x, torch.Tensor):
        t = x.to(ctx.device)
    else:
        t = torch.from_numpy(np.ascontiguousarray(x)).to(ctx.device)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def _identity4():
    return np.eye(4)


# --------------------------------------------------------------------------
# A3 / A4  projection and gate
# --------------------------------------------------------------------------

def _points4(points, ctx):
    """[N,4] float32 device tensor (a [N,3] input gets a zero fourth column)."""
    t = _dev(points, ctx, torch.float32)
    if t.ndim != 2 or t.shape[1] not in (3, 4):
        raise ValueError('points must be [N,4] (x, y, z, reflectance) or [N,3]')
    if t.shape[1] == 3:
        t = torch.cat([t, torch.zeros_like(t[:, :1])], dim=1).contiguous()
    return t


def project_points(points, TrVeloToRect, camera, max_depth=50, return_pix=False, device=None):
    """velodyne [N,4] f32 -> (u int64 [N], v int64 [N], depth f64 [N]).

    Replaces cvs_erosion.py:342-346 (V1:340-344, V4:258-262, V5:467-471,
    seg_with_pointcloud.py:132-139): homogeneous copy, 4x4 f64 transform,
    ``camera.cam2image``.  ``return_pix`` adds the gated flat pixel index
    (v*W+u, -1 when the point fails the gate of cvs_erosion.py:359)."""
    ctx = get_ctx(device)
    ctx.set_camera(TrVeloToRect, camera.K, camera.width, camera.height)
    was_np = _is_np(points)
    pts = _points4(points, ctx)
    n = pts.shape[0]
    u = torch.empty(n, dtype=torch.int64, device=ctx.device)
    v = torch.empty(n, dtype=torch.int64, device=ctx.device)
    d = torch.empty(n, dtype=torch.float64, device=ctx.device)
    pix = torch.empty(n, dtype=torch.int32, device=ctx.device) if return_pix else None
    check(ctx.lib.lf_project(ctx.h, ptr(pts), n, float(max_depth), ptr(u), ptr(v), ptr(d), ptr(pix), ctx.stream),
          ctx.h)
    out = (u, v, d) + ((pix,) if return_pix else ())
    if was_np:
        out = tuple(t.cpu().numpy() for t in out)
    return out


def gated_pixels(points, TrVeloToRect, camera, max_depth=50, device=None):
    """Flat pixel index ``v*W + u`` (int32) of every point that passes the gate of cvs_erosion.py:359,
    -1 for the others: projection + ``filter_valid`` in one pass when ``u, v, depth`` themselves are
    not needed (the kernel then lets the fp32 pre-gate answer for points that certainly fail;
    the indices are those of ``project_points(..., return_pix=True)``)."""
    ctx = get_ctx(device)
    ctx.set_camera(TrVeloToRect, camera.K, camera.width, camera.height)
    was_np = _is_np(points)
    pts = _points4(points, ctx)
    n = pts.shape[0]
    pix = torch.empty(n, dtype=torch.int32, device=ctx.device)
    check(ctx.lib.lf_project(ctx.h, ptr(pts), n, float(max_depth), None, None, None, ptr(pix), ctx.stream), ctx.h)
    return pix.cpu().numpy() if was_np else pix


def filter_valid(u, v, depth, camera, max_depth=50, device=None):
    """Boolean gate of cvs_erosion.py:359 (max_depth 30 in V4:275, V5:508,
    seg_with_pointcloud.py:154-158) on already projected arrays (one launch, ``lf_filter_valid``)."""
    ctx = get_ctx(device)
    was_np = _is_np(u)
    uu, vv, dd = _dev(u, ctx, torch.int64), _dev(v, ctx, torch.int64), _dev(depth, ctx, torch.float64)
    n = uu.shape[0]
    if vv.shape[0] != n or dd.shape[0] != n:
        raise ValueError('u, v, depth must have equal length')
    valid = torch.empty(n, dtype=torch.uint8, device=ctx.device)
    check(ctx.lib.lf_filter_valid(ctx.h, ptr(uu), ptr(vv), ptr(dd), n, int(camera.width), int(camera.height),
                                  float(max_depth), ptr(valid), ctx.stream), ctx.h)
    valid = valid.view(torch.bool)
    return valid.cpu().numpy() if was_np else valid


# --------------------------------------------------------------------------
# A5  structuring elements, erosion, packing
# --------------------------------------------------------------------------

MORPH_RECT, MORPH_CROSS, MORPH_ELLIPSE = 0, 1, 2


def structuring_element(shape, ksize):
    """k x k uint8 element with OpenCV's ``getStructuringElement`` geometry
    (cvs_erosion.py:98 uses MORPH_ELLIPSE): every row is one centred span;
    for the ellipse the half-width of row i is round(c*sqrt(1 - dy^2/r^2))."""
    k = int(ksize)
    if k < 1:
        raise ValueError('kernel size must be positive')
    se = np.zeros((k, k), np.uint8)
    r = c = k // 2
    if shape == MORPH_RECT or k == 1:
        se[:] = 1
    elif shape == MORPH_CROSS:
        se[r, :] = 1
        se[:, c] = 1
    elif shape == MORPH_ELLIPSE:
        inv_r2 = 1.0 / (r * r) if r else 0.0
        for i in range(k):
            dy = i - r
            if abs(dy) <= r:
                dx = int(np.rint(c * np.sqrt((r * r - dy * dy) * inv_r2)))
                se[i, max(c - dx, 0):min(c + dx + 1, k)] = 1
    else:
        raise ValueError('unknown structuring element shape')
    return se


class PackedMasks:
    """Bit-packed instance masks on the device: uint32 [M, H, ceil(W/32)].

    ``raw=True``: the words are binarised masks that have NOT been eroded yet (packed on the
    sending side by ``io.pack_masks`` / ``lf_pack_masks_host``); ``FrameBatch.run`` and
    ``erode_pack`` then erode them on the device from the packed words (mode 3).  Otherwise
    they are taken as the final masks the point pass reads."""

    def __init__(self, words, H, W, bbox=None, raw=False):
        self.words, self.H, self.W = words, H, W
        self.bbox = bbox        # int32 [M,4] {2^20-x0, 2^20-y0, x1+1, y1+1} (zeros = empty) or None
        self.raw = bool(raw)

    def __len__(self):
        return self.words.shape[0]

    def unpack(self, ctx=None):
        ctx = ctx or get_ctx(self.words.device)
        m = len(self)
        out = torch.empty((m, self.H, self.W), dtype=torch.uint8, device=self.words.device)
        check(ctx.lib.lf_unpack_masks(ctx.h, ptr(self.words), m, self.H, self.W, ptr(out), ctx.stream), ctx.h)
        return out


def _mask_mode(t, float_threshold):
    if t.dtype == torch.uint8:
        return _abi.MODE_U8
    if t.dtype == torch.bool:
        return _abi.MODE_U8
    if t.dtype == torch.float32:
        return _abi.MODE_F32_GT_HALF if float_threshold == 'gt_half' else _abi.MODE_F32_GE1
    raise ValueError('masks must be uint8, bool or float32, got %s' % t.dtype)


def erode_pack(masks, erosion_kernel_size=3, erosion_iterations=1, shape=MORPH_ELLIPSE, se=None,
               float_threshold='ge_one', device=None):
    """[M,H,W] masks -> PackedMasks, eroded ``erosion_iterations`` times.

    uint8/bool: non-zero is set.  float32: ``>= 1.0`` is set, which is what the
    reference's uint8 truncations leave (cvs_erosion.py:101-103, 153-154;
    V2:197-200); ``float_threshold='gt_half'`` gives seg_with_pointcloud.py:167.
    ``erosion_kernel_size=0`` or ``erosion_iterations=0`` packs without erosion."""
    ctx = get_ctx(device)
    if isinstance(masks, (list, tuple)):
        masks = np.stack(masks) if len(masks) else np.zeros((0, 1, 1), np.uint8)
    if isinstance(masks, PackedMasks):
        # packed on the sending side: erode from the words (1/8 of the uint8 traffic)
        t, mode = masks.words.contiguous(), _abi.MODE_PACKED
        M, H, W = len(masks), masks.H, masks.W
        if tuple(t.shape) != (M, H, (W + 31) // 32) or t.dtype not in (torch.int32, torch.uint32):
            raise ValueError('PackedMasks.words must be 32-bit [M, H, ceil(W/32)]')
    else:
        t = _dev(masks, ctx)
        if t.ndim != 3:
            raise ValueError('masks must be [M,H,W]')
        mode = _mask_mode(t, float_threshold)
        if t.dtype == torch.bool:
            t = t.view(torch.uint8)
        M, H, W = t.shape
    wpr = (W + 31) // 32
    words = torch.empty((M, H, wpr), dtype=torch.int32, device=ctx.device)
    iters = int(erosion_iterations) if erosion_kernel_size else 0
    if iters > 0:
        se_arr = np.ascontiguousarray(se, dtype=np.uint8) if se is not None else \
            structuring_element(shape, erosion_kernel_size)
        k = se_arr.shape[0]
        if se_arr.shape != (k, k):
            raise ValueError('structuring element must be square')
        se_ptr = se_arr.ctypes.data
    else:
        se_arr, k, se_ptr = None, 1, None
    bbox = torch.zeros((M, 4), dtype=torch.int32, device=ctx.device)
    check(ctx.lib.lf_erode_pack(ctx.h, ptr(t), mode, M, H, W, se_ptr, k, iters, ptr(words), ptr(bbox), ctx.stream),
          ctx.h)
    return PackedMasks(words, H, W, bbox)


def erode_masks(masks, erosion_kernel_size=3, erosion_iterations=1, shape=MORPH_ELLIPSE, device=None):
    """float32 [M,H,W] 0/1 masks -> float32 eroded masks, the post-YOLO half of
    ``image_segmentation_with_erosion`` (cvs_erosion.py:96-106; V3:81-97)."""
    was_np = _is_np(masks)
    ctx = get_ctx(device)
    pm = erode_pack(masks, erosion_kernel_size, erosion_iterations, shape, device=device)
    out = pm.unpack(ctx).to(torch.float32)
    return out.cpu().numpy() if was_np else out


# --------------------------------------------------------------------------
# A6  point-in-mask gather
# --------------------------------------------------------------------------

def resize_masks(masks, size, device=None):
    """``cv2.resize(mask.astype(np.uint8), (W, H))`` for a stack of masks, on the device
    and bit-exact with OpenCV's default INTER_LINEAR (cvs_erosion.py:153; V4:293; V5:531).

    masks [M,h,w]: uint8 / bool as they are, float32 truncated like ``astype(np.uint8)``
    (only values >= 1 survive; the reference's masks are 0.0 / 1.0).  ``size`` = (W, H),
    OpenCV's argument order, or a camera.  Returns uint8 [M,H,W] (numpy in -> numpy out)."""
    ctx = get_ctx(device)
    W, H = (size.width, size.height) if hasattr(size, 'width') else (int(size[0]), int(size[1]))
    was_np = _is_np(masks)
    if isinstance(masks, (list, tuple)):
        masks = np.stack(masks) if len(masks) else np.zeros((0, 1, 1), np.uint8)
    t = _dev(masks, ctx)
    if t.ndim != 3:
        raise ValueError('masks must be [M,h,w]')
    if t.dtype == torch.bool:
        t = t.view(torch.uint8)
    elif t.dtype != torch.uint8:
        t = t.clamp(0, 255).to(torch.uint8)              # astype(np.uint8) truncation of 0.0 / 1.0 masks
    M, h, w = t.shape
    out = torch.empty((M, H, W), dtype=torch.uint8, device=ctx.device)
    check(ctx.lib.lf_resize_masks(ctx.h, ptr(t), M, h, w, ptr(out), H, W, ctx.stream), ctx.h)
    return out.cpu().numpy() if was_np else out


def resize_pack(pm, size, device=None, out=None, bbox=None):
    """``PackedMasks`` at any resolution -> ``PackedMasks`` at ``size`` (camera or (W, H)) whose bits are
    ``cv2.resize(mask01, (W, H)) != 0``: the resize of cvs_erosion.py:153 followed by its ``> 0.5``,
    straight from packed words to packed words (``lf_resize_pack``).  ``out`` / ``bbox``: optional
    preallocated destination words [M,H,ceil(W/32)] and zeroed bbox records [M,4]."""
    ctx = get_ctx(device)
    W, H = (size.width, size.height) if hasattr(size, 'width') else (int(size[0]), int(size[1]))
    M = len(pm)
    words = out if out is not None else torch.empty((M, H, (W + 31) // 32), dtype=torch.int32, device=ctx.device)
    bb = bbox if bbox is not None else torch.zeros((M, 4), dtype=torch.int32, device=ctx.device)
    check(ctx.lib.lf_resize_pack(ctx.h, ptr(pm.words), M, pm.H, pm.W, ptr(words), ptr(bb), H, W, ctx.stream), ctx.h)
    return PackedMasks(words, H, W, bb)


def _as_packed(masks, camera, ctx, float_threshold='ge_one'):
    if isinstance(masks, PackedMasks):
        pm = masks
        if pm.raw:
            pm = PackedMasks(pm.words, pm.H, pm.W)       # used as they are: no erosion asked for here
        if len(pm) and (pm.H != camera.height or pm.W != camera.width):
            return resize_pack(pm, camera, device=ctx.device)
        return pm
    if isinstance(masks, (list, tuple)):
        masks = np.stack(masks) if len(masks) else np.zeros((0, camera.height, camera.width), np.uint8)
    shape = tuple(masks.shape[1:]) if len(masks) else (camera.height, camera.width)
    if len(masks) and shape != (camera.height, camera.width):
        # cvs_erosion.py:153: cv2.resize(mask.astype(np.uint8), (W, H)), then "> 0.5"
        if not isinstance(masks, torch.Tensor):
            masks = np.asarray(masks)
            if masks.dtype != np.uint8 and masks.dtype != np.bool_:
                masks = masks.astype(np.uint8)
        masks = resize_masks(_dev(masks, ctx), camera, device=ctx.device)
    return erode_pack(masks, 0, 0, float_threshold=float_threshold, device=ctx.device)


def mask_membership(u_valid, v_valid, masks, camera, device=None):
    """uint64 bit field per point: bit c = point lies in mask c (<= 64 masks per call)."""
    ctx = get_ctx(device)
    pm = _as_packed(masks, camera, ctx)
    uu, vv = _dev(u_valid, ctx, torch.int64), _dev(v_valid, ctx, torch.int64)
    n = uu.shape[0]
    bits = torch.zeros(n, dtype=torch.int64, device=ctx.device)
    M = len(pm)
    if M > _abi.MAX_CARS:
        raise ValueError('more than %d masks: split the call' % _abi.MAX_CARS)
    check(ctx.lib.lf_mask_gather(ctx.h, ptr(pm.words), M, camera.height, camera.width, ptr(uu), ptr(vv), n,
                                 ptr(bits), ctx.stream), ctx.h)
    return bits


def extract_car_points_by_mask(points_valid, u_valid, v_valid, masks, camera, device=None):
    """List of M arrays [n_c,3]: the gated points whose pixel lies in mask c
    (cvs_erosion.py:148-162; V1:186-208).  An empty selection is a (0,3)
    float64 array, as in the reference."""
    ctx = get_ctx(device)
    was_np = _is_np(points_valid)
    M = len(masks)
    if M == 0:
        return []
    if M > _abi.MAX_CARS and not isinstance(masks, PackedMasks):
        # one 64-bit membership word per point and call: more masks go in groups of 64
        out = []
        for g in range(0, M, _abi.MAX_CARS):
            out += extract_car_points_by_mask(points_valid, u_valid, v_valid, masks[g:g + _abi.MAX_CARS], camera, device)
        return out
    bits = mask_membership(u_valid, v_valid, masks, camera, device)
    pts = _dev(points_valid, ctx)
    if pts.dtype not in (torch.float32, torch.float64):
        pts = pts.to(torch.float64)
    if pts.ndim != 2 or pts.shape[1] < 3 or pts.shape[0] != bits.shape[0]:
        raise ValueError('points_valid must be [n,>=3] with one row per (u, v)')
    n = pts.shape[0]
    # counts of all masks, then ONE order-preserving compaction of every car's points (CSR), two read-backs
    scratch = torch.empty(max(int(ctx.lib.lf_car_scratch_elems(n, M)), 2), dtype=torch.int32, device=ctx.device)
    counts = torch.zeros(M, dtype=torch.int32, device=ctx.device)
    check(ctx.lib.lf_car_counts(ctx.h, ptr(bits), n, M, ptr(scratch), ptr(counts), ctx.stream), ctx.h)
    h_counts = counts.cpu().numpy()
    total = int(h_counts.sum())
    out = torch.empty((max(total, 1), 3), dtype=pts.dtype, device=ctx.device)
    if total:
        check(ctx.lib.lf_car_compact(ctx.h, ptr(bits), n, M, ptr(pts), int(pts.dtype == torch.float64), pts.shape[1],
                                     ptr(scratch), ptr(out), None, ctx.stream), ctx.h)
    data = out.cpu().numpy() if was_np else out
    sets, o = [], 0
    for c in range(M):
        k = int(h_counts[c])
        if k == 0:
            sets.append(np.array([]).reshape(0, 3) if was_np else
                        torch.zeros((0, 3), dtype=torch.float64, device=ctx.device))
        else:
            sets.append(data[o:o + k])
        o += k
    return sets


# --------------------------------------------------------------------------
# A7  boxes (host numpy float64, as in the reference)
# --------------------------------------------------------------------------

def filter_visible_bboxes(bboxes_3d, camera):
    """Keep boxes with >= 2 corners in front of the camera (depth > 0.1) and
    inside the image (cvs_erosion.py:55-74)."""
    keep = []
    for bbox in bboxes_3d:
        if 'corners_cam0' not in bbox:
            continue
        corners = np.array(bbox['corners_cam0'])
        u, v, depth = camera.cam2image(corners.T)
        ok = (depth > 0.1) & (u >= 0) & (u < camera.width) & (v >= 0) & (v < camera.height)
        if np.sum(ok) >= 2:
            keep.append(bbox)
    return keep


def transform_bboxes_to_velodyne(bboxes_3d, TrVeloToCam):
    """Adds ``corners_velo`` (list) to every box (cvs_erosion.py:41-52)."""
    cam_to_velo = np.linalg.inv(TrVeloToCam)
    for bbox in bboxes_3d:
        if 'corners_cam0' in bbox:
            c = np.array(bbox['corners_cam0'])
            homo = np.hstack([c, np.ones((c.shape[0], 1))])
            bbox['corners_velo'] = np.matmul(cam_to_velo, homo.T).T[:, :3].tolist()
    return bboxes_3d


def boxes_to_device(bboxes_3d, ctx):
    """list of dicts with corners_velo -> (lf_box tensor on device, host struct array)."""
    corners = np.array([b['corners_velo'] for b in bboxes_3d], dtype=np.float64).reshape(-1, 8, 3)
    host = _abi.prepare_boxes(corners)
    dev = torch.from_numpy(host.view(np.uint8).reshape(-1, _abi.BOX_DTYPE.itemsize).copy()).to(ctx.device)
    return dev, host


# --------------------------------------------------------------------------
# A8  containment
# --------------------------------------------------------------------------

def _in_box(points, bbox_corners, oriented, device=None):
    if len(points) == 0:
        return np.array([])
    ctx = get_ctx(device)
    was_np = _is_np(points)
    pts = _dev(points, ctx)
    if pts.dtype not in (torch.float32, torch.float64):
        pts = pts.to(torch.float64)
    if pts.ndim != 2 or pts.shape[1] < 3:
        raise ValueError('points must be [n,>=3]')
    host = _abi.prepare_boxes(np.asarray(bbox_corners, dtype=np.float64).reshape(1, 8, 3))
    box = torch.from_numpy(host.view(np.uint8).copy()).to(ctx.device)
    out = torch.empty(pts.shape[0], dtype=torch.uint8, device=ctx.device)
    check(ctx.lib.lf_points_in_box(ctx.h, ptr(pts), int(pts.dtype == torch.float64), pts.shape[0],
                                   pts.shape[1], ptr(box), int(oriented), ptr(out), ctx.stream), ctx.h)
    res = out.to(torch.bool)
    return res.cpu().numpy() if was_np else res


def oriented_point_in_bbox(points, bbox_corners, device=None):
    """Closed slab test along c1-c0, c3-c0, c4-c0 (cvs_erosion.py:114-145)."""
    return _in_box(points, bbox_corners, True, device)


def point_in_bbox(points, bbox_corners, device=None):
    """Axis-aligned test against min/max of the corners (V1:118-139)."""
    return _in_box(points, bbox_corners, False, device)


# --------------------------------------------------------------------------
# A9  per-car statistics
# --------------------------------------------------------------------------

def inside_count_matrix(car_point_sets, bboxes_3d, use_oriented=True, device=None):
    """int64 [M,B]: points of car c inside box b -- the double loop of
    calculate_car_point_statistics (cvs_erosion.py:175-195) in one launch."""
    ctx = get_ctx(device)
    M, B = len(car_point_sets), len(bboxes_3d)
    if M == 0 or B == 0:
        return np.zeros((M, B), np.int64)
    boxes_with = [i for i, b in enumerate(bboxes_3d) if 'corners_velo' in b]
    sets = [_dev(s, ctx) if len(s) else None for s in car_point_sets]
    lens = [0 if s is None else s.shape[0] for s in sets]
    out = np.zeros((M, B), np.int64)
    if sum(lens) == 0 or not boxes_with:
        return out
    f64 = any(s is not None and s.dtype == torch.float64 for s in sets)
    dt = torch.float64 if f64 else torch.float32
    pts = torch.cat([s[:, :3].to(dt) for s in sets if s is not None]).contiguous()
    set_id = torch.repeat_interleave(torch.arange(M, dtype=torch.int32, device=ctx.device),
                                     torch.tensor(lens, device=ctx.device))
    dev_boxes, _ = boxes_to_device([bboxes_3d[i] for i in boxes_with], ctx)
    Bw = len(boxes_with)
    inside = torch.zeros((M, Bw), dtype=torch.int32, device=ctx.device)
    check(ctx.lib.lf_count_in_boxes(ctx.h, ptr(pts), int(f64), pts.shape[0], 3, ptr(set_id), ptr(dev_boxes), Bw,
                                    int(use_oriented), ptr(inside), ctx.stream), ctx.h)
    out[:, boxes_with] = inside.cpu().numpy()
    return out


def _stats_from_counts(totals, cnt, colors, min_points, extra=None):
    """First strict maximum from 0; matched iff best >= min_points
    (cvs_erosion.py:181-225).  Cars without points give no row."""
    stats = []
    for ci, total in enumerate(totals):
        total = int(total)
        if total == 0:
            continue
        best, best_idx = 0, -1
        row = cnt[ci]
        if row.size:
            j = int(np.argmax(row))              # first maximum
            if row[j] > 0:
                best, best_idx = row[j], j
        if best_idx >= 0 and best >= min_points:
            outside = total - best
            s = {'car_id': ci, 'matched_bbox_id': best_idx, 'total_points': total,
                 'points_inside_bbox': best, 'points_outside_bbox': outside,
                 'inside_percentage': (best / total) * 100, 'outside_percentage': (outside / total) * 100,
                 'color': colors[ci]}
        else:
            s = {'car_id': ci, 'matched_bbox_id': -1, 'total_points': total, 'points_inside_bbox': 0,
                 'points_outside_bbox': total, 'inside_percentage': 0.0, 'outside_percentage': 100.0,
                 'color': colors[ci]}
        if extra is not None:
            s.update(extra(ci, s))
        stats.append(s)
    return stats


def calculate_car_point_statistics(car_point_sets, bboxes_3d, colors, min_points=10, use_oriented=True,
                                   device=None, return_points=False):
    """Per-car dicts {car_id, matched_bbox_id, total_points, points_inside_bbox,
    points_outside_bbox, inside_percentage, outside_percentage, color}
    (cvs_erosion.py:165-229; V3:320-428 adds ``use_oriented``).

    ``return_points=True`` gives the V2 / V3 form of the dicts (V3_point_cloud_with_erosion.py:386-398,
    416-428; V2_point_cloud_without_erosion.py has the same body): additionally ``corners_velo`` (8x3
    array of the matched box), ``inside_mask`` (bool per car point, against the matched box) and
    ``car_points``; None / None / the points for unmatched cars -- what
    ``create_colored_point_cloud_with_bbox_analysis`` (V3:471-515) and ``print_summary_statistics``
    consume."""
    if not bboxes_3d or len(car_point_sets) == 0:
        return []
    cnt = inside_count_matrix(car_point_sets, bboxes_3d, use_oriented, device)
    extra = None
    if return_points:
        def extra(ci, s):
            bi = s['matched_bbox_id']
            if bi < 0:
                return {'corners_velo': None, 'inside_mask': None, 'car_points': car_point_sets[ci]}
            corners = np.array(bboxes_3d[bi]['corners_velo'])
            return {'corners_velo': corners, 'inside_mask': _in_box(car_point_sets[ci], corners, use_oriented, device),
                    'car_points': car_point_sets[ci]}
    return _stats_from_counts([len(s) for s in car_point_sets], cnt, colors, min_points, extra)


def summary_statistics_text(car_statistics):
    """The text ``print_summary_statistics`` prints (V3_point_cloud_with_erosion.py:431-468;
    V2_point_cloud_without_erosion.py:409-446), as one string."""
    if not car_statistics:
        return '\nNo car statistics to display.\n'
    matched = [s for s in car_statistics if s['matched_bbox_id'] >= 0]
    rule = '-' * 60
    lines = ['', '=' * 60, '{:^60}'.format('SUMMARY STATISTICS'), '=' * 60,
             'Total cars detected: {}'.format(len(car_statistics)),
             'Successfully matched: {}'.format(len(matched)),
             'Unmatched: {}'.format(len(car_statistics) - len(matched))]
    if matched:
        row = '{:<8} {:<8} {:<8} {:<8} {:<8} {:<10}'
        lines += ['', row.format('Car ID', 'BBox ID', 'Total', 'Inside', 'Outside', 'Inside %'), rule]
        num = '{:<8} {:<8} {:<8} {:<8} {:<8} {:<10.1f}'
        for s in matched:
            lines.append(num.format(s['car_id'], s['matched_bbox_id'], s['total_points'], s['points_inside_bbox'],
                                    s['points_outside_bbox'], s['inside_percentage']))
        tp = sum(s['total_points'] for s in matched)
        ti = sum(s['points_inside_bbox'] for s in matched)
        to = sum(s['points_outside_bbox'] for s in matched)
        lines += [rule, num.format('TOTAL', '', tp, ti, to, (ti / tp * 100) if tp > 0 else 0)]
    return '\n'.join(lines) + '\n'


def print_summary_statistics(car_statistics):
    """Summary table of the per-car statistics, printed as the reference prints it (V3:431-468)."""
    print(summary_statistics_text(car_statistics), end='')


def match_car_points_to_bboxes(car_point_sets, bboxes_3d, colors, min_points=10, use_oriented=True,
                               device=None):
    """(corners_velo, rgb colour, count) per matched car (V1:211-265)."""
    pairs = []
    if not bboxes_3d or len(car_point_sets) == 0:
        return pairs
    cnt = inside_count_matrix(car_point_sets, bboxes_3d, use_oriented, device)
    for ci, pts in enumerate(car_point_sets):
        if len(pts) == 0:
            continue
        best, best_idx = 0, -1
        for bi in range(cnt.shape[1]):
            if cnt[ci, bi] > best and cnt[ci, bi] >= min_points:
                best, best_idx = cnt[ci, bi], bi
        if best_idx >= 0:
            corners = np.array(bboxes_3d[best_idx]['corners_velo'])
            col = np.array([colors[ci][2], colors[ci][1], colors[ci][0]]) / 255.0
            pairs.append((corners, col, best))
    return pairs


def extension_metrics(inside, total, n_valid_in_box):
    """Point-level precision / recall / IoU / match ratio (README.md:14, 38
    name them; the reference has no code for them -- SURVEY 0.1 D3)."""
    tp, fp, fn = float(inside), float(total - inside), float(n_valid_in_box - inside)
    return {'precision': tp / (tp + fp) if tp + fp > 0 else 0.0,
            'recall': tp / (tp + fn) if tp + fn > 0 else 0.0,
            'point_iou': tp / (tp + fp + fn) if tp + fp + fn > 0 else 0.0,
            'match_ratio': (tp / fp) if fp > 0 else float('inf')}


def generate_colors(n):
    """The car colours every script uses (cvs_erosion.py:107)."""
    return [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(n)]


# --------------------------------------------------------------------------
# fused batch path: A3+A4+A5+A6+A8+A9 for F frames in four launches
# --------------------------------------------------------------------------

class FrameBatch:
    """Device-resident CSR batch of frames.

    points [sum N,4] f32, masks [sum M,H,W] uint8/float32 (or PackedMasks),
    boxes lf_box [sum B]; per-frame counts give the CSR offsets.  Every buffer
    the step writes is allocated here once, so ``run`` launches kernels only
    (and can be captured in a CUDA graph)."""

    def __init__(self, points, n_points, masks, n_masks, boxes, n_boxes, camera, TrVeloToRect, device=None,
                 want_car_bits=False):
        """Any number of frames; any number of masks per frame (beyond 64 the point pass runs once per group
        of 64 masks, inside ``lf_fuse_batch``) unless ``want_car_bits`` (one 64-bit word per point)."""
        ctx = get_ctx(device)
        self.ctx, self.camera = ctx, camera
        self.TrVeloToRect = np.asarray(TrVeloToRect, dtype=np.float64)
        self.F = len(n_points)
        n_points = np.asarray(n_points, np.int64)
        n_masks = np.asarray(n_masks, np.int64)
        n_boxes = np.asarray(n_boxes, np.int64)
        if not (len(n_masks) == len(n_boxes) == self.F):
            raise ValueError('per-frame count arrays must have equal length')
        self.points = _points4(points, ctx) if not isinstance(points, torch.Tensor) or points.shape[-1] != 4 \
            or points.dtype != torch.float32 or not points.is_cuda else points.contiguous()
        if self.points.shape[0] != int(n_points.sum()):
            raise ValueError('points rows != sum of per-frame counts')
        self.masks = masks if isinstance(masks, PackedMasks) else (None if masks is None else _dev(masks, ctx))
        n_m = len(self.masks) if self.masks is not None else 0
        if n_m != int(n_masks.sum()):
            raise ValueError('mask count != sum of per-frame counts')
        self._h_corners = None
        if isinstance(boxes, torch.Tensor):
            self.boxes = boxes.to(ctx.device).contiguous()
        else:
            if not (isinstance(boxes, np.ndarray) and boxes.dtype == _abi.BOX_DTYPE):
                self._h_corners = np.asarray(boxes, dtype=np.float64).reshape(-1, 8, 3)
            host = boxes if isinstance(boxes, np.ndarray) and boxes.dtype == _abi.BOX_DTYPE else \
                _abi.prepare_boxes(boxes)
            self.boxes = torch.from_numpy(np.ascontiguousarray(host).view(np.uint8).reshape(-1, 192).copy()) \
                .to(ctx.device)
        if self.boxes.numel() // 192 != int(n_boxes.sum()):
            raise ValueError('box count != sum of per-frame counts')

        def csr(a, dt):
            return np.concatenate([[0], np.cumsum(a)]).astype(dt)
        self.h_pt_off, self.h_mask_off, self.h_box_off = csr(n_points, np.int64), csr(n_masks, np.int32), \
            csr(n_boxes, np.int32)
        self.h_cnt_off = csr(n_masks * n_boxes, np.int64)
        self.pt_off = torch.from_numpy(self.h_pt_off).to(ctx.device)
        self.mask_off = torch.from_numpy(self.h_mask_off).to(ctx.device)
        self.box_off = torch.from_numpy(self.h_box_off).to(ctx.device)
        self.cnt_off = torch.from_numpy(self.h_cnt_off).to(ctx.device)
        self.max_n = int(n_points.max()) if self.F else 0
        self.max_m = int(n_masks.max()) if self.F else 0
        self.max_b = int(n_boxes.max()) if self.F else 0
        sm, sb, sc = int(n_masks.sum()), int(n_boxes.sum()), int(self.h_cnt_off[-1])
        # one allocation for everything the step zeroes and everything it hands back:
        # mask bbox [sm,4] | total | n_in_box | n_valid | inside | best_idx | best_cnt | pad | agg(int64 x8)
        self._o = 4 * sm
        self._ob = self._o + sm + sb + self.F + sc        # best_idx, then best_cnt
        self._nz = self._ob + 2 * sm
        self._nz += self._nz & 1
        self.counters = torch.zeros(self._nz + 16, dtype=torch.int32, device=ctx.device)
        o = self._o
        self.mask_bbox = self.counters[:o].view(max(sm, 0), 4) if sm else None
        self.total = self.counters[o:o + sm]
        self.n_in_box = self.counters[o + sm:o + sm + sb]
        self.n_valid = self.counters[o + sm + sb:o + sm + sb + self.F]
        self.inside = self.counters[o + sm + sb + self.F:o + sm + sb + self.F + sc]
        self.best_idx = self.counters[self._ob:self._ob + sm]
        self.best_cnt = self.counters[self._ob + sm:self._ob + 2 * sm]
        self.agg = self.counters[self._nz:].view(torch.int64)
        self.car_bits = torch.empty(self.points.shape[0], dtype=torch.int64, device=ctx.device) \
            if want_car_bits else None
        self._native = None
        if isinstance(self.masks, PackedMasks) and self.masks.raw:
            H, W = camera.height, camera.width
            if n_m and (self.masks.H, self.masks.W) != (H, W):
                raise ValueError('raw packed masks must be at camera resolution')
            self.packed = PackedMasks(torch.empty((n_m, H, (W + 31) // 32), dtype=torch.int32,
                                                  device=ctx.device), H, W)
        elif isinstance(self.masks, PackedMasks):
            if n_m and (self.masks.H, self.masks.W) != (camera.height, camera.width):
                # final packed masks are read with the camera's row stride: another size would be read out of bounds
                raise ValueError('final PackedMasks must be at camera resolution (%dx%d); use resize_pack first'
                                 % (camera.width, camera.height))
            self.packed = self.masks
        else:
            H, W = camera.height, camera.width
            if self.masks is not None and n_m and tuple(self.masks.shape[1:]) != (H, W):
                # masks below camera resolution (YOLO without retina_masks): eroded where they
                # are, THEN resized, as cvs_erosion.py:96-106 / :153 do -- see run()
                self._native = self.masks
                h, w = self.masks.shape[1:]
                self._native_words = torch.empty((n_m, h, (w + 31) // 32), dtype=torch.int32, device=ctx.device)
            self.packed = PackedMasks(torch.empty((n_m, H, (W + 31) // 32), dtype=torch.int32,
                                                  device=ctx.device), H, W)
        self.launches_per_run = 0

    def run(self, erosion_kernel_size=3, erosion_iterations=1, max_depth=50, min_points=10, use_oriented=True,
            shape=MORPH_ELLIPSE, float_threshold='ge_one', mark=None, one_call=False, exact=False, sparse_rows=False):
        """memset + erode/pack + fused point pass + per-car decision, all async
        on the current stream.  Returns nothing; read ``results()`` / ``table()``.

        ``sparse_rows=True`` (uint8 / float32 / raw packed masks eroded by this call): the erode kernel writes only
        the rows of the packed masks that have a set pixel plus a row map, and the point pass consults the map
        (``lf_erode_pack_rows`` / ``lf_fuse_batch_rows``) -- same counters, nine tenths of the packed words never
        travel to HBM; ``self.packed.words`` is then NOT a complete mask (rows without a set pixel keep whatever
        they held), so leave it off when something else reads the packed masks afterwards.

        ``one_call=True`` issues erode + point pass + decision through the single ABI entry
        ``lf_process_batch`` instead of the three calls (``mark(name)`` between them);
        ``exact=True`` is the verification mode of the point pass (``LF_FUSE_EXACT``: no point
        is classified in fp32) -- same results, the tests compare the two."""
        ctx, lib = self.ctx, self.ctx.lib
        ctx.set_camera(self.TrVeloToRect, self.camera.K, self.camera.width, self.camera.height)
        st = ctx.stream
        launches = 1
        mark = mark or (lambda name: None)
        mark('begin')
        check(lib.lf_zero_counters(ctx.h, ptr(self.counters), self.counters.numel() * 4, st), ctx.h)
        mark('zero')
        pre = isinstance(self.masks, PackedMasks)          # masks arrived bit-packed
        raw = self.masks is not None and len(self.masks) > 0 and (not pre or self.masks.raw)
        iters, se, mode, mt = 0, None, 0, None
        if raw and getattr(self, '_native', None) is not None:
            # erode at the masks' own resolution (packed), then cv2.resize + "> 0.5" from packed words
            # to packed words: two launches, no temporaries
            nat = self._native.view(torch.uint8) if self._native.dtype == torch.bool else self._native
            n_it = int(erosion_iterations) if erosion_kernel_size else 0
            se_n = structuring_element(shape, erosion_kernel_size) if n_it else None
            h, w = nat.shape[1:]
            check(lib.lf_erode_pack(ctx.h, ptr(nat), _mask_mode(nat, float_threshold), len(nat), h, w,
                                    se_n.ctypes.data if n_it else None, se_n.shape[0] if n_it else 1, n_it,
                                    ptr(self._native_words), None, st), ctx.h)
            check(lib.lf_resize_pack(ctx.h, ptr(self._native_words), len(nat), h, w, ptr(self.packed.words),
                                     ptr(self.mask_bbox), self.camera.height, self.camera.width, st), ctx.h)
            launches += 2
            raw = False
        if raw:
            iters = int(erosion_iterations) if erosion_kernel_size else 0
            se = structuring_element(shape, erosion_kernel_size) if iters else None
            if pre:
                mode, mt = _abi.MODE_PACKED, self.masks.words
            else:
                mode = _mask_mode(self.masks, float_threshold)
                mt = self.masks.view(torch.uint8) if self.masks.dtype == torch.bool else self.masks
        bbox = self.packed.bbox if pre and not self.masks.raw else self.mask_bbox
        flags = (_abi.FUSE_ORIENTED if use_oriented else 0) | (_abi.FUSE_EXACT if exact else 0)
        if one_call and self.car_bits is None and getattr(self, '_native', None) is None:
            check(lib.lf_process_batch(
                ctx.h, ptr(self.points), ptr(self.pt_off), ptr(mt) if raw else None, mode, len(mt) if raw else 0,
                ptr(self.mask_off), se.ctypes.data if iters else None, se.shape[0] if iters else 1, iters,
                ptr(self.packed.words), ptr(bbox), ptr(self.boxes), ptr(self.box_off), ptr(self.cnt_off),
                self.F, self.max_n, self.max_m, self.max_b, float(max_depth), flags, int(min_points),
                ptr(self.total), ptr(self.inside), ptr(self.n_in_box), ptr(self.n_valid),
                ptr(self.best_idx), ptr(self.best_cnt), ptr(self.agg), None, st), ctx.h)
            mark('car_stats')
            self.launches_per_run = 1 + (1 if raw else 0) + (2 if (self.F and self.max_n) else 0)
            return
        row_map = None
        if raw and sparse_rows and not pre:                # (packed words in: the direct kernel writes every row anyway)
            if getattr(self, '_row_map', None) is None:
                self._row_map = torch.empty((max(len(mt), 1), (self.camera.height + 31) // 32), dtype=torch.int32,
                                            device=ctx.device)
            row_map = self._row_map
        if raw:
            check(lib.lf_erode_pack_rows(ctx.h, ptr(mt), mode, len(mt), self.camera.height, self.camera.width,
                                         se.ctypes.data if iters else None, se.shape[0] if iters else 1, iters,
                                         ptr(self.packed.words), ptr(self.mask_bbox), ptr(row_map), st), ctx.h)
            launches += 2 if row_map is not None else 1    # + the gap-fill kernel behind the row map
        mark('erode_pack')
        if self.F and self.max_n:
            check(lib.lf_fuse_batch_rows(ctx.h, ptr(self.points), ptr(self.pt_off), ptr(self.packed.words), ptr(bbox),
                                         ptr(self.mask_off), ptr(self.boxes), ptr(self.box_off), ptr(self.cnt_off),
                                         self.F, self.max_n, self.max_m, self.max_b, float(max_depth),
                                         flags, ptr(self.total), ptr(self.inside), ptr(self.n_in_box),
                                         ptr(self.n_valid), ptr(self.car_bits), ptr(row_map), st), ctx.h)
            mark('fuse')
            check(lib.lf_car_stats(ctx.h, ptr(self.total), ptr(self.inside), ptr(self.mask_off),
                                   ptr(self.box_off), ptr(self.cnt_off), self.F, int(min_points),
                                   ptr(self.best_idx), ptr(self.best_cnt), ptr(self.agg), st), ctx.h)
            launches += 2
        mark('car_stats')
        self.launches_per_run = launches

    def capture(self, *args, **kwargs):
        """Record one ``run(*args, **kwargs)`` into a CUDA graph and return it (``g.replay()``
        re-executes memset + kernels with one launch).  Every ABI call enqueues on the
        caller's stream without allocating or synchronising, so the whole step is capturable;
        worth it for small batches (cfg1: one frame), where launch latency is the step."""
        kwargs.pop('mark', None)
        self.run(*args, **kwargs)                      # warm-up: function attributes, lazy module load
        torch.cuda.synchronize(self.ctx.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.run(*args, **kwargs)
        return g

    def upload(self, points=None, masks=None, threads=0, slices=16, float_threshold='ge_one'):
        """Refill the batch from HOST buffers of the shapes it was built with, asynchronously on
        the current stream (pinned CPU tensors copy without staging).  A batch built on
        ``PackedMasks(raw=True)`` takes uint8 / bool / float32 host masks [sum M,H,W] and sends
        them at one bit per pixel: ``lf_pack_masks_host`` packs slice i+1 into pinned staging
        while slice i is on the wire; erosion still runs on the device.  A batch built on
        full-resolution device masks copies them as they are."""
        if points is not None:
            src = points if isinstance(points, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(points))
            self.points.copy_(src.view(self.points.shape), non_blocking=True)
        if masks is None:
            return
        if isinstance(self.masks, PackedMasks):
            if not self.masks.raw:
                raise ValueError('this batch holds final packed masks; build it on PackedMasks(raw=True) to refill')
            from . import io as IO
            m = masks.numpy() if isinstance(masks, torch.Tensor) else np.asarray(masks)
            words = self.masks.words
            if tuple(m.shape) != (len(self.masks), self.masks.H, self.masks.W):
                raise ValueError('masks must be [%d, %d, %d]' % (len(self.masks), self.masks.H, self.masks.W))
            if getattr(self, '_h_words', None) is None:
                self._h_words = torch.empty(words.shape, dtype=words.dtype, pin_memory=True)
                self._h_free = torch.cuda.Event()
            else:
                self._h_free.synchronize()              # the previous upload has left the staging buffer
            hw = self._h_words.numpy()
            n = len(m)
            step = max(1, (n + slices - 1) // max(1, slices))
            for a in range(0, n, step):
                b = min(n, a + step)
                IO.pack_masks(m[a:b], float_threshold=float_threshold, threads=threads, out=hw[a:b])
                words[a:b].copy_(self._h_words[a:b], non_blocking=True)
            self._h_free.record(torch.cuda.current_stream(self.ctx.device))
        else:
            if getattr(self, '_native', None) is not None:
                raise ValueError('low-resolution masks: rebuild the batch instead')
            src = masks if isinstance(masks, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(masks))
            dst = self.masks.view(torch.uint8) if self.masks.dtype == torch.bool else self.masks
            if src.dtype == torch.bool:
                src = src.view(torch.uint8)
            if src.dtype != dst.dtype or tuple(src.shape) != tuple(dst.shape):
                raise ValueError('masks must be %s %s' % (dst.dtype, tuple(dst.shape)))
            dst.copy_(src, non_blocking=True)

    def download(self, out=None):
        """Counter block -> host, asynchronously on the current stream; ``out``: pinned int32 tensor
        of ``self.counters.shape`` (allocated once when omitted).  Synchronise the stream before
        reading it; ``results()`` is the blocking, decoded form."""
        if out is None:
            if getattr(self, '_h_out', None) is None:
                self._h_out = torch.empty(self.counters.shape, dtype=self.counters.dtype, pin_memory=True)
            out = self._h_out
        out.copy_(self.counters, non_blocking=True)
        return out

    def results(self, host=None):
        """Host copies: dict of numpy arrays (one D2H of the counter block, or the block
        ``download()`` already brought over)."""
        c = self.counters.cpu().numpy() if host is None else (host.numpy() if isinstance(host, torch.Tensor) else host)
        sm, sb, sc = int(self.h_mask_off[-1]), int(self.h_box_off[-1]), int(self.h_cnt_off[-1])
        o = self._o
        total = c[o:o + sm].astype(np.int64); o += sm
        n_in_box = c[o:o + sb].astype(np.int64); o += sb
        n_valid = c[o:o + self.F].astype(np.int64); o += self.F
        inside = c[o:o + sc].astype(np.int64)
        agg = c[self._nz:].view(np.int64).copy()
        return {'total': total, 'n_in_box': n_in_box, 'n_valid': n_valid, 'inside': inside, 'agg': agg,
                'best_idx': c[self._ob:self._ob + sm].copy(), 'best_cnt': c[self._ob + sm:self._ob + 2 * sm].copy()}

    def table(self, host=None, min_points=10):
        """The per-car table of the whole batch, vectorised on the host from ONE copy of the counter block
        (``download()`` / ``results``): a dict of equally long arrays {frame (index in the batch), car_id,
        matched_bbox_id, total_points, points_inside_bbox, points_outside_bbox, inside_percentage,
        outside_percentage} with one row per car that has points in a frame that has boxes and gated
        points -- the rows ``calculate_car_point_statistics`` returns (cvs_erosion.py:197-225), in order."""
        c = self.counters.cpu().numpy() if host is None else (host.numpy() if isinstance(host, torch.Tensor) else host)
        sm, sb = int(self.h_mask_off[-1]), int(self.h_box_off[-1])
        if getattr(self, '_h_car_frame', None) is None:
            nm = np.diff(self.h_mask_off)
            self._h_car_frame = np.repeat(np.arange(self.F, dtype=np.int64), nm)
            self._h_car_id = np.arange(sm, dtype=np.int64) - np.repeat(self.h_mask_off[:-1].astype(np.int64), nm)
            self._h_car_hasbox = np.repeat(np.diff(self.h_box_off) > 0, nm)
        o = self._o
        total = c[o:o + sm]
        n_valid = c[o + sm + sb:o + sm + sb + self.F]
        bi = c[self._ob:self._ob + sm]
        bc = c[self._ob + sm:self._ob + 2 * sm]
        keep = (total > 0) & self._h_car_hasbox & (n_valid[self._h_car_frame] > 0)
        total = total[keep].astype(np.int64)
        bi, bc = bi[keep], bc[keep].astype(np.int64)
        matched = (bi >= 0) & (bc >= min_points)
        inside = np.where(matched, bc, 0)
        outside = total - inside
        return {'frame': self._h_car_frame[keep], 'car_id': self._h_car_id[keep],
                'matched_bbox_id': np.where(matched, bi, -1).astype(np.int64), 'total_points': total,
                'points_inside_bbox': inside, 'points_outside_bbox': outside,
                'inside_percentage': np.where(matched, (inside / total) * 100, 0.0),
                'outside_percentage': np.where(matched, (outside / total) * 100, 100.0)}

    def frame_stats(self, res, f, min_points=10, colors=None, extended=False, return_points=False,
                    use_oriented=True):
        """The list of per-car dicts of frame f, as calculate_car_point_statistics
        returns them (cvs_erosion.py:165-229).  ``extended=True`` adds the point-level precision /
        recall / point_iou / match_ratio (README.md:14, 38; ``extension_metrics``).
        ``return_points=True`` (batch built with ``want_car_bits=True``) adds the V2 / V3 keys
        ``corners_velo``, ``inside_mask``, ``car_points`` (V3:386-398) from the membership bits."""
        m0, m1 = self.h_mask_off[f], self.h_mask_off[f + 1]
        b0, b1 = self.h_box_off[f], self.h_box_off[f + 1]
        M, B = m1 - m0, b1 - b0
        if B == 0 or M == 0:
            return []
        cnt = res['inside'][self.h_cnt_off[f]:self.h_cnt_off[f + 1]].reshape(M, B)
        colors = colors or generate_colors(M)
        extra = None
        if extended:
            nin = res['n_in_box'][b0:b1]

            def extra(ci, s):
                if s['matched_bbox_id'] < 0:
                    return extension_metrics(0, s['total_points'], 0)
                return extension_metrics(s['points_inside_bbox'], s['total_points'], nin[s['matched_bbox_id']])
        if return_points:
            if self.car_bits is None:
                raise ValueError('return_points needs a batch built with want_car_bits=True')
            p0, p1 = int(self.h_pt_off[f]), int(self.h_pt_off[f + 1])
            bits, fpts = self.car_bits[p0:p1], self.points[p0:p1, :3]
            corners = self.host_corners(f)
            metrics = extra

            def extra(ci, s):
                out = metrics(ci, s) if metrics is not None else {}
                cp = fpts[((bits >> ci) & 1).to(torch.bool)].cpu().numpy()
                bi = s['matched_bbox_id']
                out.update({'corners_velo': corners[bi] if bi >= 0 else None,
                            'inside_mask': _in_box(cp, corners[bi], use_oriented, self.ctx.device) if bi >= 0 else None,
                            'car_points': cp})
                return out
        return _stats_from_counts(res['total'][m0:m1], cnt, colors, min_points, extra)

    def host_corners(self, f):
        """[B_f,8,3] float64 corners_velo of frame f's boxes, when the batch was built from corners."""
        if self._h_corners is None:
            raise ValueError('this batch was built from prepared lf_box records: corners are not kept')
        return self._h_corners[int(self.h_box_off[f]):int(self.h_box_off[f + 1])]


def process_frames_batched(points, n_points, masks, n_masks, bboxes_per_frame, camera, TrVeloToRect,
                           erosion_kernel_size=3, erosion_iterations=1, max_depth=50, min_points=10,
                           use_oriented=True, device=None, extended=False, return_points=False):
    """Body of the per-frame loop of cvs_erosion.process_frames (:342-372) for F
    frames at once.  ``bboxes_per_frame``: list (per frame) of lists of box
    dicts carrying ``corners_velo``.  Returns a list (per frame) of per-car
    statistics lists; ``extended`` / ``return_points`` as in ``FrameBatch.frame_stats``."""
    corners = [np.array([b['corners_velo'] for b in fr], dtype=np.float64).reshape(-1, 8, 3)
               for fr in bboxes_per_frame]
    n_boxes = [len(c) for c in corners]
    allc = np.concatenate(corners) if corners else np.zeros((0, 8, 3))
    batch = FrameBatch(points, n_points, masks, n_masks, allc, n_boxes, camera, TrVeloToRect, device,
                       want_car_bits=return_points)
    batch.run(erosion_kernel_size, erosion_iterations, max_depth, min_points, use_oriented,
              sparse_rows=True)                        # the packed masks are this call's own intermediate
    res = batch.results()
    out = []
    for f in range(batch.F):
        out.append(batch.frame_stats(res, f, min_points, extended=extended, return_points=return_points,
                                     use_oriented=use_oriented) if res['n_valid'][f] > 0 else [])
    return out


# --------------------------------------------------------------------------
# A10  depth maps
# --------------------------------------------------------------------------

def per_car_depth_maps(points, TrVeloToRect, masks, camera, max_depth=30, policy='last', full=False,
                       device=None):
    """[(car_id, depthMap f64 [H,W])] with car_id = i+1, as the loop of
    seg_with_pointcloud.py:153-170 builds them: gate at ``max_depth``, mask test
    ``mask[v,u] > 0.5`` on the raw float masks, ``policy='last'`` = the last point
    in index order wins (the reference); ``'nearest'`` = z-buffer.  ``full=True``
    also returns the unmasked frame map (seg_with_pointcloud.py:198-200)."""
    ctx = get_ctx(device)
    ctx.set_camera(TrVeloToRect, camera.K, camera.width, camera.height)
    was_np = _is_np(points)
    pts = _points4(points, ctx)
    pm = _as_packed(masks, camera, ctx, float_threshold='gt_half')
    M = len(pm)
    H, W = camera.height, camera.width
    scratch = torch.empty(H * W, dtype=torch.int64, device=ctx.device)
    per = torch.empty((M, H, W), dtype=torch.float64, device=ctx.device) if M else None
    fl = torch.empty((H, W), dtype=torch.float64, device=ctx.device) if full else None
    pol = {'last': _abi.POLICY_LAST, 'nearest': _abi.POLICY_NEAREST}[policy]
    check(ctx.lib.lf_depth_maps(ctx.h, ptr(pts), pts.shape[0], float(max_depth), pol, ptr(pm.words) if M else None,
                                M, ptr(scratch), ptr(fl), ptr(per), ctx.stream), ctx.h)
    if was_np:
        maps = [(i + 1, per[i].cpu().numpy()) for i in range(M)]
        return (maps, fl.cpu().numpy()) if full else maps
    maps = [(i + 1, per[i]) for i in range(M)]
    return (maps, fl) if full else maps


def per_car_depth_maps_batch(points, n_points, masks, n_masks, TrVeloToRect, camera, max_depth=30, policy='last',
                             full=False, per_car=True, device=None):
    """``per_car_depth_maps`` for F frames at once (``lf_depth_maps_batch``: three launches for the
    whole batch instead of three per frame).  points: [sum N,4] (frames concatenated, ``n_points``
    per frame); masks: [sum M,H,W] or ``PackedMasks`` (``n_masks`` per frame).  Returns a list with
    one entry per frame, each the [(car_id, depthMap)] list ``per_car_depth_maps`` returns for that
    frame alone (seg_with_pointcloud.py:153-170); with ``full=True`` also the [F,H,W] frame maps.
    numpy in -> numpy out, tensors in -> views of one device buffer out."""
    ctx = get_ctx(device)
    ctx.set_camera(TrVeloToRect, camera.K, camera.width, camera.height)
    was_np = _is_np(points)
    pts = _points4(points, ctx)
    n_points = [int(x) for x in n_points]
    n_masks = [int(x) for x in n_masks]
    F = len(n_points)
    if len(n_masks) != F or sum(n_points) != pts.shape[0]:
        raise ValueError('n_points / n_masks must describe the %d concatenated frames' % F)
    pm = _as_packed(masks, camera, ctx, float_threshold='gt_half')
    if sum(n_masks) != len(pm):
        raise ValueError('n_masks must sum to the number of masks')
    H, W = camera.height, camera.width
    pt_off = torch.tensor(np.concatenate([[0], np.cumsum(n_points)]), dtype=torch.int64, device=ctx.device)
    mask_off = torch.tensor(np.concatenate([[0], np.cumsum(n_masks)]), dtype=torch.int32, device=ctx.device)
    scratch = torch.empty((F, H * W), dtype=torch.int64, device=ctx.device)
    M = len(pm)
    per = torch.empty((M, H, W), dtype=torch.float64, device=ctx.device) if (per_car and M) else None
    fl = torch.empty((F, H, W), dtype=torch.float64, device=ctx.device) if full else None
    pol = {'last': _abi.POLICY_LAST, 'nearest': _abi.POLICY_NEAREST}[policy]
    check(ctx.lib.lf_depth_maps_batch(ctx.h, ptr(pts), ptr(pt_off), F, max(n_points) if F else 0, float(max_depth), pol,
                                      ptr(pm.words) if M else None, ptr(mask_off), ptr(scratch), ptr(fl), ptr(per),
                                      ctx.stream), ctx.h)
    out, m0 = [], 0
    host = per.cpu().numpy() if (was_np and per is not None) else per
    for f in range(F):
        out.append([(i + 1, host[m0 + i]) for i in range(n_masks[f])] if host is not None else [])
        m0 += n_masks[f]
    if full:
        return out, (fl.cpu().numpy() if was_np else fl)
    return out


# --------------------------------------------------------------------------
# A11  CSV (host; byte-compatible with the pandas writer of the reference)
# --------------------------------------------------------------------------

CSV_COLUMNS = ['frame', 'car_id', 'matched_bbox_id', 'total_points', 'points_inside_bbox', 'points_outside_bbox',
               'inside_percentage', 'outside_percentage', 'is_matched', 'timestamp']


def csv_rows(car_statistics, frame_number, timestamp=None):
    """Row dicts of append_to_master_csv (cvs_erosion.py:240-255)."""
    import datetime
    rows = []
    for s in car_statistics:
        rows.append({'frame': frame_number, 'car_id': s['car_id'], 'matched_bbox_id': s['matched_bbox_id'],
                     'total_points': s['total_points'], 'points_inside_bbox': s['points_inside_bbox'],
                     'points_outside_bbox': s['points_outside_bbox'],
                     'inside_percentage': round(s['inside_percentage'], 2),
                     'outside_percentage': round(s['outside_percentage'], 2),
                     'is_matched': s['matched_bbox_id'] >= 0,
                     'timestamp': timestamp if timestamp is not None else datetime.datetime.now().isoformat()})
    return rows


def _csv_cell(v):
    if isinstance(v, (bool, np.bool_)):
        return 'True' if v else 'False'
    if isinstance(v, (float, np.floating)):
        return repr(float(v))
    return str(v)


def append_to_master_csv(car_statistics, frame_number, master_csv_path, timestamp=None):
    """Append the frame's rows; the header is written only when the file is
    created (cvs_erosion.py:232-265).  ``timestamp`` overrides now().isoformat()."""
    import os
    if not car_statistics:
        return
    d = os.path.dirname(master_csv_path)
    if d:
        os.makedirs(d, exist_ok=True)
    new = not os.path.exists(master_csv_path)
    with open(master_csv_path, 'a', newline='') as fh:
        if new:
            fh.write(','.join(CSV_COLUMNS) + '\n')
        for row in csv_rows(car_statistics, frame_number, timestamp):
            fh.write(','.join(_csv_cell(row[c]) for c in CSV_COLUMNS) + '\n')


def analyze_master_csv(master_csv_path, verbose=False):
    """Aggregates cvs_erosion.py:268-295 prints, returned as a dict."""
    import csv
    import os
    if not os.path.exists(master_csv_path):
        return None
    with open(master_csv_path, newline='') as fh:
        rows = list(csv.DictReader(fh))
    matched = [r for r in rows if r['is_matched'] == 'True']
    out = {'frames': len({r['frame'] for r in rows}), 'detections': len(rows), 'matched': len(matched),
           'unmatched': len(rows) - len(matched),
           'matching_rate': (len(matched) / len(rows) * 100) if rows else float('nan')}
    if matched:
        pct = np.array([float(r['inside_percentage']) for r in matched])
        out.update({'avg_points_per_car': float(np.mean([int(r['total_points']) for r in matched])),
                    'avg_inside_pct': float(pct.mean()), 'min_inside_pct': float(pct.min()),
                    'max_inside_pct': float(pct.max())})
    if verbose:
        print('Total frames processed:', out['frames'])
        print('Total car detections:', out['detections'])
        print('Matched cars: %d  Unmatched cars: %d  Matching rate: %.1f%%' %
              (out['matched'], out['unmatched'], out['matching_rate']))
    return out
