"""2D detection <-> projected 3D box matching of the V4 / V5 scripts.

O(#detections x #boxes) scalar work (<= 12 x 314 per frame): kept on the host
in float64 so that the IoUs, scores and the Hungarian assignment are the
reference's, value for value.  The data-parallel part of those scripts
(projection, gate, mask gather) is served by ``lidarfuse``.
"""
import numpy as np


def calculate_iou_2d(box1, box2):
    """IoU of [x1,y1,x2,y2] boxes, 0 for touching / disjoint boxes
    (V4_BBox_IoU_filtering.py:118-137; V5_ProjectingBBoxes.py:255-274)."""
    ax1, ay1, ax2, ay2 = box1
    bx1, by1, bx2, by2 = box2
    left, top = max(ax1, bx1), max(ay1, by1)
    right, bottom = min(ax2, bx2), min(ay2, by2)
    if right <= left or bottom <= top:
        return 0.0
    inter = (right - left) * (bottom - top)
    union = (ax2 - ax1) * (ay2 - ay1) + (bx2 - bx1) * (by2 - by1) - inter
    return inter / union if union > 0 else 0.0


def project_3d_bbox_to_2d(bbox_3d, camera):
    """({bbox, center, size, area, avg_depth}, corners_cam0) of the corners in
    front of the camera, or (None, None) (V5_ProjectingBBoxes.py:215-252)."""
    if 'corners_cam0' not in bbox_3d:
        return None, None
    corners = np.array(bbox_3d['corners_cam0'])
    u, v, depth = camera.cam2image(corners.T)
    front = depth > 0
    if not np.any(front):
        return None, None
    uu, vv = u[front], v[front]
    x0, x1, y0, y1 = np.min(uu), np.max(uu), np.min(vv), np.max(vv)
    w, h = x1 - x0, y1 - y0
    info = {'bbox': [x0, y0, x1, y1], 'center': [(x0 + x1) / 2, (y0 + y1) / 2], 'size': [w, h],
            'area': w * h, 'avg_depth': np.mean(depth[front])}
    return info, corners


def match_detections_to_bboxes(boxes_2d, bboxes_3d, colors, camera, min_iou=0.25):
    """[(corners_velo, rgb)] -- per detection the box with the highest IoU above
    ``min_iou`` (strict, first wins) (V4_BBox_IoU_filtering.py:140-183)."""
    pairs = []
    if not bboxes_3d or len(boxes_2d) == 0:
        return pairs
    projected = []
    for bbox in bboxes_3d:
        info, _ = project_3d_bbox_to_2d(bbox, camera)
        projected.append(None if info is None else info['bbox'])
    for di, det in enumerate(boxes_2d):
        x1, y1, x2, y2 = det
        best, best_idx = 0, -1
        for bi, pb in enumerate(projected):
            if pb is None:
                continue
            iou = calculate_iou_2d([x1, y1, x2, y2], pb)
            if iou > best and iou > min_iou:
                best, best_idx = iou, bi
        if best_idx >= 0 and 'corners_velo' in bboxes_3d[best_idx]:
            rgb = np.array([colors[di][2], colors[di][1], colors[di][0]]) / 255.0
            pairs.append((np.array(bboxes_3d[best_idx]['corners_velo']), rgb))
    return pairs


def calculate_matching_score(detection_info, bbox_3d_info, weight_iou=0.5, weight_center=0.3, weight_size=0.2):
    """0.5 IoU + 0.3 max(0, 1 - |dcentre|/1000) + 0.2 min/max area
    (V5_ProjectingBBoxes.py:277-304)."""
    iou = calculate_iou_2d(detection_info['bbox'], bbox_3d_info['bbox'])
    dist = np.linalg.norm(np.array(detection_info['center']) - np.array(bbox_3d_info['center']))
    center_score = max(0, 1 - dist / 1000)
    det_area = detection_info['size'][0] * detection_info['size'][1]
    box_area = bbox_3d_info['area']
    size_score = min(det_area, box_area) / max(det_area, box_area) if det_area > 0 and box_area > 0 else 0
    total = weight_iou * iou + weight_center * center_score + weight_size * size_score
    return total, {'iou': iou, 'center_score': center_score, 'size_score': size_score, 'total_score': total}


def assign_detections(boxes_2d, bboxes_3d, camera, min_score_threshold=0.3, min_iou_threshold=0.15):
    """Accepted (det_idx, bbox_idx) pairs of the Hungarian assignment on
    cost = 1 - score (V5_ProjectingBBoxes.py:319-368)."""
    from scipy.optimize import linear_sum_assignment
    dets = []
    for b in boxes_2d:
        if len(b) == 4:
            x1, y1, x2, y2 = b
            dets.append({'bbox': [x1, y1, x2, y2], 'center': [(x1 + x2) / 2, (y1 + y2) / 2],
                         'size': [x2 - x1, y2 - y1], 'area': (x2 - x1) * (y2 - y1)})
    infos, index = [], []
    for j, bbox in enumerate(bboxes_3d):
        info, _ = project_3d_bbox_to_2d(bbox, camera)
        if info is not None:
            infos.append(info)
            index.append(j)
    if not infos or not dets:
        return []
    cost = np.zeros((len(dets), len(infos)))
    detail = {}
    for i, d in enumerate(dets):
        for j, p in enumerate(infos):
            s, info = calculate_matching_score(d, p)
            cost[i, j] = 1 - s
            detail[(i, j)] = info
    out = []
    for di, bj in zip(*linear_sum_assignment(cost)):
        info = detail[(di, bj)]
        if info['total_score'] >= min_score_threshold and info['iou'] >= min_iou_threshold:
            out.append((int(di), int(index[bj])))
    return out


def improved_match_detections_to_bboxes(boxes_2d, bboxes_3d, mask_colors, camera, min_score_threshold=0.3,
                                        min_iou_threshold=0.15):
    """[(corners_velo, rgb)]: accepted assignments in detection order, then every
    unmatched box in light grey (V5_ProjectingBBoxes.py:307-416)."""
    pairs = []
    if not bboxes_3d or len(boxes_2d) == 0:
        return pairs
    if all(project_3d_bbox_to_2d(b, camera)[0] is None for b in bboxes_3d):
        return pairs                                  # "No valid 3D bbox projections" (:343-345)
    accepted = assign_detections(boxes_2d, bboxes_3d, camera, min_score_threshold, min_iou_threshold)
    taken = set()
    for di, bi in accepted:
        taken.add(bi)
        if 'corners_velo' not in bboxes_3d[bi]:
            continue
        if di < len(mask_colors):
            rgb = np.array([mask_colors[di][2], mask_colors[di][1], mask_colors[di][0]], dtype=float) / 255.0
        else:
            rgb = np.array([1.0, 0.0, 0.0])
        pairs.append((np.array(bboxes_3d[bi]['corners_velo']), rgb))
    for i, bbox in enumerate(bboxes_3d):
        if i not in taken and 'corners_velo' in bbox:
            pairs.append((np.array(bbox['corners_velo']), [0.7, 0.7, 0.7]))
    return pairs


# --------------------------------------------------------------------------
# device versions: batched projection + IoU / score matrices
# --------------------------------------------------------------------------

def _ctx_with_camera(camera, device=None):
    """lidarfuse context whose K / image size are ``camera``'s (the box kernels do not
    use TrVeloToRect: the reference projects cam0 corners with K alone)."""
    from . import lidarfuse as LF
    ctx = LF.get_ctx(device)
    Kc = np.ascontiguousarray(np.asarray(camera.K, dtype=np.float64).reshape(3, 4))
    key = ctx._cam_key
    if key is None or key[1] != Kc.tobytes() or key[2] != int(camera.width) or key[3] != int(camera.height):
        ctx.set_camera(np.eye(4), camera.K, camera.width, camera.height)
    return ctx


def project_boxes_batched(corners_cam0, camera, depth_range=(0.1, 100), min_points_in_view=4,
                          return_corners=False, device=None):
    """corners_cam0 [n,8,3] f64 (any number of frames' boxes, concatenated) ->
    structured array (``_abi.BOX2D_DTYPE``): the projected rectangle of
    ``project_3d_bbox_to_2d`` (V5:215-252), the corner count of
    ``filter_visible_bboxes`` (cvs_erosion.py:55-74) and the verdict of
    ``is_bbox_in_camera_view`` (secondtest.py:277-359), one record per box.
    ``return_corners`` adds (u, v, depth) [n,8] as ``camera.cam2image`` gives them."""
    import torch
    from . import _abi
    from ._abi import check, ptr
    ctx = _ctx_with_camera(camera, device)
    c = np.ascontiguousarray(np.asarray(corners_cam0, dtype=np.float64).reshape(-1, 8, 3))
    n = len(c)
    out = torch.zeros((max(n, 1), 72), dtype=torch.uint8, device=ctx.device)
    u = v = d = None
    if return_corners:
        u = torch.empty((max(n, 1), 8), dtype=torch.int64, device=ctx.device)
        v = torch.empty((max(n, 1), 8), dtype=torch.int64, device=ctx.device)
        d = torch.empty((max(n, 1), 8), dtype=torch.float64, device=ctx.device)
    if n:
        dc = torch.from_numpy(c).to(ctx.device)
        check(ctx.lib.lf_project_boxes(ctx.h, ptr(dc), n, float(depth_range[0]), float(depth_range[1]),
                                       int(min_points_in_view), ptr(out), ptr(u), ptr(v), ptr(d), ctx.stream), ctx.h)
    rec = out.cpu().numpy()[:n].reshape(-1).view(_abi.BOX2D_DTYPE).copy()
    if return_corners:
        return rec, u.cpu().numpy()[:n], v.cpu().numpy()[:n], d.cpu().numpy()[:n]
    return rec


def _corner_array(bboxes_3d):
    idx = [i for i, b in enumerate(bboxes_3d) if 'corners_cam0' in b]
    c = np.array([bboxes_3d[i]['corners_cam0'] for i in idx], dtype=np.float64).reshape(-1, 8, 3)
    return idx, c


def filter_visible_bboxes_device(bboxes_3d, camera, device=None):
    """``filter_visible_bboxes`` (cvs_erosion.py:55-74) with the projection on the device."""
    idx, c = _corner_array(bboxes_3d)
    rec = project_boxes_batched(c, camera, device=device)
    return [bboxes_3d[i] for i, r in zip(idx, rec) if r['n_visible'] >= 2]


def filter_bboxes_in_camera_view(bboxes_3d, camera, min_points_in_view=4, depth_range=(0.1, 100), device=None):
    """``filter_bboxes_in_camera_view`` (secondtest.py:362-...): keeps the boxes for which
    ``is_bbox_in_camera_view`` (:277-359) answers True."""
    from . import _abi
    idx, c = _corner_array(bboxes_3d)
    rec = project_boxes_batched(c, camera, depth_range, min_points_in_view, device=device)
    return [bboxes_3d[i] for i, r in zip(idx, rec) if r['view_code'] == _abi.VIEW_VALID]


def match_matrices(boxes_2d_per_frame, rects, n_boxes_per_frame, weights=(0.5, 0.3, 0.2), device=None):
    """Per frame the (IoU [D,B], score [D,B]) float64 matrices of
    ``calculate_iou_2d`` (V4:118-137) and ``calculate_matching_score`` (V5:277-304)
    for every (detection, projected box) pair, computed in ONE launch for all frames.
    ``rects``: records of ``project_boxes_batched`` for the concatenated boxes."""
    import torch
    from . import lidarfuse as LF
    from ._abi import check, ptr
    ctx = LF.get_ctx(device)
    F = len(boxes_2d_per_frame)
    dets = [np.asarray(b, dtype=np.float64).reshape(-1, 4) for b in boxes_2d_per_frame]
    nd = np.array([len(d) for d in dets], dtype=np.int64)
    nb = np.asarray(n_boxes_per_frame, dtype=np.int64)
    if len(nb) != F or int(nb.sum()) != len(rects):
        raise ValueError('n_boxes_per_frame does not match rects')
    det_off = np.concatenate([[0], np.cumsum(nd)]).astype(np.int32)
    box_off = np.concatenate([[0], np.cumsum(nb)]).astype(np.int32)
    out_off = np.concatenate([[0], np.cumsum(nd * nb)]).astype(np.int64)
    total = int(out_off[-1])
    if F == 0 or total == 0:
        return [(np.zeros((int(a), int(b))), np.zeros((int(a), int(b)))) for a, b in zip(nd, nb)]
    d_dets = torch.from_numpy(np.ascontiguousarray(np.concatenate(dets))).to(ctx.device)
    d_rects = torch.from_numpy(np.ascontiguousarray(rects).view(np.uint8).reshape(-1, 72).copy()).to(ctx.device)
    iou = torch.empty(total, dtype=torch.float64, device=ctx.device)
    score = torch.empty(total, dtype=torch.float64, device=ctx.device)
    out = []
    for f0 in range(0, F, 65535):                       # gridDim.y
        f1 = min(F, f0 + 65535)
        do = torch.from_numpy(det_off[f0:f1 + 1].copy()).to(ctx.device)
        bo = torch.from_numpy(box_off[f0:f1 + 1].copy()).to(ctx.device)
        oo = torch.from_numpy(out_off[f0:f1 + 1].copy()).to(ctx.device)
        check(ctx.lib.lf_match_matrices(ctx.h, ptr(d_dets), ptr(do), ptr(d_rects), ptr(bo), ptr(oo), f1 - f0,
                                        int((nd[f0:f1] * nb[f0:f1]).max()), float(weights[0]), float(weights[1]),
                                        float(weights[2]), ptr(iou), ptr(score), ctx.stream), ctx.h)
    h_iou, h_score = iou.cpu().numpy(), score.cpu().numpy()
    for f in range(F):
        sl = slice(int(out_off[f]), int(out_off[f + 1]))
        out.append((h_iou[sl].reshape(int(nd[f]), int(nb[f])), h_score[sl].reshape(int(nd[f]), int(nb[f]))))
    return out


def assign_detections_batched(boxes_2d_per_frame, bboxes_3d_per_frame, camera, min_score_threshold=0.3,
                              min_iou_threshold=0.15, device=None):
    """``assign_detections`` for many frames: projection and the score matrices on the
    device (two launches for the whole batch), the Hungarian step per frame on the host
    (scipy, as V5:360).  Returns a list (per frame) of accepted (det_idx, bbox_idx)."""
    from scipy.optimize import linear_sum_assignment
    idxs, cs = zip(*[_corner_array(b) for b in bboxes_3d_per_frame]) if bboxes_3d_per_frame else ((), ())
    allc = np.concatenate(cs) if cs else np.zeros((0, 8, 3))
    rects = project_boxes_batched(allc, camera, device=device)
    dets = [[b for b in fr if len(b) == 4] for fr in boxes_2d_per_frame]
    mats = match_matrices(dets, rects, [len(c) for c in cs], device=device)
    out, o = [], 0
    for f, (iou, score) in enumerate(mats):
        nb = len(cs[f])
        front = rects[o:o + nb]['n_front'] > 0
        o += nb
        cols = np.nonzero(front)[0]
        if len(cols) == 0 or iou.shape[0] == 0:
            out.append([])
            continue
        cost = 1 - score[:, cols]
        acc = []
        for di, bj in zip(*linear_sum_assignment(cost)):
            if score[di, cols[bj]] >= min_score_threshold and iou[di, cols[bj]] >= min_iou_threshold:
                acc.append((int(di), int(idxs[f][cols[bj]])))
        out.append(acc)
    return out
