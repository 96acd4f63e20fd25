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
