"""numpy / OpenCV restatement of the reference's per-point fusion + evaluation path.

ORACLE / TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``) -- the product
package must never import this file.

Every function cites the reference lines it restates (paths relative to
/root/reference).  Pinned by ``tests/test_oracle_golden.py`` against vectors the
reference's own unmodified functions produced here (``oracle/gen_golden.py``).
"""
import os

import cv2
import numpy as np

# --------------------------------------------------------------------------
# A3  projection                         Coding_testes/cvs_erosion.py:342-346
# --------------------------------------------------------------------------

def project_points(points, TrVeloToRect, camera):
    """velodyne [N,4] f32 -> (u int64 [N], v int64 [N], depth f64 [N]).

    cvs_erosion.py:342-346 (same in V1:340-344, V4:258-262, V5:467-471,
    seg_with_pointcloud.py:132-139): overwrite the reflectance column with 1,
    multiply by the 4x4 f64 velo->rect transform (float32 promotes to f64
    inside matmul), keep x,y,z and hand the [3,N] view to ``cam2image``.
    """
    homo = np.array(points, dtype=np.float32, copy=True)
    homo[:, 3] = 1
    cam = np.matmul(TrVeloToRect, homo.T).T[:, :3]
    u, v, depth = camera.cam2image(cam.T)
    return u.astype(int), v.astype(int), depth


# --------------------------------------------------------------------------
# A4  validity gate                      Coding_testes/cvs_erosion.py:359-367
# --------------------------------------------------------------------------

def filter_valid(u, v, depth, camera, max_depth=50):
    """Frustum + range gate.  max_depth 50 in V1/V2/V3/cvs_erosion (:359),
    30 in V4:275, V5:508, seg_with_pointcloud.py:154-158."""
    return ((u >= 0) & (u < camera.width) & (v >= 0) & (v < camera.height)
            & (depth > 0) & (depth < max_depth))


# --------------------------------------------------------------------------
# A5  mask erosion                       Coding_testes/cvs_erosion.py:96-106
# --------------------------------------------------------------------------

def erode_masks(masks, erosion_kernel_size=3, erosion_iterations=1, shape=cv2.MORPH_ELLIPSE):
    """float32 [M,H,W] 0/1 masks -> float32 eroded masks (cvs_erosion.py:98-106;
    V3_point_cloud_with_erosion.py:81-97): scale to uint8 by truncation,
    ``cv2.erode`` with an elliptical structuring element, scale back."""
    se = cv2.getStructuringElement(shape, (erosion_kernel_size, erosion_kernel_size))
    out = []
    for m in masks:
        m8 = (np.asarray(m) * 255).astype(np.uint8)
        e = cv2.erode(m8, se, iterations=erosion_iterations)
        out.append(e.astype(np.float32) / 255.0)
    if len(out) == 0:
        return np.zeros((0,) + tuple(np.shape(masks)[1:]), np.float32)
    return np.array(out)


def binary_masks(masks, camera):
    """Effective boolean mask the gather sees (cvs_erosion.py:153-154):
    ``cv2.resize(mask.astype(uint8), (W,H))`` then ``> 0.5``."""
    out = []
    for m in masks:
        r = cv2.resize(np.asarray(m).astype(np.uint8), (camera.width, camera.height))
        out.append(r > 0.5)
    if not out:
        return np.zeros((0, camera.height, camera.width), bool)
    return np.stack(out)


# --------------------------------------------------------------------------
# A6  point-in-mask gather               Coding_testes/cvs_erosion.py:148-162
# --------------------------------------------------------------------------

def mask_membership(u_valid, v_valid, masks, camera):
    """bool [M,Nv]: row c is ``mask_resized_c[v_valid, u_valid] > 0.5``."""
    b = binary_masks(masks, camera)
    if b.shape[0] == 0:
        return np.zeros((0, len(u_valid)), bool)
    return b[:, v_valid, u_valid]


def extract_car_points_by_mask(points_valid, u_valid, v_valid, masks, camera):
    """List of M arrays [n_c,3]; an empty selection becomes a (0,3) float64 array
    (cvs_erosion.py:156-160)."""
    sel = mask_membership(u_valid, v_valid, masks, camera)
    sets = []
    for row in sel:
        sets.append(points_valid[row] if row.any() else np.array([]).reshape(0, 3))
    return sets


# --------------------------------------------------------------------------
# A7  box preparation                    Coding_testes/cvs_erosion.py:31-74
# --------------------------------------------------------------------------

def filter_visible_bboxes(bboxes_3d, camera):
    """Keep a box when >= 2 of its 8 cam0 corners have depth > 0.1 and land in
    the image (cvs_erosion.py:55-74).  Corners go straight through
    ``cam2image`` -- no rectifying rotation."""
    keep = []
    for bbox in bboxes_3d:
        if 'corners_cam0' not in bbox:
            continue
        c = np.array(bbox['corners_cam0'])
        u, v, depth = camera.cam2image(c.T)
        ok = (depth > 0.1) & (u >= 0) & (u < camera.width) & (v >= 0) & (v < camera.height)
        if np.sum(ok) >= 2:
            keep.append(bbox)
    return keep


def transform_bboxes_to_velodyne(bboxes_3d, TrVeloToCam):
    """corners_velo = (inv(TrVeloToCam) @ [corners,1]^T)^T[:, :3] stored as a
    Python list on the same dict (cvs_erosion.py:41-52)."""
    inv = np.linalg.inv(TrVeloToCam)
    for bbox in bboxes_3d:
        if 'corners_cam0' in bbox:
            c = np.array(bbox['corners_cam0'])
            ch = np.hstack([c, np.ones((c.shape[0], 1))])
            bbox['corners_velo'] = np.matmul(inv, ch.T).T[:, :3].tolist()
    return bboxes_3d


# --------------------------------------------------------------------------
# A8  containment                        Coding_testes/cvs_erosion.py:114-145
# --------------------------------------------------------------------------

def point_in_bbox(points, bbox_corners):
    """Axis-aligned test (V1_BBox_Pointwise_filtering.py:118-139)."""
    if len(points) == 0:
        return np.array([])
    lo = np.min(bbox_corners, axis=0)
    hi = np.max(bbox_corners, axis=0)
    return np.all((points >= lo) & (points <= hi), axis=1)


def oriented_point_in_bbox(points, bbox_corners):
    """Closed-interval slab test along c1-c0, c3-c0, c4-c0
    (cvs_erosion.py:114-138).  f32 points minus f64 corner promotes to f64."""
    if len(points) == 0:
        return np.array([])
    c0 = bbox_corners[0]
    rel = points - c0
    inside = np.ones(len(points), bool)
    for j in (1, 3, 4):
        e = bbox_corners[j] - c0
        t = np.dot(rel, e) / np.dot(e, e)
        inside &= (t >= 0) & (t <= 1)
    return inside


# --------------------------------------------------------------------------
# A9  per-car statistics                 Coding_testes/cvs_erosion.py:165-229
# --------------------------------------------------------------------------

def inside_count_matrix(car_point_sets, bboxes_3d, use_oriented=True):
    """int64 [M,B]: points of car c inside box b."""
    M, B = len(car_point_sets), len(bboxes_3d)
    cnt = np.zeros((M, B), np.int64)
    test = oriented_point_in_bbox if use_oriented else point_in_bbox
    for ci, pts in enumerate(car_point_sets):
        if len(pts) == 0:
            continue
        for bi, bbox in enumerate(bboxes_3d):
            if 'corners_velo' not in bbox:
                continue
            cnt[ci, bi] = np.sum(test(pts, np.array(bbox['corners_velo'])))
    return cnt


def calculate_car_point_statistics(car_point_sets, bboxes_3d, colors, min_points=10, use_oriented=True,
                                   return_points=False):
    """Per-car dicts (cvs_erosion.py:165-229; V3:320-428 adds use_oriented).

    best box = first strict maximum of the inside count starting from 0; a
    match needs best_count >= min_points; cars with no points are skipped but
    keep their index.  ``return_points=True``: the V2 / V3 form of the dicts with
    ``corners_velo``, ``inside_mask`` (of the best box) and ``car_points``
    (V3_point_cloud_with_erosion.py:386-398, 416-428); None / None for unmatched cars."""
    stats = []
    if not bboxes_3d or len(car_point_sets) == 0:
        return stats
    cnt = inside_count_matrix(car_point_sets, bboxes_3d, use_oriented)
    for ci, pts in enumerate(car_point_sets):
        total = len(pts)
        if total == 0:
            continue
        best, best_idx = 0, -1
        for bi in range(cnt.shape[1]):
            if cnt[ci, bi] > best:
                best, best_idx = cnt[ci, bi], bi
        if best_idx >= 0 and best >= min_points:
            outside = total - best
            stats.append({'car_id': ci, 'matched_bbox_id': best_idx, 'total_points': total,
                          'points_inside_bbox': best, 'points_outside_bbox': outside,
                          'inside_percentage': (best / total) * 100,
                          'outside_percentage': (outside / total) * 100,
                          'color': colors[ci]})
        else:
            stats.append({'car_id': ci, 'matched_bbox_id': -1, 'total_points': total,
                          'points_inside_bbox': 0, 'points_outside_bbox': total,
                          'inside_percentage': 0.0, 'outside_percentage': 100.0,
                          'color': colors[ci]})
        if return_points:
            test = oriented_point_in_bbox if use_oriented else point_in_bbox
            bi = stats[-1]['matched_bbox_id']
            corners = np.array(bboxes_3d[bi]['corners_velo']) if bi >= 0 else None
            stats[-1].update({'corners_velo': corners, 'inside_mask': test(pts, corners) if bi >= 0 else None,
                              'car_points': pts})
    return stats


def summary_statistics_text(car_statistics):
    """What print_summary_statistics prints (V3_point_cloud_with_erosion.py:431-468; V2:409-446), as one string."""
    if not car_statistics:
        return '\nNo car statistics to display.\n'
    matched = [s for s in car_statistics if s['matched_bbox_id'] >= 0]
    lines = ['\n' + '=' * 60, format('SUMMARY STATISTICS', '^60'), '=' * 60,
             'Total cars detected: %d' % len(car_statistics), 'Successfully matched: %d' % len(matched),
             'Unmatched: %d' % (len(car_statistics) - len(matched))]
    if matched:
        lines.append('\n%-8s %-8s %-8s %-8s %-8s %-10s' % ('Car ID', 'BBox ID', 'Total', 'Inside', 'Outside', 'Inside %'))
        lines.append('-' * 60)
        for s in matched:
            lines.append('%-8s %-8s %-8s %-8s %-8s %s' % (s['car_id'], s['matched_bbox_id'], s['total_points'],
                                                          s['points_inside_bbox'], s['points_outside_bbox'],
                                                          format(s['inside_percentage'], '<10.1f')))
        tp = sum(s['total_points'] for s in matched)
        ti = sum(s['points_inside_bbox'] for s in matched)
        to = sum(s['points_outside_bbox'] for s in matched)
        avg = (ti / tp * 100) if tp > 0 else 0
        lines.append('-' * 60)
        lines.append('%-8s %-8s %-8s %-8s %-8s %s' % ('TOTAL', '', tp, ti, to, format(avg, '<10.1f')))
    return '\n'.join(lines) + '\n'


def match_car_points_to_bboxes(car_point_sets, bboxes_3d, colors, min_points=10, use_oriented=True):
    """(corners_velo, rgb colour, count) per matched car
    (V1_BBox_Pointwise_filtering.py:211-265).  Unlike the statistics function
    the threshold is applied *inside* the running maximum."""
    pairs = []
    if not bboxes_3d or len(car_point_sets) == 0:
        return pairs
    cnt = inside_count_matrix(car_point_sets, bboxes_3d, use_oriented)
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
    """Point-level precision / recall / IoU / match ratio.  NOT in the
    reference's code (README.md:14, 38 only name them) -- defined in
    SURVEY.md section 8(a) row A9; parity for these is unpinned."""
    tp = float(inside)
    fp = float(total - inside)
    fn = float(n_valid_in_box - inside)
    return {
        'precision': tp / (tp + fp) if tp + fp > 0 else 0.0,
        'recall': tp / (tp + fn) if tp + fn > 0 else 0.0,
        'point_iou': tp / (tp + fp + fn) if tp + fp + fn > 0 else 0.0,
        'match_ratio': (tp / fp) if fp > 0 else float('inf'),
    }


# --------------------------------------------------------------------------
# A10 per-car depth maps                 Coding_testes/seg_with_pointcloud.py:153-170
# --------------------------------------------------------------------------

def per_car_depth_maps_loop(u, v, depth, masks, camera, max_depth=30):
    """Literal restatement of the interpreted double loop (last write wins in
    point-index order).  Slow: small cases only."""
    valid = filter_valid(u, v, depth, camera, max_depth)
    maps = []
    for i, mask in enumerate(masks):
        dm = np.zeros((camera.height, camera.width))
        for idx in np.where(valid)[0]:
            x, y = u[idx], v[idx]
            if mask[y, x] > 0.5:
                dm[y, x] = depth[idx]
        maps.append((i + 1, dm))
    return maps


def per_car_depth_maps(u, v, depth, masks, camera, max_depth=30, policy='last'):
    """Vectorised equivalent.  ``policy='last'`` is the reference's behaviour
    (numpy fancy assignment applies repeated indices in order, so the last
    point wins, like the loop); ``policy='nearest'`` is the north-star
    z-buffer (minimum depth per pixel)."""
    valid = filter_valid(u, v, depth, camera, max_depth)
    idx = np.where(valid)[0]
    H, W = camera.height, camera.width
    full = np.zeros((H, W))
    if policy == 'last':
        full[v[idx], u[idx]] = depth[idx]
    else:
        order = idx[np.argsort(-depth[idx], kind='stable')]
        full[v[order], u[order]] = depth[order]
    hit = np.zeros((H, W), bool)
    hit[v[idx], u[idx]] = True
    maps = []
    for i, mask in enumerate(masks):
        sel = hit & (np.asarray(mask) > 0.5)
        maps.append((i + 1, np.where(sel, full, 0.0)))
    return maps


# --------------------------------------------------------------------------
# A11 CSV                                Coding_testes/cvs_erosion.py:232-295
# --------------------------------------------------------------------------

CSV_COLUMNS = ['frame', 'car_id', 'matched_bbox_id', 'total_points', 'points_inside_bbox',
               'points_outside_bbox', 'inside_percentage', 'outside_percentage', 'is_matched',
               'timestamp']


def csv_rows(car_statistics, frame_number, timestamp):
    rows = []
    for s in car_statistics:
        rows.append({
            'frame': frame_number, 'car_id': s['car_id'], 'matched_bbox_id': s['matched_bbox_id'],
            'total_points': s['total_points'], 'points_inside_bbox': s['points_inside_bbox'],
            'points_outside_bbox': s['points_outside_bbox'],
            'inside_percentage': round(s['inside_percentage'], 2),
            'outside_percentage': round(s['outside_percentage'], 2),
            'is_matched': s['matched_bbox_id'] >= 0, 'timestamp': timestamp})
    return rows


def append_to_master_csv(car_statistics, frame_number, master_csv_path, timestamp):
    """pandas append with a header only on creation (cvs_erosion.py:232-265);
    the timestamp is injected instead of ``datetime.now().isoformat()``."""
    import pandas as pd
    if not car_statistics:
        return
    d = os.path.dirname(master_csv_path)
    if d:
        os.makedirs(d, exist_ok=True)
    df = pd.DataFrame(csv_rows(car_statistics, frame_number, timestamp))
    if os.path.exists(master_csv_path):
        df.to_csv(master_csv_path, mode='a', header=False, index=False)
    else:
        df.to_csv(master_csv_path, index=False)


def analyze_master_csv(master_csv_path):
    """Aggregates printed by cvs_erosion.py:268-295, returned as a dict."""
    import pandas as pd
    if not os.path.exists(master_csv_path):
        return None
    df = pd.read_csv(master_csv_path)
    out = {'frames': int(df['frame'].nunique()), 'detections': int(len(df)),
           'matched': int(df['is_matched'].sum()), 'unmatched': int((~df['is_matched']).sum()),
           'matching_rate': float(df['is_matched'].mean() * 100)}
    m = df[df['is_matched'] == True]  # noqa: E712
    if len(m) > 0:
        out.update({'avg_points_per_car': float(m['total_points'].mean()),
                    'avg_inside_pct': float(m['inside_percentage'].mean()),
                    'min_inside_pct': float(m['inside_percentage'].min()),
                    'max_inside_pct': float(m['inside_percentage'].max())})
    return out


# --------------------------------------------------------------------------
# A12 2D matching (host-side surface)    V4_BBox_IoU_filtering.py:118-183, V5:215-416
# --------------------------------------------------------------------------

def calculate_iou_2d(box1, box2):
    """V4_BBox_IoU_filtering.py:118-137 (same in V5:255-274)."""
    ax0, ay0, ax1, ay1 = box1
    bx0, by0, bx1, by1 = box2
    ix0, iy0 = max(ax0, bx0), max(ay0, by0)
    ix1, iy1 = min(ax1, bx1), min(ay1, by1)
    if ix1 <= ix0 or iy1 <= iy0:
        return 0.0
    inter = (ix1 - ix0) * (iy1 - iy0)
    union = (ax1 - ax0) * (ay1 - ay0) + (bx1 - bx0) * (by1 - by0) - inter
    return inter / union if union > 0 else 0.0


def projected_bbox_2d(bbox_3d, camera):
    """min/max pixel box over corners with depth > 0 (V4:157-168, V5:215-252).
    Returns None when no corner is in front of the camera."""
    c = np.array(bbox_3d['corners_cam0'])
    u, v, depth = camera.cam2image(c.T)
    front = depth > 0
    if np.sum(front) == 0:
        return None
    uu, vv = u[front], v[front]
    return [np.min(uu), np.min(vv), np.max(uu), np.max(vv)], float(np.mean(depth[front]))


def match_detections_to_bboxes(boxes_2d, bboxes_3d, colors, camera, min_iou=0.25):
    """V4_BBox_IoU_filtering.py:140-183: best IoU > min_iou wins (strict)."""
    pairs = []
    if not bboxes_3d or len(boxes_2d) == 0:
        return pairs
    for di, b2 in enumerate(boxes_2d):
        x1, y1, x2, y2 = b2
        best, best_idx = 0, -1
        for bi, bbox in enumerate(bboxes_3d):
            if 'corners_cam0' not in bbox:
                continue
            pb = projected_bbox_2d(bbox, camera)
            if pb is None:
                continue
            iou = calculate_iou_2d([x1, y1, x2, y2], pb[0])
            if iou > best and iou > min_iou:
                best, best_idx = iou, bi
        if best_idx >= 0 and 'corners_velo' in bboxes_3d[best_idx]:
            col = np.array([colors[di][2], colors[di][1], colors[di][0]]) / 255.0
            pairs.append((np.array(bboxes_3d[best_idx]['corners_velo']), col))
    return pairs


def calculate_matching_score(det, proj, weight_iou=0.5, weight_center=0.3, weight_size=0.2):
    """V5_ProjectingBBoxes.py:277-304."""
    iou = calculate_iou_2d(det['bbox'], proj['bbox'])
    dist = np.linalg.norm(np.array(det['center']) - np.array(proj['center']))
    center_score = max(0, 1 - dist / 1000)
    da = det['size'][0] * det['size'][1]
    ba = proj['area']
    size_ratio = min(da, ba) / max(da, ba) if (da > 0 and ba > 0) else 0
    total = weight_iou * iou + weight_center * center_score + weight_size * size_ratio
    return total, {'iou': iou, 'center_score': center_score, 'size_score': size_ratio,
                   'total_score': total}


def improved_match(boxes_2d, bboxes_3d, camera, min_score_threshold=0.3, min_iou_threshold=0.15):
    """Assignment part of V5_ProjectingBBoxes.py:307-416: returns the list of
    accepted (det_idx, original_bbox_idx) pairs."""
    from scipy.optimize import linear_sum_assignment
    if not bboxes_3d or len(boxes_2d) == 0:
        return []
    dets = []
    for b in boxes_2d:
        if len(b) == 4:
            x1, y1, x2, y2 = b
            dets.append({'bbox': [x1, y1, x2, y2], 'center': [(x1 + x2) / 2, (y1 + y2) / 2],
                         'size': [x2 - x1, y2 - y1], 'area': (x2 - x1) * (y2 - y1)})
    projs, valid_idx = [], []
    for j, bbox in enumerate(bboxes_3d):
        pb = projected_bbox_2d(bbox, camera)
        if pb is None:
            continue
        x0, y0, x1, y1 = pb[0]
        projs.append({'bbox': pb[0], 'center': [(x0 + x1) / 2, (y0 + y1) / 2],
                      'size': [x1 - x0, y1 - y0], 'area': (x1 - x0) * (y1 - y0)})
        valid_idx.append(j)
    if not projs:
        return []
    cost = np.zeros((len(dets), len(projs)))
    det_ = {}
    for i, d in enumerate(dets):
        for j, p in enumerate(projs):
            s, info = calculate_matching_score(d, p)
            cost[i, j] = 1 - s
            det_[(i, j)] = info
    out = []
    for di, bj in zip(*linear_sum_assignment(cost)):
        info = det_[(di, bj)]
        if info['total_score'] >= min_score_threshold and info['iou'] >= min_iou_threshold:
            out.append((int(di), int(valid_idx[bj])))
    return out


# --------------------------------------------------------------------------
# one frame, end to end                  Coding_testes/cvs_erosion.py:320-376
# --------------------------------------------------------------------------

def process_frame(points, bboxes_3d_raw, masks, camera, TrVeloToCam, TrVeloToRect,
                  erosion_kernel_size=3, erosion_iterations=1, max_depth=50, min_points=10,
                  use_oriented=True):
    """Body of the per-frame loop of cvs_erosion.process_frames (:331-376) with
    masks supplied instead of YOLO.  ``erosion_kernel_size=0`` skips erosion
    (the V2 variant).  Returns every intermediate the CUDA path is checked
    against."""
    import copy
    boxes = transform_bboxes_to_velodyne(filter_visible_bboxes(copy.deepcopy(bboxes_3d_raw), camera),
                                         TrVeloToCam)
    u, v, depth = project_points(points, TrVeloToRect, camera)
    valid = filter_valid(u, v, depth, camera, max_depth)
    vidx = np.where(valid)[0]
    u_valid, v_valid = u[valid], v[valid]
    points_valid = points[vidx, :3]
    if erosion_kernel_size and len(masks) > 0:
        masks_e = erode_masks(masks, erosion_kernel_size, erosion_iterations)
    else:
        masks_e = masks
    colors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks_e))]
    member = mask_membership(u_valid, v_valid, masks_e, camera)
    sets = [points_valid[r] if r.any() else np.array([]).reshape(0, 3) for r in member]
    inside = inside_count_matrix(sets, boxes, use_oriented)
    n_in_box = np.zeros(len(boxes), np.int64)
    test = oriented_point_in_bbox if use_oriented else point_in_bbox
    for bi, b in enumerate(boxes):
        if len(points_valid):
            n_in_box[bi] = np.sum(test(points_valid, np.array(b['corners_velo'])))
    stats = calculate_car_point_statistics(sets, boxes, colors, min_points, use_oriented) \
        if len(vidx) else []
    return {'u': u, 'v': v, 'depth': depth, 'valid': valid, 'boxes': boxes, 'masks_eroded': masks_e,
            'member': member, 'total': member.sum(axis=1).astype(np.int64) if len(member) else
            np.zeros(0, np.int64), 'inside': inside, 'n_in_box': n_in_box, 'stats': stats,
            'car_point_sets': sets}


# --------------------------------------------------------------------------
# colour buffers (restated; ORACLE / TEST INFRASTRUCTURE ONLY)
# --------------------------------------------------------------------------

def colored_cloud_reference(points_valid, u_valid, v_valid, masks, colors, camera):
    """The geometry list of V4_BBox_IoU_filtering.py:290-307 (same loop in
    V5_ProjectingBBoxes.py:524-550) without Open3D: [(points [n,3], rgb [n,3])] per mask
    with at least one point, then the grey background cloud; plus ``bg_assigned``."""
    clouds = []
    bg_assigned = np.zeros(len(points_valid), dtype=bool)
    for mask, color in zip(masks, colors):
        mask_resized = cv2.resize(np.asarray(mask).astype(np.uint8), (camera.width, camera.height))
        mask_indices = mask_resized[v_valid, u_valid] > 0.5
        if np.count_nonzero(mask_indices) > 0:
            car_points = points_valid[mask_indices]
            bg_assigned[mask_indices] = True
            car_color = np.array([color[2], color[1], color[0]]) / 255.0          # BGR to RGB
            clouds.append((mask_indices, car_points, np.tile(car_color, (len(car_points), 1))))
    remaining = points_valid[~bg_assigned]
    if len(remaining) > 0:
        clouds.append((~bg_assigned, remaining, np.tile([0.5, 0.5, 0.5], (len(remaining), 1))))
    return clouds, bg_assigned


def same_color_reference(points, u, v, depth, masks, mask_colors, camera, max_depth=30):
    """The per-point loop of Same_color.py:106-132 (inline in projectVeloToImage, so restated):
    for every valid point in index order the FIRST mask with ``mask[y, x] > 0.5`` gives the colour
    ``mask_colors[i] / 255`` (stored order, no BGR swap); points no mask claims are background.
    -> (colored_points [n,3], colored_colors [n,3], full_points [k,3]) as the script builds them."""
    valid = (u >= 0) & (u < camera.width) & (v >= 0) & (v < camera.height) & (depth > 0) & (depth < max_depth)
    colored_points, colored_colors, full_points = [], [], []
    for idx in np.where(valid)[0]:
        x, y = u[idx], v[idx]
        pt = points[idx, :3]
        matched = False
        for i, mask in enumerate(masks):
            if y < mask.shape[0] and x < mask.shape[1] and mask[y, x] > 0.5:
                colored_points.append(pt)
                colored_colors.append(np.array(mask_colors[i]) / 255.0)
                matched = True
                break
        if not matched:
            full_points.append(pt)
    return (np.array(colored_points).reshape(-1, 3), np.array(colored_colors).reshape(-1, 3),
            np.array(full_points).reshape(-1, 3))


def vis_cloud_reference(points_valid, u_valid, v_valid, masks, mask_colors, camera):
    """vis.py:175-205 (inline, restated): per mask the indices with ``mask[vv, uu] > 0.5`` found by
    a Python loop over the valid points, colour ``[c[2], c[1], c[0]] / 255``; ``bg_assigned``.
    -> ([(indices, rgb)] per mask with points, bg_assigned)."""
    clouds = []
    bg_assigned = np.zeros(len(points_valid), dtype=bool)
    for mask, color in zip(masks, mask_colors):
        normalized_color = np.array([color[2], color[1], color[0]], dtype=float) / 255.0
        idxs = [i for i, (uu, vv) in enumerate(zip(u_valid, v_valid))
                if 0 <= uu < camera.width and 0 <= vv < camera.height and mask[vv, uu] > 0.5]
        idxs = np.array(idxs, dtype=np.int64)
        if len(idxs) > 0:
            clouds.append((idxs, normalized_color))
            bg_assigned[idxs] = True
    return clouds, bg_assigned


def jet_lut_restated(N=256):
    """matplotlib's 'jet' (``_cm._jet_data`` through ``colors._create_lookup_table``,
    gamma 1).  matplotlib is absent from this image: PARITY UNPINNED for these values."""
    data = {'red': ((0.00, 0, 0), (0.35, 0, 0), (0.66, 1, 1), (0.89, 1, 1), (1.00, 0.5, 0.5)),
            'green': ((0.000, 0, 0), (0.125, 0, 0), (0.375, 1, 1), (0.640, 1, 1), (0.910, 0, 0), (1.000, 0, 0)),
            'blue': ((0.00, 0.5, 0.5), (0.11, 1, 1), (0.34, 1, 1), (0.65, 0, 0), (1.00, 0, 0))}
    out = []
    for ch in ('red', 'green', 'blue'):
        a = np.array(data[ch], dtype=float)
        x, y0, y1 = a[:, 0] * (N - 1), a[:, 1], a[:, 2]
        xind = (N - 1) * np.linspace(0, 1, N)
        ind = np.searchsorted(x, xind)[1:-1]
        dist = (xind[1:-1] - x[ind - 1]) / (x[ind] - x[ind - 1])
        out.append(np.clip(np.concatenate([[y1[0]], dist * (y0[ind] - y1[ind - 1]) + y1[ind - 1], [y0[-1]]]), 0, 1))
    return np.stack(out, axis=1)


def depth_overlay_reference(depthMap, masking_image, lut):
    """seg_with_pointcloud.py:176-180 with ``cm`` = the lookup table ``lut`` [N,3]:
    Colormap.__call__ on floats is ``lut[int(x * N)]`` with ``x * N == N -> N - 1``."""
    N = len(lut)
    x = depthMap / np.max(depthMap)
    xa = x * N
    xa[xa == N] = N - 1
    depthImage = lut[np.clip(xa.astype(int), 0, N - 1)]
    image_withseg = np.array(masking_image) / 255.
    image_withseg[depthMap > 0] = depthImage[depthMap > 0]
    image_withseg = np.uint8(image_withseg * 255)
    return cv2.cvtColor(image_withseg, cv2.COLOR_RGB2BGR)
