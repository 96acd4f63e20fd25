"""The frame loop of cvs_erosion.process_frames (cvs_erosion.py:298-379), batched.

The reference walks the sweeps one by one; here every frame that survives the
reference's skip rules is packed into one CSR batch and the whole
sequence goes through the GPU in four launches.  YOLO inference is out of
scope: ``detections(frame) -> (masks [M,H,W], boxes_xyxy [M,4], conf [M])`` is
an input (any order; it is sorted by confidence here as cvs_erosion.py:86-94
does).
"""
import os

import numpy as np

from . import kitti360 as K
from . import lidarfuse as LF


def sort_by_confidence(masks, boxes, conf):
    """cvs_erosion.py:90-94: argsort(confidences)[::-1] applied to boxes, masks, confidences."""
    order = np.argsort(conf)[::-1]
    return masks[order], boxes[order], conf[order]


class FrameSource:
    """KITTI-360 layout reader (cvs_erosion.py:18-38, 320-335)."""

    def __init__(self, root, seq=0, bbox_dir=None):
        self.root = root
        self.velo = K.Kitti360Viewer3DRaw(seq=seq, mode='velodyne', root=root)
        self.bbox_dir = bbox_dir or os.path.join(root, 'bboxes_3D_cam0')

    def frames(self):
        return self.velo.frames()

    def points(self, frame):
        return self.velo.loadVelodyneData(frame)

    def boxes(self, frame):
        return K.load_bounding_boxes(os.path.join(self.bbox_dir, 'BBoxes_%d.json' % frame))


def collect_frames(source, detections, camera, TrVeloToCam, log=None):
    """Apply the reference's per-frame skip rules and gather what is left.
    Returns a list of dicts {frame, points, masks, bboxes}."""
    kept = []
    for frame in source.frames():
        try:
            points = source.points(frame)
        except Exception as e:                       # cvs_erosion.py:325-329
            if log:
                log('failed to load frame %d: %s' % (frame, e))
            continue
        raw = source.boxes(frame)
        if not raw:                                  # :334-335 (missing / empty JSON)
            continue
        boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw, camera), TrVeloToCam)
        det = detections(frame)
        if det is None or len(det[0]) == 0:          # no masks -> nothing to extract (:361-362)
            continue
        masks, b2d, conf = sort_by_confidence(np.asarray(det[0]), np.asarray(det[1]), np.asarray(det[2]))
        kept.append({'frame': frame, 'points': points, 'masks': masks, 'bboxes': boxes})
    return kept


def process_frames(source, detections, camera, TrVeloToCam, TrVeloToRect, master_csv_path=None,
                   erosion_kernel_size=3, erosion_iterations=1, max_depth=50, min_points=10, use_oriented=True,
                   timestamp=None, device=None, log=None):
    """Whole-sequence equivalent of cvs_erosion.process_frames: returns
    {frame: car_statistics} and appends the rows to ``master_csv_path``."""
    kept = collect_frames(source, detections, camera, TrVeloToCam, log)
    if not kept:
        return {}
    pts = np.concatenate([k['points'] for k in kept]).astype(np.float32, copy=False)
    masks = np.concatenate([k['masks'] for k in kept])
    stats = LF.process_frames_batched(pts, [len(k['points']) for k in kept], masks,
                                      [len(k['masks']) for k in kept], [k['bboxes'] for k in kept], camera,
                                      TrVeloToRect, erosion_kernel_size, erosion_iterations, max_depth,
                                      min_points, use_oriented, device)
    out = {}
    for k, st in zip(kept, stats):
        out[k['frame']] = st
        if st and master_csv_path:                   # cvs_erosion.py:375-376
            LF.append_to_master_csv(st, k['frame'], master_csv_path, timestamp)
    return out


def process_depth_maps(source, detections, camera, TrVeloToRect, max_depth=30, policy='last', frames_per_batch=32,
                       device=None, log=None):
    """Whole-sequence equivalent of seg_with_pointcloud.py:120-170 (``projectVeloToImage`` per frame):
    {frame: [(car_id, depthMap f64 [H,W])]} for every frame with at least one mask.  The masks are
    used as the detector returned them -- no confidence sort, no erosion, ``mask[v,u] > 0.5`` -- and
    the gate is ``depth < max_depth`` (30 m there).  ``frames_per_batch`` frames go through
    ``lf_depth_maps_batch`` in one call (a per-car map is 4.2 MB: 32 frames x 12 cars = 1.6 GB)."""
    todo = []
    for frame in source.frames():
        try:
            points = source.points(frame)
        except Exception as e:
            if log:
                log('failed to load frame %d: %s' % (frame, e))
            continue
        det = detections(frame)
        if det is None or len(det[0]) == 0:
            continue
        todo.append((frame, points, np.asarray(det[0])))
    out = {}
    for i in range(0, len(todo), max(1, frames_per_batch)):
        part = todo[i:i + max(1, frames_per_batch)]
        pts = np.concatenate([p for _, p, _ in part]).astype(np.float32, copy=False)
        maps = LF.per_car_depth_maps_batch(pts, [len(p) for _, p, _ in part], np.concatenate([m for _, _, m in part]),
                                           [len(m) for _, _, m in part], TrVeloToRect, camera, max_depth=max_depth,
                                           policy=policy, device=device)
        for (frame, _, _), mp in zip(part, maps):
            out[frame] = mp
    return out


def process_sweeps_streamed(paths, frame_inputs, camera, TrVeloToRect, frames_per_batch=64, erosion_kernel_size=3,
                            erosion_iterations=1, max_depth=50, min_points=10, use_oriented=True, device=None,
                            frame_ids=None):
    """Real-data loop for sequences that do not fit one batch: the ``.bin`` sweeps ``paths`` are read
    by ``io.SweepStreamer`` (reader thread -> pinned staging -> copy stream) while the previous
    batch is on the GPU; ``frame_inputs(i) -> (masks [M,H,W] uint8/float32 numpy or CUDA tensor,
    corners_velo [B,8,3] float64)`` supplies what the detector and the box loader produced for
    sweep i.  Returns the table rows ``(frame, car_id, matched_bbox_id, total, inside)`` of every
    car with at least one point, as ``append_to_master_csv`` writes them (cvs_erosion.py:232-265).
    One ``results()`` read-back per batch."""
    import torch
    from . import io as IO
    ctx = LF.get_ctx(device)
    frame_ids = list(range(len(paths))) if frame_ids is None else list(frame_ids)
    rows = []
    i0 = 0
    pending = None

    def drain(p):
        batch, ids = p
        res = batch.results()
        for f, fid in enumerate(ids):
            m0, m1 = batch.h_mask_off[f], batch.h_mask_off[f + 1]
            has_boxes = batch.h_box_off[f + 1] > batch.h_box_off[f]
            for c in range(m0, m1):
                t = int(res['total'][c])
                if t == 0 or not has_boxes:                # no points -> no row; no boxes -> no statistics
                    continue
                bi, bc = int(res['best_idx'][c]), int(res['best_cnt'][c])
                matched = bi >= 0 and bc >= min_points
                rows.append((fid, c - m0, bi if matched else -1, t, bc if matched else 0))

    for points, counts, _ in IO.SweepStreamer(paths, frames_per_batch, device=ctx.device.index):
        n = len(counts)
        masks, nm, corners, nb = [], [], [], []
        for i in range(i0, i0 + n):
            m, c = frame_inputs(i)
            m = m if isinstance(m, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(m))
            masks.append(m.to(ctx.device, non_blocking=True))
            nm.append(len(m))
            c = np.asarray(c, np.float64).reshape(-1, 8, 3)
            corners.append(c)
            nb.append(len(c))
        mt = torch.cat(masks) if masks else torch.zeros((0, camera.height, camera.width), dtype=torch.uint8,
                                                        device=ctx.device)
        batch = LF.FrameBatch(points, counts, mt, nm, np.concatenate(corners) if corners else np.zeros((0, 8, 3)), nb,
                              camera, TrVeloToRect, device=ctx.device.index)
        batch.run(erosion_kernel_size, erosion_iterations, max_depth, min_points, use_oriented, sparse_rows=True)
        if pending is not None:
            drain(pending)                                 # read batch i-1 back while batch i runs
        pending = (batch, frame_ids[i0:i0 + n])
        i0 += n
    if pending is not None:
        drain(pending)
    return rows
