"""Deterministic stand-in instance masks (YOLOv11 inference is out of scope).

ORACLE / TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference obtains its car masks from ``model.predict(image, classes=2,
retina_masks=True)`` (Coding_testes/cvs_erosion.py:79-94): float32 [M,H,W]
arrays of 0.0/1.0 at camera resolution, sorted by confidence descending.  The
weights are absent, so tests feed masks rendered from the frame's own 3D boxes:
the filled convex hull of each box's projected corners, dilated by a few pixels
to imitate the "bleed-out" the erosion step exists to remove.  They are inputs,
not results, so their exact shape carries no parity claim.
"""
import cv2
import numpy as np


def _hull_mask(u, v, W, H, dilate_px):
    pts = np.stack([u, v], axis=1).astype(np.int32)
    hull = cv2.convexHull(pts.reshape(-1, 1, 2))
    m = np.zeros((H, W), np.uint8)
    cv2.fillConvexPoly(m, hull, 1)
    if dilate_px > 0:
        k = 2 * dilate_px + 1
        m = cv2.dilate(m, cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k)))
    return m


def standin_detections(bboxes_3d_raw, camera, max_masks=16, dilate_px=4,
                       max_depth=60.0, min_area=150):
    """Render masks for the boxes of one frame.

    Returns ``(masks f32 [M,H,W] of 0/1, boxes_xyxy f32 [M,4], conf f32 [M])``
    in descending-area (= descending-confidence) order.
    """
    W, H = camera.width, camera.height
    cands = []
    for bbox in bboxes_3d_raw:
        c = np.array(bbox['corners_cam0'], dtype=np.float64)
        u, v, depth = camera.cam2image(c.T.copy())
        if not np.all(depth > 0.5) or float(np.mean(depth)) > max_depth:
            continue
        if u.max() < 0 or u.min() >= W or v.max() < 0 or v.min() >= H:
            continue
        u = np.clip(u, -2 * W, 3 * W)
        v = np.clip(v, -2 * H, 3 * H)
        m = _hull_mask(u, v, W, H, dilate_px)
        area = int(m.sum())
        if area < min_area:
            continue
        ys, xs = np.nonzero(m)
        box = np.array([xs.min(), ys.min(), xs.max() + 1, ys.max() + 1], np.float32)
        cands.append((area, m, box))
    cands.sort(key=lambda t: -t[0])
    cands = cands[:max_masks]
    if not cands:
        return (np.zeros((0, H, W), np.float32), np.zeros((0, 4), np.float32),
                np.zeros((0,), np.float32))
    masks = np.stack([c[1] for c in cands]).astype(np.float32)
    boxes = np.stack([c[2] for c in cands])
    conf = (0.97 - 0.03 * np.arange(len(cands))).astype(np.float32)
    return masks, boxes, conf


def random_blob_masks(rng, M, H, W, n_blobs=3, max_r=60):
    """Random unions of ellipses + salt noise, uint8 0/1 -- for property tests."""
    masks = np.zeros((M, H, W), np.uint8)
    for m in range(M):
        for _ in range(n_blobs):
            cx, cy = int(rng.integers(0, W)), int(rng.integers(0, H))
            ax, ay = int(rng.integers(2, max_r)), int(rng.integers(2, max_r))
            cv2.ellipse(masks[m], (cx, cy), (ax, ay), float(rng.uniform(0, 180)), 0, 360, 1, -1)
        holes = rng.random((H, W)) < 0.01
        masks[m][holes] = 0
        specks = rng.random((H, W)) < 0.002
        masks[m][specks] = 1
    return masks
