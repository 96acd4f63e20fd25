"""One frame through the REFERENCE'S OWN per-frame body (Coding_testes/cvs_erosion.py:338-372), with disk
I/O and YOLO inference taken out.  ORACLE / TEST INFRASTRUCTURE ONLY: the checker for the port
(oracle/pipeline.py) and the thing ``bench.py --impl reference`` / ``cpu_baseline`` time -- never the
product.

Every step that is a function in the reference is CALLED from the unmodified script (loaded by
oracle/reference_loader.py from /root/reference or its staged copy oracle/_ref):
    filter_visible_bboxes (:55-74), transform_bboxes_to_velodyne (:41-52),
    image_segmentation_with_erosion (:77-111; ``model.predict`` answers from the stub with the masks that
    are this frame's input, the confidence sort and the erosion loop are the reference's),
    extract_car_points_by_mask (:148-162), calculate_car_point_statistics (:165-229).
The two pieces the reference has inline in ``process_frames`` are spelled here statement for statement:
the projection (:342-346) and the gate + compaction (:359-367).  ``camera.cam2image`` is the restated
kitti360scripts method (the submodule is absent from the reference tree).
"""
import contextlib
import copy
import io

import numpy as np

from . import reference_loader as RL

_state = {'det': None}


def load():
    """The unmodified cvs_erosion module, its YOLO stub answering with the masks set by ``set_masks``."""
    cvs = RL.load('cvs_erosion')
    RL.set_detection_provider(lambda image: _state['det'])
    return cvs


def process_frame(cvs, points, bboxes_3d_raw, masks, camera, TrVeloToCam, TrVeloToRect, erosion_kernel_size=3,
                  quiet=True):
    """masks: float32 [M,H,W] 0/1 in confidence order (what ``result.masks.data`` holds).  Returns the list
    of per-car dicts ``calculate_car_point_statistics`` returned ([] where the reference ``continue``s)."""
    M = len(masks)
    conf = np.linspace(0.99, 0.5, M).astype(np.float32) if M else np.zeros(0, np.float32)
    _state['det'] = (masks, np.zeros((M, 4), np.float32), conf) if M else None
    sink = io.StringIO()
    with (contextlib.redirect_stdout(sink) if quiet else contextlib.nullcontext()):
        bboxes_3d_filtered = cvs.filter_visible_bboxes(copy.deepcopy(bboxes_3d_raw), camera)        # :338
        bboxes_3d = cvs.transform_bboxes_to_velodyne(bboxes_3d_filtered, TrVeloToCam)               # :339
        points_homo = points.copy()                                                                 # :342
        points_homo[:, 3] = 1
        pointsCam = np.matmul(TrVeloToRect, points_homo.T).T[:, :3]
        u, v, depth = camera.cam2image(pointsCam.T)
        u, v = u.astype(int), v.astype(int)                                                         # :346
        image = np.zeros((camera.height, camera.width, 3), np.uint8)                                # cv2.imread stand-in
        seg_image, masks_e, colors, boxes, confidences = cvs.image_segmentation_with_erosion(
            image, erosion_kernel_size=erosion_kernel_size)                                         # :356
        valid = (u >= 0) & (u < camera.width) & (v >= 0) & (v < camera.height) & (depth > 0) & (depth < 50)   # :359
        valid_indices = np.where(valid)[0]
        if len(valid_indices) == 0:
            return []
        u_valid = u[valid]
        v_valid = v[valid]
        points_valid = points[valid_indices, :3]                                                    # :367
        if masks_e is not None and len(masks_e) > 0:
            car_point_sets = cvs.extract_car_points_by_mask(points_valid, u_valid, v_valid, masks_e, camera)   # :371
            return cvs.calculate_car_point_statistics(car_point_sets, bboxes_3d, colors, min_points=10)      # :372
    return []
