"""Golden vectors for the 'next' rows of SURVEY 8(f), from the UNMODIFIED reference.

ORACLE / TEST INFRASTRUCTURE ONLY (build container: needs /root/reference).

    python -m oracle.gen_golden_next      ->  tests/golden/golden_next.json

For every KITTI360_sample frame, with the raw boxes of ``BBoxes_<frame>.json``:

* ``secondtest.is_bbox_in_camera_view`` (secondtest.py:277-359): verdict, reason and the
  counters it reports, per box;
* ``V5_ProjectingBBoxes.project_3d_bbox_to_2d`` (V5:215-252): bbox / area / avg_depth;
* ``cvs_erosion.filter_visible_bboxes`` (cvs_erosion.py:55-74): the kept indices;
* ``V5.calculate_matching_score`` (V5:277-304) and ``V4.calculate_iou_2d`` (V4:118-137)
  for every (stand-in detection, raw box) pair, detections handed over as Python floats
  (float64), plus the accepted pairs of the Hungarian step restated from
  V5:319-368 on the reference's own score function.
"""
import contextlib
import io
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import kitti360_restated as k360          # noqa: E402
from oracle import mask_provider as MP                # noqa: E402
from oracle import reference_loader as RL             # noqa: E402

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def main():
    if not RL.available():
        raise SystemExit('reference not present; golden vectors can only be generated in the build container')
    with contextlib.redirect_stdout(io.StringIO()):
        st = RL.load('secondtest')
        v5 = RL.load('V5_ProjectingBBoxes')
        v4 = RL.load('V4_BBox_IoU_filtering')
        cvs = RL.load('cvs_erosion')
    camera, TrVeloToCam, TrVeloToRect = k360.calibration_chain(RL.SAMPLE)
    with open(os.path.join(GOLDEN, 'golden.json')) as fh:
        frames = json.load(fh)['_meta']['frames']
    gold = {'_meta': {'generator': 'oracle/gen_golden_next.py running /root/reference code unmodified',
                      'numpy': np.__version__}}
    for fr in frames:
        path = os.path.join(RL.SAMPLE, 'bboxes_3D_cam0', 'BBoxes_%d.json' % fr)
        if not os.path.isfile(path):
            path = os.path.join(GOLDEN, 'kitti360_sample', 'bboxes_3D_cam0', 'BBoxes_%d.json' % fr)
        if not os.path.isfile(path):
            continue
        with open(path) as fh:
            raw = json.load(fh)
        g = {'view': [], 'rect': []}
        with contextlib.redirect_stdout(io.StringIO()):
            for b in raw:
                keep, info = st.is_bbox_in_camera_view(b, camera)
                g['view'].append({'keep': bool(keep), 'reason': info['reason'],
                                  'corners_in_view': int(info.get('corners_in_view', -1)),
                                  'corners_with_valid_depth': int(info.get('corners_with_valid_depth', -1)),
                                  'avg_depth': float(info.get('avg_depth', 0.0))})
                pinfo, _ = v5.project_3d_bbox_to_2d(b, camera)
                g['rect'].append(None if pinfo is None else
                                 {'bbox': [int(x) for x in pinfo['bbox']], 'area': int(pinfo['area']),
                                  'avg_depth': float(pinfo['avg_depth'])})
            kept = cvs.filter_visible_bboxes(raw, camera)
        g['visible_idx'] = [i for i, b in enumerate(raw) if any(b is k for k in kept)]
        # stand-in detections (deterministic; masks are inputs, YOLO is out of scope)
        masks, boxes2d, conf = MP.standin_detections(raw, camera)
        dets = [[float(x) for x in b] for b in np.asarray(boxes2d, dtype=np.float64).reshape(-1, 4)]
        g['dets'] = dets
        iou, score = [], []
        with contextlib.redirect_stdout(io.StringIO()):
            for d in dets:
                x1, y1, x2, y2 = d
                di = {'bbox': [x1, y1, x2, y2], 'center': [(x1 + x2) / 2, (y1 + y2) / 2],
                      'size': [x2 - x1, y2 - y1], 'area': (x2 - x1) * (y2 - y1)}
                ri, rs = [], []
                for b in raw:
                    pinfo, _ = v5.project_3d_bbox_to_2d(b, camera)
                    if pinfo is None:
                        ri.append(0.0)
                        rs.append(None)
                        continue
                    ri.append(float(v4.calculate_iou_2d(d, pinfo['bbox'])))
                    rs.append(float(v5.calculate_matching_score(di, pinfo)[0]))
                iou.append(ri)
                score.append(rs)
        g['iou'] = iou
        g['score'] = score
        gold[str(fr)] = g
    with open(os.path.join(GOLDEN, 'golden_next.json'), 'w') as fh:
        json.dump(gold, fh, sort_keys=True)
    print('frames:', len(gold) - 1, 'boxes:', sum(len(v['view']) for k, v in gold.items() if k != '_meta'))


if __name__ == '__main__':
    main()
