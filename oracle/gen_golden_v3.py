#!/usr/bin/env python
"""Golden vectors for the V2 / V3 return shape of calculate_car_point_statistics and for
print_summary_statistics, produced by the UNMODIFIED reference functions
(Coding_testes/V3_point_cloud_with_erosion.py:320-428, 431-468; V2 has the same bodies).

ORACLE / TEST INFRASTRUCTURE ONLY; runs only where /root/reference exists (the build container).
Inputs are the car point sets and boxes of oracle.pipeline.process_frame on the 20 sample frames with the
stand-in masks (those sets are themselves pinned to the reference's by tests/test_oracle_golden.py).
Writes tests/golden/golden_v3.json:
  per frame and variant (erosion3 / noerosion, oriented / aabb): for every dict the reference returns
  {car_id, matched_bbox_id, total, inside, outside, inside_mask sha (packbits), corners_velo sha,
  car_points sha}, and the exact text print_summary_statistics printed.
"""
import contextlib
import hashlib
import io
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

from oracle import kitti360_restated as k360, mask_provider as MP, pipeline as P, reference_loader as RL  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def digest(stats):
    out = []
    for s in stats:
        out.append({'car_id': int(s['car_id']), 'matched_bbox_id': int(s['matched_bbox_id']),
                    'total_points': int(s['total_points']), 'points_inside_bbox': int(s['points_inside_bbox']),
                    'points_outside_bbox': int(s['points_outside_bbox']),
                    'inside_percentage': float(s['inside_percentage']),
                    'outside_percentage': float(s['outside_percentage']),
                    'inside_mask_sha': None if s['inside_mask'] is None else
                    sha(np.packbits(np.asarray(s['inside_mask'], bool))),
                    'inside_mask_len': None if s['inside_mask'] is None else int(len(s['inside_mask'])),
                    'corners_velo_sha': None if s['corners_velo'] is None else sha(np.asarray(s['corners_velo'], np.float64)),
                    'car_points_sha': sha(np.asarray(s['car_points'], np.float32)),
                    'keys': sorted(s.keys())})
    return out


def main():
    if not RL.available():
        raise SystemExit('reference not present; golden vectors can only be generated in the build container')
    from conftest import load_boxes, load_frame
    v3 = RL.load('V3_point_cloud_with_erosion')
    cam, TVC, TVR = k360.calibration_chain(RL.SAMPLE)
    with open(os.path.join(ROOT, 'tests', 'golden', 'golden.json')) as f:
        frames = json.load(f)['_meta']['frames']
    gold = {'_meta': {'generator': 'oracle/gen_golden_v3.py running /root/reference V3 functions unmodified',
                      'frames': frames}}
    for fr in frames:
        pts, raw = load_frame(fr), load_boxes(fr)
        masks, _, _ = MP.standin_detections(raw, cam)
        g = {}
        for name, k in (('erosion3', 3), ('noerosion', 0)):
            w = P.process_frame(pts, raw, masks, cam, TVC, TVR, erosion_kernel_size=k)
            colors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks))]
            for oname, oriented in (('oriented', True), ('aabb', False)):
                buf = io.StringIO()
                with contextlib.redirect_stdout(buf):
                    st = v3.calculate_car_point_statistics(w['car_point_sets'], w['boxes'], colors, min_points=10,
                                                           use_oriented=oriented)
                calc_text = buf.getvalue()
                buf = io.StringIO()
                with contextlib.redirect_stdout(buf):
                    v3.print_summary_statistics(st)
                g['%s_%s' % (name, oname)] = {'stats': digest(st), 'summary_text': buf.getvalue(),
                                              'calc_text_sha': sha(np.frombuffer(calc_text.encode(), np.uint8))}
        gold[str(fr)] = g
    with open(os.path.join(ROOT, 'tests', 'golden', 'golden_v3.json'), 'w') as f:
        json.dump(gold, f, indent=1, sort_keys=True)
    n = sum(len(v['erosion3_oriented']['stats']) for k, v in gold.items() if k != '_meta')
    print('frames', len(frames), 'dicts (erosion3, oriented)', n)


if __name__ == '__main__':
    main()
