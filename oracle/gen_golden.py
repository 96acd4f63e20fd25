"""Generate ``tests/golden/`` by running the REFERENCE'S OWN, UNMODIFIED code.

ORACLE / TEST INFRASTRUCTURE ONLY.  Run in the build container (needs
/root/reference):

    python -m oracle.gen_golden [--inputs]

What is executed from /root/reference (through ``oracle/reference_loader.py``):

* ``cvs_erosion.process_frames()`` end to end -- its inline projection and
  gate (:342-367), ``image_segmentation_with_erosion`` (:77-111, with the stub
  YOLO feeding stand-in masks), ``extract_car_points_by_mask``,
  ``calculate_car_point_statistics``, ``append_to_master_csv`` (timestamp frozen
  by swapping the module's ``datetime``).  Recording wrappers around the
  module's functions capture every intermediate.
* V1/V2 functions on the captured inputs (no erosion, AABB variant,
  ``match_car_points_to_bboxes``).
* V4 ``match_detections_to_bboxes`` and V5 ``improved_match_detections_to_bboxes``.
* ``seg_with_pointcloud.projectVeloToImage()`` end to end with a recording
  matplotlib stub (the per-car depth-map loop :160-170).

``--inputs`` additionally (re)writes the input fixtures (velodyne frames as
compressed npz, calibration text, box JSON) copied as DATA from
KITTI360_sample; without it only the golden outputs are refreshed.
"""
import argparse
import contextlib
import copy
import datetime as _dt
import glob
import hashlib
import io
import json
import os
import shutil
import sys
import tempfile
import types

import cv2
import numpy as np

from . import kitti360_restated as k360
from . import mask_provider as MP
from . import reference_loader as RL

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), 'tests', 'golden')
SAMPLE_OUT = os.path.join(GOLDEN, 'kitti360_sample')
SEQ = '2013_05_28_drive_0000_sync'
FROZEN_TS = '2025-01-01T00:00:00'
DEPTH_STRIDE = 16


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def frames_available():
    fs = sorted(glob.glob(os.path.join(RL.SAMPLE, 'data_3d_raw', SEQ, 'velodyne_points', 'data', '*.bin')))
    return [int(os.path.basename(f).split('.')[0]) for f in fs]


def write_inputs():
    os.makedirs(os.path.join(SAMPLE_OUT, 'velodyne'), exist_ok=True)
    os.makedirs(os.path.join(SAMPLE_OUT, 'calibration'), exist_ok=True)
    os.makedirs(os.path.join(SAMPLE_OUT, 'bboxes_3D_cam0'), exist_ok=True)
    for name in ['perspective.txt', 'calib_cam_to_velo.txt', 'calib_cam_to_pose.txt']:
        shutil.copyfile(os.path.join(RL.SAMPLE, 'calibration', name),
                        os.path.join(SAMPLE_OUT, 'calibration', name))
    for f in glob.glob(os.path.join(RL.SAMPLE, 'bboxes_3D_cam0', '*.json')):
        shutil.copyfile(f, os.path.join(SAMPLE_OUT, 'bboxes_3D_cam0', os.path.basename(f)))
    for fr in frames_available():
        p = np.fromfile(os.path.join(RL.SAMPLE, 'data_3d_raw', SEQ, 'velodyne_points', 'data',
                                     '%010d.bin' % fr), dtype=np.float32).reshape(-1, 4)
        np.savez_compressed(os.path.join(SAMPLE_OUT, 'velodyne', '%010d.npz' % fr),
                            xyzi=np.ascontiguousarray(p.T))


# ---------------------------------------------------------------- detections

class Detections:
    """Stand-in YOLO output per frame, keyed by the image content."""

    def __init__(self, camera):
        self.by_hash = {}
        self.by_frame = {}
        for fr in frames_available():
            img_path = os.path.join(RL.SAMPLE, 'data_2d_raw', SEQ, 'image_00', 'data_rect', '%010d.png' % fr)
            jpath = os.path.join(RL.SAMPLE, 'bboxes_3D_cam0', 'BBoxes_%d.json' % fr)
            if not os.path.isfile(img_path):
                continue
            img = cv2.imread(img_path)
            self.by_hash[sha(img)] = fr
            if os.path.isfile(jpath):
                raw = json.load(open(jpath))
                masks, boxes, conf = MP.standin_detections(raw, camera)
            else:
                masks = np.zeros((0, camera.height, camera.width), np.float32)
                boxes = np.zeros((0, 4), np.float32)
                conf = np.zeros((0,), np.float32)
            perm = np.random.default_rng(fr).permutation(len(conf))
            self.by_frame[fr] = (masks, boxes, conf, perm)
        self.last_frame = None

    def __call__(self, image):
        fr = self.by_hash[sha(image)]
        self.last_frame = fr
        masks, boxes, conf, perm = self.by_frame[fr]
        if len(conf) == 0:
            return None
        return masks[perm], boxes[perm], conf[perm]        # "YOLO order" = shuffled

    def sorted_for(self, fr):
        masks, boxes, conf, _ = self.by_frame[fr]
        return masks, boxes, conf                           # confidence-descending


class FrozenDatetime:
    @staticmethod
    def now():
        return _dt.datetime.fromisoformat(FROZEN_TS)


def jsonable_stats(stats):
    out = []
    for s in stats:
        out.append({'car_id': int(s['car_id']), 'matched_bbox_id': int(s['matched_bbox_id']),
                    'total_points': int(s['total_points']),
                    'points_inside_bbox': int(s['points_inside_bbox']),
                    'points_outside_bbox': int(s['points_outside_bbox']),
                    'inside_percentage': float(s['inside_percentage']),
                    'outside_percentage': float(s['outside_percentage']),
                    'color': [int(c) for c in s['color']]})
    return out


# ---------------------------------------------------------------- cvs_erosion

def run_cvs_erosion(det, gold):
    ref = RL.load('cvs_erosion')
    RL.set_detection_provider(det)
    rec = {'frame': None}
    per = {}

    cam_cls = k360.CameraPerspective
    orig_c2i = cam_cls.cam2image

    def c2i(self, points):
        u, v, d = orig_c2i(self, points)
        if points.shape[-1] > 64 and rec['frame'] is not None:
            per.setdefault(rec['frame'], {})['proj'] = (u.copy(), v.copy(), d.copy())
        return u, v, d

    orig_load = ref.Kitti360Viewer3DRaw.loadVelodyneData

    def load(self, frame=0):
        rec['frame'] = frame
        return orig_load(self, frame)

    o_extract, o_stats, o_obb = ref.extract_car_points_by_mask, ref.calculate_car_point_statistics, \
        ref.oriented_point_in_bbox

    def extract(points_valid, u_valid, v_valid, masks, camera):
        r = o_extract(points_valid, u_valid, v_valid, masks, camera)
        per[rec['frame']]['extract'] = (points_valid.copy(), u_valid.copy(), v_valid.copy(),
                                        np.array(masks), [np.array(x) for x in r])
        return r

    def obb(points, corners):
        r = o_obb(points, corners)
        per[rec['frame']].setdefault('obb_counts', []).append(int(np.sum(r)))
        return r

    def stats(car_point_sets, bboxes_3d, colors, min_points=10):
        r = o_stats(car_point_sets, bboxes_3d, colors, min_points=min_points)
        per[rec['frame']]['stats'] = (copy.deepcopy(bboxes_3d), r)
        return r

    cam_cls.cam2image = c2i
    ref.Kitti360Viewer3DRaw.loadVelodyneData = load
    ref.extract_car_points_by_mask, ref.calculate_car_point_statistics = extract, stats
    ref.oriented_point_in_bbox = obb
    ref.datetime = FrozenDatetime
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp()
    try:
        os.chdir(tmp)
        with contextlib.redirect_stdout(io.StringIO()):
            ref.process_frames(seq=0, cam_id=0)
        csv_text = open(os.path.join(tmp, 'results', 'master_car_statistics.csv')).read()
    finally:
        os.chdir(cwd)
        cam_cls.cam2image = orig_c2i
        ref.Kitti360Viewer3DRaw.loadVelodyneData = orig_load
        ref.extract_car_points_by_mask, ref.calculate_car_point_statistics = o_extract, o_stats
        ref.oriented_point_in_bbox = o_obb
        shutil.rmtree(tmp, ignore_errors=True)

    with open(os.path.join(GOLDEN, 'master_car_statistics_erosion3.csv'), 'w') as f:
        f.write(csv_text)

    arrays = {}
    for fr, d in sorted(per.items()):
        g = gold.setdefault(str(fr), {})
        if 'proj' in d:
            u, v, depth = d['proj']
            W, H = 1408, 376
            base = (u >= 0) & (u < W) & (v >= 0) & (v < H) & (depth > 0)
            g['n_points'] = int(len(u))
            g['u_sha'], g['v_sha'] = sha(u.astype(np.int64)), sha(v.astype(np.int64))
            g['valid50_sha'] = sha((base & (depth < 50)).astype(np.uint8))
            g['valid30_sha'] = sha((base & (depth < 30)).astype(np.uint8))
            g['n_valid50'] = int((base & (depth < 50)).sum())
            g['n_valid30'] = int((base & (depth < 30)).sum())
            arrays['depth_sub_%d' % fr] = depth[::DEPTH_STRIDE].copy()
            arrays['u_sub_%d' % fr] = u[::DEPTH_STRIDE].astype(np.int64)
            arrays['v_sub_%d' % fr] = v[::DEPTH_STRIDE].astype(np.int64)
        if 'extract' in d:
            pv, uv, vv, masks, sets = d['extract']
            g['u_valid_sha'], g['v_valid_sha'] = sha(uv.astype(np.int64)), sha(vv.astype(np.int64))
            g['points_valid_sha'] = sha(pv.astype(np.float32))
            g['eroded_mask_sha'] = [sha(np.packbits(m > 0.5)) for m in masks]
            g['eroded_mask_area'] = [int((m > 0.5).sum()) for m in masks]
            g['eroded_total_points'] = [int(len(s)) for s in sets]
            g['eroded_car_points_sha'] = [sha(np.asarray(s, np.float32)) for s in sets]
        if 'stats' in d:
            boxes, st = d['stats']
            g['n_boxes_visible'] = len(boxes)
            arrays['corners_velo_%d' % fr] = np.array([b['corners_velo'] for b in boxes]).reshape(-1, 8, 3)
            g['stats_erosion3'] = jsonable_stats(st)
            nonempty = [i for i, n in enumerate(g['eroded_total_points']) if n > 0]
            cnt = np.array(d.get('obb_counts', []), np.int64).reshape(len(nonempty), len(boxes))
            full = np.zeros((len(g['eroded_total_points']), len(boxes)), np.int64)
            full[nonempty] = cnt
            arrays['inside_erosion3_%d' % fr] = full
    return per, arrays


# ---------------------------------------------------------------- V1/V2/V4/V5 function surfaces

def run_function_surfaces(det, per, gold, arrays, camera):
    v1 = RL.load('V1_BBox_Pointwise_filtering')
    v2 = RL.load('V2_point_cloud_without_erosion')
    v4 = RL.load('V4_BBox_IoU_filtering')
    v5 = RL.load('V5_ProjectingBBoxes')
    for fr, d in sorted(per.items()):
        if 'stats' not in d:
            continue
        g = gold[str(fr)]
        pv, uv, vv, _, _ = d['extract']
        boxes = d['stats'][0]
        masks, boxes2d, conf = det.sorted_for(fr)
        colors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks))]
        with contextlib.redirect_stdout(io.StringIO()):
            sets = v2.extract_car_points_by_mask(pv, uv, vv, masks, camera)
            st_obb = v2.calculate_car_point_statistics(sets, boxes, colors, min_points=10, use_oriented=True)
            st_aabb = v2.calculate_car_point_statistics(sets, boxes, colors, min_points=10, use_oriented=False)
            pairs = v1.match_car_points_to_bboxes(sets, boxes, colors, min_points=10, use_oriented=True)
            m4 = v4.match_detections_to_bboxes(boxes2d, boxes, colors, camera)
            m5 = v5.improved_match_detections_to_bboxes(boxes2d, boxes, colors, camera)
        g['raw_total_points'] = [int(len(s)) for s in sets]
        g['raw_car_points_sha'] = [sha(np.asarray(s, np.float32)) for s in sets]
        g['stats_noerosion'] = jsonable_stats(st_obb)
        g['stats_noerosion_aabb'] = jsonable_stats(st_aabb)
        g['v1_matched_counts'] = [int(p[2]) for p in pairs]
        arrays['v1_matched_corners_%d' % fr] = np.array([p[0] for p in pairs]).reshape(-1, 8, 3)
        arrays['v4_matched_corners_%d' % fr] = np.array([p[0] for p in m4]).reshape(-1, 8, 3)
        arrays['v4_matched_colors_%d' % fr] = np.array([p[1] for p in m4], np.float64).reshape(-1, 3)
        arrays['v5_matched_corners_%d' % fr] = np.array([p[0] for p in m5]).reshape(-1, 8, 3)
        arrays['v5_matched_colors_%d' % fr] = np.array([np.asarray(p[1], np.float64) for p in m5]).reshape(-1, 3)
    # a few scalar IoU known answers from the reference's own function
    rng = np.random.default_rng(7)
    kat = []
    for _ in range(64):
        a = np.sort(rng.uniform(0, 100, 4).reshape(2, 2), axis=0).T.reshape(-1)[[0, 2, 1, 3]]
        b = np.sort(rng.uniform(0, 100, 4).reshape(2, 2), axis=0).T.reshape(-1)[[0, 2, 1, 3]]
        kat.append([a.tolist(), b.tolist(), float(v4.calculate_iou_2d(a, b))])
    gold['iou2d_kat'] = kat


# ---------------------------------------------------------------- seg_with_pointcloud depth maps

class _Axes:
    def __init__(self, sink):
        self.sink = sink

    def imshow(self, arr, **kw):
        self.sink.append(np.array(arr))

    def set_title(self, *a, **k):
        pass

    def axis(self, *a, **k):
        pass


class _Fig:
    def savefig(self, path, **k):
        _Fig.saved.append(os.path.basename(path))

    saved = []


def run_depth_maps(det, gold, arrays):
    sink = []
    plt = types.ModuleType('matplotlib.pyplot')
    plt.get_cmap = lambda name: (lambda x: np.zeros(np.shape(x) + (4,)))
    plt.subplots = lambda *a, **k: (_Fig(), [_Axes(sink), _Axes([])])
    plt.close = lambda *a, **k: None
    plt.show = lambda *a, **k: None
    mpl = types.ModuleType('matplotlib')
    mpl.cm = types.ModuleType('matplotlib.cm')
    mpl.pyplot = plt
    saved = {k: sys.modules.get(k) for k in ['matplotlib', 'matplotlib.cm', 'matplotlib.pyplot']}
    sys.modules['matplotlib'], sys.modules['matplotlib.cm'], sys.modules['matplotlib.pyplot'] = mpl, mpl.cm, plt
    try:
        ref = RL.load('seg_with_pointcloud')
        ref.cam_id = 0            # the function reads the __main__ global (seg_with_pointcloud.py:121,229)
        RL.set_detection_provider(det)
        _Fig.saved = []
        with contextlib.redirect_stdout(io.StringIO()):
            ref.projectVeloToImage(seq=0)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    # file names encode (frame, car id): '<frame:010d>,depth_map_car_<id:02d>_.png'
    assert len(sink) == len(_Fig.saved)
    for name, dm in zip(_Fig.saved, sink):
        fr = int(name.split(',')[0])
        car = int(name.split('_')[3])
        nz = np.flatnonzero(dm)
        arrays['depthmap_pix_%d_%d' % (fr, car)] = nz.astype(np.int32)
        arrays['depthmap_val_%d_%d' % (fr, car)] = dm.reshape(-1)[nz]
        gold.setdefault(str(fr), {}).setdefault('depthmap_cars', []).append(car)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--inputs', action='store_true')
    args = ap.parse_args()
    if not RL.available():
        raise SystemExit('reference not present; golden vectors can only be generated in the build container')
    os.makedirs(GOLDEN, exist_ok=True)
    if args.inputs or not os.path.isdir(SAMPLE_OUT):
        write_inputs()
    camera, TrVeloToCam, TrVeloToRect = k360.calibration_chain(RL.SAMPLE)
    det = Detections(camera)
    gold = {'_meta': {'frozen_timestamp': FROZEN_TS, 'depth_stride': DEPTH_STRIDE,
                      'frames': frames_available(),
                      'generator': 'oracle/gen_golden.py running /root/reference code unmodified',
                      'numpy': np.__version__, 'cv2': cv2.__version__}}
    per, arrays = run_cvs_erosion(det, gold)
    run_function_surfaces(det, per, gold, arrays, camera)
    run_depth_maps(det, gold, arrays)
    arrays['TrVeloToRect'] = TrVeloToRect
    arrays['TrVeloToCam'] = TrVeloToCam
    arrays['K'] = camera.K
    with open(os.path.join(GOLDEN, 'golden.json'), 'w') as f:
        json.dump(gold, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(GOLDEN, 'golden_arrays.npz'), **arrays)
    print('frames with projections:', sum(1 for k, v in gold.items() if k != '_meta' and 'u_sha' in v))
    print('frames with stats:', sum(1 for k, v in gold.items() if k != '_meta' and 'stats_erosion3' in v))
    print('depth maps:', sum(len(v.get('depthmap_cars', [])) for k, v in gold.items() if k not in ('_meta', 'iou2d_kat')))


if __name__ == '__main__':
    main()
