"""The numpy restatement (oracle/pipeline.py) against vectors produced by the
reference's own unmodified code (oracle/gen_golden.py).  CPU only."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_boxes, load_frame, standin
from oracle import pipeline as P


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


FRAMES_FAST = [100, 947, 2098, 2449, 1866, 570]


def _frames(golden):
    return [f for f in golden['_meta']['frames'] if str(f) in golden and 'u_sha' in golden[str(f)]]


def test_fixture_inventory(golden):
    assert len(golden['_meta']['frames']) == 20
    # frame 2717 has no box file (BBoxes_2711.json is misnamed) -> reference skips it (cvs_erosion.py:332-335)
    assert '2717' not in golden
    assert len(_frames(golden)) == 19


def test_projection_matches_reference(golden, golden_arrays, oracle_calib):
    cam, TVC, TVR = oracle_calib
    assert np.array_equal(TVR, golden_arrays['TrVeloToRect'])
    stride = golden['_meta']['depth_stride']
    for fr in _frames(golden):
        g = golden[str(fr)]
        pts = load_frame(fr)
        assert len(pts) == g['n_points']
        u, v, d = P.project_points(pts, TVR, cam)
        assert sha(u.astype(np.int64)) == g['u_sha']
        assert sha(v.astype(np.int64)) == g['v_sha']
        assert sha(P.filter_valid(u, v, d, cam, 50).astype(np.uint8)) == g['valid50_sha']
        assert sha(P.filter_valid(u, v, d, cam, 30).astype(np.uint8)) == g['valid30_sha']
        np.testing.assert_allclose(d[::stride], golden_arrays['depth_sub_%d' % fr], rtol=1e-12, atol=0)


@pytest.mark.parametrize('fr', FRAMES_FAST)
def test_frame_pipeline_matches_reference(fr, golden, golden_arrays, oracle_calib):
    cam, TVC, TVR = oracle_calib
    g = golden.get(str(fr))
    if g is None or 'stats_erosion3' not in g:
        pytest.skip('reference produced no statistics for this frame')
    masks, boxes2d, conf = standin(fr, cam)
    out = P.process_frame(load_frame(fr), load_boxes(fr), masks, cam, TVC, TVR)
    assert len(out['boxes']) == g['n_boxes_visible']
    cv = np.array([b['corners_velo'] for b in out['boxes']]).reshape(-1, 8, 3)
    assert np.array_equal(cv, golden_arrays['corners_velo_%d' % fr])
    assert [sha(np.packbits(m > 0.5)) for m in out['masks_eroded']] == g['eroded_mask_sha']
    assert out['total'].tolist() == g['eroded_total_points']
    assert [sha(np.asarray(s, np.float32)) for s in out['car_point_sets']] == g['eroded_car_points_sha']
    assert np.array_equal(out['inside'], golden_arrays['inside_erosion3_%d' % fr])
    got = [{k: (float(v) if 'percentage' in k else (list(v) if k == 'color' else int(v)))
            for k, v in s.items()} for s in out['stats']]
    assert got == g['stats_erosion3']
    # no-erosion (V2) and AABB variants
    out2 = P.process_frame(load_frame(fr), load_boxes(fr), masks, cam, TVC, TVR, erosion_kernel_size=0)
    assert out2['total'].tolist() == g['raw_total_points']
    got2 = [{k: (float(v) if 'percentage' in k else (list(v) if k == 'color' else int(v)))
             for k, v in s.items()} for s in out2['stats']]
    assert got2 == g['stats_noerosion']
    out3 = P.process_frame(load_frame(fr), load_boxes(fr), masks, cam, TVC, TVR, erosion_kernel_size=0,
                           use_oriented=False)
    got3 = [{k: (float(v) if 'percentage' in k else (list(v) if k == 'color' else int(v)))
             for k, v in s.items()} for s in out3['stats']]
    assert got3 == g['stats_noerosion_aabb']
    colors = [s['color'] for s in out2['stats']]
    allcolors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks))]
    pairs = P.match_car_points_to_bboxes(out2['car_point_sets'], out2['boxes'], allcolors)
    assert [int(p[2]) for p in pairs] == g['v1_matched_counts']
    assert np.array_equal(np.array([p[0] for p in pairs]).reshape(-1, 8, 3),
                          golden_arrays['v1_matched_corners_%d' % fr])


def test_csv_bytes_match_reference(golden, oracle_calib, tmp_path):
    cam, TVC, TVR = oracle_calib
    path = str(tmp_path / 'results' / 'master_car_statistics.csv')
    ts = golden['_meta']['frozen_timestamp']
    for fr in golden['_meta']['frames']:
        raw = load_boxes(fr)
        if not raw:
            continue
        masks, _, _ = standin(fr, cam)
        if len(masks) == 0:
            continue
        out = P.process_frame(load_frame(fr), raw, masks, cam, TVC, TVR)
        P.append_to_master_csv(out['stats'], fr, path, ts)
    want = open(os.path.join(GOLDEN, 'master_car_statistics_erosion3.csv')).read()
    assert open(path).read() == want
    agg = P.analyze_master_csv(path)
    assert agg['detections'] == want.count('\n') - 1


def test_depth_maps_match_reference(golden, golden_arrays, oracle_calib):
    cam, TVC, TVR = oracle_calib
    checked = 0
    for fr in [100, 250, 1866]:
        g = golden[str(fr)]
        masks, _, conf = standin(fr, cam)
        perm = np.random.default_rng(fr).permutation(len(conf))   # the "YOLO order" gen_golden used
        u, v, d = P.project_points(load_frame(fr), TVR, cam)
        maps = P.per_car_depth_maps(u, v, d, masks[perm], cam, max_depth=30, policy='last')
        for car, dm in maps:
            nz = np.flatnonzero(dm)
            if car in g.get('depthmap_cars', []):
                assert np.array_equal(nz.astype(np.int32), golden_arrays['depthmap_pix_%d_%d' % (fr, car)])
                assert np.array_equal(dm.reshape(-1)[nz], golden_arrays['depthmap_val_%d_%d' % (fr, car)])
                checked += 1
            else:
                assert len(nz) == 0      # reference skipped it: max == 0 (seg_with_pointcloud.py:174-175)
    assert checked >= 5


def test_depth_map_loop_equals_vectorised(oracle_calib):
    cam, TVC, TVR = oracle_calib
    masks, _, _ = standin(1866, cam)
    u, v, d = P.project_points(load_frame(1866)[:40000], TVR, cam)
    a = P.per_car_depth_maps_loop(u, v, d, masks, cam)
    b = P.per_car_depth_maps(u, v, d, masks, cam)
    for (i, x), (j, y) in zip(a, b):
        assert i == j and np.array_equal(x, y)


def test_2d_matching_matches_reference(golden, golden_arrays, oracle_calib):
    cam, TVC, TVR = oracle_calib
    for a, b, want in golden['iou2d_kat']:
        assert P.calculate_iou_2d(a, b) == want
    for fr in [100, 2098, 947]:
        masks, boxes2d, conf = standin(fr, cam)
        out = P.process_frame(load_frame(fr), load_boxes(fr), masks, cam, TVC, TVR, erosion_kernel_size=0)
        colors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks))]
        m4 = P.match_detections_to_bboxes(boxes2d, out['boxes'], colors, cam)
        assert np.array_equal(np.array([p[0] for p in m4]).reshape(-1, 8, 3),
                              golden_arrays['v4_matched_corners_%d' % fr])
        pairs = P.improved_match(boxes2d, out['boxes'], cam)
        n_matched = len(pairs)
        got = np.array([out['boxes'][bi]['corners_velo'] for _, bi in pairs]).reshape(-1, 8, 3)
        assert np.array_equal(got, golden_arrays['v5_matched_corners_%d' % fr][:n_matched])


def test_live_reference_if_present(oracle_calib):
    """When /root/reference is mounted (build container) re-run two of its
    functions directly; skipped on the GPU box."""
    from oracle import reference_loader as RL
    if not RL.available():
        pytest.skip('reference tree not present')
    ref = RL.load('cvs_erosion')
    cam, TVC, TVR = oracle_calib
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(500, 3)).astype(np.float32) * 3
    corners = np.array(load_boxes(100)[0]['corners_cam0'])
    corners = corners - corners.mean(0)
    assert np.array_equal(ref.oriented_point_in_bbox(pts, corners), P.oriented_point_in_bbox(pts, corners))
    assert len(ref.filter_visible_bboxes(load_boxes(100), cam)) == len(P.filter_visible_bboxes(load_boxes(100), cam))


def test_c_restatement_matches_numpy_bit_for_bit(oracle_calib):
    """oracle/oracle_c.c (explicit fma order) == numpy/OpenBLAS on the sample frames: depth bit for bit,
    u / v exactly; containment flags equal (they are only ULP-sensitive on a face)."""
    import ctypes as C
    import subprocess
    root = os.path.join(os.path.dirname(GOLDEN), '..', 'oracle')
    subprocess.run(['make', '-s', '-C', root], check=True)
    lib = C.CDLL(os.path.join(root, '_build', 'liboracle_c.so'))
    lib.oracle_project.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.oracle_oriented_in.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    cam, TVC, TVR = oracle_calib
    for fr in (100, 2449):
        pts = np.ascontiguousarray(load_frame(fr))
        n = len(pts)
        u, v, d = np.empty(n, np.int64), np.empty(n, np.int64), np.empty(n, np.float64)
        T, K = np.ascontiguousarray(TVR), np.ascontiguousarray(cam.K)
        lib.oracle_project(T.ctypes.data, K.ctypes.data, pts.ctypes.data, n, u.ctypes.data, v.ctypes.data,
                           d.ctypes.data)
        ou, ov, od = P.project_points(pts, TVR, cam)
        assert np.array_equal(u, ou) and np.array_equal(v, ov)
        assert np.array_equal(d, od), 'numpy/OpenBLAS on this host does not accumulate k=0..3 with fma'
        boxes = P.transform_bboxes_to_velodyne(P.filter_visible_bboxes(load_boxes(fr), cam), TVC)
        xyz = np.ascontiguousarray(pts[P.filter_valid(ou, ov, od, cam, 50), :3])
        for b in boxes[:8]:
            c = np.ascontiguousarray(np.array(b['corners_velo']))
            got = np.empty(len(xyz), np.uint8)
            lib.oracle_oriented_in(xyz.ctypes.data, len(xyz), c.ctypes.data, got.ctypes.data)
            assert np.array_equal(got.astype(bool), P.oriented_point_in_bbox(xyz, c))


def test_v3_dicts_and_summary_text_match_the_reference_vectors():
    """Oracle restatement of the V2 / V3 form of calculate_car_point_statistics (V3:320-428) and of
    print_summary_statistics (V3:431-468) against tests/golden/golden_v3.json, which the unmodified
    reference functions produced (oracle/gen_golden_v3.py); the product's text formatter is host code and
    is checked here too."""
    import hashlib
    import json
    import os
    from conftest import GOLDEN, SAMPLE, load_boxes, load_frame
    from lidar_object_detection_b200 import lidarfuse as LF
    from oracle import kitti360_restated as k360, mask_provider as MP, pipeline as P

    def sha(a):
        return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]
    with open(os.path.join(GOLDEN, 'golden_v3.json')) as f:
        gold = json.load(f)
    cam, TVC, TVR = k360.calibration_chain(SAMPLE)
    checked = 0
    for fr in gold['_meta']['frames'][::3]:
        pts, raw = load_frame(fr), load_boxes(fr)
        masks, _, _ = MP.standin_detections(raw, cam)
        for name, k in (('erosion3', 3), ('noerosion', 0)):
            w = P.process_frame(pts, raw, masks, cam, TVC, TVR, erosion_kernel_size=k)
            colors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks))]
            for oname, oriented in (('oriented', True), ('aabb', False)):
                want = gold[str(fr)]['%s_%s' % (name, oname)]
                st = P.calculate_car_point_statistics(w['car_point_sets'], w['boxes'], colors, 10, oriented,
                                                      return_points=True)
                assert len(st) == len(want['stats'])
                for s, g in zip(st, want['stats']):
                    assert sorted(s.keys()) == g['keys']
                    assert [int(s[k]) for k in ('car_id', 'matched_bbox_id', 'total_points', 'points_inside_bbox',
                                                'points_outside_bbox')] == \
                        [g[k] for k in ('car_id', 'matched_bbox_id', 'total_points', 'points_inside_bbox',
                                        'points_outside_bbox')]
                    assert s['inside_percentage'] == g['inside_percentage']
                    assert sha(np.asarray(s['car_points'], np.float32)) == g['car_points_sha']
                    if g['inside_mask_sha'] is None:
                        assert s['inside_mask'] is None and s['corners_velo'] is None
                    else:
                        assert sha(np.packbits(np.asarray(s['inside_mask'], bool))) == g['inside_mask_sha']
                        assert sha(np.asarray(s['corners_velo'], np.float64)) == g['corners_velo_sha']
                    checked += 1
                assert P.summary_statistics_text(st) == want['summary_text']
                assert LF.summary_statistics_text(st) == want['summary_text']
    assert checked > 50
    assert LF.summary_statistics_text([]) == P.summary_statistics_text([]) == '\nNo car statistics to display.\n'


def test_port_and_reference_functions_give_identical_rows_on_synthetic_frames():
    """bench.py --impl reference times oracle/reference_frame.py (the reference's own functions from the
    unmodified / staged scripts); the oracle port must give the same per-car rows on the frames the bench
    uses (synthetic HDL-64 frames, float32 masks), for both erosion sizes of the workloads."""
    from conftest import SAMPLE
    from lidar_object_detection_b200 import synthetic as S
    from oracle import kitti360_restated as k360, pipeline as P, reference_loader as RL
    if not RL.available():
        pytest.skip('reference scripts neither mounted nor staged')
    from oracle import reference_frame as RF
    cvs = RF.load()
    cam, TVC, TVR = k360.calibration_chain(SAMPLE)
    cars = S.sort_cars_by_image_area(S.make_cars(2, 16, 77), cam, TVR)
    boxes = S.reference_boxes(cars, cam, TVC)
    key = ('car_id', 'matched_bbox_id', 'total_points', 'points_inside_bbox', 'points_outside_bbox')
    rows = 0
    for f in range(2):
        pts = S.numpy_frame(cars[f], 77 + f, 64, 2048)
        masks = S.numpy_masks(cars[f], cam, TVR).astype(np.float32)
        raw = [{'corners_cam0': b['corners_cam0']} for b in boxes[f]]
        for k in (3, 7):
            ref = RF.process_frame(cvs, pts, raw, masks, cam, TVC, TVR, erosion_kernel_size=k)
            port = P.process_frame(pts, raw, masks, cam, TVC, TVR, erosion_kernel_size=k)['stats']
            assert [[int(s[c]) for c in key] for s in ref] == [[int(s[c]) for c in key] for s in port]
            assert [s['inside_percentage'] for s in ref] == [s['inside_percentage'] for s in port]
            rows += len(ref)
    assert rows > 20
