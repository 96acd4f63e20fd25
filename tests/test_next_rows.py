"""SURVEY 8(f) 'next' rows: box projection / 2D matching matrices (rank 2), mask resize
(rank 1), colour buffers (rank 3), I/O formats (rank 4).

Golden vectors: tests/golden/golden_next.json, produced by oracle/gen_golden_next.py from
the UNMODIFIED reference (secondtest.py, V4, V5, cvs_erosion.py) in the build container.
CPU tests pin the host restatements against them; ``-m gpu`` tests pin the CUDA path
against both.  Bars: integers (pixel rectangles, corner counts, verdicts, kept indices)
bit-exact; IoU bit-exact; scores and mean depths within 1e-12 relative (one fp64
rounding of a different but fixed summation order, stated here).
"""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, SAMPLE, load_boxes

REASONS = {'valid': 0, 'all_behind_camera': 1, 'no_intersection': 2, 'too_small': 3}


@pytest.fixture(scope='module')
def gnext():
    with open(os.path.join(GOLDEN, 'golden_next.json')) as f:
        return json.load(f)


@pytest.fixture(scope='module')
def cam():
    from lidar_object_detection_b200 import kitti360 as K
    return K.calibration_chain(SAMPLE)[0]


def _frames(gnext):
    return sorted(int(k) for k in gnext if k != '_meta')


# ------------------------------------------------------------------ CPU: host restatements

def test_host_projection_and_scores_match_reference(gnext, cam):
    from lidar_object_detection_b200 import lidarfuse as LF, matching2d as M
    for fr in _frames(gnext):
        g = gnext[str(fr)]
        raw = load_boxes(fr)
        assert len(raw) == len(g['view'])
        kept = LF.filter_visible_bboxes(raw, cam)
        assert [i for i, b in enumerate(raw) if any(b is k for k in kept)] == g['visible_idx']
        infos = [M.project_3d_bbox_to_2d(b, cam)[0] for b in raw]
        for info, want in zip(infos, g['rect']):
            assert (info is None) == (want is None)
            if want is not None:
                assert [int(x) for x in info['bbox']] == want['bbox']
                assert int(info['area']) == want['area']
                assert float(info['avg_depth']) == want['avg_depth']
        for d, wi, ws in zip(g['dets'], g['iou'], g['score']):
            x1, y1, x2, y2 = d
            di = {'bbox': d, 'center': [(x1 + x2) / 2, (y1 + y2) / 2], 'size': [x2 - x1, y2 - y1],
                  'area': (x2 - x1) * (y2 - y1)}
            for info, a, b in zip(infos, wi, ws):
                if info is None:
                    assert b is None
                    continue
                assert M.calculate_iou_2d(d, info['bbox']) == a
                assert M.calculate_matching_score(di, info)[0] == b


# ------------------------------------------------------------------ GPU: rank 2

@pytest.mark.gpu
def test_device_box_projection_matches_reference(gnext, cam):
    from lidar_object_detection_b200 import matching2d as M
    frames = _frames(gnext)
    raws = [load_boxes(fr) for fr in frames]
    allc = np.array([b['corners_cam0'] for raw in raws for b in raw], dtype=np.float64).reshape(-1, 8, 3)
    rec, u, v, d = M.project_boxes_batched(allc, cam, return_corners=True)     # ONE launch, 992 boxes
    o = 0
    for fr, raw in zip(frames, raws):
        g = gnext[str(fr)]
        r = rec[o:o + len(raw)]
        for i, b in enumerate(raw):
            uu, vv, dd = cam.cam2image(np.array(b['corners_cam0'], dtype=np.float64).T)
            assert np.array_equal(u[o + i], uu) and np.array_equal(v[o + i], vv)
            assert np.array_equal(d[o + i], dd)                                   # bit-exact depths
        assert [i for i in range(len(raw)) if r['n_visible'][i] >= 2] == g['visible_idx']
        for i, (view, rect) in enumerate(zip(g['view'], g['rect'])):
            assert int(r['view_code'][i]) == REASONS[view['reason']], (fr, i)
            if view['reason'] == 'valid':
                assert int(r['corners_in_view'][i]) == view['corners_in_view']
                assert int(r['n_in_range'][i]) == view['corners_with_valid_depth']
                assert abs(float(r['range_avg_depth'][i]) - view['avg_depth']) <= 1e-12 * abs(view['avg_depth'])
            assert (int(r['n_front'][i]) == 0) == (rect is None)
            if rect is not None:
                assert [int(r[k][i]) for k in ('x0', 'y0', 'x1', 'y1')] == rect['bbox']
                assert abs(float(r['avg_depth'][i]) - rect['avg_depth']) <= 1e-12 * abs(rect['avg_depth'])
        # the two list-level filters, as the reference's callers use them
        assert [raw.index(b) for b in M.filter_visible_bboxes_device(raw, cam)] == g['visible_idx']
        assert [i for i, b in enumerate(raw) if any(b is k for k in M.filter_bboxes_in_camera_view(raw, cam))] == \
            [i for i, w in enumerate(g['view']) if w['keep']]
        o += len(raw)


@pytest.mark.gpu
def test_device_match_matrices_and_assignment(gnext, cam):
    from lidar_object_detection_b200 import matching2d as M
    frames = _frames(gnext)
    raws = [load_boxes(fr) for fr in frames]
    dets = [gnext[str(fr)]['dets'] for fr in frames]
    allc = np.array([b['corners_cam0'] for raw in raws for b in raw], dtype=np.float64).reshape(-1, 8, 3)
    rec = M.project_boxes_batched(allc, cam)
    mats = M.match_matrices(dets, rec, [len(r) for r in raws])                 # ONE launch, all frames
    for fr, raw, (iou, score) in zip(frames, raws, mats):
        g = gnext[str(fr)]
        assert iou.shape == (len(g['dets']), len(raw))
        for di in range(iou.shape[0]):
            for bi in range(iou.shape[1]):
                assert iou[di, bi] == g['iou'][di][bi], (fr, di, bi)              # bit-exact
                ws = g['score'][di][bi]
                if ws is None:
                    assert score[di, bi] == -np.inf
                else:
                    assert abs(score[di, bi] - ws) <= 1e-12 * max(1.0, abs(ws))
    acc = M.assign_detections_batched(dets, raws, cam)
    for fr, raw, d, a in zip(frames, raws, dets, acc):
        assert a == M.assign_detections(d, raw, cam), fr


@pytest.mark.gpu
def test_device_box_projection_edge_cases(cam):
    from lidar_object_detection_b200 import matching2d as M
    assert len(M.project_boxes_batched(np.zeros((0, 8, 3)), cam)) == 0
    c = np.zeros((3, 8, 3))
    c[0, :, 2] = -5.0                                   # all behind the camera
    c[1] = np.array([[x, y, z] for z in (10.0, 12.0) for y in (-1.0, 1.0) for x in (-1.0, 1.0)])
    c[2, :, 2] = 0.0                                    # depth exactly 0 -> -1e-6 (cam2image)
    rec, u, v, d = M.project_boxes_batched(c, cam, return_corners=True)
    assert rec['n_front'][0] == 0 and rec['view_code'][0] == 1
    assert rec['n_front'][1] == 8 and rec['view_code'][1] == 0 and rec['n_visible'][1] == 8
    assert np.all(d[2] == -1e-6) and rec['n_front'][2] == 0
    for i in range(3):
        uu, vv, dd = cam.cam2image(c[i].T.copy())
        assert np.array_equal(u[i], uu) and np.array_equal(v[i], vv) and np.array_equal(d[i], dd)
    assert M.match_matrices([], rec[:0], []) == []


# ------------------------------------------------------------------ GPU: rank 1 (mask resize)

@pytest.mark.gpu
@pytest.mark.parametrize('src', [(376, 1408), (384, 1408), (160, 608), (104, 384), (188, 704), (752, 2816),
                                 (100, 333), (3, 5), (1, 1), (20, 5000), (192, 704), (48, 176), (700, 32), (12, 1408)])
def test_device_resize_is_bit_exact_with_cv2(src):
    """cv2.resize(u8, (W, H)) (default INTER_LINEAR) against lf_resize_masks: identity,
    the 2x2-average substitution at exactly 2x, fixed-point bilinear up and down;
    binary masks (the reference's case) and full-range grey values."""
    import cv2
    from lidar_object_detection_b200 import lidarfuse as LF
    rng = np.random.default_rng(src[0] * 7919 + src[1])
    h, w = src
    imgs = [(rng.random((h, w)) < 0.3).astype(np.uint8), rng.integers(0, 256, (h, w), dtype=np.uint8),
            np.zeros((h, w), np.uint8)]
    cv2.circle(imgs[2], (w // 2, h // 2), max(1, min(h, w) // 3), 1, -1)
    stack = np.stack(imgs)
    for (W, H) in [(1408, 376), (1000, 123), (33, 17)]:
        got = LF.resize_masks(stack, (W, H))
        assert got.shape == (3, H, W) and got.dtype == np.uint8
        for i in range(3):
            assert np.array_equal(got[i], cv2.resize(imgs[i], (W, H))), (src, (W, H), i)


@pytest.mark.gpu
@pytest.mark.parametrize('src', [(376, 1408), (192, 704), (752, 2816), (96, 320), (160, 608), (50, 77), (1000, 3000),
                                 (17, 33), (1, 1)])
def test_packed_resize_is_bit_exact_with_cv2(src):
    """lf_resize_pack: packed 0/1 masks -> packed ``cv2.resize(mask01, (W, H)) != 0`` (the resize of
    cvs_erosion.py:153 and its ``> 0.5``), every branch: identity, the 2x2-average substitution at
    exactly 2x, enlarging (64-bit window), shrinking by more than 2 (multi-word spans), ragged last
    words; bounding boxes as lf_erode_pack writes them."""
    import cv2
    import torch
    from lidar_object_detection_b200 import io as IO, lidarfuse as LF
    rng = np.random.default_rng(src[0] * 31 + src[1])
    h, w = src
    imgs = [(rng.random((h, w)) < 0.5).astype(np.uint8), (rng.random((h, w)) < 0.02).astype(np.uint8),
            np.zeros((h, w), np.uint8), np.ones((h, w), np.uint8), np.zeros((h, w), np.uint8)]
    cv2.circle(imgs[2], (w // 2, h // 2), max(1, min(h, w) // 3), 1, -1)
    imgs[4][h // 3:h // 3 + max(1, h // 5), w // 4:w // 4 + max(1, w // 7)] = 1
    stack = np.stack(imgs)
    pm = LF.PackedMasks(torch.from_numpy(IO.pack_masks(stack).view(np.int32)).cuda(), h, w)
    for (W, H) in [(1408, 376), (1000, 123), (33, 17), (w, h), (max(1, w // 2), max(1, h // 2))]:
        out = LF.resize_pack(pm, (W, H))
        got = out.unpack().cpu().numpy()
        bb = out.bbox.cpu().numpy()
        for i in range(len(imgs)):
            want = cv2.resize(imgs[i], (W, H)) != 0
            assert np.array_equal(got[i] != 0, want), (src, (W, H), i)
            ys, xs = np.nonzero(want)
            if len(ys) == 0:
                assert (bb[i] == 0).all()
            else:
                assert tuple(bb[i]) == ((1 << 20) - xs.min(), (1 << 20) - ys.min(), xs.max() + 1, ys.max() + 1)
        # same answer as the uint8 resize followed by the pack
        ref = LF.erode_pack(LF.resize_masks(torch.from_numpy(stack).cuda(), (W, H)), 0, 0)
        assert torch.equal(out.words, ref.words)


@pytest.mark.gpu
def test_low_resolution_masks_follow_the_reference_order(cam):
    """Masks below camera resolution: the reference erodes them where they are
    (cvs_erosion.py:96-106) and resizes inside extract_car_points_by_mask (:153).  The
    batch path and the function surface do the same; all counters bit-exact with the oracle."""
    import cv2
    from conftest import load_frame, standin
    from lidar_object_detection_b200 import kitti360 as K, lidarfuse as LF
    from oracle import pipeline as P
    camera, TVC, TVR = K.calibration_chain(SAMPLE)
    for fr, size in ((100, (704, 188)), (947, (608, 160))):
        pts, raw = load_frame(fr), load_boxes(fr)
        full = standin(fr, camera)[0]
        low = np.stack([cv2.resize(m.astype(np.uint8), size, interpolation=cv2.INTER_NEAREST) for m in full]) \
            .astype(np.float32)
        want = P.process_frame(pts, raw, low, camera, TVC, TVR)
        boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw, camera), TVC)
        got = LF.process_frames_batched(pts, [len(pts)], low, [len(low)], [boxes], camera, TVR)[0]
        key = ('car_id', 'matched_bbox_id', 'total_points', 'points_inside_bbox', 'points_outside_bbox')
        assert [[int(s[k]) for k in key] for s in got] == [[int(s[k]) for k in key] for s in want['stats']]
        assert sum(s['total_points'] for s in got) > 100
        # function surface: extract_car_points_by_mask resizes like the reference
        u, v, d = LF.project_points(pts, TVR, camera)
        valid = LF.filter_valid(u, v, d, camera)
        er = LF.erode_masks(low)
        sets = LF.extract_car_points_by_mask(pts[valid, :3], u[valid], v[valid], er, camera)
        wsets = P.extract_car_points_by_mask(pts[valid, :3], u[valid], v[valid], P.erode_masks(low), camera)
        assert [len(s) for s in sets] == [len(s) for s in wsets]
        for a, b in zip(sets, wsets):
            assert np.array_equal(np.asarray(a, np.float32).reshape(-1, 3), np.asarray(b, np.float32).reshape(-1, 3))


# ------------------------------------------------------------------ rank 3 (colour buffers)

def test_jet_lut_known_answers():
    """Published properties of matplotlib's jet: dark blue -> blue -> cyan -> yellow -> red
    -> dark red; 256 rows in [0, 1].  (The table itself is an unpinned restatement:
    matplotlib is not in this image.)"""
    from lidar_object_detection_b200 import colour as C
    from oracle import pipeline as P
    lut = C.jet_lut()
    assert lut.shape == (256, 3) and lut.min() >= 0 and lut.max() <= 1
    assert np.allclose(lut[0], [0, 0, 0.5]) and np.allclose(lut[-1], [0.5, 0, 0])
    assert np.allclose(lut[int(0.375 * 255)], [0.0588, 1.0, 0.9108], atol=5e-3) or lut[96, 1] == 1.0
    assert lut[128, 1] == 1.0 and abs(lut[128, 0] - lut[128, 2]) < 0.1            # green-yellow middle
    assert np.array_equal(lut, P.jet_lut_restated())


@pytest.mark.gpu
def test_device_point_colours_match_the_v4_v5_loop():
    from conftest import load_frame, standin
    from lidar_object_detection_b200 import colour as C, kitti360 as K
    from oracle import pipeline as P
    camera, TVC, TVR = K.calibration_chain(SAMPLE)
    for fr in (100, 2098, 1866):
        pts = load_frame(fr)
        masks = standin(fr, camera)[0]
        colors = [(int(i * 60) % 255, int(i * 120) % 255, int(i * 180) % 255) for i in range(len(masks))]
        pv, rgb = C.colored_points(pts, TVR, masks, colors, camera, max_depth=30)
        want = P.process_frame(pts, load_boxes(fr), masks, camera, TVC, TVR, erosion_kernel_size=0, max_depth=30)
        valid = want['valid']
        assert np.array_equal(pv, pts[valid, :3])
        clouds, assigned = P.colored_cloud_reference(pts[valid, :3], want['u'][valid], want['v'][valid], masks,
                                                     colors, camera)
        expect = np.zeros((int(valid.sum()), 3))
        for sel, cloud_pts, cloud_rgb in clouds:          # draw order: later clouds overdraw earlier ones
            expect[sel] = cloud_rgb
        assert np.array_equal(rgb, expect), fr
        assert np.array_equal(np.any(rgb != 0.5, axis=1), assigned | np.any(expect != 0.5, axis=1))
        assert assigned.sum() > 50


@pytest.mark.gpu
def test_device_point_colours_match_the_same_color_and_vis_loops():
    """Same_color.py:106-132 (first matching mask wins, colour in stored order) and vis.py:175-205
    (one cloud per mask, BGR -> RGB, bg_assigned): the scripts' per-point Python loops, restated in
    the oracle, against lf_point_colors with policy first / last."""
    from conftest import load_frame, standin
    from lidar_object_detection_b200 import colour as C, kitti360 as K, lidarfuse as LF
    from oracle import pipeline as P
    camera, TVC, TVR = K.calibration_chain(SAMPLE)
    for fr in (100, 2098):
        pts = load_frame(fr)
        masks = standin(fr, camera)[0].astype(np.float32)
        masks[1:] = np.maximum(masks[1:], masks[:-1] * (np.arange(masks.shape[2]) % 3 == 0))   # overlaps between masks
        colors = [(int(i * 60) % 255, int(i * 120 + 7) % 255, int(i * 180 + 3) % 255) for i in range(len(masks))]
        u, v, d = P.project_points(pts, TVR, camera)
        cp, cc, fp = P.same_color_reference(pts, u, v, d, masks, colors, camera)
        pv, rgb = C.colored_points(pts, TVR, masks, colors, camera, max_depth=30, policy='first', swap_bgr=False)
        car = np.any(rgb != 0.5, axis=1)
        bits = LF.mask_membership(*[a[P.filter_valid(u, v, d, camera, 30)] for a in (u, v)], masks, camera)
        car = (bits.cpu().numpy() if hasattr(bits, 'cpu') else np.asarray(bits)) != 0   # a grey car colour would not show in rgb
        assert np.array_equal(pv[car], cp) and np.array_equal(rgb[car], cc), fr
        assert np.array_equal(pv[~car], fp) and np.all(rgb[~car] == 0.5)
        assert len(cp) > 50
        # vis.py: last cloud drawn wins, colours swapped to RGB
        valid = P.filter_valid(u, v, d, camera, 30)
        clouds, assigned = P.vis_cloud_reference(pts[valid, :3], u[valid], v[valid], masks, colors, camera)
        expect = np.full((int(valid.sum()), 3), 0.5)
        for idxs, col in clouds:
            expect[idxs] = col
        pv2, rgb2 = C.colored_points(pts, TVR, masks, colors, camera, max_depth=30)
        assert np.array_equal(rgb2, expect) and np.array_equal(assigned, car)
        # where masks overlap the two policies must differ
        assert np.any(np.any(rgb2 != C.colored_points(pts, TVR, masks, colors, camera, max_depth=30, policy='first')[1],
                             axis=1))
    with pytest.raises(ValueError):
        C.point_colors(np.zeros(3, np.int64), [(1, 2, 3)], policy='nearest')


@pytest.mark.gpu
def test_device_depth_overlay_matches_numpy(golden_arrays):
    from conftest import load_frame, standin
    from lidar_object_detection_b200 import colour as C, kitti360 as K, lidarfuse as LF
    from oracle import pipeline as P
    camera, TVC, TVR = K.calibration_chain(SAMPLE)
    rng = np.random.default_rng(5)
    pts = load_frame(100)
    masks = standin(100, camera)[0]
    maps = LF.per_car_depth_maps(pts, TVR, masks, camera, max_depth=30, policy='last')
    image = rng.integers(0, 256, (camera.height, camera.width, 3), dtype=np.uint8)
    n = 0
    for car_id, dm in maps:
        dm = np.asarray(dm)
        if dm.max() == 0:
            continue
        got = C.depth_overlay(dm, image)
        want = P.depth_overlay_reference(dm.copy(), image, P.jet_lut_restated())
        assert got.dtype == np.uint8 and np.array_equal(got, want)
        n += 1
    assert n >= 3
    # all cars in one launch (shared image), and a batch with one image per map; an empty map gives the image
    stack = np.stack([np.asarray(dm) for _, dm in maps] + [np.zeros_like(np.asarray(maps[0][1]))])
    got = C.depth_overlays(stack, image)
    for i, dm in enumerate(stack):
        if dm.max() > 0:
            assert np.array_equal(got[i], P.depth_overlay_reference(dm.copy(), image, P.jet_lut_restated())), i
        else:
            assert np.array_equal(got[i], image[:, :, ::-1])          # np.uint8(x / 255 * 255) round trip, RGB -> BGR
    images = np.stack([np.roll(image, 7 * i, axis=1) for i in range(len(stack))])
    got = C.depth_overlays(stack, images)
    for i, dm in enumerate(stack):
        if dm.max() > 0:
            assert np.array_equal(got[i], P.depth_overlay_reference(dm.copy(), images[i], P.jet_lut_restated())), i
    assert C.depth_overlays(stack[:0], image).shape == (0,) + image.shape
    with pytest.raises(ValueError):
        C.depth_overlays(stack, image[:-1])
    # a pixel count that is not a multiple of four takes the one-pixel-per-thread kernel
    small = rng.random((3, 5, 7)) * (rng.random((3, 5, 7)) < 0.5)
    small[2] = 0.0
    simg = rng.integers(0, 256, (3, 5, 7, 3), dtype=np.uint8)
    got = C.depth_overlays(small, simg)
    for i in range(2):
        assert np.array_equal(got[i], P.depth_overlay_reference(small[i].copy(), simg[i], P.jet_lut_restated())), i
    assert np.array_equal(got[2], simg[2][:, :, ::-1])


# ------------------------------------------------------------------ rank 4 (I/O formats)

def test_csv_formatter_is_byte_identical(tmp_path):
    """lf_format_csv_rows (plain C, host) against the committed reference CSV
    (tests/golden/master_car_statistics_erosion3.csv, written by the unmodified
    cvs_erosion.append_to_master_csv through pandas) and against the Python writer on
    rounding edge cases (x.xx5 ties, 100.0, 0.0, one-decimal values)."""
    import csv
    from lidar_object_detection_b200 import io as IO, lidarfuse as LF
    path = os.path.join(GOLDEN, 'master_car_statistics_erosion3.csv')
    with open(path, 'rb') as fh:
        want = fh.read()
    with open(path, newline='') as fh:
        recs = list(csv.DictReader(fh))
    rows = [(int(r['frame']), int(r['car_id']), int(r['matched_bbox_id']), int(r['total_points']),
             int(r['points_inside_bbox'])) for r in recs]
    ts = recs[0]['timestamp']
    assert all(r['timestamp'] == ts for r in recs)
    assert IO.format_csv_rows(rows, ts, with_header=True) == want
    # edge cases against the Python writer
    rng = np.random.default_rng(3)
    stats, rows = [], []
    cases = [(8, 1), (8, 3), (8, 5), (8, 7), (1000, 125), (200, 1), (3, 1), (3, 2), (7, 7), (9, 0), (16, 1), (32, 1),
             (64, 1), (400, 1), (800, 1), (1600, 1), (2000, 1)]
    cases += [(int(t), int(rng.integers(0, t + 1))) for t in rng.integers(1, 5000, 3000)]
    for i, (t, ins) in enumerate(cases):
        bid = -1 if ins == 0 else i % 7
        stats.append({'car_id': i, 'matched_bbox_id': bid, 'total_points': t, 'points_inside_bbox': ins,
                      'points_outside_bbox': t - ins, 'inside_percentage': (ins / t) * 100,
                      'outside_percentage': ((t - ins) / t) * 100})
        rows.append((42, i, bid, t, ins))
    p1, p2 = str(tmp_path / 'a.csv'), str(tmp_path / 'b.csv')
    LF.append_to_master_csv(stats, 42, p1, timestamp='2025-01-01T00:00:00')
    IO.append_rows_to_master_csv(rows, p2, timestamp='2025-01-01T00:00:00')
    IO.append_rows_to_master_csv(rows[:5], p2, timestamp='t2')          # appended without a second header
    LF.append_to_master_csv(stats[:5], 42, p1, timestamp='t2')
    assert open(p1, 'rb').read() == open(p2, 'rb').read()


def test_sweep_reader_and_mask_packer_host(tmp_path):
    from conftest import load_frame
    from lidar_object_detection_b200 import io as IO
    paths = []
    for fr in (100, 250):
        p = str(tmp_path / ('%010d.bin' % fr))
        load_frame(fr).astype(np.float32).tofile(p)
        paths.append(p)
    n = [IO.sweep_points(p) for p in paths]
    buf = np.zeros((sum(n), 4), np.float32)
    assert IO.read_sweeps_into(paths, buf) == n
    assert np.array_equal(buf[:n[0]], load_frame(100)) and np.array_equal(buf[n[0]:], load_frame(250))
    with pytest.raises(ValueError):
        IO.read_sweeps_into(paths, np.zeros((10, 4), np.float32))
    rng = np.random.default_rng(1)
    m = rng.random((3, 5, 70)) < 0.4
    w = IO.pack_masks_host(m)
    assert w.shape == (3, 5, 3) and w.dtype == np.uint32
    for x in (0, 31, 32, 69):
        assert np.array_equal((w[:, :, x >> 5] >> np.uint32(x & 31)) & 1, m[:, :, x])
    assert np.all(w[:, :, 2] >> np.uint32(6) == 0)                      # padding bits beyond W stay clear


@pytest.mark.gpu
def test_sweep_streamer_and_packed_mask_container(tmp_path):
    import torch
    from conftest import load_frame, standin
    from lidar_object_detection_b200 import io as IO, kitti360 as K, lidarfuse as LF
    camera, TVC, TVR = K.calibration_chain(SAMPLE)
    frames = [100, 250, 360, 644, 729, 947, 1134]
    paths = []
    for fr in frames:
        p = str(tmp_path / ('%010d.bin' % fr))
        load_frame(fr).astype(np.float32).tofile(p)
        paths.append(p)
    seen = 0
    st = IO.SweepStreamer(paths, frames_per_batch=2)
    assert len(st) == 4
    for pts, counts, ps in st:                            # 4 batches over 2 slots: both get reused
        assert pts.is_cuda and pts.shape == (sum(counts), 4)
        want = np.concatenate([load_frame(frames[seen + i]) for i in range(len(counts))])
        u, v, d = LF.project_points(pts, TVR, camera)    # consumer work on the current stream
        assert np.array_equal(pts.cpu().numpy(), want)
        uw, vw, dw = LF.project_points(want, TVR, camera)
        assert torch.equal(u.cpu(), torch.from_numpy(uw))
        seen += len(counts)
    assert seen == len(frames)
    # packed mask container: device words -> .npz -> device words, usable by the gather
    masks, _, conf = standin(100, camera)
    pm = LF.erode_pack(masks, 3, 1)
    f = str(tmp_path / 'masks_100.npz')
    IO.save_packed_masks(f, pm.words, pm.H, pm.W, conf=conf, bbox=pm.bbox)
    pm2, conf2 = IO.load_packed_masks(f)
    assert torch.equal(pm.words, pm2.words) and torch.equal(pm.bbox, pm2.bbox) and np.array_equal(conf, conf2)
    assert os.path.getsize(f) < masks.size // 8
    assert np.array_equal(pm2.words.cpu().numpy().view(np.uint32), IO.pack_masks_host(pm.unpack().cpu().numpy()))
