"""CUDA path (through the C ABI) against the oracle and the golden vectors the
reference's own code produced.  Everything here needs a B200: `-m gpu`."""
import hashlib
import os

import cv2
import numpy as np
import pytest

from conftest import GOLDEN, SAMPLE, load_boxes, load_frame, standin
from oracle import mask_provider as MP
from oracle import pipeline as P

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope='module')
def LF():
    import torch
    assert torch.cuda.is_available()
    from lidar_object_detection_b200 import lidarfuse
    return lidarfuse


@pytest.fixture(scope='module')
def calib():
    from lidar_object_detection_b200 import kitti360 as K
    return K.calibration_chain(SAMPLE)


def _frames(golden):
    return [f for f in golden['_meta']['frames'] if str(f) in golden and 'u_sha' in golden[str(f)]]


# ------------------------------------------------------------------ A3 / A4

def test_projection_bit_exact_on_all_sample_frames(LF, calib, golden, golden_arrays, oracle_calib):
    """u, v, gate: bit-exact with the reference's numpy (sha of the int64 arrays the
    reference produced); depth: bit-exact with the oracle here, <= 1e-6 relative
    (north_star tolerance) against the committed reference samples."""
    cam, TVC, TVR = calib
    ocam, _, oTVR = oracle_calib
    assert np.array_equal(TVR, golden_arrays['TrVeloToRect'])
    stride = golden['_meta']['depth_stride']
    margin = 1.0
    for fr in golden['_meta']['frames']:
        pts = load_frame(fr)
        u, v, d, pix = LF.project_points(pts, TVR, cam, max_depth=50, return_pix=True)
        assert u.dtype == np.int64 and v.dtype == np.int64 and d.dtype == np.float64
        ou, ov, od = P.project_points(pts, oTVR, ocam)
        assert np.array_equal(u, ou) and np.array_equal(v, ov)
        assert np.array_equal(d, od), 'depth not bit-identical with numpy/OpenBLAS on this host'
        valid = P.filter_valid(ou, ov, od, ocam, 50)
        assert np.array_equal(pix >= 0, valid)
        assert np.array_equal(pix[valid], (ov[valid] * cam.width + ou[valid]).astype(np.int32))
        assert np.array_equal(LF.filter_valid(u, v, d, cam, 30), P.filter_valid(ou, ov, od, ocam, 30))
        g = golden.get(str(fr))
        if g and 'u_sha' in g:
            assert sha(u) == g['u_sha'] and sha(v) == g['v_sha']
            assert sha((pix >= 0).astype(np.uint8)) == g['valid50_sha']
            np.testing.assert_allclose(d[::stride], golden_arrays['depth_sub_%d' % fr], rtol=1e-6, atol=0)
            assert np.array_equal(d[::stride], golden_arrays['depth_sub_%d' % fr])
        # distance of the pre-rounding quotients to a .5 tie
        homo = pts.astype(np.float64).copy(); homo[:, 3] = 1
        pc = (oTVR @ homo.T)
        q = (ocam.K[:3, :3] @ pc[:3])
        with np.errstate(all='ignore'):
            fu = q[0] / np.abs(q[2])
        fu = fu[valid]
        margin = min(margin, float(np.min(np.abs(np.abs(fu - np.floor(fu)) - 0.5))))
    assert margin > 1e-10, 'a projected point sits on a rounding tie: pixel parity is ULP-sensitive'


def test_projection_special_values(LF, calib, oracle_calib):
    """NaN / inf / zero-depth / huge inputs follow numpy's int64 cast (SURVEY A.2 notes)."""
    cam, TVC, TVR = calib
    ocam, _, oTVR = oracle_calib
    pts = np.array([[np.nan, 0, 0, 0], [np.inf, 1, 1, 0], [1e30, -1e30, 5, 0], [0, 0, 0, 0],
                    [1e-30, 1e-30, 1e-30, 1], [10, 0.5, -0.5, 0.2], [-10, 0.5, -0.5, 0.2],
                    [3e38, 3e38, 3e38, 0]], np.float32)
    u, v, d = LF.project_points(pts, TVR, cam)
    with np.errstate(all='ignore'):
        ou, ov, od = P.project_points(pts, oTVR, ocam)
    assert np.array_equal(u, ou) and np.array_equal(v, ov)
    assert np.array_equal(np.isnan(d), np.isnan(od))
    assert np.array_equal(d[~np.isnan(d)], od[~np.isnan(od)])
    e = LF.project_points(np.zeros((0, 4), np.float32), TVR, cam)
    assert all(len(x) == 0 for x in e)


def test_pixel_only_projection_equals_the_full_call(LF, calib):
    """lf_project asked for the gated pixel alone lets the fp32 pre-gate answer for points that
    certainly fail the gate: the pixel indices must be those of the full fp64 call, on real sweeps
    (both range limits) and on special values."""
    import torch
    from lidar_object_detection_b200._abi import check, ptr
    cam, TVC, TVR = calib
    ctx = LF.get_ctx(0)
    special = np.array([[np.nan, 0, 0, 0], [np.inf, 1, 1, 0], [1e30, -1e30, 5, 0], [0, 0, 0, 0], [3e38, 3e38, 3e38, 0],
                        [10, 0.5, -0.5, 0.2], [-10, 0.5, -0.5, 0.2], [49.99, 1, 0, 0], [50.3, 1, 0, 0]], np.float32)
    for fr in (100, 2449):
        pts = np.concatenate([load_frame(fr), special])
        for max_depth in (50.0, 30.0, 1e9):
            _, _, _, want = LF.project_points(pts, TVR, cam, max_depth=max_depth, return_pix=True)
            t = torch.from_numpy(pts).to(ctx.device)
            pix = torch.full((len(pts),), 7, dtype=torch.int32, device=ctx.device)
            check(ctx.lib.lf_project(ctx.h, ptr(t), len(pts), max_depth, None, None, None, ptr(pix), ctx.stream), ctx.h)
            got = pix.cpu().numpy()
            assert np.array_equal(got, np.asarray(want)), (fr, max_depth)
            assert np.array_equal(LF.gated_pixels(pts, TVR, cam, max_depth=max_depth), got)
            assert (got >= 0).sum() > 20000


# ------------------------------------------------------------------ A5

@pytest.mark.parametrize('W,H', [(1408, 376), (1000, 123), (33, 17), (64, 8), (128, 40), (32, 20), (16, 9)])
def test_erosion_matches_cv2(LF, W, H):
    rng = np.random.default_rng(W * 1000 + H)
    masks = MP.random_blob_masks(rng, 5, H, W, n_blobs=4, max_r=max(4, min(H, W) // 3))
    masks[3] = 1                                   # all ones: borders must not erode
    masks[4, :, :] = 0
    masks[4, H // 4:, W // 5:] = 1                 # touches two borders
    for shape, cvshape in [(LF.MORPH_ELLIPSE, cv2.MORPH_ELLIPSE), (LF.MORPH_RECT, cv2.MORPH_RECT),
                           (LF.MORPH_CROSS, cv2.MORPH_CROSS)]:
        for k in (1, 3, 5, 7, 9, 15):
            se = cv2.getStructuringElement(cvshape, (k, k))
            assert np.array_equal(LF.structuring_element(shape, k), se), (shape, k)
            for it in (1, 2, 3):
                if k * it > 21:
                    continue
                want = np.stack([cv2.erode(m * 255, se, iterations=it) == 255 for m in masks])
                got = LF.erode_pack(masks, k, it, shape).unpack().cpu().numpy().astype(bool)
                assert np.array_equal(got, want), (W, H, shape, k, it)
    # pack only, float modes, the float round trip of cvs_erosion.py:101-103
    f = masks.astype(np.float32) * rng.choice([1.0, 0.999, 0.6, 0.4], size=masks.shape).astype(np.float32)
    assert np.array_equal(LF.erode_pack(f, 0, 0).unpack().cpu().numpy() != 0, f.astype(np.uint8) != 0)
    assert np.array_equal(LF.erode_pack(f, 0, 0, float_threshold='gt_half').unpack().cpu().numpy() != 0, f > 0.5)
    ref = P.erode_masks(f, 3, 1)
    got = LF.erode_masks(f, 3, 1)
    assert got.dtype == np.float32
    assert np.array_equal(got != 0, ref.astype(np.uint8) != 0)     # what the gather sees (cvs_erosion.py:153-154)
    # float32 masks through every compiled-in element (bulk-copy path when W % 32 == 0), special values
    f[0, 0, :8] = [np.nan, -0.0, 2.0, np.inf, -np.inf, 1.0, 0.5, 0.50001]
    for mode, binar in (('ge_one', lambda a: a >= 1.0), ('gt_half', lambda a: a > 0.5)):
        with np.errstate(invalid='ignore'):
            b = binar(f).astype(np.uint8)
        for shape, cvshape in [(LF.MORPH_ELLIPSE, cv2.MORPH_ELLIPSE), (LF.MORPH_RECT, cv2.MORPH_RECT)]:
            for k, it in ((3, 1), (5, 2), (7, 1)):
                se = cv2.getStructuringElement(cvshape, (k, k))
                want = np.stack([cv2.erode(m * 255, se, iterations=it) == 255 for m in b])
                got = LF.erode_pack(f, k, it, shape, float_threshold=mode).unpack().cpu().numpy().astype(bool)
                assert np.array_equal(got, want), (W, H, mode, shape, k, it)
    assert len(LF.erode_pack(np.zeros((0, H, W), np.uint8))) == 0


def test_eroded_sample_masks_match_reference_sha(LF, calib, golden):
    cam, TVC, TVR = calib
    for fr in [100, 947, 2098]:
        g = golden[str(fr)]
        masks, _, _ = standin(fr, cam)
        got = LF.erode_masks(masks, 3, 1)
        assert [sha(np.packbits(m > 0.5)) for m in got] == g['eroded_mask_sha']
        assert [int((m > 0.5).sum()) for m in got] == g['eroded_mask_area']


# ------------------------------------------------------------------ A6 / A8 / A9 function surface

def _norm_stats(stats):
    return [{k: (float(v) if 'percentage' in k else (list(v) if k == 'color' else int(v))) for k, v in s.items()}
            for s in stats]


@pytest.mark.parametrize('fr', [100, 947, 2098, 2449, 1866, 570])
def test_function_surface_matches_oracle_and_reference(LF, calib, golden, golden_arrays, fr):
    """extract_car_points_by_mask / oriented_point_in_bbox / calculate_car_point_statistics
    called one by one, the way the V1-V3 scripts call them."""
    cam, TVC, TVR = calib
    g = golden.get(str(fr), {})
    pts, raw = load_frame(fr), load_boxes(fr)
    masks, _, _ = standin(fr, cam)
    boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw, cam), TVC)
    if 'n_boxes_visible' in g:
        assert len(boxes) == g['n_boxes_visible']
        cv = np.array([b['corners_velo'] for b in boxes]).reshape(-1, 8, 3)
        assert np.array_equal(cv, golden_arrays['corners_velo_%d' % fr])
    u, v, d = LF.project_points(pts, TVR, cam)
    valid = LF.filter_valid(u, v, d, cam, 50)
    pv, uv, vv = pts[valid, :3], u[valid], v[valid]
    want = P.process_frame(pts, raw, masks, cam, TVC, TVR)
    em = LF.erode_masks(masks, 3, 1)
    sets = LF.extract_car_points_by_mask(pv, uv, vv, em, cam)
    assert len(sets) == len(want['car_point_sets'])
    for a, b in zip(sets, want['car_point_sets']):
        assert a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a, b)
    if 'eroded_car_points_sha' in g:
        assert [sha(np.asarray(s, np.float32)) for s in sets] == g['eroded_car_points_sha']
    # containment, box by box (first 12 boxes), oriented and axis-aligned
    min_margin = np.inf
    for b in boxes[:12]:
        c = np.array(b['corners_velo'])
        assert np.array_equal(LF.oriented_point_in_bbox(pv, c), P.oriented_point_in_bbox(pv, c))
        assert np.array_equal(LF.point_in_bbox(pv, c), P.point_in_bbox(pv, c))
        rel = pv - c[0]
        for j in (1, 3, 4):
            e = c[j] - c[0]
            t = np.dot(rel, e) / np.dot(e, e)
            min_margin = min(min_margin, float(np.min(np.minimum(np.abs(t), np.abs(t - 1)))))
    assert min_margin > 1e-12, 'a point sits on a box face: counts would be ULP-sensitive (SURVEY 7.3 item 2)'
    assert LF.oriented_point_in_bbox(np.zeros((0, 3)), np.zeros((8, 3))).shape == (0,)
    colors = LF.generate_colors(len(masks))
    for oriented in (True, False):
        got = LF.calculate_car_point_statistics(sets, boxes, colors, 10, oriented)
        ref = P.calculate_car_point_statistics(want['car_point_sets'], want['boxes'], colors, 10, oriented)
        assert _norm_stats(got) == _norm_stats(ref)
        assert all(type(a['points_inside_bbox']) is type(b['points_inside_bbox']) for a, b in zip(got, ref))
    if 'stats_erosion3' in g:
        assert _norm_stats(LF.calculate_car_point_statistics(sets, boxes, colors)) == g['stats_erosion3']
        assert np.array_equal(LF.inside_count_matrix(sets, boxes), golden_arrays['inside_erosion3_%d' % fr])
    if 'v1_matched_counts' in g:
        raw_sets = LF.extract_car_points_by_mask(pv, uv, vv, masks, cam)
        assert [len(s) for s in raw_sets] == g['raw_total_points']
        pairs = LF.match_car_points_to_bboxes(raw_sets, boxes, colors)
        assert [int(p[2]) for p in pairs] == g['v1_matched_counts']
        assert np.array_equal(np.array([p[0] for p in pairs]).reshape(-1, 8, 3),
                              golden_arrays['v1_matched_corners_%d' % fr])
    assert LF.calculate_car_point_statistics(sets, [], colors) == []
    assert LF.calculate_car_point_statistics([], boxes, colors) == []


# ------------------------------------------------------------------ fused batch path

def _batch_inputs(frames, cam, TVC, LF):
    pts, masks, boxes, raws = [], [], [], []
    for fr in frames:
        raw = load_boxes(fr)
        pts.append(load_frame(fr))
        masks.append(standin(fr, cam)[0])
        raws.append(raw)
        boxes.append(LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw, cam), TVC))
    return pts, masks, boxes, raws


@pytest.mark.parametrize('variant', ['erosion3', 'noerosion', 'aabb', 'erosion5x2_30m'])
def test_fused_batch_matches_oracle_on_all_sample_frames(LF, calib, golden, variant):
    """All 20 sample frames (ragged N, M = 0..16, B = 0..300) in ONE batch:
    every counter bit-exact with the oracle, statistics equal to the reference's."""
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames']
    pts, masks, boxes, raws = _batch_inputs(frames, cam, TVC, LF)
    kw = {'erosion3': dict(erosion_kernel_size=3, erosion_iterations=1, max_depth=50, use_oriented=True),
          'noerosion': dict(erosion_kernel_size=0, erosion_iterations=1, max_depth=50, use_oriented=True),
          'aabb': dict(erosion_kernel_size=0, erosion_iterations=1, max_depth=50, use_oriented=False),
          'erosion5x2_30m': dict(erosion_kernel_size=5, erosion_iterations=2, max_depth=30, use_oriented=True)}[variant]
    corners = np.concatenate([np.array([b['corners_velo'] for b in bs], dtype=np.float64).reshape(-1, 8, 3)
                              for bs in boxes])
    batch = LF.FrameBatch(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks),
                          [len(m) for m in masks], corners, [len(b) for b in boxes], cam, TVR, want_car_bits=True)
    batch.run(min_points=10, **kw)
    res = batch.results()
    bits = batch.car_bits.cpu().numpy()
    agg = np.zeros(8, np.int64)
    for i, fr in enumerate(frames):
        want = P.process_frame(pts[i], raws[i], masks[i], cam, TVC, TVR, min_points=10, **kw)
        m0, m1 = batch.h_mask_off[i], batch.h_mask_off[i + 1]
        b0, b1 = batch.h_box_off[i], batch.h_box_off[i + 1]
        assert res['n_valid'][i] == want['valid'].sum()
        assert np.array_equal(res['total'][m0:m1], want['total'])
        assert np.array_equal(res['n_in_box'][b0:b1], want['n_in_box'])
        got_inside = res['inside'][batch.h_cnt_off[i]:batch.h_cnt_off[i + 1]].reshape(m1 - m0, b1 - b0)
        assert np.array_equal(got_inside, want['inside'])
        fb = bits[batch.h_pt_off[i]:batch.h_pt_off[i + 1]]
        member = np.zeros((m1 - m0, len(pts[i])), bool)
        if m1 > m0:
            member[:, want['valid']] = want['member']
        for c in range(m1 - m0):
            assert np.array_equal(((fb >> c) & 1).astype(bool), member[c])
        stats = batch.frame_stats(res, i) if res['n_valid'][i] else []
        assert _norm_stats(stats) == _norm_stats(want['stats'])
        g = golden.get(str(fr), {})
        key = {'erosion3': 'stats_erosion3', 'noerosion': 'stats_noerosion', 'aabb': 'stats_noerosion_aabb'}.get(variant)
        if key and key in g:
            assert _norm_stats(stats) == g[key]
        # K6 outputs and aggregates
        for s in want['stats']:
            c = s['car_id']
            if s['matched_bbox_id'] >= 0:
                assert res['best_idx'][m0 + c] == s['matched_bbox_id']
                assert res['best_cnt'][m0 + c] == s['points_inside_bbox']
        if want['stats']:
            matched = [s for s in want['stats'] if s['matched_bbox_id'] >= 0]
            agg += np.array([1, len(want['stats']), len(matched), sum(s['total_points'] for s in matched),
                             sum(s['points_inside_bbox'] for s in matched),
                             sum(s['total_points'] for s in want['stats']), 0, 0])
    assert np.array_equal(res['agg'], agg)
    # the vectorised per-car table (what an end-to-end caller reads back) = the per-frame dicts, in order
    tab = batch.table(None, 10)
    rows = [(i, s) for i in range(len(frames)) for s in (batch.frame_stats(res, i) if res['n_valid'][i] else [])]
    assert len(tab['frame']) == len(rows)
    for r, (i, s) in enumerate(rows):
        assert (tab['frame'][r], tab['car_id'][r], tab['matched_bbox_id'][r], tab['total_points'][r],
                tab['points_inside_bbox'][r], tab['points_outside_bbox'][r]) == \
            (i, s['car_id'], s['matched_bbox_id'], s['total_points'], s['points_inside_bbox'], s['points_outside_bbox'])
        assert tab['inside_percentage'][r] == s['inside_percentage']
        assert tab['outside_percentage'][r] == s['outside_percentage']
    # running the same batch again gives the same answer (counters are re-zeroed)
    batch.run(min_points=10, **kw)
    res2 = batch.results()
    for k in res:
        assert np.array_equal(res[k], res2[k]), k


@pytest.mark.parametrize('variant', ['erosion3', 'noerosion', 'erosion7', 'rect5x2_aabb'])
def test_persistent_batch_kernel_equals_separate_launches(LF, calib, golden, variant):
    """lf_process_batch (one persistent kernel: erode items and fuse items claimed from a
    dependency-ordered queue) against the oracle and against the stand-alone launches,
    on the ragged 20-frame sample batch (M = 0..16 per frame, frames without boxes)."""
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames']
    pts, masks, boxes, raws = _batch_inputs(frames, cam, TVC, LF)
    kw = {'erosion3': dict(erosion_kernel_size=3, erosion_iterations=1, max_depth=50, use_oriented=True),
          'noerosion': dict(erosion_kernel_size=0, erosion_iterations=1, max_depth=50, use_oriented=True),
          'erosion7': dict(erosion_kernel_size=7, erosion_iterations=1, max_depth=50, use_oriented=True),
          'rect5x2_aabb': dict(erosion_kernel_size=5, erosion_iterations=2, max_depth=30, use_oriented=False,
                               shape=LF.MORPH_RECT)}[variant]
    corners = np.concatenate([np.array([b['corners_velo'] for b in bs], dtype=np.float64).reshape(-1, 8, 3)
                              for bs in boxes])
    u8 = (np.concatenate(masks) != 0).astype(np.uint8)
    batch = LF.FrameBatch(np.concatenate(pts), [len(p) for p in pts], u8, [len(m) for m in masks], corners,
                          [len(b) for b in boxes], cam, TVR)
    batch.run(min_points=10, one_call=False, **kw)
    sep = batch.results()
    sep_packed = batch.packed.words.cpu().numpy().copy()
    sep_bbox = batch.mask_bbox.cpu().numpy().copy()
    batch.packed.words.zero_()
    for rep in range(3):                       # the queue is re-armed by the counter memset
        batch.run(min_points=10, one_call=True, **kw)
        got = batch.results()
        for k in sep:
            assert np.array_equal(sep[k], got[k]), (k, rep)
        assert np.array_equal(batch.packed.words.cpu().numpy(), sep_packed)
        assert np.array_equal(batch.mask_bbox.cpu().numpy(), sep_bbox)
    if 'shape' not in kw:
        for i in (0, 5, 19):
            want = P.process_frame(pts[i], raws[i], masks[i], cam, TVC, TVR, min_points=10, **kw)
            m0, m1 = batch.h_mask_off[i], batch.h_mask_off[i + 1]
            assert got['n_valid'][i] == want['valid'].sum()
            assert np.array_equal(got['total'][m0:m1], want['total'])
            gi = got['inside'][batch.h_cnt_off[i]:batch.h_cnt_off[i + 1]].reshape(want['inside'].shape)
            assert np.array_equal(gi, want['inside'])


def test_persistent_batch_kernel_many_frames(LF, calib):
    """More frames than resident CTAs' worth of work: 300 synthetic frames, persistent
    kernel == stand-alone launches on every counter."""
    from lidar_object_detection_b200 import synthetic as S
    cam, TVC, TVR = calib
    sb = S.SyntheticBatch(300, cam, TVC, TVR, ncars=16, seed=99)
    sb.batch.run(3, 1, 50, 10, True, one_call=False)
    sep = sb.batch.results()
    sb.batch.run(3, 1, 50, 10, True, one_call=True)
    got = sb.batch.results()
    for k in sep:
        assert np.array_equal(sep[k], got[k]), k
    assert got['total'].sum() > 0 and got['inside'].sum() > 0


# ------------------------------------------------------------------ A11 CSV, end to end

class _NpzSource:
    """The committed sample (sweeps stored as npz) behind the driver's FrameSource interface."""

    def __init__(self, frames):
        self._frames = frames

    def frames(self):
        return list(self._frames)

    def points(self, frame):
        return load_frame(frame)

    def boxes(self, frame):
        return load_boxes(frame)


def test_master_csv_is_byte_identical_with_the_reference(LF, calib, golden, tmp_path):
    """cfg2: all sample frames -> master_car_statistics.csv, compared byte for byte
    with the file the unmodified cvs_erosion.process_frames wrote (timestamp frozen)."""
    from lidar_object_detection_b200 import driver
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames']

    def detections(frame):
        m, b, c = standin(frame, cam)
        if len(c) == 0:
            return None
        perm = np.random.default_rng(frame).permutation(len(c))      # "YOLO order", as gen_golden fed it
        return m[perm], b[perm], c[perm]

    path = str(tmp_path / 'results' / 'master_car_statistics.csv')
    out = driver.process_frames(_NpzSource(frames), detections, cam, TVC, TVR, master_csv_path=path,
                                timestamp=golden['_meta']['frozen_timestamp'])
    want = open(os.path.join(GOLDEN, 'master_car_statistics_erosion3.csv')).read()
    assert open(path).read() == want
    assert 2717 not in out                                             # missing box file -> frame skipped
    agg = LF.analyze_master_csv(path)
    ref = P.analyze_master_csv(path)
    assert agg['frames'] == ref['frames'] and agg['detections'] == ref['detections']
    assert agg['matched'] == ref['matched'] and agg['unmatched'] == ref['unmatched']
    for k in ('matching_rate', 'avg_points_per_car', 'avg_inside_pct', 'min_inside_pct', 'max_inside_pct'):
        assert abs(agg[k] - ref[k]) <= 1e-9 * max(1.0, abs(ref[k]))


# ------------------------------------------------------------------ A10 depth maps

def test_depth_maps_match_reference_and_oracle(LF, calib, golden, golden_arrays):
    cam, TVC, TVR = calib
    checked = 0
    for fr in [100, 250, 1866]:
        g = golden[str(fr)]
        masks, _, conf = standin(fr, cam)
        perm = np.random.default_rng(fr).permutation(len(conf))
        pts = load_frame(fr)
        maps, full = LF.per_car_depth_maps(pts, TVR, masks[perm], cam, max_depth=30, policy='last', full=True)
        u, v, d = P.project_points(pts, TVR, cam)
        want = P.per_car_depth_maps(u, v, d, masks[perm], cam, max_depth=30, policy='last')
        for (car, dm), (wcar, wdm) in zip(maps, want):
            assert car == wcar and dm.dtype == np.float64 and np.array_equal(dm, wdm)
            nz = np.flatnonzero(dm)
            if car in g.get('depthmap_cars', []):
                assert np.array_equal(nz.astype(np.int32), golden_arrays['depthmap_pix_%d_%d' % (fr, car)])
                got, ref = dm.reshape(-1)[nz], golden_arrays['depthmap_val_%d_%d' % (fr, car)]
                np.testing.assert_allclose(got, ref, rtol=1e-6, atol=0)
                assert np.array_equal(got, ref)
                checked += 1
            else:
                assert len(nz) == 0
        valid = P.filter_valid(u, v, d, cam, 30)
        ref_full = np.zeros((cam.height, cam.width))
        ref_full[v[valid], u[valid]] = d[valid]
        assert np.array_equal(full, ref_full)
        near = LF.per_car_depth_maps(pts, TVR, masks[perm], cam, max_depth=30, policy='nearest')
        wnear = P.per_car_depth_maps(u, v, d, masks[perm], cam, max_depth=30, policy='nearest')
        for (car, dm), (wcar, wdm) in zip(near, wnear):
            assert np.array_equal(dm, wdm)
            assert np.all(dm[dm > 0] <= maps[car - 1][1][dm > 0])     # nearest <= last, same support
            assert np.array_equal(dm > 0, maps[car - 1][1] > 0)
    assert checked >= 5
    assert LF.per_car_depth_maps(load_frame(100), TVR, np.zeros((0, cam.height, cam.width), np.float32), cam) == []


@pytest.mark.parametrize('policy', ['last', 'nearest'])
def test_depth_maps_batch_equals_per_frame_and_oracle(LF, calib, golden, policy):
    """lf_depth_maps_batch over the 20 sample frames in one call (ragged: 0..12 masks per frame, an
    empty frame in the middle) = the oracle's per-frame maps, bit for bit (seg_with_pointcloud.py:153-170)."""
    cam, TVC, TVR = calib
    frames = list(golden['_meta']['frames'])
    pts, masks, n_pts, n_masks = [], [], [], []
    for fr in frames:
        pts.append(load_frame(fr)); masks.append(standin(fr, cam)[0])
        n_pts.append(len(pts[-1])); n_masks.append(len(masks[-1]))
    pts.insert(7, np.zeros((0, 4), np.float32)); masks.insert(7, np.zeros((0, cam.height, cam.width), np.float32))
    n_pts.insert(7, 0); n_masks.insert(7, 0)
    got, full = LF.per_car_depth_maps_batch(np.concatenate(pts), n_pts, np.concatenate(masks), n_masks, TVR, cam,
                                            max_depth=30, policy=policy, full=True)
    assert len(got) == len(n_pts) and full.shape == (len(n_pts), cam.height, cam.width)
    cars = 0
    for f in range(len(n_pts)):
        u, v, d = P.project_points(pts[f], TVR, cam) if n_pts[f] else (np.zeros(0, np.int64),) * 2 + (np.zeros(0),)
        want = P.per_car_depth_maps(u, v, d, masks[f], cam, max_depth=30, policy=policy)
        assert [c for c, _ in got[f]] == [c for c, _ in want]
        for (_, dm), (_, wdm) in zip(got[f], want):
            assert dm.dtype == np.float64 and np.array_equal(dm, wdm)
            cars += 1
        valid = P.filter_valid(u, v, d, cam, 30)
        ref_full = np.zeros((cam.height, cam.width))
        if policy == 'last':
            ref_full[v[valid], u[valid]] = d[valid]
        else:
            ref_full[:] = np.inf
            np.minimum.at(ref_full, (v[valid], u[valid]), d[valid])
            ref_full[np.isinf(ref_full)] = 0.0
        assert np.array_equal(full[f], ref_full), f
    assert cars == sum(n_masks) >= 100
    # one frame through the batch entry point = the single-frame entry point
    one = LF.per_car_depth_maps_batch(pts[0], [n_pts[0]], masks[0], [n_masks[0]], TVR, cam, policy=policy)[0]
    ref = LF.per_car_depth_maps(pts[0], TVR, masks[0], cam, policy=policy)
    assert all(np.array_equal(a[1], b[1]) for a, b in zip(one, ref)) and len(one) == len(ref)
    assert LF.per_car_depth_maps_batch(np.zeros((0, 4), np.float32), [], np.zeros((0, cam.height, cam.width), np.uint8),
                                       [], TVR, cam) == []


def test_sequence_depth_maps_match_the_reference_vectors(LF, calib, golden, golden_arrays):
    """cfg2: driver.process_depth_maps over all sample frames (batches of 8) against the vectors the
    unmodified seg_with_pointcloud.projectVeloToImage produced (depth <= 1e-6 relative by contract,
    bit-identical in fact) and against the oracle for the frames without a recorded vector."""
    from lidar_object_detection_b200 import driver
    cam, TVC, TVR = calib
    frames = list(golden['_meta']['frames'])

    def det(fr):                     # the detector order the golden vectors were recorded with
        m, b, c = standin(fr, cam)
        perm = np.random.default_rng(fr).permutation(len(c))
        return m[perm], b[perm], c[perm]

    got = driver.process_depth_maps(_NpzSource(frames), det, cam, TVR, frames_per_batch=8)
    checked = 0
    for fr in frames:
        masks = det(fr)[0]
        if len(masks) == 0:
            assert fr not in got
            continue
        assert [c for c, _ in got[fr]] == list(range(1, len(masks) + 1))
        for car, dm in got[fr]:
            key = 'depthmap_pix_%d_%d' % (fr, car)
            nz = np.flatnonzero(dm)
            if car not in golden[str(fr)].get('depthmap_cars', []):
                assert len(nz) == 0
            elif key in golden_arrays:
                assert np.array_equal(nz.astype(np.int32), golden_arrays[key])
                ref = golden_arrays['depthmap_val_%d_%d' % (fr, car)]
                np.testing.assert_allclose(dm.reshape(-1)[nz], ref, rtol=1e-6, atol=0)
                assert np.array_equal(dm.reshape(-1)[nz], ref)
                checked += 1
    assert checked >= 90
    fr = [f for f in frames if f in got][3]
    u, v, d = P.project_points(load_frame(fr), TVR, cam)
    for (car, dm), (wcar, wdm) in zip(got[fr], P.per_car_depth_maps(u, v, d, det(fr)[0], cam, max_depth=30)):
        assert car == wcar and np.array_equal(dm, wdm)


# ------------------------------------------------------------------ synthetic batches (cfg3-shaped)

def test_synthetic_batch_matches_oracle_and_invariants(LF, calib):
    """A small cfg3-shaped batch generated on the device: a few frames against the
    oracle (all counters bit-exact), the rest through size-independent properties."""
    from lidar_object_detection_b200 import synthetic as S
    cam, TVC, TVR = calib
    sb = S.SyntheticBatch(24, cam, TVC, TVR, ncars=16, seed=1234)
    sb.batch.run(3, 1, 50, 10, True)
    res = sb.batch.results()
    b = sb.batch
    for f in (0, 7, 23):
        pts, masks, boxes = sb.host_frame(f)
        want = P.process_frame(pts, boxes, masks.astype(np.float32), cam, TVC, TVR)
        assert len(want['boxes']) == sb.n_boxes[f]
        assert res['n_valid'][f] == want['valid'].sum()
        assert np.array_equal(res['total'][b.h_mask_off[f]:b.h_mask_off[f + 1]], want['total'])
        assert np.array_equal(res['n_in_box'][b.h_box_off[f]:b.h_box_off[f + 1]], want['n_in_box'])
        got = res['inside'][b.h_cnt_off[f]:b.h_cnt_off[f + 1]].reshape(16, sb.n_boxes[f])
        assert np.array_equal(got, want['inside'])
        assert _norm_stats(b.frame_stats(res, f)) == _norm_stats(want['stats'])
    tot = res['total'].reshape(24, 16)
    assert np.all(res['n_valid'] > 0.05 * sb.n) and np.all(res['n_valid'] < 0.5 * sb.n)
    assert np.all(tot <= res['n_valid'][:, None])
    for f in range(24):
        ins = res['inside'][b.h_cnt_off[f]:b.h_cnt_off[f + 1]].reshape(16, sb.n_boxes[f])
        assert np.all(ins <= tot[f][:, None])
        assert np.all(ins <= res['n_in_box'][b.h_box_off[f]:b.h_box_off[f + 1]][None, :])
    # erosion is anti-extensive: totals can only shrink (SURVEY Appendix B observation)
    sb.batch.run(0, 0, 50, 10, True)
    res0 = sb.batch.results()
    assert np.all(res0['total'] >= res['total']) and np.any(res0['total'] > res['total'])
    sb.batch.run(7, 1, 50, 10, True)
    assert np.all(sb.batch.results()['total'] <= res['total'])
    assert np.array_equal(res0['n_valid'], res['n_valid']) and np.array_equal(res0['n_in_box'], res['n_in_box'])


def test_dense_stress_shape_cfg5(LF, calib):
    """BASELINE configs[4] in small: 2 097 152-point frames x 64 instances, 7x7 ellipse
    erosion.  Frame 0 against the oracle (every counter bit-exact); persistent kernel ==
    stand-alone launches on all frames."""
    from lidar_object_detection_b200 import synthetic as S
    cam, TVC, TVR = calib
    sb = S.SyntheticBatch(3, cam, TVC, TVR, ncars=64, seed=4242, rings=64, azimuths=32768)
    assert sb.n == 2097152
    b = sb.batch
    b.run(7, 1, 50, 10, True)
    res = b.results()
    pts, masks, boxes = sb.host_frame(0)
    want = P.process_frame(pts, boxes, masks.astype(np.float32), cam, TVC, TVR, erosion_kernel_size=7)
    assert res['n_valid'][0] == want['valid'].sum()
    assert np.array_equal(res['total'][:64], want['total'])
    assert np.array_equal(res['n_in_box'][:sb.n_boxes[0]], want['n_in_box'])
    assert np.array_equal(res['inside'][:b.h_cnt_off[1]].reshape(64, sb.n_boxes[0]), want['inside'])
    assert _norm_stats(b.frame_stats(res, 0)) == _norm_stats(want['stats'])
    assert res['total'].sum() > 10000
    b.run(7, 1, 50, 10, True, one_call=True)
    got = b.results()
    for k in res:
        assert np.array_equal(res[k], got[k]), k


def test_mask_bounding_boxes(LF):
    """lf_erode_pack's optional per-mask bounding box (the fused pass culls masks with it)."""
    rng = np.random.default_rng(11)
    for (H, W) in [(376, 1408), (50, 70)]:
        masks = MP.random_blob_masks(rng, 6, H, W, n_blobs=2, max_r=20)
        masks[0] = 0
        masks[1] = 0
        masks[1, H - 1, W - 1] = 1
        masks[2] = 1
        for k, it in [(0, 0), (3, 1), (5, 2)]:
            pm = LF.erode_pack(masks, k, it)
            got = pm.unpack().cpu().numpy()
            bb = pm.bbox.cpu().numpy()
            for m in range(len(masks)):
                ys, xs = np.nonzero(got[m])
                if len(ys) == 0:
                    assert bb[m].tolist() == [0, 0, 0, 0]
                else:
                    assert bb[m].tolist() == [(1 << 20) - xs.min(), (1 << 20) - ys.min(), xs.max() + 1, ys.max() + 1]


def test_point_pass_fp32_classification_equals_the_exact_mode(LF, calib, golden):
    """The point pass decides most points from an fp32 residual test (fuse_body.cuh, stage 3) and
    sends the rest through the float64 chain; LF_FUSE_EXACT sends everything through it.  Both
    must give identical counters and membership bits: on the ragged real batch, on a synthetic
    batch, and on an adversarial frame whose points sit on pixel-rounding ties, on the image
    border and on the range gate (those must all land in the exact path and agree with numpy)."""
    from lidar_object_detection_b200 import synthetic as S
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames']
    pts, masks, boxes, raws = _batch_inputs(frames, cam, TVC, LF)
    corners = np.concatenate([np.array([b['corners_velo'] for b in bs], dtype=np.float64).reshape(-1, 8, 3)
                              for bs in boxes])
    batch = LF.FrameBatch(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks),
                          [len(m) for m in masks], corners, [len(b) for b in boxes], cam, TVR, want_car_bits=True)
    for md in (50, 30):
        batch.run(3, 1, md, 10, True)
        fast, fbits = batch.results(), batch.car_bits.cpu().numpy().copy()
        batch.car_bits.fill_(-1)
        batch.run(3, 1, md, 10, True, exact=True)
        ex = batch.results()
        for k in fast:
            assert np.array_equal(fast[k], ex[k]), (k, md)
        assert np.array_equal(fbits, batch.car_bits.cpu().numpy())
    sb = S.SyntheticBatch(24, cam, TVC, TVR, ncars=16, seed=99)
    sb.batch.run(3, 1, 50, 10, True)
    fast = sb.batch.results()
    sb.batch.run(3, 1, 50, 10, True, exact=True)
    ex = sb.batch.results()
    for k in fast:
        assert np.array_equal(fast[k], ex[k]), k
    # adversarial frame: back-project pixel positions u + 0.5 + delta (ties and near-ties), the image
    # border +-1 px and depths around the gate into the velodyne frame, round to float32
    rng = np.random.default_rng(5)
    n = 200000
    K = np.asarray(cam.K)[:3, :3]
    delta = rng.choice([0.0, 1e-7, -1e-7, 1e-5, -1e-5, 1e-3, -1e-3, 0.25], n)
    u = rng.integers(-2, cam.width + 2, n) + 0.5 + delta
    v = rng.integers(-2, cam.height + 2, n) + 0.5 + rng.choice([0.0, 1e-6, -1e-6, 0.3], n)
    d = rng.choice([0.05, 0.5, 3.0, 12.0, 29.9999, 30.0, 30.0001, 49.99999, 50.0, 50.00001, 70.0, -5.0], n) \
        * (1 + rng.choice([0.0, 1e-7, -1e-7], n))
    cam_pts = np.stack([(u - K[0, 2]) * d / K[0, 0], (v - K[1, 2]) * d / K[1, 1], d, np.ones(n)])
    velo = (np.linalg.inv(TVR) @ cam_pts).T.astype(np.float32)
    velo[:, 3] = 0
    full = np.ones((2, cam.height, cam.width), np.uint8)
    full[1, ::2] = 0
    c0 = np.array([b['corners_velo'] for b in boxes[0]], dtype=np.float64).reshape(-1, 8, 3)
    adv = LF.FrameBatch(velo, [n], full, [2], c0, [len(c0)], cam, TVR, want_car_bits=True)
    for md in (50, 30):
        uu, vv, dd = P.project_points(velo, TVR, cam)
        valid = P.filter_valid(uu, vv, dd, cam, md)
        adv.run(0, 0, md, 10, True)
        res = adv.results()
        bits = adv.car_bits.cpu().numpy()
        assert res['n_valid'][0] == valid.sum()
        assert np.array_equal((bits & 1).astype(bool), valid)
        assert np.array_equal(((bits >> 1) & 1).astype(bool), valid & (vv % 2 == 1))
        assert res['total'].tolist() == [valid.sum(), (valid & (vv % 2 == 1)).sum()]
        adv.run(0, 0, md, 10, True, exact=True)
        ex = adv.results()
        for k in res:
            assert np.array_equal(res[k], ex[k]), (k, md)


def test_point_pass_segment_tails_and_many_tiny_frames(LF, calib):
    """Frame lengths around the CTA geometry (whole iterations, one point more, one less, a few
    points) and a batch of more than 65 535 frames (several launches inside lf_fuse_batch):
    totals equal the per-point membership bits and the numpy projection."""
    cam, TVC, TVR = calib
    pts0 = load_frame(100)
    masks0 = standin(100, cam)[0][:3]
    lens = [1, 31, 32, 33, 255, 256, 257, 2047, 2048, 2049, 4097, 16384, 16385, 20000]
    pts = np.concatenate([pts0[20000:20000 + n] for n in lens])
    masks = np.concatenate([masks0] * len(lens))
    b = LF.FrameBatch(pts, lens, masks, [3] * len(lens), np.zeros((0, 8, 3)), [0] * len(lens), cam, TVR,
                      want_car_bits=True)
    b.run(0, 0, 50, 10, True)
    res = b.results()
    bits = b.car_bits.cpu().numpy()
    o = 0
    for f, n in enumerate(lens):
        p = pts0[20000:20000 + n]
        uu, vv, dd = P.project_points(p, TVR, cam)
        valid = P.filter_valid(uu, vv, dd, cam, 50)
        assert res['n_valid'][f] == valid.sum(), n
        for c in range(3):
            member = np.zeros(n, bool)
            member[valid] = masks0[c][vv[valid], uu[valid]] > 0.5
            assert np.array_equal(((bits[o:o + n] >> c) & 1).astype(bool), member), (n, c)
            assert res['total'][3 * f + c] == member.sum()
        o += n
    # 70 000 frames of 40 points, one shared tiny mask set per frame is too much memory: no masks,
    # one box per frame; n_valid and n_in_box per frame against numpy
    F, n = 70000, 40
    tiny = pts0[10300:10300 + n]
    big = np.tile(tiny, (F, 1))
    c0 = np.array(LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(load_boxes(100), cam), TVC)[19]
                  ['corners_velo']).reshape(1, 8, 3)
    bb = LF.FrameBatch(big, [n] * F, None, [0] * F, np.tile(c0, (F, 1, 1)), [1] * F, cam, TVR)
    bb.run(0, 0, 50, 10, True)
    r = bb.results()
    uu, vv, dd = P.project_points(tiny, TVR, cam)
    valid = P.filter_valid(uu, vv, dd, cam, 50)
    assert np.all(r['n_valid'] == valid.sum()) and valid.sum() > 0
    nin = P.oriented_point_in_bbox(tiny[valid, :3], c0[0]).sum()
    assert nin > 0 and np.all(r['n_in_box'] == nin)


def _sha16(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def _v3_digest(stats):
    out = []
    for s in stats:
        out.append({'car_id': int(s['car_id']), 'matched_bbox_id': int(s['matched_bbox_id']),
                    'total_points': int(s['total_points']), 'points_inside_bbox': int(s['points_inside_bbox']),
                    'points_outside_bbox': int(s['points_outside_bbox']),
                    'inside_percentage': float(s['inside_percentage']),
                    'outside_percentage': float(s['outside_percentage']),
                    'inside_mask_sha': None if s['inside_mask'] is None else
                    _sha16(np.packbits(np.asarray(s['inside_mask'], bool))),
                    'inside_mask_len': None if s['inside_mask'] is None else int(len(s['inside_mask'])),
                    'corners_velo_sha': None if s['corners_velo'] is None else
                    _sha16(np.asarray(s['corners_velo'], np.float64)),
                    'car_points_sha': _sha16(np.asarray(s['car_points'], np.float32)),
                    'keys': sorted(s.keys())})
    return out


@pytest.fixture(scope='module')
def golden_v3():
    import json
    import os
    from conftest import GOLDEN
    with open(os.path.join(GOLDEN, 'golden_v3.json')) as f:
        return json.load(f)


def test_v3_return_shape_matches_the_reference_vectors(LF, calib, golden, golden_v3, capsys):
    """calculate_car_point_statistics(return_points=True) -- the V2 / V3 dicts with corners_velo,
    inside_mask and car_points (V3:386-398, 416-428) -- and print_summary_statistics (V3:431-468)
    against what the UNMODIFIED reference functions returned / printed (oracle/gen_golden_v3.py):
    the function surface on every sample frame, the batch path (membership bits -> car points)
    on all frames in one batch."""
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames']
    pts, masks, boxes, raws = _batch_inputs(frames, cam, TVC, LF)
    for variant, k in (('erosion3', 3), ('noerosion', 0)):
        batch = LF.process_frames_batched(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks),
                                          [len(m) for m in masks], boxes, cam, TVR, erosion_kernel_size=k,
                                          return_points=True)
        for i, fr in enumerate(frames):
            g = golden_v3[str(fr)]
            assert _v3_digest(batch[i]) == g[variant + '_oriented']['stats'], (fr, variant, 'batch')
            assert LF.summary_statistics_text(batch[i]) == g[variant + '_oriented']['summary_text']
            if fr not in (100, 947, 2449, 570):
                continue
            u, v, d = LF.project_points(pts[i], TVR, cam)
            valid = LF.filter_valid(u, v, d, cam, 50)
            em = LF.erode_masks(masks[i], 3, 1) if k else masks[i]
            sets = LF.extract_car_points_by_mask(pts[i][valid, :3], u[valid], v[valid], em, cam)
            colors = LF.generate_colors(len(masks[i]))
            for oname, oriented in (('oriented', True), ('aabb', False)):
                st = LF.calculate_car_point_statistics(sets, boxes[i], colors, 10, oriented, return_points=True)
                want = g['%s_%s' % (variant, oname)]
                assert _v3_digest(st) == want['stats'], (fr, variant, oname)
                capsys.readouterr()
                LF.print_summary_statistics(st)
                assert capsys.readouterr().out == want['summary_text']
                ref = P.calculate_car_point_statistics(sets, boxes[i], colors, 10, oriented, return_points=True)
                assert _v3_digest(ref) == want['stats']


def test_extension_metrics_match_the_oracle(LF, calib, golden):
    """Point-level precision / recall / IoU / match ratio (north-star item 5; README.md:14, 38 -- the
    reference names them and has no code for them, SURVEY 0.1 D3, so parity is against the oracle's
    definition): FrameBatch.frame_stats(extended=True) and process_frames_batched(extended=True) on the 20
    sample frames and on a synthetic batch, every value equal to oracle.pipeline.extension_metrics of the
    oracle's own counters; match_ratio is inf when no point of the car lies outside the box."""
    from lidar_object_detection_b200 import synthetic as S
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames']
    pts, masks, boxes, raws = _batch_inputs(frames, cam, TVC, LF)
    keys = ('precision', 'recall', 'point_iou', 'match_ratio')

    def want_metrics(w):
        out = []
        for s in w['stats']:
            if s['matched_bbox_id'] < 0:
                out.append(P.extension_metrics(0, s['total_points'], 0))
            else:
                out.append(P.extension_metrics(s['points_inside_bbox'], s['total_points'],
                                               w['n_in_box'][s['matched_bbox_id']]))
        return out

    got = LF.process_frames_batched(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks),
                                    [len(m) for m in masks], boxes, cam, TVR, extended=True)
    n = 0
    for i in range(len(frames)):
        w = P.process_frame(pts[i], raws[i], masks[i], cam, TVC, TVR)
        wm = want_metrics(w)
        assert _norm_stats([{k: v for k, v in s.items() if k not in keys} for s in got[i]]) == _norm_stats(w['stats'])
        assert len(got[i]) == len(wm)
        for s, m in zip(got[i], wm):
            for k in keys:
                assert s[k] == m[k], (frames[i], s['car_id'], k)       # same float64 expression: exact
            assert 0.0 <= s['precision'] <= 1.0 and 0.0 <= s['recall'] <= 1.0 and 0.0 <= s['point_iou'] <= 1.0
            if s['matched_bbox_id'] >= 0:
                assert s['precision'] == s['inside_percentage'] / 100 or abs(s['precision'] - s['inside_percentage'] / 100) < 1e-12
            n += 1
    assert n > 80
    # synthetic frames: tight stand-in masks make cars whose every point is inside the box -> inf ratio
    sb = S.SyntheticBatch(12, cam, TVC, TVR, ncars=16, seed=321)
    sb.batch.run(3, 1, 50, 10, True)
    res = sb.batch.results()
    infs = 0
    for f in range(12):
        p, m, b = sb.host_frame(f)
        w = P.process_frame(p, b, m.astype(np.float32), cam, TVC, TVR)
        st = sb.batch.frame_stats(res, f, extended=True)
        for s, mm in zip(st, want_metrics(w)):
            for k in keys:
                assert s[k] == mm[k], (f, s['car_id'], k)
            infs += s['match_ratio'] == float('inf')
    assert LF.extension_metrics(5, 5, 9) == {'precision': 1.0, 'recall': 5 / 9, 'point_iou': 5 / 9,
                                             'match_ratio': float('inf')}
    assert LF.extension_metrics(0, 0, 0) == {'precision': 0.0, 'recall': 0.0, 'point_iou': 0.0,
                                             'match_ratio': float('inf')}


def test_device_compaction_and_gate_kernels(LF, calib):
    """lf_filter_valid against the numpy gate (cvs_erosion.py:359) including the INT64_MIN pixels numpy
    gives NaN / inf points, and lf_car_counts + lf_car_compact against boolean indexing
    (cvs_erosion.py:156-160): 64 overlapping random membership words over several block sizes, float32
    rows of stride 4 and float64 rows of stride 3, point indices returned in order."""
    import ctypes as C
    import torch
    from lidar_object_detection_b200 import _abi
    cam, TVC, TVR = calib
    rng = np.random.default_rng(3)
    n = 50000
    u = rng.integers(-50, cam.width + 50, n)
    v = rng.integers(-50, cam.height + 50, n)
    d = rng.uniform(-5, 60, n)
    u[:5] = np.iinfo(np.int64).min
    d[5:9] = [0.0, 50.0, np.nan, np.inf]
    for md in (50, 30):
        want = (u >= 0) & (u < cam.width) & (v >= 0) & (v < cam.height) & (d > 0) & (d < md)
        got = LF.filter_valid(u, v, d, cam, md)
        assert got.dtype == np.bool_ and np.array_equal(got, want)
    assert LF.filter_valid(u[:0], v[:0], d[:0], cam).shape == (0,)
    ctx = LF.get_ctx()
    lib = ctx.lib
    p = lambda t: C.c_void_p(t.data_ptr())
    for n, m, f64 in ((1, 1, False), (31, 3, False), (2048, 64, True), (2049, 17, False), (70001, 64, False), (5000, 40, True)):
        dens = rng.uniform(0, 0.3, m)
        member = rng.random((m, n)) < dens[:, None]
        member[m // 2] = False                                  # an empty car in the middle
        bits = np.zeros(n, np.uint64)
        for c in range(m):
            bits |= member[c].astype(np.uint64) << np.uint64(c)
        stride = 3 if f64 else 4
        pts = rng.normal(size=(n, stride)).astype(np.float64 if f64 else np.float32)
        tb = torch.from_numpy(bits.view(np.int64)).cuda()
        tp = torch.from_numpy(pts).cuda()
        scratch = torch.empty(int(lib.lf_car_scratch_elems(n, m)), dtype=torch.int32, device='cuda')
        counts = torch.full((m,), -1, dtype=torch.int32, device='cuda')
        assert lib.lf_car_counts(ctx.h, p(tb), n, m, p(scratch), p(counts), ctx.stream) == _abi.LF_OK
        hc = counts.cpu().numpy()
        assert np.array_equal(hc, member.sum(axis=1))
        tot = int(hc.sum())
        out = torch.full((max(tot, 1), 3), 7.0, dtype=tp.dtype, device='cuda')
        idx = torch.full((max(tot, 1),), -1, dtype=torch.int32, device='cuda')
        assert lib.lf_car_compact(ctx.h, p(tb), n, m, p(tp), int(f64), stride, p(scratch), p(out), p(idx),
                                  ctx.stream) == _abi.LF_OK
        ho, hi, o = out.cpu().numpy(), idx.cpu().numpy(), 0
        for c in range(m):
            sel = np.nonzero(member[c])[0]
            assert np.array_equal(hi[o:o + len(sel)], sel), (n, m, c)
            assert np.array_equal(ho[o:o + len(sel)], pts[sel, :3])
            o += len(sel)
        assert o == tot


# ------------------------------------------------------------------ ragged / degenerate batches

@pytest.mark.parametrize('persistent', [False, True])
def test_degenerate_frames_in_one_batch(LF, calib, persistent):
    """Frames without points, without masks, without boxes and with NaN / inf / behind-camera
    points next to a real frame: per-frame results equal the single-frame results, empty
    frames produce no rows, nothing reads out of bounds (both launch modes)."""
    cam, TVC, TVR = calib
    pts0, raw0 = load_frame(100), load_boxes(100)
    masks0 = standin(100, cam)[0]
    boxes0 = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw0, cam), TVC)
    c0 = np.array([b['corners_velo'] for b in boxes0], dtype=np.float64).reshape(-1, 8, 3)
    weird = np.array([[np.nan, 0, 0, 0], [np.inf, 1, 1, 0], [-np.inf, 2, 2, 0], [0, 0, 0, 0], [-10, 0, 0, 0],
                      [1e30, 1e30, 1e30, 0], [5, 0.1, -0.5, 0]], dtype=np.float32)
    pts = np.concatenate([pts0, weird, pts0[:50000]])
    n_pts = [len(pts0), 0, len(weird), 50000, 0]
    masks = np.concatenate([masks0, masks0[:2], masks0[:3]])
    n_masks = [len(masks0), 2, 3, 0, 0]
    corners = np.concatenate([c0, c0[:4], c0])
    n_boxes = [len(c0), 4, 0, len(c0), 0]
    batch = LF.FrameBatch(pts, n_pts, masks, n_masks, corners, n_boxes, cam, TVR)
    batch.run(3, 1, 50, 10, True, one_call=persistent)
    res = batch.results()
    want = P.process_frame(pts0, raw0, masks0, cam, TVC, TVR)
    M0 = len(masks0)
    assert res['n_valid'][0] == want['valid'].sum() and res['n_valid'][1] == 0 and res['n_valid'][4] == 0
    assert np.array_equal(res['total'][:M0], want['total'])
    assert np.array_equal(res['inside'][:batch.h_cnt_off[1]].reshape(want['inside'].shape), want['inside'])
    assert not res['total'][M0:M0 + 2].any()                       # frame 1: masks but no points
    wv = P.process_frame(weird, [], masks0[:3], cam, TVC, TVR)['valid']
    assert res['n_valid'][2] == wv.sum() == 1                      # only the last weird point is in view
    assert res['n_valid'][3] == P.process_frame(pts0[:50000], raw0, masks0[:0], cam, TVC, TVR)['valid'].sum()
    assert _norm_stats(batch.frame_stats(res, 0)) == _norm_stats(want['stats'])
    for f in (1, 2, 3, 4):
        assert batch.frame_stats(res, f) == [] or all(s['total_points'] == 0 for s in batch.frame_stats(res, f))
    # empty batch and oversize frames
    e = LF.FrameBatch(np.zeros((0, 4), np.float32), [], np.zeros((0, cam.height, cam.width), np.uint8), [],
                      np.zeros((0, 8, 3)), [], cam, TVR)
    e.run(one_call=persistent)
    assert e.results()['total'].size == 0
    # more than 64 masks per frame: the point pass runs once per group of 64 masks; per-point membership
    # words cannot hold them
    big = LF.FrameBatch(pts0[:1000], [1000], np.zeros((65, cam.height, cam.width), np.uint8), [65], c0[:1], [1], cam,
                        TVR, want_car_bits=True)
    with pytest.raises(ValueError):
        big.run(one_call=persistent)
    rng = np.random.default_rng(17)
    many = np.concatenate([masks0] * 14)[:130]                    # 130 masks: groups of 64 + 64 + 2
    many = many[rng.permutation(130)]
    b130 = LF.FrameBatch(np.concatenate([pts0, pts0[:30000]]), [len(pts0), 30000], np.concatenate([many, many[:70]]),
                         [130, 70], np.concatenate([c0, c0[:5]]), [len(c0), 5], cam, TVR)
    b130.run(3, 1, 50, 10, True, one_call=persistent)
    r130 = b130.results()
    w130 = P.process_frame(pts0, raw0, many, cam, TVC, TVR)
    assert np.array_equal(r130['total'][:130], w130['total'])
    assert np.array_equal(r130['inside'][:b130.h_cnt_off[1]].reshape(130, len(c0)), w130['inside'])
    assert np.array_equal(r130['n_in_box'][:len(c0)], w130['n_in_box']) and r130['n_valid'][0] == w130['valid'].sum()
    assert _norm_stats(b130.frame_stats(r130, 0)) == _norm_stats(w130['stats'])
    w70 = P.process_frame(pts0[:30000], raw0, many[:70], cam, TVC, TVR)
    assert np.array_equal(r130['total'][130:], w70['total']) and r130['n_valid'][1] == w70['valid'].sum()
    assert np.array_equal(r130['inside'][b130.h_cnt_off[1]:].reshape(70, 5), w70['inside'][:, :5])
    u, v, d = LF.project_points(pts0, TVR, cam)
    valid = LF.filter_valid(u, v, d, cam, 50)
    sets = LF.extract_car_points_by_mask(pts0[valid, :3], u[valid], v[valid], LF.erode_masks(many, 3, 1), cam)
    assert len(sets) == 130 and [len(x) for x in sets] == w130['total'].tolist()


def test_argument_errors_surface_as_python_exceptions(LF, calib):
    cam, TVC, TVR = calib
    m = np.zeros((1, cam.height, cam.width), np.uint8)
    with pytest.raises(ValueError):
        LF.erode_pack(m, se=np.ones((4, 4), np.uint8), erosion_kernel_size=4)      # even element
    with pytest.raises(ValueError):
        LF.erode_pack(m, se=np.zeros((3, 3), np.uint8))                            # empty element
    with pytest.raises(ValueError):
        LF.erode_pack(m.astype(np.int32))                                          # unsupported dtype
    with pytest.raises(ValueError):
        LF.project_points(np.zeros((5, 2), np.float32), TVR, cam)
    with pytest.raises(ValueError):
        LF.FrameBatch(np.zeros((5, 4), np.float32), [4], m, [1], np.zeros((0, 8, 3)), [0], cam, TVR)


def test_cuda_graph_replay_of_the_whole_step(LF, calib):
    """FrameBatch.capture: memset + erode + fuse + decision recorded once, replayed with one
    launch; results identical with direct launches (cfg1-sized batch, short segments)."""
    import torch
    cam, TVC, TVR = calib
    pts, raw = load_frame(100), load_boxes(100)
    masks = (standin(100, cam)[0] != 0).astype(np.uint8)
    boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw, cam), TVC)
    c = np.array([b['corners_velo'] for b in boxes], dtype=np.float64).reshape(-1, 8, 3)
    b = LF.FrameBatch(pts, [len(pts)], masks, [len(masks)], c, [len(c)], cam, TVR)
    b.run(3, 1, 50, 10, True)
    ref = b.results()
    g = b.capture(3, 1, 50, 10, True)
    b.counters.fill_(-7)                                   # the graph must re-zero everything itself
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    got = b.results()
    for k in ref:
        assert np.array_equal(ref[k], got[k]), k
    want = P.process_frame(pts, raw, masks.astype(np.float32), cam, TVC, TVR)
    assert np.array_equal(got['total'], want['total'])


# ------------------------------------------------------------------ packed masks in (SURVEY 8(d) secondary accounting)

@pytest.mark.parametrize('W,H', [(1408, 376), (128, 40), (1000, 123), (33, 17), (64, 8), (96, 300), (32, 20)])
def test_erosion_from_packed_words_matches_cv2(LF, W, H):
    """Masks packed on the sending side (lf_pack_masks_host) and eroded on the device from the words
    (lf_erode_pack mode 3): bulk-copy path (whole quads per row), generic path (any W), pad bits."""
    import torch
    from lidar_object_detection_b200 import io as IO
    rng = np.random.default_rng(W * 77 + H)
    masks = MP.random_blob_masks(rng, 6, H, W, n_blobs=4, max_r=max(4, min(H, W) // 3))
    masks[3] = 1
    masks[4, :, :] = 0
    masks[4, H // 4:, W // 5:] = 1
    masks[5] = 0
    words = IO.pack_masks(masks * rng.integers(1, 256, masks.shape, dtype=np.uint8))
    assert np.array_equal(words, IO.pack_masks_host(masks))
    raw = LF.PackedMasks(torch.from_numpy(words.view(np.int32)).cuda(), H, W, raw=True)
    for shape, cvshape in [(LF.MORPH_ELLIPSE, cv2.MORPH_ELLIPSE), (LF.MORPH_RECT, cv2.MORPH_RECT),
                           (LF.MORPH_CROSS, cv2.MORPH_CROSS)]:
        for k, it in ((1, 1), (3, 1), (3, 3), (5, 1), (5, 2), (7, 1), (9, 1), (15, 1)):
            se = cv2.getStructuringElement(cvshape, (k, k))
            want = np.stack([cv2.erode(m * 255, se, iterations=it) == 255 for m in masks])
            pm = LF.erode_pack(raw, k, it, shape)
            assert np.array_equal(pm.unpack().cpu().numpy().astype(bool), want), (W, H, shape, k, it)
            ref = LF.erode_pack(masks, k, it, shape)
            assert torch.equal(pm.words, ref.words) and torch.equal(pm.bbox, ref.bbox), (W, H, shape, k, it)
    # pack only (iterations 0): a copy with the pad bits cleared
    assert torch.equal(LF.erode_pack(raw, 0, 0).words, LF.erode_pack(masks, 0, 0).words)
    # a run-time element that is no row span (generic element path)
    se = np.array([[1, 0, 1], [0, 1, 0], [1, 0, 1]], np.uint8)
    want = np.stack([cv2.erode(m * 255, se, iterations=1) == 255 for m in masks])
    assert np.array_equal(LF.erode_pack(raw, 3, 1, se=se).unpack().cpu().numpy().astype(bool), want)
    # in place is refused (strips read their neighbours' rows)
    ctx = LF.get_ctx(0)
    from lidar_object_detection_b200 import _abi
    se3 = LF.structuring_element(LF.MORPH_ELLIPSE, 3)
    rc = ctx.lib.lf_erode_pack(ctx.h, _abi.ptr(raw.words), 3, len(raw), H, W, se3.ctypes.data, 3, 1,
                               _abi.ptr(raw.words), None, ctx.stream)
    assert rc == _abi.LF_E_ARG


def test_batch_from_packed_masks_equals_batch_from_uint8_masks(LF, calib):
    """The whole step with masks that crossed PCIe bit-packed: identical counters and decisions."""
    import torch
    from lidar_object_detection_b200 import io as IO
    cam, TVC, TVR = calib
    frames = [100, 250, 2449]
    pts, npts, masks, nm, corners, nb = [], [], [], [], [], []
    for fr in frames:
        p = load_frame(fr)
        m = (standin(fr, cam)[0] != 0).astype(np.uint8)
        boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(load_boxes(fr), cam), TVC)
        c = np.array([b['corners_velo'] for b in boxes], dtype=np.float64).reshape(-1, 8, 3)
        pts.append(p); npts.append(len(p)); masks.append(m); nm.append(len(m)); corners.append(c); nb.append(len(c))
    pts, masks, corners = np.concatenate(pts), np.concatenate(masks), np.concatenate(corners)
    a = LF.FrameBatch(pts, npts, masks, nm, corners, nb, cam, TVR)
    a.run(3, 1, 50, 10, True)
    ra = a.results()
    words = torch.from_numpy(IO.pack_masks(masks).view(np.int32)).cuda()
    for persistent in (False, True):
        b = LF.FrameBatch(pts, npts, LF.PackedMasks(words, cam.height, cam.width, raw=True), nm, corners, nb, cam, TVR)
        b.run(3, 1, 50, 10, True, one_call=persistent)
        rb = b.results()
        for k in ra:
            assert np.array_equal(ra[k], rb[k]), (k, persistent)
        assert torch.equal(a.packed.words, b.packed.words)
    g = b.capture(5, 1, 50, 10, True)                      # 5x5 from the packed words, as a graph
    g.replay()
    a.run(5, 1, 50, 10, True)
    ra, rb = a.results(), b.results()
    for k in ra:
        assert np.array_equal(ra[k], rb[k]), k
    # packed masks that are NOT raw are final: used as they are
    c = LF.FrameBatch(pts, npts, LF.PackedMasks(a.packed.words.clone(), cam.height, cam.width, a.mask_bbox.clone()),
                      nm, corners, nb, cam, TVR)
    c.run(3, 1, 50, 10, True)
    rc = c.results()
    for k in ra:
        assert np.array_equal(ra[k], rc[k]), k


def test_erosion_shape_sweep_with_canaries(LF):
    """Bounds check of our own (no sanitizer on this pool): every mask format over a sweep of image
    shapes -- one-word rows, one-quad rows, strips of 1..97 rows, ragged last words -- through the C ABI
    into output buffers fenced by canary words; results against cv2, canaries untouched."""
    import torch
    from lidar_object_detection_b200 import _abi, io as IO
    ctx = LF.get_ctx(0)
    rng = np.random.default_rng(99)
    CAN = 0x5A5A5A5A
    widths = [1, 2, 31, 32, 33, 63, 64, 65, 96, 127, 128, 129, 160, 256, 257, 512, 1408]
    heights = [1, 2, 3, 15, 16, 17, 33, 95, 96, 97, 193]
    n_cases = 0
    for W in widths:
        for H in heights:
            if W * H > 80000 and (W, H) != (1408, 97):
                continue
            M = 3
            masks = (rng.random((M, H, W)) < 0.85).astype(np.uint8)
            masks[1] = 1
            masks[2, : H // 2] = 0
            wpr = (W + 31) // 32
            words_in = torch.from_numpy(IO.pack_masks_host(masks).view(np.int32)).cuda()
            srcs = {0: torch.from_numpy(masks).cuda(), 1: torch.from_numpy(masks.astype(np.float32)).cuda(),
                    2: torch.from_numpy(masks.astype(np.float32) * 0.75).cuda(), 3: words_in}
            for k, it, shape, cvshape in ((3, 1, LF.MORPH_ELLIPSE, cv2.MORPH_ELLIPSE), (7, 2, LF.MORPH_ELLIPSE, cv2.MORPH_ELLIPSE),
                                          (5, 1, LF.MORPH_RECT, cv2.MORPH_RECT)):
                se = LF.structuring_element(shape, k)
                want = np.stack([cv2.erode(m * 255, cv2.getStructuringElement(cvshape, (k, k)), iterations=it) == 255
                                 for m in masks])
                want_words = IO.pack_masks_host(want)
                for mode, src in srcs.items():
                    n = M * H * wpr
                    buf = torch.full((n + 64,), CAN, dtype=torch.int32, device='cuda')
                    bb = torch.full((4 * M + 16,), 0, dtype=torch.int32, device='cuda')
                    bb[:8] = CAN
                    bb[8 + 4 * M:] = CAN
                    rc = ctx.lib.lf_erode_pack(ctx.h, _abi.ptr(src), mode, M, H, W, se.ctypes.data, k, it,
                                               buf.data_ptr() + 32 * 4, bb.data_ptr() + 8 * 4, ctx.stream)
                    assert rc == 0, (W, H, k, it, mode, rc)
                    got = buf.cpu().numpy()
                    assert (got[:32] == CAN).all() and (got[32 + n:] == CAN).all(), ('canary', W, H, k, it, mode)
                    assert np.array_equal(got[32:32 + n].view(np.uint32).reshape(M, H, wpr), want_words), (W, H, k, it, mode)
                    gb = bb.cpu().numpy()
                    assert (gb[:8] == CAN).all() and (gb[8 + 4 * M:] == CAN).all(), ('bbox canary', W, H, k, it, mode)
                    for c in range(M):                                  # bounding box of the eroded mask
                        ys, xs = np.nonzero(want[c])
                        rec = gb[8 + 4 * c:12 + 4 * c]
                        if len(ys) == 0:
                            assert (rec == 0).all(), (W, H, k, it, mode, c)
                        else:
                            assert tuple(rec) == ((1 << 20) - xs.min(), (1 << 20) - ys.min(), xs.max() + 1, ys.max() + 1), \
                                (W, H, k, it, mode, c)
                    n_cases += 1
    assert n_cases > 1500


def test_streamed_sweeps_give_the_rows_of_the_single_batch(LF, calib, golden, tmp_path):
    """Real-data loop (I/O row): .bin sweeps -> SweepStreamer (reader thread, pinned staging, copy
    stream) -> one FrameBatch per 7 sweeps, read back one batch behind.  Rows identical with the whole
    sequence in ONE batch and with the reference's CSV (same counts, same decisions)."""
    import csv
    from lidar_object_detection_b200 import driver
    cam, TVC, TVR = calib
    frames = [f for f in golden['_meta']['frames'] if load_boxes(f)]
    paths, inputs = [], []
    for fr in frames:
        p = str(tmp_path / ('%010d.bin' % fr))
        load_frame(fr).astype(np.float32).tofile(p)
        paths.append(p)
        m, b, c = standin(fr, cam)                                   # confidence-descending already
        boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(load_boxes(fr), cam), TVC)
        inputs.append(((m != 0).astype(np.uint8), np.array([bx['corners_velo'] for bx in boxes]).reshape(-1, 8, 3)))
    rows = driver.process_sweeps_streamed(paths, lambda i: inputs[i], cam, TVR, frames_per_batch=7, frame_ids=frames)
    # the reference's CSV for the same detections (written by the unmodified cvs_erosion.process_frames)
    want = []
    with open(os.path.join(GOLDEN, 'master_car_statistics_erosion3.csv')) as fh:
        for r in csv.DictReader(fh):
            want.append((int(r['frame']), int(r['car_id']), int(r['matched_bbox_id']), int(r['total_points']),
                         int(r['points_inside_bbox'])))
    assert sorted(rows) == sorted(want)
    assert [r[0] for r in rows] == sorted(r[0] for r in rows)        # batches come back in order


# ------------------------------------------------------------------ 8(e) exchange through peer memory

def test_peer_store_exchange_equals_the_table_all_reduce(LF, calib):
    """lf_publish_table (rows stored into every rank's table copy) against distributed.local_table /
    CounterExchange on a ragged batch; world size 1 here (the N = 2 comparison runs inside bench.py
    under torchrun on every multi-GPU launch)."""
    import torch
    import torch.distributed as dist
    from lidar_object_detection_b200 import distributed as D
    cam, TVC, TVR = calib
    frames = [100, 250, 1866, 2449]
    pts, masks, boxes, _ = _batch_inputs(frames, cam, TVC, LF)
    masks[1] = masks[1][:0]                                    # a frame without cars
    boxes[2] = []                                              # a frame without boxes
    nb = [len(b) for b in boxes]
    corners = np.array([b['corners_velo'] for fr in boxes for b in fr], np.float64).reshape(-1, 8, 3)
    batch = LF.FrameBatch(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks), [len(m) for m in masks],
                          corners, nb, cam, TVR, device=0)
    batch.run(3, 1, 50, 10, True)
    max_cars = max(len(m) for m in masks) + 2
    own = not dist.is_initialized()
    if own:
        dist.init_process_group('nccl', init_method='tcp://127.0.0.1:29577', rank=0, world_size=1,
                                device_id=torch.device('cuda', 0))
    try:
        assert D.PeerCounterExchange.available()
        ex = D.PeerCounterExchange(len(frames), max_cars, 0, 1, batch.ctx.device)
        ref = D.CounterExchange(len(frames), max_cars, 0, 1, batch.ctx.device)
        ref.bind(batch.h_mask_off, nb)
        for _ in range(3):                                     # both halves of the double buffer, twice
            table, slots = ex.exchange_batch(batch)
            want, wagg = ref.exchange_batch(batch)
            torch.cuda.synchronize()
            assert table.dtype == torch.int32 and torch.equal(table.to(torch.int64), want)
            assert torch.equal(ex.agg, wagg) and slots.shape == (1, 8)
        sm = int(batch.h_mask_off[-1])
        loc = D.local_table(batch.total[:sm], batch.best_cnt[:sm], batch.best_idx[:sm], nb, batch.h_mask_off, max_cars)
        assert torch.equal(loc, table)
        assert int(table[1].abs().sum()) == 0 and int(table[2, :, 3].sum()) == 0 and int(table[0, :, 0].sum()) > 0
        assert D.rows_from_table(table, frames) == D.rows_from_table(want, frames)
    finally:
        if own:
            dist.destroy_process_group()


# ------------------------------------------------------------------ A6 / A8 stand-alone kernels, every layout

@pytest.mark.parametrize('dtype,stride,B', [(np.float32, 4, 70), (np.float32, 3, 70), (np.float64, 3, 70), (np.float64, 5, 70),
                                            (np.float32, 4, 530)])
@pytest.mark.parametrize('oriented', [True, False])
def test_count_and_containment_kernels_every_layout(LF, dtype, stride, B, oriented):
    """lf_points_in_box / lf_count_in_boxes straight through the C ABI: float32 records of 4 (one 128-bit load
    per point), packed rows of 3, float64 rows, 70 boxes (more than one lane block, staged in shared memory) and
    530 (read from global memory), interleaved point sets, NaN / inf / huge coordinates -- against the oracle's per-box test
    (cvs_erosion.py:114-145, 175-195; V1:118-139)."""
    import torch
    from lidar_object_detection_b200 import _abi
    from lidar_object_detection_b200._abi import check, ptr
    ctx = LF.get_ctx(None)
    rng = np.random.default_rng(11 + stride + B)
    S, n = 5, 20011
    centres = rng.uniform(-20, 20, (B, 3))
    corners = np.zeros((B, 8, 3))
    signs = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], float)
    for b in range(B):
        yaw = rng.uniform(0, 2 * np.pi)
        half = rng.uniform(0.5, 3.0, 3) * signs
        c, s_ = np.cos(yaw), np.sin(yaw)
        corners[b] = np.stack([c * half[:, 0] - s_ * half[:, 1], s_ * half[:, 0] + c * half[:, 1], half[:, 2]], 1) + centres[b]
    # points: sweeps of neighbours around the boxes (so that whole warps see few boxes), some exactly on corners
    pick = rng.integers(0, B, n)
    xyz = (centres[pick] + rng.normal(0, 2.0, (n, 3))).astype(dtype)
    order = np.argsort(pick, kind='stable')
    xyz = xyz[order]
    if not oriented:
        xyz[:8] = corners[3].astype(dtype)                  # on the faces: exact for min / max comparisons (no arithmetic)
    xyz[100] = (np.nan, 0, 0); xyz[101] = (np.inf, 1, 1); xyz[102] = (1e30, -1e30, 0); xyz[103] = (0, np.nan, np.nan)
    rows = np.zeros((n, stride), dtype)
    rows[:, :3] = xyz
    rows[:, 3:] = 7.0
    set_id = rng.integers(0, S, n).astype(np.int32)
    d_rows = torch.from_numpy(rows).to(ctx.device)
    d_sid = torch.from_numpy(set_id).to(ctx.device)
    host = _abi.prepare_boxes(corners)
    d_boxes = torch.from_numpy(host.view(np.uint8).reshape(-1, _abi.BOX_DTYPE.itemsize).copy()).to(ctx.device)
    test = P.oriented_point_in_bbox if oriented else P.point_in_bbox
    with np.errstate(all='ignore'):
        want_in = np.stack([test(xyz.astype(np.float64), corners[b]) for b in range(B)], 1)      # [n, B]
    if oriented:                                            # counts must not hinge on the last bit of a dot product
        fin = np.isfinite(xyz).all(1) & (np.abs(xyz) < 1e6).all(1)
        for b in range(B):
            rel = xyz[fin].astype(np.float64) - corners[b, 0]
            for j in (1, 3, 4):
                e = corners[b, j] - corners[b, 0]
                t = rel @ e / (e @ e)
                assert np.min(np.minimum(np.abs(t), np.abs(t - 1))) > 1e-12
    want = np.zeros((S, B), np.int32)
    for s_ in range(S):
        want[s_] = want_in[set_id == s_].sum(0)
    inside = torch.zeros((S, B), dtype=torch.int32, device=ctx.device)
    check(ctx.lib.lf_count_in_boxes(ctx.h, ptr(d_rows), int(dtype == np.float64), n, stride, ptr(d_sid), ptr(d_boxes), B,
                                    int(oriented), ptr(inside), ctx.stream), ctx.h)
    assert np.array_equal(inside.cpu().numpy(), want)
    assert want.sum() > 1000
    for b in (0, 3, B - 1):
        flags = torch.empty(n, dtype=torch.uint8, device=ctx.device)
        check(ctx.lib.lf_points_in_box(ctx.h, ptr(d_rows), int(dtype == np.float64), n, stride, ptr(d_boxes[b]),
                                       int(oriented), ptr(flags), ctx.stream), ctx.h)
        assert np.array_equal(flags.cpu().numpy().astype(bool), want_in[:, b])


def test_mask_gather_more_than_eight_masks(LF):
    """lf_mask_gather with 1, 8, 9 and 64 masks (the words are requested eight at a time) against numpy indexing."""
    import torch
    from lidar_object_detection_b200 import io as IO
    from lidar_object_detection_b200._abi import check, ptr
    ctx = LF.get_ctx(None)
    rng = np.random.default_rng(3)
    H, W, n = 37, 100, 5000
    u = rng.integers(-3, W + 3, n).astype(np.int64)
    v = rng.integers(-3, H + 3, n).astype(np.int64)
    ok = (u >= 0) & (u < W) & (v >= 0) & (v < H)
    d_u, d_v = torch.from_numpy(u).to(ctx.device), torch.from_numpy(v).to(ctx.device)
    for m in (1, 8, 9, 64):
        masks = (rng.random((m, H, W)) < 0.4).astype(np.uint8)
        words = torch.from_numpy(IO.pack_masks(masks).view(np.int32)).to(ctx.device)
        bits = torch.empty(n, dtype=torch.int64, device=ctx.device)
        check(ctx.lib.lf_mask_gather(ctx.h, ptr(words), m, H, W, ptr(d_u), ptr(d_v), n, ptr(bits), ctx.stream), ctx.h)
        got = bits.cpu().numpy().view(np.uint64)
        want = np.zeros(n, np.uint64)
        for c in range(m):
            want[ok] |= masks[c, v[ok], u[ok]].astype(np.uint64) << np.uint64(c)
        assert np.array_equal(got, want), m


# ------------------------------------------------------------------ row map: only the rows with a set pixel are written

@pytest.mark.parametrize('W,H', [(1408, 376), (128, 40), (1000, 123), (64, 8)])
def test_erode_pack_rows_writes_the_map_and_only_needs_the_mapped_rows(LF, W, H):
    """lf_erode_pack_rows against lf_erode_pack for every mask format and kernel family (compile-time elements
    with the in-kernel map, the direct packed kernel and the generic paths with the map read off afterwards):
    the map is exactly "row has a set pixel", the mapped rows hold the dense result, the bounding boxes agree;
    the packed buffer is poisoned first, so a row that is skipped shows."""
    import torch
    from lidar_object_detection_b200 import _abi, io as IO
    from lidar_object_detection_b200._abi import check, ptr
    ctx = LF.get_ctx(None)
    rng = np.random.default_rng(W + 3 * H)
    masks = MP.random_blob_masks(rng, 6, H, W, n_blobs=2, max_r=max(3, min(H, W) // 5))
    masks[3] = 1
    masks[4] = 0
    masks[5, :, :] = 0
    masks[5, H // 2, W // 3] = 1                             # a single pixel: gone after any erosion
    wpr, rmw = (W + 31) // 32, (H + 31) // 32
    inputs = [('u8', torch.from_numpy(masks * 255).to(ctx.device), _abi.MODE_U8 if hasattr(_abi, 'MODE_U8') else 0),
              ('f32', torch.from_numpy(masks.astype(np.float32)).to(ctx.device), 1),
              ('packed', torch.from_numpy(IO.pack_masks_host(masks).view(np.int32)).to(ctx.device), 3)]
    for name, t, mode in inputs:
        for shape in (LF.MORPH_ELLIPSE, LF.MORPH_RECT):
            for k, it in ((3, 1), (5, 2), (1, 0)):
                se = LF.structuring_element(shape, k) if it else None
                dense = torch.zeros((6, H, wpr), dtype=torch.int32, device=ctx.device)
                bb_d = torch.zeros((6, 4), dtype=torch.int32, device=ctx.device)
                check(ctx.lib.lf_erode_pack(ctx.h, ptr(t), mode, 6, H, W, se.ctypes.data if it else None, k, it,
                                            ptr(dense), ptr(bb_d), ctx.stream), ctx.h)
                sparse = torch.full((6, H, wpr), 0x5A5A5A5A, dtype=torch.int32, device=ctx.device)
                bb_s = torch.zeros((6, 4), dtype=torch.int32, device=ctx.device)
                rm = torch.full((6, rmw), -1, dtype=torch.int32, device=ctx.device)    # the call zeroes it
                check(ctx.lib.lf_erode_pack_rows(ctx.h, ptr(t), mode, 6, H, W, se.ctypes.data if it else None, k, it,
                                                 ptr(sparse), ptr(bb_s), ptr(rm), ctx.stream), ctx.h)
                d, s_, m = dense.cpu().numpy(), sparse.cpu().numpy(), rm.cpu().numpy().view(np.uint32)
                rows = (d != 0).any(axis=2)                                           # [6, H]
                got_rows = ((m[:, np.arange(H) >> 5] >> (np.arange(H) & 31).astype(np.uint32)) & 1).astype(bool)
                key = (name, W, H, shape, k, it)
                assert np.array_equal(got_rows, rows), key
                assert np.array_equal(s_[rows], d[rows]), key
                assert torch.equal(bb_s, bb_d), key
                assert np.all((s_[~rows] == 0) | (s_[~rows] == 0x5A5A5A5A)), key      # skipped, or written as zeros
                # every row from the first to the last row with a set pixel is valid (skipped rows in between are
                # zero-filled); outside that range the kernels with the in-kernel map really skip
                ys = np.arange(H)
                for i in range(6):
                    if not rows[i].any():
                        continue
                    y0, y1 = ys[rows[i]].min(), ys[rows[i]].max()
                    assert np.array_equal(s_[i, y0:y1 + 1], d[i, y0:y1 + 1]), key
                    if name != 'packed' and W % 32 == 0 and k in (3, 5):
                        outside = np.ones(H, bool)
                        outside[y0:y1 + 1] = False
                        assert np.all(s_[i, outside] == 0x5A5A5A5A), key


def test_sparse_rows_batch_equals_the_dense_batch(LF, calib, golden):
    """FrameBatch.run(sparse_rows=True): same counters and decisions as the dense step, with the packed buffer
    poisoned beforehand (a word of an unmapped row that reached a counter would show) -- the real sample frames
    (up to 314 boxes), a synthetic batch, and frames with 130 masks (two passes of 64)."""
    import torch
    from lidar_object_detection_b200 import synthetic as S
    cam, TVC, TVR = calib
    frames = golden['_meta']['frames'][:6]
    pts, masks, boxes, raws = _batch_inputs(frames, cam, TVC, LF)
    corners = [np.array([b['corners_velo'] for b in bs], dtype=np.float64).reshape(-1, 8, 3) for bs in boxes]
    batches = [LF.FrameBatch(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks), [len(m) for m in masks],
                             np.concatenate(corners), [len(b) for b in boxes], cam, TVR)]
    batches.append(S.SyntheticBatch(12, cam, TVC, TVR, ncars=16, device=0).batch)
    rng = np.random.default_rng(8)
    many = MP.random_blob_masks(rng, 130, cam.height, cam.width, n_blobs=1, max_r=40)
    batches.append(LF.FrameBatch(pts[0], [len(pts[0])], many, [130], corners[0], [len(boxes[0])], cam, TVR))
    for b in batches:
        for args in ((3, 1, 50, 10, True), (5, 2, 30, 5, False), (0, 0, 50, 10, True)):
            b.run(*args)
            want = {k: v.copy() for k, v in b.results().items()}
            b.packed.words.fill_(-1)
            b.run(*args, sparse_rows=True)
            got = b.results()
            assert set(got) == set(want)
            for k in want:
                assert np.array_equal(got[k], want[k]), (k, args)
            assert sum(int(v.sum()) for k, v in want.items() if k == 'total') > 0
