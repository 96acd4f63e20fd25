"""CPU-side checks of the boundary: the library builds, loads and exports every
symbol include/lidarfuse.h declares; host-only entry points behave; the product
package never imports the oracle."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, SAMPLE


@pytest.fixture(scope='module')
def lib():
    from lidar_object_detection_b200 import build, _abi
    build.build()
    return _abi.load()


def test_every_header_symbol_is_exported(lib):
    from lidar_object_detection_b200 import _abi
    hdr = open(os.path.join(ROOT, 'include', 'lidarfuse.h')).read()
    names = set(re.findall(r'^\s*(?:int|int64_t|const char\*)\s+(lf_\w+)\s*\(', hdr, flags=re.M))
    assert len(names) >= 18
    assert names == set(_abi.SIGNATURES), names ^ set(_abi.SIGNATURES)
    for n in names:
        assert getattr(lib, n) is not None
    assert lib.lf_version() == 200
    assert lib.lf_strerror(0) == b'ok' and b'argument' in lib.lf_strerror(-1)


def test_box_struct_layout_and_host_preparation(lib):
    from lidar_object_detection_b200 import _abi
    assert _abi.BOX_DTYPE.itemsize == 192
    rng = np.random.default_rng(3)
    # a rotated 4 x 2 x 1.5 box
    yaw = 0.7
    R = np.array([[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]])
    signs = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]])
    corners = (signs * np.array([2.0, 1.0, 0.75])) @ R.T + np.array([10.0, 3.0, -1.0])
    b = _abi.prepare_boxes(corners[None])[0]
    assert np.array_equal(b['c0'], corners[0])
    assert np.array_equal(b['e1'], corners[1] - corners[0])
    assert np.array_equal(b['e2'], corners[3] - corners[0])
    assert np.array_equal(b['e3'], corners[4] - corners[0])
    assert np.allclose(b['ee'], [16.0, 4.0, 2.25], rtol=1e-14)
    assert np.array_equal(b['amin'], corners.min(0)) and np.array_equal(b['amax'], corners.max(0))
    assert np.all(b['lo'] < corners.min(0)) and np.all(b['hi'] > corners.max(0))
    assert np.all(b['lo'] > corners.min(0) - 0.01) and np.all(b['hi'] < corners.max(0) + 0.01)
    # the early-out bounds contain every point the exact test accepts, also for skewed edge sets
    from oracle import pipeline as P
    for _ in range(50):
        c = rng.normal(size=(8, 3)) * 3
        bb = _abi.prepare_boxes(c[None])[0]
        pts = (rng.normal(size=(4000, 3)) * 6).astype(np.float32)
        inside = P.oriented_point_in_bbox(pts, c)
        near = np.all((pts >= bb['lo']) & (pts <= bb['hi']), axis=1)
        assert not np.any(inside & ~near)
    deg = _abi.prepare_boxes(np.zeros((1, 8, 3)))[0]                    # degenerate: bounds disabled
    assert np.all(np.isinf(deg['lo'])) and np.all(np.isinf(deg['hi']))
    assert len(_abi.prepare_boxes(np.zeros((0, 8, 3)))) == 0


def test_no_cpu_fallback_without_a_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip('a device is present')
    from lidar_object_detection_b200 import _abi, lidarfuse, kitti360
    cam, TVC, TVR = kitti360.calibration_chain(SAMPLE)
    with pytest.raises(_abi.LidarFuseError):
        lidarfuse.project_points(np.zeros((4, 4), np.float32), TVR, cam)
    h = ctypes.c_void_p()
    assert lib.lf_create(ctypes.byref(h), 0) == _abi.LF_E_CUDA


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'lidar_object_detection_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), f
                assert 'oracle.' not in src.replace('oracle.py', ''), f


def test_host_loaders_match_the_oracle_restatement(oracle_calib):
    from lidar_object_detection_b200 import kitti360 as K, lidarfuse as LF
    from oracle import pipeline as P
    from conftest import load_boxes
    cam, TVC, TVR = K.calibration_chain(SAMPLE)
    ocam, oTVC, oTVR = oracle_calib
    assert np.array_equal(TVR, oTVR) and np.array_equal(TVC, oTVC) and np.array_equal(cam.K, ocam.K)
    assert (cam.width, cam.height) == (1408, 376)
    raw = load_boxes(2449)
    a = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(load_boxes(2449), cam), TVC)
    b = P.transform_bboxes_to_velodyne(P.filter_visible_bboxes(raw, ocam), oTVC)
    assert len(a) == len(b) and all(x['corners_velo'] == y['corners_velo'] for x, y in zip(a, b))
    pts = np.random.default_rng(0).normal(size=(3, 50)) * 20
    for x, y in zip(cam.cam2image(pts.copy()), ocam.cam2image(pts.copy())):
        assert np.array_equal(x, y)


def test_2d_matching_matches_reference_vectors(golden, golden_arrays):
    from lidar_object_detection_b200 import kitti360 as K, lidarfuse as LF, matching2d as M2
    from conftest import load_boxes, standin
    cam, TVC, TVR = K.calibration_chain(SAMPLE)
    for a, b, want in golden['iou2d_kat']:
        assert M2.calculate_iou_2d(a, b) == want
    for fr in [100, 2098, 947, 2449]:
        masks, boxes2d, conf = standin(fr, cam)
        boxes = LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(load_boxes(fr), cam), TVC)
        colors = LF.generate_colors(len(masks))
        m4 = M2.match_detections_to_bboxes(boxes2d, boxes, colors, cam)
        assert np.array_equal(np.array([p[0] for p in m4]).reshape(-1, 8, 3), golden_arrays['v4_matched_corners_%d' % fr])
        assert np.array_equal(np.array([p[1] for p in m4], np.float64).reshape(-1, 3),
                              golden_arrays['v4_matched_colors_%d' % fr])
        m5 = M2.improved_match_detections_to_bboxes(boxes2d, boxes, colors, cam)
        assert np.array_equal(np.array([p[0] for p in m5]).reshape(-1, 8, 3), golden_arrays['v5_matched_corners_%d' % fr])
        assert np.array_equal(np.array([np.asarray(p[1], np.float64) for p in m5]).reshape(-1, 3),
                              golden_arrays['v5_matched_colors_%d' % fr])


def test_csv_writer_bytes_without_pandas(golden, tmp_path):
    """The hand-written CSV writer reproduces the pandas bytes of the reference file."""
    from lidar_object_detection_b200 import lidarfuse as LF
    from conftest import GOLDEN
    want = open(os.path.join(GOLDEN, 'master_car_statistics_erosion3.csv')).read()
    path = str(tmp_path / 'm.csv')
    ts = golden['_meta']['frozen_timestamp']
    for fr in golden['_meta']['frames']:
        g = golden.get(str(fr))
        if not g or not g.get('stats_erosion3'):
            continue
        stats = [dict(s, points_inside_bbox=np.int64(s['points_inside_bbox'])) for s in g['stats_erosion3']]
        LF.append_to_master_csv(stats, fr, path, ts)
    assert open(path).read() == want


def test_argument_checks_need_no_device(lib):
    """Null contexts / bad sizes are rejected before any CUDA call (status codes, no abort)."""
    from lidar_object_detection_b200 import _abi
    assert lib.lf_process_batch(None, *([None] * 3), 0, 0, None, None, 3, 1, *([None] * 5), 1, 1, 1, 1, 50.0, 1, 10,
                                *([None] * 9)) == _abi.LF_E_ARG
    assert lib.lf_resize_masks(None, None, 1, 4, 4, None, 8, 8, None) == _abi.LF_E_ARG
    assert lib.lf_project_boxes(None, None, 1, 0.1, 100.0, 4, None, None, None, None, None) == _abi.LF_E_ARG
    assert lib.lf_point_colors(None, None, 1, None, 1, None, None, None, 0, None) == _abi.LF_E_ARG
    assert lib.lf_format_csv_rows(None, None, None, None, None, -1, b't', 0, None, 0) == _abi.LF_E_ARG
    assert lib.lf_format_csv_rows(None, None, None, None, None, 0, b't', 1, None, 0) == \
        len('frame,car_id,matched_bbox_id,total_points,points_inside_bbox,points_outside_bbox,'
            'inside_percentage,outside_percentage,is_matched,timestamp\n')
    assert b'argument' in lib.lf_strerror(_abi.LF_E_ARG).lower() or lib.lf_strerror(_abi.LF_E_ARG)


def test_host_mask_packer_matches_numpy():
    """lf_pack_masks_host (plain C++, no device): every binarisation rule, ragged widths, threads,
    and every instruction-set path the host offers (scalar, AVX2, AVX-512BW)."""
    from lidar_object_detection_b200 import _abi, io as IO
    lib = _abi.load()
    top = lib.lf_host_simd(2)
    try:
        for cap in range(top, -1, -1):
            assert lib.lf_host_simd(cap) == cap
            _check_host_packer(IO)
    finally:
        lib.lf_host_simd(2)


def _check_host_packer(IO):
    rng = np.random.default_rng(5)
    for W in (1408, 1000, 37, 64, 96, 160, 31):
        m = (rng.random((7, 29, W)) < 0.3).astype(np.uint8) * rng.integers(1, 256, (7, 29, W), dtype=np.uint8)
        want = IO.pack_masks_host(m)
        for threads in (1, 3, 0):
            assert np.array_equal(IO.pack_masks(m, threads=threads), want), (W, threads)
        assert np.array_equal(IO.pack_masks(m.astype(bool)), want)
        f = (rng.random((7, 29, W)) * 1.3).astype(np.float32)
        f[0, 0, :6] = [np.nan, -0.0, 1.0, 0.5, np.inf, -np.inf]
        with np.errstate(invalid='ignore'):
            assert np.array_equal(IO.pack_masks(f), IO.pack_masks_host(f >= 1.0)), W
            assert np.array_equal(IO.pack_masks(f, 'gt_half'), IO.pack_masks_host(f > 0.5)), W
    out = np.full((7, 29, 1), 0xFFFFFFFF, np.uint32)
    assert IO.pack_masks(m, out=out) is out and np.array_equal(out, want)      # pad bits come out zero
    assert IO.pack_masks(np.zeros((0, 4, 4), np.uint8)).shape == (0, 4, 1)
    with pytest.raises(ValueError):
        IO.pack_masks(np.zeros((2, 4, 4), np.int16))
    with pytest.raises(ValueError):
        IO.pack_masks(m, out=np.zeros((7, 29, 2), np.uint32))
