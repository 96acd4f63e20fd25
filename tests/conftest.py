import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')
SAMPLE = os.path.join(GOLDEN, 'kitti360_sample')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason='no CUDA device')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope='session')
def golden():
    with open(os.path.join(GOLDEN, 'golden.json')) as f:
        return json.load(f)


@pytest.fixture(scope='session')
def golden_arrays():
    return np.load(os.path.join(GOLDEN, 'golden_arrays.npz'))


@pytest.fixture(scope='session')
def sample_frames(golden):
    return list(golden['_meta']['frames'])


def load_frame(frame):
    z = np.load(os.path.join(SAMPLE, 'velodyne', '%010d.npz' % frame))
    return np.ascontiguousarray(z['xyzi'].T)


def load_boxes(frame):
    p = os.path.join(SAMPLE, 'bboxes_3D_cam0', 'BBoxes_%d.json' % frame)
    if not os.path.isfile(p):
        return []
    with open(p) as f:
        return json.load(f)


@pytest.fixture(scope='session')
def oracle_calib():
    from oracle import kitti360_restated as k360
    return k360.calibration_chain(SAMPLE)


_det_cache = {}


def standin(frame, camera):
    """Stand-in detections (confidence-descending order) for a sample frame."""
    from oracle import mask_provider as MP
    if frame not in _det_cache:
        _det_cache[frame] = MP.standin_detections(load_boxes(frame), camera)
    return _det_cache[frame]
