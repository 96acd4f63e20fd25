"""lf_depth_maps_noalloc (static device scratch, no allocation; no counterpart in the reference)
against lf_depth_maps, through the C ABI.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_frame, standin

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def test_noalloc_depth_maps_equal_the_scratch_taking_entry_point():
    """Two calls in a row per policy (different frames, so stale keys of the first call would show),
    outputs bit-identical with lf_depth_maps (the depth-map tests of this repo compare for equality);
    an image beyond LF_NOALLOC_MAX_PIXELS is refused with LF_E_LIMIT."""
    import torch
    from lidar_object_detection_b200 import _abi, kitti360 as K, lidarfuse as LF
    from conftest import SAMPLE
    lib = _abi.load()
    cam, TVC, TVR = K.calibration_chain(SAMPLE)
    ctx = LF.get_ctx()
    ctx.set_camera(TVR, cam.K, cam.width, cam.height)
    H, W = cam.height, cam.width
    assert lib.lf_depth_maps_noalloc_max_pixels() == 1 << 20 >= H * W
    dev = ctx.device
    st = ctx.stream
    for policy in (_abi.POLICY_LAST, _abi.POLICY_NEAREST):
        for fr in (100, 947):                               # two calls in a row on the shared static array
            pts = torch.from_numpy(load_frame(fr)).to(dev)
            masks = standin(fr, cam)[0]
            pm = LF.erode_pack(masks, 0, 0, float_threshold='gt_half')
            m = len(pm)
            scratch = torch.empty(H * W, dtype=torch.int64, device=dev)
            want_full = torch.empty((H, W), dtype=torch.float64, device=dev)
            want_per = torch.empty((m, H, W), dtype=torch.float64, device=dev)
            got_full = torch.full((H, W), -1.0, dtype=torch.float64, device=dev)
            got_per = torch.full((m, H, W), -1.0, dtype=torch.float64, device=dev)
            rc = lib.lf_depth_maps(ctx.h, _ptr(pts), pts.shape[0], 30.0, policy, _ptr(pm.words), m, _ptr(scratch),
                                   _ptr(want_full), _ptr(want_per), st)
            assert rc == _abi.LF_OK
            rc = lib.lf_depth_maps_noalloc(ctx.h, _ptr(pts), pts.shape[0], 30.0, policy, _ptr(pm.words), m,
                                           _ptr(got_full), _ptr(got_per), st)
            assert rc == _abi.LF_OK
            torch.cuda.synchronize(dev)
            assert torch.equal(got_full, want_full) and torch.equal(got_per, want_per), (policy, fr)
            assert float(want_full.max()) > 0                 # the comparison is not between two empty maps
    # no points: the static array must still be cleared (maps all zero), not keep the last call's keys
    empty = torch.zeros((0, 4), dtype=torch.float32, device=dev)
    z = torch.full((H, W), -1.0, dtype=torch.float64, device=dev)
    assert lib.lf_depth_maps_noalloc(ctx.h, None, 0, 30.0, _abi.POLICY_LAST, None, 0, _ptr(z), None, st) == _abi.LF_OK
    torch.cuda.synchronize(dev)
    assert float(z.abs().max()) == 0.0 and empty.numel() == 0
    # beyond the fixed capacity: a clear error, nothing launched
    ctx.set_camera(TVR, cam.K, 2048, 1024)
    try:
        rc = lib.lf_depth_maps_noalloc(ctx.h, _ptr(pts), pts.shape[0], 30.0, _abi.POLICY_LAST, None, 0, None, None, st)
        assert rc == _abi.LF_E_LIMIT
        assert b'capacity' in lib.lf_strerror(rc)
    finally:
        ctx.set_camera(TVR, cam.K, cam.width, cam.height)
