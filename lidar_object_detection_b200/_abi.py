"""ctypes binding of include/lidarfuse.h.

There is no CPU fallback: if the CUDA library cannot be loaded, or no CUDA
device is present when a context is requested, this module raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('LIDARFUSE_LIB') or os.path.join(HERE, '_lib', 'liblidarfuse.so')   # override: kernel experiments

LF_OK, LF_E_ARG, LF_E_CUDA, LF_E_UNSUPPORTED, LF_E_NOCAMERA, LF_E_LIMIT = 0, -1, -2, -3, -4, -5
POLICY_NEAREST, POLICY_LAST = 0, 1
MODE_U8, MODE_F32_GE1, MODE_F32_GT_HALF, MODE_PACKED = 0, 1, 2, 3
MAX_CARS = 64
FUSE_ORIENTED, FUSE_EXACT = 1, 2          # flag word `use_oriented` of lf_fuse_batch / lf_process_batch

# numpy mirror of struct lf_box (192 bytes)
BOX_DTYPE = np.dtype([('c0', '<f8', 3), ('e1', '<f8', 3), ('e2', '<f8', 3), ('e3', '<f8', 3),
                      ('ee', '<f8', 3), ('amin', '<f8', 3), ('amax', '<f8', 3),
                      ('lo', '<f4', 3), ('hi', '<f4', 3)])
assert BOX_DTYPE.itemsize == 192
# numpy mirror of struct lf_box2d (72 bytes)
BOX2D_DTYPE = np.dtype([('x0', '<i8'), ('y0', '<i8'), ('x1', '<i8'), ('y1', '<i8'), ('avg_depth', '<f8'),
                        ('range_avg_depth', '<f8'), ('n_front', '<i4'), ('n_visible', '<i4'), ('n_in_range', '<i4'),
                        ('corners_in_view', '<i4'), ('view_code', '<i4'), ('pad_', '<i4')])
assert BOX2D_DTYPE.itemsize == 72
VIEW_VALID, VIEW_ALL_BEHIND, VIEW_NO_INTERSECTION, VIEW_TOO_SMALL = 0, 1, 2, 3

_vp, _i32, _i64, _f64, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_uint64

SIGNATURES = {
    'lf_version': (C.c_int, []),
    'lf_strerror': (C.c_char_p, [_i32]),
    'lf_create': (C.c_int, [C.POINTER(_vp), _i32]),
    'lf_destroy': (C.c_int, [_vp]),
    'lf_last_cuda_error': (C.c_int, [_vp]),
    'lf_set_camera': (C.c_int, [_vp, _vp, _vp, _i32, _i32]),
    'lf_project': (C.c_int, [_vp, _vp, _i64, _f64, _vp, _vp, _vp, _vp, _vp]),
    'lf_erode_pack': (C.c_int, [_vp, _vp, _i32, _i64, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp]),
    'lf_erode_pack_rows': (C.c_int, [_vp, _vp, _i32, _i64, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    'lf_unpack_masks': (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _vp]),
    'lf_pack_masks_host': (C.c_int, [_vp, _i32, _i64, _i32, _i32, _vp, _i32]),
    'lf_host_simd': (C.c_int, [_i32]),
    'lf_publish_table': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i64, _vp, _i32, _i32, _i64, _i64,
                                   _vp]),
    'lf_prepare_boxes': (C.c_int, [_vp, _i64, _vp]),
    'lf_fuse_batch': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _i32, _f64, _i32,
                                _vp, _vp, _vp, _vp, _vp, _vp]),
    'lf_fuse_batch_rows': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _i32, _f64, _i32,
                                     _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    'lf_process_batch': (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp,
                                   _i32, _i64, _i32, _i32, _f64, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                   _vp, _vp]),
    'lf_mask_gather': (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _i64, _vp, _vp]),
    'lf_filter_valid': (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _f64, _vp, _vp]),
    'lf_car_scratch_elems': (C.c_int64, [_i64, _i32]),
    'lf_car_counts': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _vp, _vp]),
    'lf_car_compact': (C.c_int, [_vp, _vp, _i64, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    'lf_points_in_box': (C.c_int, [_vp, _vp, _i32, _i64, _i32, _vp, _i32, _vp, _vp]),
    'lf_count_in_boxes': (C.c_int, [_vp, _vp, _i32, _i64, _i32, _vp, _vp, _i32, _i32, _vp, _vp]),
    'lf_car_stats': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    'lf_zero_counters': (C.c_int, [_vp, _vp, _i64, _vp]),
    'lf_depth_maps': (C.c_int, [_vp, _vp, _i64, _f64, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    'lf_depth_maps_noalloc_max_pixels': (C.c_int64, []),
    'lf_depth_maps_noalloc': (C.c_int, [_vp, _vp, _i64, _f64, _i32, _vp, _i32, _vp, _vp, _vp]),
    'lf_depth_maps_batch': (C.c_int, [_vp, _vp, _vp, _i32, _i64, _f64, _i32, _vp, _vp, _vp, _vp, _vp, _vp]),
    'lf_resize_masks': (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _i32, _i32, _vp]),
    'lf_resize_pack': (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _vp, _i32, _i32, _vp]),
    'lf_project_boxes': (C.c_int, [_vp, _vp, _i64, _f64, _f64, _i32, _vp, _vp, _vp, _vp, _vp]),
    'lf_match_matrices': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _f64, _f64, _f64, _vp, _vp, _vp]),
    'lf_point_colors': (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _vp, _vp, _i32, _vp]),
    'lf_depth_overlay': (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    'lf_depth_overlay_batch': (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    'lf_format_csv_rows': (C.c_int64, [_vp, _vp, _vp, _vp, _vp, _i64, C.c_char_p, _i32, _vp, _i64]),
    'lf_synth_points': (C.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _u64, _vp, _vp]),
    'lf_synth_masks': (C.c_int, [_vp, _vp, _i32, _i32, _vp, _f64, _vp, _vp]),
}

_lib = None


class LidarFuseError(RuntimeError):
    pass


def load():
    """dlopen the in-tree library and bind every symbol of the header."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise LidarFuseError('%s is missing: run `python -m lidar_object_detection_b200.build` '
                             '(there is no CPU fallback)' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code, ctx=None):
    if code == LF_OK:
        return
    lib = load()
    msg = lib.lf_strerror(code).decode()
    if code == LF_E_CUDA and ctx is not None:
        msg += ' [cudaError_t %d]' % lib.lf_last_cuda_error(ctx)
    if code in (LF_E_ARG, LF_E_UNSUPPORTED, LF_E_LIMIT):
        raise ValueError('lidarfuse: ' + msg)
    raise LidarFuseError('lidarfuse: ' + msg)


def ptr(t):
    """Device / host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


def prepare_boxes(corners_velo):
    """corners [n,8,3] f64 (host) -> structured array of lf_box via lf_prepare_boxes."""
    corners = np.ascontiguousarray(np.asarray(corners_velo, dtype=np.float64).reshape(-1, 8, 3))
    out = np.zeros(len(corners), dtype=BOX_DTYPE)
    if len(corners):
        check(load().lf_prepare_boxes(corners.ctypes.data, len(corners), out.ctypes.data))
    return out
