"""Host-side loaders with the names the reference scripts use.

These stay in numpy float64 on purpose: the sixteen doubles
of TrVeloToRect and the box corners must be bit-identical with what the
reference's numpy calls produce, and they are O(1) / O(#boxes) work per frame.

``CameraPerspective`` / ``loadCalibration*`` follow the interface of the
kitti360scripts package the reference imports (cvs_erosion.py:9-10) -- the
submodule is empty in the reference tree, so these are written against its call
sites (cvs_erosion.py:64, 303, 308-309, 314, 345).
"""
import glob
import json
import os

import numpy as np

_BOTTOM = np.array([[0.0, 0.0, 0.0, 1.0]])


def loadCalibrationRigid(filename):
    """One line of 12 numbers -> 4x4 f64 (cvs_erosion.py:308)."""
    if not os.path.isfile(filename):
        raise RuntimeError('%s does not exist!' % filename)
    return np.vstack([np.loadtxt(filename).reshape(3, 4), _BOTTOM])


def loadCalibrationCameraToPose(filename):
    """``image_0k: 12 numbers`` lines -> {name: 4x4 f64} (cvs_erosion.py:309)."""
    if not os.path.isfile(filename):
        raise RuntimeError('%s does not exist!' % filename)
    out = {}
    with open(filename) as fh:
        for line in fh:
            if ':' not in line:
                continue
            name, rest = line.split(':', 1)
            name = name.strip()
            if name.startswith('image_'):
                vals = np.array([float(t) for t in rest.split()])
                out[name] = np.vstack([vals.reshape(3, 4), _BOTTOM])
    for cam in ('image_00', 'image_01', 'image_02', 'image_03'):
        if cam not in out:
            raise RuntimeError('%s has no entry for %s' % (filename, cam))
    return out


class CameraPerspective:
    """Rectified pinhole camera 0/1: attributes ``K`` (3x4), ``R_rect`` (4x4),
    ``width``, ``height`` and the host ``cam2image`` used for the handful of box
    corners (cvs_erosion.py:64).  Whole point clouds go through
    ``lidarfuse.project_points`` (CUDA) instead."""

    def __init__(self, root_dir, seq='2013_05_28_drive_0009_sync', cam_id=0):
        if cam_id not in (0, 1):
            raise AssertionError('perspective cameras are 0 and 1')
        self.cam_id = cam_id
        self.intrinsic_file = os.path.join(root_dir, 'calibration', 'perspective.txt')
        tag = '%02d:' % cam_id
        K = R = size = None
        with open(self.intrinsic_file) as fh:
            for raw in fh.read().splitlines():
                tok = [t for t in raw.split(' ') if t != '']
                if not tok:
                    continue
                if tok[0] == 'P_rect_' + tag:
                    K = np.array([float(t) for t in tok[1:]]).reshape(3, 4)
                elif tok[0] == 'R_rect_' + tag:
                    R = np.eye(4)
                    R[:3, :3] = np.array([float(t) for t in tok[1:]]).reshape(3, 3)
                elif tok[0] == 'S_rect_' + tag:
                    size = (int(float(tok[1])), int(float(tok[2])))
        if K is None or R is None or size is None:
            raise AssertionError('incomplete intrinsics in %s' % self.intrinsic_file)
        self.K, self.R_rect = K, R
        self.width, self.height = size

    def cam2image(self, points):
        """[3,N] (or [B,3,N]) f64 -> (u int64, v int64, depth f64); depth==0 -> -1e-6."""
        single = points.ndim == 2
        pts = points[None] if single else points
        proj = np.matmul(self.K[:3, :3].reshape(1, 3, 3), pts)
        depth = proj[:, 2, :]
        depth[depth == 0] = -1e-6
        with np.errstate(all='ignore'):
            u = np.round(proj[:, 0, :] / np.abs(depth)).astype(int)
            v = np.round(proj[:, 1, :] / np.abs(depth)).astype(int)
        return (u[0], v[0], depth[0]) if single else (u, v, depth)


def calibration_chain(root_dir, seq=0, cam_id=0):
    """(camera, TrVeloToCam, TrVeloToRect), the preamble of every reference script
    (cvs_erosion.py:300-314)."""
    sequence = '2013_05_28_drive_%04d_sync' % seq
    camera = CameraPerspective(root_dir, sequence, cam_id)
    cam0_to_velo = loadCalibrationRigid(os.path.join(root_dir, 'calibration', 'calib_cam_to_velo.txt'))
    cam_to_pose = loadCalibrationCameraToPose(os.path.join(root_dir, 'calibration', 'calib_cam_to_pose.txt'))
    camk_to_cam0 = np.linalg.inv(cam_to_pose['image_00']) @ cam_to_pose['image_%02d' % cam_id]
    cam_to_velo = cam0_to_velo @ camk_to_cam0
    velo_to_cam = np.linalg.inv(cam_to_velo)
    velo_to_rect = np.matmul(camera.R_rect, velo_to_cam)
    return camera, velo_to_cam, velo_to_rect


class Kitti360Viewer3DRaw:
    """Raw velodyne sweeps (cvs_erosion.py:18-28)."""

    def __init__(self, seq=0, mode='velodyne', root=None):
        root = root or os.environ.get('KITTI360_DATASET', '')
        if mode != 'velodyne':
            raise ValueError('only the velodyne mode is on the fusion path')
        self.sensor_dir = 'velodyne_points'
        self.raw3DPcdPath = os.path.join(root, 'data_3d_raw', '2013_05_28_drive_%04d_sync' % seq, self.sensor_dir,
                                         'data')

    def frames(self):
        files = sorted(glob.glob(os.path.join(self.raw3DPcdPath, '*.bin')))
        return [int(os.path.basename(f).split('.')[0]) for f in files]

    def loadVelodyneData(self, frame=0):
        path = os.path.join(self.raw3DPcdPath, '%010d.bin' % frame)
        if not os.path.isfile(path):
            raise RuntimeError('%s does not exist!' % path)
        return np.fromfile(path, dtype=np.float32).reshape(-1, 4)


def load_bounding_boxes(json_path):
    """List of {index, corners_cam0}; a missing file yields [] (cvs_erosion.py:31-38)."""
    if not os.path.exists(json_path):
        return []
    with open(json_path) as fh:
        return json.load(fh)
