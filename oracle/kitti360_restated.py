"""Restatement of the kitti360Scripts pieces the reference imports.

ORACLE / TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference does ``from kitti360scripts.devkits.commons.loadCalibration import
loadCalibrationCameraToPose, loadCalibrationRigid`` and ``from
kitti360scripts.helpers.project import CameraPerspective``
(Coding_testes/cvs_erosion.py:9-10) but the submodule directory is empty
(.SUBMODULES.json: commit 40ee015de8277b2d907d32a30cdbd97072c7d7c3, status
"missing").  This file restates the published algorithm of those functions.
PARITY UNPINNED for this file: the pinned source is not available here, so the
restatement is anchored on the reference's call sites (cvs_erosion.py:64, 303,
308-309, 314, 345) and on the calibration files it parses.
"""
import os

import numpy as np


def loadCalibrationRigid(filename):
    """4x4 rigid transform from a one-line 3x4 text file.

    Call site: Coding_testes/cvs_erosion.py:308 (``TrCam0ToVelo``).
    """
    if not os.path.isfile(filename):
        raise RuntimeError('%s does not exist!' % filename)
    lastrow = np.array([0, 0, 0, 1]).reshape(1, 4)
    return np.concatenate((np.loadtxt(filename).reshape(3, 4), lastrow))


def _read_variable(lines, name, M, N):
    for line in lines:
        if line.startswith(name):
            vals = line.replace('%s:' % name, '').split()
            assert len(vals) == M * N
            return np.array([float(x) for x in vals]).reshape(M, N)
    return None


def loadCalibrationCameraToPose(filename):
    """dict image_00..image_03 -> 4x4 camera-to-pose transforms.

    Call site: Coding_testes/cvs_erosion.py:309.
    """
    if not os.path.isfile(filename):
        raise RuntimeError('%s does not exist!' % filename)
    with open(filename, 'r') as fid:
        lines = fid.readlines()
    Tr = {}
    lastrow = np.array([0, 0, 0, 1]).reshape(1, 4)
    for camera in ['image_00', 'image_01', 'image_02', 'image_03']:
        Tr[camera] = np.concatenate((_read_variable(lines, camera, 3, 4), lastrow))
    return Tr


class CameraPerspective:
    """Rectified perspective camera (cam 0 / 1) of KITTI-360.

    Attributes used by the reference: ``K`` (3x4), ``R_rect`` (4x4), ``width``,
    ``height`` and ``cam2image`` (Coding_testes/cvs_erosion.py:64, 68, 314,
    345, 359).  The pose file (``cam2world``) is never touched by the hot path
    and is loaded only when present.
    """

    def __init__(self, root_dir, seq='2013_05_28_drive_0009_sync', cam_id=0):
        assert cam_id == 0 or cam_id == 1
        self.cam_id = cam_id
        calib_dir = os.path.join(root_dir, 'calibration')
        self.intrinsic_file = os.path.join(calib_dir, 'perspective.txt')
        self.load_intrinsics(self.intrinsic_file)

    def load_intrinsics(self, intrinsic_file):
        intrinsic_loaded = False
        width = height = -1
        with open(intrinsic_file) as f:
            intrinsics = f.read().splitlines()
        for line in intrinsics:
            line = line.split(' ')
            if line[0] == 'P_rect_%02d:' % self.cam_id:
                K = np.reshape([float(x) for x in line[1:] if x != ''], [3, 4])
                intrinsic_loaded = True
            elif line[0] == 'R_rect_%02d:' % self.cam_id:
                R_rect = np.eye(4)
                R_rect[:3, :3] = np.array([float(x) for x in line[1:] if x != '']).reshape(3, 3)
            elif line[0] == 'S_rect_%02d:' % self.cam_id:
                width = int(float(line[1]))
                height = int(float(line[2]))
        assert intrinsic_loaded
        assert width > 0 and height > 0
        self.K = K
        self.width, self.height = width, height
        self.R_rect = R_rect

    def cam2image(self, points):
        """points [3,N] (or [B,3,N]) f64 -> (u int64, v int64, depth f64).

        ``depth == 0`` is replaced by -1e-6 IN PLACE of the projected array;
        u, v are divided by |depth| and rounded half-to-even (np.round).
        """
        ndim = points.ndim
        if ndim == 2:
            points = np.expand_dims(points, 0)
        points_proj = np.matmul(self.K[:3, :3].reshape([1, 3, 3]), points)
        depth = points_proj[:, 2, :]
        depth[depth == 0] = -1e-6
        with np.errstate(invalid='ignore', divide='ignore', over='ignore'):
            u = np.round(points_proj[:, 0, :] / np.abs(depth)).astype(int)
            v = np.round(points_proj[:, 1, :] / np.abs(depth)).astype(int)
        if ndim == 2:
            u = u[0]
            v = v[0]
            depth = depth[0]
        return u, v, depth


def calibration_chain(root_dir, seq=0, cam_id=0):
    """The calibration preamble every reference script repeats.

    Follows Coding_testes/cvs_erosion.py:300-314 line by line.  Returns
    ``(camera, TrVeloToCam, TrVeloToRect)``.
    """
    sequence = '2013_05_28_drive_%04d_sync' % seq
    camera = CameraPerspective(root_dir, sequence, cam_id)
    fileCameraToVelo = os.path.join(root_dir, 'calibration', 'calib_cam_to_velo.txt')
    fileCameraToPose = os.path.join(root_dir, 'calibration', 'calib_cam_to_pose.txt')
    TrCam0ToVelo = loadCalibrationRigid(fileCameraToVelo)
    TrCamToPose = loadCalibrationCameraToPose(fileCameraToPose)
    TrCamkToCam0 = np.linalg.inv(TrCamToPose['image_00']) @ TrCamToPose[f'image_{cam_id:02d}']
    TrCamToVelo = TrCam0ToVelo @ TrCamkToCam0
    TrVeloToCam = np.linalg.inv(TrCamToVelo)
    TrVeloToRect = np.matmul(camera.R_rect, TrVeloToCam)
    return camera, TrVeloToCam, TrVeloToRect
