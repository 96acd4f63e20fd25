"""Import the reference's scripts UNMODIFIED so their own functions can be run.

ORACLE / TEST INFRASTRUCTURE ONLY.  Reads the scripts from ``/root/reference``
where it exists (the build container) and otherwise from ``oracle/_ref/``,
the byte-for-byte copies ``oracle/stage_reference.py`` makes at ``build()`` time
(git-ignored, shipped to the GPU box with the snapshot) -- nothing at run time
on the GPU box reads ``/root/reference``.  Used by ``oracle/gen_golden*.py``
(fixture generation), by the CPU tests that re-check the restatement against the
live reference, and by ``bench.py --impl reference`` / the ``cpu_baseline`` leg
(``oracle/reference_frame.py``), which time the reference's own functions.

The scripts cannot be imported as they are: at import time they build
``YOLO('yolo11x-seg.pt')`` (weights absent, ultralytics not installed), import
``open3d`` / ``matplotlib`` (absent) and ``kitti360scripts`` (empty submodule)
-- Coding_testes/cvs_erosion.py:8-15.  We therefore plant stub modules in
``sys.modules``:

* ``ultralytics.YOLO``  -> an object whose ``predict`` returns detections from a
  pluggable provider (masks are inputs; see ``oracle/mask_provider.py``),
  shaped like the attributes the scripts read (``result.masks.data.cpu().numpy()``,
  ``result.boxes.xyxy`` / ``.conf``, ``result.orig_img``, ``len(result.boxes)``).
* ``kitti360scripts.*`` -> ``oracle/kitti360_restated.py``.
* ``open3d``, ``matplotlib``, ``PIL``, ``mpl_toolkits`` -> inert placeholders.

Not one line of the reference is copied: the files are executed from where
they lie.
"""
import importlib.util
import os
import sys
import types

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = os.environ.get('LIDAR_REFERENCE_ROOT', '/root/reference')
STAGED = False
if not os.path.isfile(os.path.join(REFERENCE_ROOT, 'Coding_testes', 'cvs_erosion.py')) and \
        os.path.isfile(os.path.join(_HERE, '_ref', 'Coding_testes', 'cvs_erosion.py')):
    # the GPU box: byte-for-byte copies staged by oracle/stage_reference.py at build() time
    REFERENCE_ROOT, STAGED = os.path.join(_HERE, '_ref'), True
SCRIPTS = os.path.join(REFERENCE_ROOT, 'Coding_testes')
SAMPLE = os.path.join(REFERENCE_ROOT, 'KITTI360_sample')
if not os.path.isdir(SAMPLE):
    # only the scripts are staged; the committed fixtures hold the sample's calibration and boxes
    SAMPLE = os.path.join(os.path.dirname(_HERE), 'tests', 'golden', 'kitti360_sample')

_provider = None


def available():
    return os.path.isfile(os.path.join(SCRIPTS, 'cvs_erosion.py'))


def set_detection_provider(fn):
    """fn(bgr_image) -> (masks f32 [M,H,W], boxes_xyxy f32 [M,4], conf f32 [M]) or None."""
    global _provider
    _provider = fn


class _T:
    """Just enough of a torch tensor: ``.cpu().numpy()``."""

    def __init__(self, a):
        self._a = np.asarray(a)

    def cpu(self):
        return self

    def numpy(self):
        return self._a

    def __len__(self):
        return len(self._a)


class _Boxes:
    def __init__(self, xyxy, conf):
        self.xyxy = _T(xyxy)
        self.conf = _T(conf)

    def __len__(self):
        return len(self.conf)


class _Masks:
    def __init__(self, data):
        self.data = _T(data)


class _Result:
    def __init__(self, img, det):
        self.orig_img = img
        if det is None or len(det[0]) == 0:
            self.masks = None
            self.boxes = _Boxes(np.zeros((0, 4), np.float32), np.zeros((0,), np.float32))
        else:
            self.masks = _Masks(det[0])
            self.boxes = _Boxes(det[1], det[2])


class _YOLO:
    def __init__(self, *a, **k):
        pass

    def predict(self, image, **kwargs):
        det = _provider(image) if _provider is not None else None
        return [_Result(image, det)]


class _Anything(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        m = _Anything(self.__name__ + '.' + name)
        setattr(self, name, m)
        return m

    def __call__(self, *a, **k):
        return self


def _install_stubs():
    from . import kitti360_restated as k360
    if 'ultralytics' not in sys.modules or not hasattr(sys.modules['ultralytics'], '_lf_stub'):
        ul = types.ModuleType('ultralytics')
        ul.YOLO = _YOLO
        ul._lf_stub = True
        sys.modules['ultralytics'] = ul
    pk = types.ModuleType('kitti360scripts')
    dk = types.ModuleType('kitti360scripts.devkits')
    cm = types.ModuleType('kitti360scripts.devkits.commons')
    lc = types.ModuleType('kitti360scripts.devkits.commons.loadCalibration')
    lc.loadCalibrationCameraToPose = k360.loadCalibrationCameraToPose
    lc.loadCalibrationRigid = k360.loadCalibrationRigid
    hp = types.ModuleType('kitti360scripts.helpers')
    pj = types.ModuleType('kitti360scripts.helpers.project')
    pj.CameraPerspective = k360.CameraPerspective
    pj.CameraFisheye = object
    for name, mod in [('kitti360scripts', pk), ('kitti360scripts.devkits', dk),
                      ('kitti360scripts.devkits.commons', cm),
                      ('kitti360scripts.devkits.commons.loadCalibration', lc),
                      ('kitti360scripts.helpers', hp), ('kitti360scripts.helpers.project', pj)]:
        sys.modules.setdefault(name, mod)
    for name in ['open3d', 'matplotlib', 'matplotlib.cm', 'matplotlib.pyplot', 'PIL', 'PIL.Image',
                 'mpl_toolkits', 'mpl_toolkits.mplot3d']:
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                sys.modules[name] = _Anything(name)


_cache = {}


def load(script):
    """Execute ``Coding_testes/<script>.py`` as a module and point its dataset
    globals at the sample that ships with the reference."""
    if script in _cache:
        return _cache[script]
    if not available():
        raise RuntimeError('reference not present at %s' % REFERENCE_ROOT)
    _install_stubs()
    path = os.path.join(SCRIPTS, script + '.py')
    spec = importlib.util.spec_from_file_location('_lfref_' + script, path)
    mod = importlib.util.module_from_spec(spec)
    saved = os.environ.get('KITTI360_DATASET')
    spec.loader.exec_module(mod)
    os.environ['KITTI360_DATASET'] = SAMPLE
    if hasattr(mod, 'bbox_dir'):
        mod.bbox_dir = os.path.join(SAMPLE, 'bboxes_3D_cam0')
    mod._lf_saved_env = saved
    _cache[script] = mod
    return mod
