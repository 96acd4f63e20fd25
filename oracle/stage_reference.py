#!/usr/bin/env python
"""Stage the reference's scripts, byte for byte, into oracle/_ref/ so that the GPU box -- which has no
/root/reference -- can still time and run the reference's OWN functions (bench.py --impl reference, the
cpu_baseline leg, tests).  ORACLE / TEST INFRASTRUCTURE ONLY.

oracle/_ref/ is git-ignored (no reference source enters the history) and NOT gpurun-ignored (it travels
with the snapshot like the built .so files).  Called by __graft_entry__.build() wherever /root/reference
is mounted; a no-op elsewhere.  A manifest with the sha256 of every staged file is written beside them.
"""
import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get('LIDAR_REFERENCE_ROOT', '/root/reference')
DST = os.path.join(HERE, '_ref')
SCRIPTS = ['cvs_erosion.py', 'seg_with_pointcloud.py', 'V1_BBox_Pointwise_filtering.py',
           'V2_point_cloud_without_erosion.py', 'V3_point_cloud_with_erosion.py', 'V4_BBox_IoU_filtering.py',
           'V5_ProjectingBBoxes.py', 'secondtest.py']


def stage(verbose=False):
    src_dir = os.path.join(REF, 'Coding_testes')
    if not os.path.isfile(os.path.join(src_dir, 'cvs_erosion.py')):
        return None
    dst_dir = os.path.join(DST, 'Coding_testes')
    os.makedirs(dst_dir, exist_ok=True)
    manifest = {}
    for name in SCRIPTS:
        src = os.path.join(src_dir, name)
        if not os.path.isfile(src):
            continue
        shutil.copyfile(src, os.path.join(dst_dir, name))
        with open(src, 'rb') as f:
            manifest[name] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(DST, 'MANIFEST.json'), 'w') as f:
        json.dump({'source': src_dir, 'sha256': manifest}, f, indent=1, sort_keys=True)
    if verbose:
        print('staged %d reference scripts into %s' % (len(manifest), dst_dir))
    return DST


if __name__ == '__main__':
    stage(verbose=True)
