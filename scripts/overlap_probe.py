#!/usr/bin/env python
"""Can the point pass of batch i run under the erode kernel of batch i + 1?  Two resident cfg3 batches on two
streams, the steps issued alternately, against the same steps on one stream (LIDARFUSE_LIB selects a library
whose point pass leaves CTA slots free: -DLF_FUSE_GRID_CAP=k).  One JSON line per mode."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from lidar_object_detection_b200 import kitti360 as K, synthetic as S
cam, TVC, TVR = K.calibration_chain(os.path.join(ROOT, 'tests', 'golden', 'kitti360_sample'))
sbs = [S.SyntheticBatch(1024, cam, TVC, TVR, ncars=16, device=0, first_frame=i * 1024) for i in range(2)]
for sb in sbs:
    sb.batch.run(3, 1, 50, 10, True)
torch.cuda.synchronize()
ref = [sb.batch.counters.clone() for sb in sbs]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]
K_STEPS = 40
SPARSE = os.environ.get("LF_PROBE_DENSE") is None
for mode in ('one stream', 'two streams'):
    for rep in range(2):
        torch.cuda.synchronize()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        for s in streams:
            s.wait_event(e0)
        for i in range(K_STEPS):
            st = streams[i % 2] if mode == 'two streams' else streams[0]
            with torch.cuda.stream(st):
                sbs[i % 2].batch.run(3, 1, 50, 10, True, sparse_rows=SPARSE)
        e1.record(streams[0]); e2.record(streams[1])
        torch.cuda.synchronize()
        ms = max(e0.elapsed_time(e1), e0.elapsed_time(e2)) / K_STEPS
    ok = all(torch.equal(sb.batch.counters, r) for sb, r in zip(sbs, ref))
    print(json.dumps({'lib': os.path.basename(os.environ.get('LIDARFUSE_LIB', 'default')), 'mode': mode, 'ms_per_step': ms, 'counters_equal': ok}))
