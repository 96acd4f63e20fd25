"""Where the end-to-end step spends its time: host pack loop vs DMA / kernels, for several slice and thread counts."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_object_detection_b200 import kitti360 as K, lidarfuse as LF, synthetic as S
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
cam, TVC, TVR = K.calibration_chain(os.path.join(ROOT, 'tests', 'golden', 'kitti360_sample'))
F, NC = 1024, 16
sb = S.SyntheticBatch(F, cam, TVC, TVR, ncars=NC, device=0)
b = sb.batch
h_pts = torch.empty(b.points.shape, dtype=b.points.dtype, pin_memory=True); h_pts.copy_(b.points)
h_masks = torch.empty(b.masks.shape, dtype=b.masks.dtype, pin_memory=True); h_masks.copy_(b.masks)
h_out = torch.empty(b.counters.shape, dtype=b.counters.dtype, pin_memory=True)
words = LF.erode_pack(b.masks, 0, 0, device=0).words
pb = LF.FrameBatch(b.points, [sb.n] * F, LF.PackedMasks(words, cam.height, cam.width, raw=True), [NC] * F, b.boxes, sb.n_boxes,
                   cam, TVR, device=0)
st = torch.cuda.current_stream()
torch.cuda.synchronize()
# DMA alone
for name, fn in (('points H2D alone', lambda: pb.points.copy_(h_pts, non_blocking=True)),):
    fn(); st.synchronize(); t0 = time.perf_counter(); fn(); st.synchronize(); print(name, round((time.perf_counter() - t0) * 1e3, 2), 'ms')
for slices in (8, 16, 32, 64):
    for threads in (8, 12, 16):
        for pts_first in (True, False):
            def step():
                t0 = time.perf_counter()
                if pts_first:
                    pb.upload(h_pts, h_masks, threads=threads, slices=slices)
                else:
                    pb.upload(None, h_masks, threads=threads, slices=slices)
                    pb.upload(h_pts, None)
                t1 = time.perf_counter()
                pb.run(3, 1, 50, 10, True); pb.download(h_out); st.synchronize()
                return t1 - t0, time.perf_counter() - t0
            step()
            r = [step() for _ in range(3)]
            print(json.dumps({'slices': slices, 'threads': threads, 'points_first': pts_first,
                              'pack_loop_ms': round(min(x[0] for x in r) * 1e3, 1), 'step_ms': round(min(x[1] for x in r) * 1e3, 1)}))
