"""cfg1 / cfg2 latency on the real KITTI360_sample frames (BASELINE configs[0], [1]):
ms per call of the whole per-frame body, device-resident inputs, direct launches vs one
CUDA-graph replay.  Not a throughput claim."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from conftest import SAMPLE, load_boxes, load_frame, standin
from lidar_object_detection_b200 import kitti360 as K, lidarfuse as LF
cam, TVC, TVR = K.calibration_chain(SAMPLE)
golden = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'golden.json')))
frames_all = golden['_meta']['frames']
for name, frames in (('cfg1 (frame 100)', [100]), ('cfg2 (20 sample frames)', frames_all)):
    pts, masks, boxes = [], [], []
    for fr in frames:
        raw = load_boxes(fr)
        pts.append(load_frame(fr)); masks.append((standin(fr, cam)[0] != 0).astype(np.uint8))
        boxes.append(LF.transform_bboxes_to_velodyne(LF.filter_visible_bboxes(raw, cam), TVC))
    corners = np.concatenate([np.array([b['corners_velo'] for b in bs], dtype=np.float64).reshape(-1, 8, 3) for bs in boxes])
    b = LF.FrameBatch(np.concatenate(pts), [len(p) for p in pts], np.concatenate(masks), [len(m) for m in masks],
                      corners, [len(x) for x in boxes], cam, TVR)
    g = b.capture(3, 1, 50, 10, True)
    ref = None
    out = {}
    for mode in ('direct', 'graph'):
        fn = (lambda: b.run(3, 1, 50, 10, True)) if mode == 'direct' else g.replay
        for _ in range(5): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        for _ in range(200): fn()
        e1.record(); torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / 200 * 1e3
        res = b.results()
        if ref is None: ref = res
        assert all(np.array_equal(ref[k], res[k]) for k in ref)
        out[mode] = {'device_ms_per_call': e0.elapsed_time(e1) / 200, 'wall_ms_per_call': wall}
    print(json.dumps({'config': name, 'frames': len(frames), 'points': int(sum(len(p) for p in pts)),
                      'masks': int(sum(len(m) for m in masks)), **out}))
