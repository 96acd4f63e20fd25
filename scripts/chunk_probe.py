#!/usr/bin/env python
"""Experiment: can the packed-mask round trip through HBM be stopped?  The step is cut
into chunks of C frames -- erode(chunk i) then point pass(chunk i) on ONE stream -- and the packed words of a
chunk live in a ring of two C-frame buffers that every chunk reuses, so that (if the L2 keeps them) the
erode kernel's 1.07 GB of packed output is overwritten in the L2 before it is evicted and the point pass
reads it from there.  Counters must equal the plain step's.  Prints one JSON line per chunk size.
    python scripts/chunk_probe.py [--frames 1024] [--chunks 16,32,64,128]"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=1024)
    ap.add_argument('--chunks', default='16,32,64,128,256')
    ap.add_argument('--steps', type=int, default=20)
    args = ap.parse_args()
    import numpy as np
    import torch
    from lidar_object_detection_b200 import _abi, kitti360 as K, lidarfuse as LF, synthetic as S
    cam, TVC, TVR = K.calibration_chain(os.path.join(ROOT, 'tests', 'golden', 'kitti360_sample'))
    sb = S.SyntheticBatch(args.frames, cam, TVC, TVR, ncars=16)
    b = sb.batch
    ctx, lib = b.ctx, b.ctx.lib
    stream = torch.cuda.current_stream()
    H, W = cam.height, cam.width
    plane = H * ((W + 31) // 32)
    se = LF.structuring_element(LF.MORPH_ELLIPSE, 3)

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps

    gp = b.capture(3, 1, 50, 10, True)
    plain = timed(gp.replay)
    want = b.counters.clone()
    print(json.dumps({'variant': 'plain step (one erode launch, one point-pass launch), CUDA graph replay', 'ms_per_step': plain}))
    p = lambda t, off=0: C.c_void_p(t.data_ptr() + off)
    for cf in [int(x) for x in args.chunks.split(',')]:
        if cf > args.frames:
            continue
        ring = torch.empty((2, cf * 16, plane), dtype=torch.int32, device='cuda')
        mo, F = b.h_mask_off, b.F

        def step():
            st = ctx.stream
            _abi.check(lib.lf_zero_counters(ctx.h, p(b.counters), b.counters.numel() * 4, st))
            for ci, f0 in enumerate(range(0, F, cf)):
                f1 = min(F, f0 + cf)
                m0, m1 = int(mo[f0]), int(mo[f1])
                buf = ring[ci & 1]
                _abi.check(lib.lf_erode_pack(ctx.h, p(b.masks, m0 * H * W), 0, m1 - m0, H, W, se.ctypes.data, 3, 1,
                                             p(buf), p(b.mask_bbox, 16 * m0), st), ctx.h)
                _abi.check(lib.lf_fuse_batch(ctx.h, p(b.points), p(b.pt_off, 8 * f0),
                                             C.c_void_p(buf.data_ptr() - m0 * plane * 4), p(b.mask_bbox),
                                             p(b.mask_off, 4 * f0), p(b.boxes), p(b.box_off, 4 * f0), p(b.cnt_off, 8 * f0),
                                             f1 - f0, b.max_n, b.max_m, b.max_b, 50.0, 1, p(b.total), p(b.inside),
                                             p(b.n_in_box), p(b.n_valid, 4 * f0), None, st), ctx.h)
            _abi.check(lib.lf_car_stats(ctx.h, p(b.total), p(b.inside), p(b.mask_off), p(b.box_off), p(b.cnt_off), F, 10,
                                        p(b.best_idx), p(b.best_cnt), p(b.agg), st), ctx.h)

        step()                                                  # warm-up outside the capture
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):                               # launch overhead out of the picture: one graph replay per step
            step()
        ms = timed(g.replay)
        same = bool(torch.equal(b.counters, want))
        print(json.dumps({'variant': 'chunks of %d frames, packed words in a 2-slot ring (%.0f MB per slot), CUDA graph replay' % (cf, cf * 16 * plane * 4 / 1e6),
                          'ms_per_step': ms, 'counters_equal_plain_step': same, 'launches_per_step': 1 + 3 * ((F + cf - 1) // cf)}))
        del ring


if __name__ == '__main__':
    main()
