#!/usr/bin/env python
"""Device timings of the SURVEY 8(f) 'next' rows (not the BASELINE metric: bench.py is).

    python bench_next.py [--reps 20]

One JSON line per kernel: launch time from CUDA events on the launching stream (after
3 warm-ups, inputs larger than L2 or the buffer re-read cold where it matters), the
algorithmic bytes of the launch and the resulting GB/s against MEASURED_PEAKS.json.
Sizes are cfg3-shaped (1024 frames x 16 masks / boxes, 131072 points per frame).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
CALIB_ROOT = os.path.join(ROOT, 'tests', 'golden', 'kitti360_sample')


def surface_kernels(args, timed, ctx, cam, TVC, TVR, LF, S, check, ptr, torch, np):
    """The stand-alone kernels behind the reference's one-call-at-a-time function surface (A3 / A4
    projection + gate, A6 gather, A8 containment, A10 depth maps) on 256 synthetic frames = 33.5 M points
    (537 MB, larger than L2)."""
    from lidar_object_detection_b200 import _abi
    H, W, st, dev = cam.height, cam.width, ctx.stream, ctx.device
    F = 256
    sb = S.SyntheticBatch(F, cam, TVC, TVR, ncars=16, device=0)
    sb.batch.run(3, 0, 50, 10, True)                        # packs the (un-eroded) masks
    torch.cuda.synchronize()
    pts, n = sb.points, sb.points.shape[0]
    u = torch.empty(n, dtype=torch.int64, device=dev)
    v = torch.empty(n, dtype=torch.int64, device=dev)
    d = torch.empty(n, dtype=torch.float64, device=dev)
    pix = torch.empty(n, dtype=torch.int32, device=dev)
    timed(lambda: check(ctx.lib.lf_project(ctx.h, ptr(pts), n, 50.0, ptr(u), ptr(v), ptr(d), ptr(pix), st), ctx.h),
          n * (16 + 28), 'lf_project (u, v int64 + depth f64 + gated pixel out)', '%d points' % n)
    timed(lambda: check(ctx.lib.lf_project(ctx.h, ptr(pts), n, 50.0, None, None, None, ptr(pix), st), ctx.h),
          n * (16 + 4), 'lf_project (gated pixel index only)', '%d points' % n)
    words = sb.batch.packed.words
    bits = torch.empty(n, dtype=torch.int64, device=dev)
    timed(lambda: check(ctx.lib.lf_mask_gather(ctx.h, ptr(words), 16, H, W, ptr(u), ptr(v), n, ptr(bits), st), ctx.h),
          n * (16 + 8), 'lf_mask_gather', '%d pixel pairs (int64 u, v) x 16 packed masks -> 64-bit membership' % n)
    inb = torch.empty(n, dtype=torch.uint8, device=dev)
    boxes = sb.batch.boxes
    timed(lambda: check(ctx.lib.lf_points_in_box(ctx.h, ptr(pts), 0, n, 4, ptr(boxes), 1, ptr(inb), st), ctx.h),
          n * (16 + 1), 'lf_points_in_box (oriented)', '%d float32 points x 1 box -> uint8 flags' % n)
    sid = (torch.arange(n, device=dev, dtype=torch.int32) & 15)
    B = int(sb.n_boxes[0])
    cnt = torch.zeros(16 * max(B, 1), dtype=torch.int32, device=dev)
    timed(lambda: check(ctx.lib.lf_count_in_boxes(ctx.h, ptr(pts), 0, n, 4, ptr(sid), ptr(boxes), B, 1, ptr(cnt), st),
                        ctx.h), n * (16 + 4), 'lf_count_in_boxes (oriented)', '%d points x %d boxes, 16 point sets' % (n, B))
    del u, v, d, pix, bits, inb, sid
    # A10 depth maps: one frame per call (the reference's loop) and the batch entry point
    npix = H * W
    scr = torch.empty(npix, dtype=torch.int64, device=dev)
    per = torch.empty((16, H, W), dtype=torch.float64, device=dev)
    n1 = sb.n
    for pol, name in ((_abi.POLICY_LAST, 'last'), (_abi.POLICY_NEAREST, 'nearest')):
        timed(lambda: check(ctx.lib.lf_depth_maps(ctx.h, ptr(pts), n1, 30.0, pol, ptr(words), 16, ptr(scr), None, ptr(per),
                                                  st), ctx.h),
              n1 * 16 + npix * 8 * 17, 'lf_depth_maps (%s)' % name, 'one frame, %d points, 16 per-car f64 maps' % n1)
    FB = 64
    scrb = torch.empty((FB, npix), dtype=torch.int64, device=dev)
    perb = torch.empty((FB * 16, H, W), dtype=torch.float64, device=dev)
    fullb = torch.empty((FB, H, W), dtype=torch.float64, device=dev)
    for pol, name in ((_abi.POLICY_LAST, 'last'), (_abi.POLICY_NEAREST, 'nearest')):
        timed(lambda: check(ctx.lib.lf_depth_maps_batch(ctx.h, ptr(pts), ptr(sb.batch.pt_off), FB, n1, 30.0, pol, ptr(words),
                                                        ptr(sb.batch.mask_off), ptr(scrb), ptr(fullb), ptr(perb), st), ctx.h),
              FB * (n1 * 16 + npix * 8 * (1 + 1 + 1 + 16)), 'lf_depth_maps_batch (%s)' % name,
              '%d frames in one call: key image cleared + read, frame map + 16 per-car f64 maps each' % FB)
        timed(lambda: check(ctx.lib.lf_depth_maps_batch(ctx.h, ptr(pts), ptr(sb.batch.pt_off), FB, n1, 30.0, pol, None,
                                                        None, ptr(scrb), ptr(fullb), None, st), ctx.h),
              FB * (n1 * 16 + npix * 8 * 3), 'lf_depth_maps_batch (%s, frame maps only)' % name, '%d frames' % FB)
    del sb, scrb, perb, fullb
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--reps', type=int, default=20)
    ap.add_argument('--only', default='', help="'surface' = only the stand-alone 8(a) kernels and the depth maps")
    args = ap.parse_args()
    import numpy as np
    import torch
    if not torch.cuda.is_available():
        raise SystemExit('bench_next.py needs a CUDA device')
    from lidar_object_detection_b200 import _abi, io as IO, kitti360 as K, lidarfuse as LF, synthetic as S
    from lidar_object_detection_b200._abi import check, ptr
    try:
        peak = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs'])
    except Exception:
        peak = 6650.0
    cam, TVC, TVR = K.calibration_chain(CALIB_ROOT)
    ctx = LF.get_ctx(0)
    ctx.set_camera(TVR, cam.K, cam.width, cam.height)
    dev, st = ctx.device, ctx.stream
    H, W = cam.height, cam.width

    def timed(fn, nbytes, name, note):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = []
        for _ in range(args.reps):
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        med = ms[len(ms) // 2]
        gbs = nbytes / (med * 1e-3) / 1e9
        print(json.dumps({'kernel': name, 'ms': med, 'ms_min': ms[0], 'algorithmic_bytes': int(nbytes),
                          'achieved_gbs': gbs, 'peak_gbs': peak, 'frac': gbs / peak, 'note': note}))

    surface_kernels(args, timed, ctx, cam, TVC, TVR, LF, S, check, ptr, torch, np)
    if args.only == 'surface':
        return
    # rank 1: resize 16384 half-resolution masks to camera resolution (bilinear), and 2x area
    M = 1024 * 16
    g = torch.Generator(device=dev).manual_seed(1)
    for (h, w), label in (((192, 704), 'bilinear 192x704 -> 376x1408'), ((752, 2816), '2x2 average 752x2816 -> 376x1408')):
        m = M if h < H else M // 8
        src = (torch.rand((m, h, w), device=dev, generator=g) < 0.3).to(torch.uint8)
        dst = torch.empty((m, H, W), dtype=torch.uint8, device=dev)
        timed(lambda: check(ctx.lib.lf_resize_masks(ctx.h, ptr(src), m, h, w, ptr(dst), H, W, st), ctx.h),
              m * (h * w + H * W), 'lf_resize_masks', '%d masks, %s' % (m, label))
        del src, dst
    # the same resize from packed words to packed words (what the batch path does for low-resolution masks)
    for (h, w), label in (((192, 704), '192x704 -> 376x1408'), ((752, 2816), '752x2816 -> 376x1408 (2x reduction)')):
        m = M if h < H else M // 8
        srcw = torch.zeros((m, h, (w + 31) // 32), dtype=torch.int32, device=dev)
        srcw[:, h // 4:h // 2, 3:9] = -1                     # one blob per mask, like an instance mask
        dstw = torch.empty((m, H, (W + 31) // 32), dtype=torch.int32, device=dev)
        bbr = torch.zeros((m, 4), dtype=torch.int32, device=dev)
        timed(lambda: check(ctx.lib.lf_resize_pack(ctx.h, ptr(srcw), m, h, w, ptr(dstw), ptr(bbr), H, W, st), ctx.h),
              (srcw.numel() + dstw.numel()) * 4, 'lf_resize_pack', '%d packed masks, %s' % (m, label))
        del srcw, dstw
    # the reference's own mask format: float32 0.0 / 1.0 masks through lf_erode_pack (bulk-copy fast path)
    mf = 2048
    f32 = torch.zeros((mf, H, W), dtype=torch.float32, device=dev)
    f32[:, 100:250, 300:900] = 1.0
    pk = torch.empty((mf, H, (W + 31) // 32), dtype=torch.int32, device=dev)
    bb = torch.zeros((mf, 4), dtype=torch.int32, device=dev)
    se = LF.structuring_element(LF.MORPH_ELLIPSE, 3)
    timed(lambda: check(ctx.lib.lf_erode_pack(ctx.h, ptr(f32), 1, mf, H, W, se.ctypes.data, 3, 1, ptr(pk), ptr(bb), st),
                        ctx.h), mf * H * W * 4, 'lf_erode_pack (float32 masks)', '%d masks 1408x376 float32, 3x3 ellipse' % mf)
    del f32
    # masks that arrive bit-packed (lf_pack_masks_host on the sending side): erosion from the words
    mp = 16384
    raw = torch.zeros((mp, H, (W + 31) // 32), dtype=torch.int32, device=dev)
    raw[:, 100:250, 10:28] = -1
    pk = torch.empty_like(raw)
    bb = torch.zeros((mp, 4), dtype=torch.int32, device=dev)
    timed(lambda: check(ctx.lib.lf_erode_pack(ctx.h, ptr(raw), 3, mp, H, W, se.ctypes.data, 3, 1, ptr(pk), ptr(bb), st),
                        ctx.h), 2 * raw.numel() * 4, 'lf_erode_pack (packed words in)',
          '%d masks 1408x376 at 1 bit per pixel, 3x3 ellipse; bytes = words in + words out' % mp)
    del raw, pk
    # rank 2: project 1M boxes; IoU/score matrices for 65536 frames x 16 x 16
    nb = 65536 * 16
    cars = S.make_cars(4096, 16, 7)
    cv = S.car_corners_velo(cars).reshape(-1, 3)
    cam0 = (np.hstack([cv, np.ones((len(cv), 1))]) @ TVC.T)[:, :3].reshape(-1, 8, 3)
    corners = torch.from_numpy(np.ascontiguousarray(np.tile(cam0, (16, 1, 1)))).to(dev)
    rects = torch.zeros((nb, 72), dtype=torch.uint8, device=dev)
    timed(lambda: check(ctx.lib.lf_project_boxes(ctx.h, ptr(corners), nb, 0.1, 100.0, 4, ptr(rects), None, None,
                                                 None, st), ctx.h),
          nb * (192 + 72), 'lf_project_boxes', '%d boxes (8 corners f64 in, 72-byte record out)' % nb)
    F = 65536
    dets = torch.rand((F * 16, 4), dtype=torch.float64, device=dev) * 300
    dets[:, 2:] += dets[:, :2] + 20
    off = torch.arange(0, F * 16 + 1, 16, dtype=torch.int32, device=dev)
    ooff = torch.arange(0, F * 256 + 1, 256, dtype=torch.int64, device=dev)
    iou = torch.empty(F * 256, dtype=torch.float64, device=dev)
    score = torch.empty(F * 256, dtype=torch.float64, device=dev)
    timed(lambda: check(ctx.lib.lf_match_matrices(ctx.h, ptr(dets), ptr(off), ptr(rects), ptr(off), ptr(ooff), 65535,
                                                  256, 0.5, 0.3, 0.2, ptr(iou), ptr(score), st), ctx.h),
          65535 * (16 * 32 + 16 * 72 + 2 * 256 * 8), 'lf_match_matrices', '65535 frames x 16 detections x 16 boxes')
    del corners, rects, dets, iou, score
    # rank 3: per-point colours for 64M points; depth overlay of 256 frames
    n = 64 * 1024 * 1024
    bits = torch.randint(0, 1 << 16, (n,), dtype=torch.int64, device=dev, generator=g)
    bits[torch.rand(n, device=dev, generator=g) < 0.8] = 0
    cols = torch.rand((16, 3), dtype=torch.float64, device=dev)
    rgb = torch.empty((n, 3), dtype=torch.float64, device=dev)
    bg = np.array([0.5, 0.5, 0.5])
    timed(lambda: check(ctx.lib.lf_point_colors(ctx.h, ptr(bits), n, ptr(cols), 16, bg.ctypes.data, ptr(rgb), None, 0, st),
                        ctx.h), n * (8 + 24), 'lf_point_colors', '%d points, 16 masks, f64 RGB out' % n)
    del bits, rgb
    from lidar_object_detection_b200 import colour as C
    lut = torch.from_numpy(C.jet_lut()).to(dev)
    nf = 64
    dm = torch.rand((nf, H, W), dtype=torch.float64, device=dev, generator=g) * 30
    dm[torch.rand((nf, H, W), device=dev, generator=g) < 0.9] = 0
    img = torch.randint(0, 256, (nf, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
    out = torch.empty_like(img)
    mx = dm.reshape(nf, -1).amax(dim=1)

    def overlay():
        for i in range(nf):
            check(ctx.lib.lf_depth_overlay(ctx.h, ptr(dm[i]), ptr(mx[i:]), ptr(img[i]), ptr(lut), 256, H, W,
                                           ptr(out[i]), st), ctx.h)
    timed(overlay, nf * H * W * (8 + 3 + 3), 'lf_depth_overlay', '%d launches of one %dx%d frame each' % (nf, W, H))
    timed(lambda: check(ctx.lib.lf_depth_overlay_batch(ctx.h, ptr(dm), ptr(mx), ptr(img), H * W * 3, ptr(lut), 256, nf, H, W,
                                                       ptr(out), st), ctx.h),
          nf * H * W * (8 + 3 + 3), 'lf_depth_overlay_batch', '%d frames of %dx%d in one launch' % (nf, W, H))
    # rank 4: real sweeps from disk through SweepStreamer + one FrameBatch per 64 sweeps (page-cache reads)
    import glob
    import tempfile
    from lidar_object_detection_b200 import driver, matching2d as M2
    npz = sorted(glob.glob(os.path.join(CALIB_ROOT, 'velodyne', '*.npz')))
    with tempfile.TemporaryDirectory() as tmp:
        uniq = []
        for pth in npz:
            fr = int(os.path.basename(pth)[:-4])
            raw = K.load_bounding_boxes(os.path.join(CALIB_ROOT, 'bboxes_3D_cam0', 'BBoxes_%d.json' % fr))
            if not raw:
                continue
            binp = os.path.join(tmp, '%010d.bin' % fr)
            np.ascontiguousarray(np.load(pth)['xyzi'].T, dtype=np.float32).tofile(binp)
            vis = LF.filter_visible_bboxes(raw, cam)
            boxes = LF.transform_bboxes_to_velodyne(vis, TVC)
            # stand-in detector output: the pixel rectangles of the (up to 12) nearest visible boxes, on the device
            rects = M2.project_boxes_batched(np.array([b['corners_cam0'] for b in vis]).reshape(-1, 8, 3), cam) \
                if vis else None
            mk = torch.zeros((min(len(vis), 12), H, W), dtype=torch.uint8, device=dev)
            for c in range(len(mk)):
                r = rects[c]
                x0, y0, x1, y1 = max(int(r['x0']), 0), max(int(r['y0']), 0), min(int(r['x1']), W - 1), min(int(r['y1']), H - 1)
                if x1 > x0 and y1 > y0:
                    mk[c, y0:y1, x0:x1] = 1
            uniq.append((binp, mk, np.array([b['corners_velo'] for b in boxes]).reshape(-1, 8, 3)))
        reps = 50
        seq = uniq * reps
        paths = [u[0] for u in seq]
        driver.process_sweeps_streamed(paths[:128], lambda i: seq[i][1:], cam, TVR, frames_per_batch=64)   # warm-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rows = driver.process_sweeps_streamed(paths, lambda i: seq[i][1:], cam, TVR, frames_per_batch=64)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        npts = sum(IO.sweep_points(p) for p in paths)
        print(json.dumps({'kernel': 'driver.process_sweeps_streamed (disk -> pinned -> H2D -> erode + fuse -> rows on host)',
                          'ms': dt * 1e3, 'sweeps': len(paths), 'sweeps_per_s': len(paths) / dt, 'points_per_s': npts / dt,
                          'read_gbs': npts * 16 / dt / 1e9, 'rows': len(rows),
                          'note': '%d real sample sweeps x %d from the page cache, 64 per batch, rectangle stand-in masks '
                                  'resident on the device' % (len(uniq), reps)}))
    # rank 4: CSV rows on the host
    nrows = 1_000_000
    rng = np.random.default_rng(0)
    total = rng.integers(1, 5000, nrows)
    rows = np.stack([rng.integers(0, 65536, nrows), rng.integers(0, 16, nrows), rng.integers(-1, 16, nrows), total,
                     (total * rng.random(nrows)).astype(np.int64)], axis=1)
    t0 = time.perf_counter()
    text = IO.format_csv_rows(rows, '2025-01-01T00:00:00', with_header=True)
    dt = time.perf_counter() - t0
    print(json.dumps({'kernel': 'lf_format_csv_rows (host, 1 thread)', 'ms': dt * 1e3, 'rows': nrows,
                      'rows_per_s': nrows / dt, 'bytes': len(text)}))


if __name__ == '__main__':
    main()
