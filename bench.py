#!/usr/bin/env python
"""points/s through project + mask (erode + gather) + bbox evaluation on synthetic
HDL-64-shaped frames (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step = one pass of the hot path over one batch per GPU: counter memset,
erode+pack of the uint8 masks, the fused point pass and the per-car decision
(4 launches),
plus -- at N > 1 -- the single all-reduce of the per-car table.
N = 1 runs BASELINE configs[2]: 1024 frames x 131072 points x 16 masks
(1408x376) x 16 boxes, all resident in HBM (10.8 GB >> 126 MB L2, so no L2
flush is needed between steps).  N > 1 is weak scaling: every rank holds its
own 1024-frame shard (cfg4's frames-per-chunk).

--impl reference times the reference's CPU path (the numpy/OpenCV oracle port
of cvs_erosion.py's per-frame body) on the host cores, on the same kind of
frames, and prints the same JSON line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CALIB_ROOT = os.path.join(ROOT, 'tests', 'golden', 'kitti360_sample')
METRIC = 'points/sec (project+mask+bbox eval)'
FRAMES, NCARS, RINGS, AZ = 1024, 16, 64, 2048
KSIZE = 3
# BASELINE.json configs[2] (the metric's configuration) and configs[4] (dense stress)
WORKLOADS = {'cfg3': dict(frames=1024, ncars=16, rings=64, az=2048, ksize=3),
             'cfg5': dict(frames=64, ncars=64, rings=64, az=32768, ksize=7)}


def set_workload(name, frames=None):
    global FRAMES, NCARS, RINGS, AZ, KSIZE
    w = WORKLOADS[name]
    FRAMES, NCARS, RINGS, AZ, KSIZE = frames or w['frames'], w['ncars'], w['rings'], w['az'], w['ksize']
W_IMG, H_IMG = 1408, 376
SEED = 0x5EED0000


def algorithmic_bytes_per_frame(n_points, n_masks, n_boxes):
    """compulsory reads + writes, each counted once."""
    return 16 * n_points + n_masks * H_IMG * W_IMG + 192 * n_boxes + 8 * n_masks * (n_boxes + 3)


# --------------------------------------------------------------------------- CPU arm

def _cpu_worker(rank, nframes, warm, steps, seed, barrier, queue, workload='cfg3', kind='port'):
    """One host process: generate ``nframes`` synthetic frames with numpy (untimed),
    then run ``warm`` + ``steps`` frames of the CPU path (cycling over its frames); only the
    ``steps`` part is timed.  kind 'reference': the reference's OWN functions, from the unmodified
    scripts (oracle/reference_frame.py: cvs_erosion.py:338-372 once each, nothing added);
    'port': the oracle's numpy restatement (only where the scripts are not staged)."""
    os.environ['OPENBLAS_NUM_THREADS'] = '1'
    os.environ['OMP_NUM_THREADS'] = '1'
    import numpy as np
    import cv2
    cv2.setNumThreads(1)
    from lidar_object_detection_b200 import kitti360 as K, synthetic as S
    from oracle import pipeline as P
    set_workload(workload)
    cam, TVC, TVR = K.calibration_chain(CALIB_ROOT)
    cars = S.sort_cars_by_image_area(S.make_cars(nframes, NCARS, seed + 7919 * rank), cam, TVR)
    boxes = S.reference_boxes(cars, cam, TVC)
    frames = []
    for f in range(nframes):
        pts = S.numpy_frame(cars[f], seed + 7919 * rank + f, RINGS, AZ)
        masks = S.numpy_masks(cars[f], cam, TVR).astype(np.float32)     # YOLO hands float32 masks over
        raw = [{'corners_cam0': b['corners_cam0']} for b in boxes[f]]
        frames.append((pts, raw, masks))
    rows = 0
    if kind == 'reference':
        from oracle import reference_frame as RF
        cvs = RF.load()

        def one(fr):
            return len(RF.process_frame(cvs, *fr, cam, TVC, TVR, erosion_kernel_size=KSIZE))
    else:
        def one(fr):
            return len(P.process_frame(*fr, cam, TVC, TVR, erosion_kernel_size=KSIZE)['stats'])
    for i in range(warm):
        one(frames[i % nframes])
    barrier.wait()
    t0 = time.time()
    for i in range(steps):
        rows += one(frames[i % nframes])
    t1 = time.time()
    queue.put((rank, t0, t1, steps * len(frames[0][0]), rows))


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_kind():
    """'reference' when the reference's scripts are present (/root/reference or staged oracle/_ref)."""
    try:
        from oracle import reference_loader as RL
        return 'reference' if RL.available() else 'port'
    except Exception:
        return 'port'


def cpu_reference_throughput(steps, warm=1, distinct_frames=2, workers=None, workload='cfg3', kind=None):
    """points/s of the CPU path, one process per host core, each timing ``steps``
    frames (a bounded sample of the workload)."""
    import multiprocessing as mp
    workers = workers or host_cores()
    kind = kind or cpu_kind()
    ctx = mp.get_context('spawn')
    barrier, queue = ctx.Barrier(workers), ctx.Queue()
    procs = [ctx.Process(target=_cpu_worker, args=(r, distinct_frames, warm, steps, SEED, barrier, queue, workload, kind))
             for r in range(workers)]
    for p in procs:
        p.start()
    res = [queue.get() for _ in procs]
    for p in procs:
        p.join()
    t0 = min(r[1] for r in res)
    t1 = max(r[2] for r in res)
    points = sum(r[3] for r in res)
    return {'value': points / (t1 - t0), 'seconds': t1 - t0, 'points': points, 'cores': workers,
            'frames': steps * workers, 'rows': sum(r[4] for r in res), 'kind': kind}


CPU_KIND_NOTE = {'reference': "the reference's own functions from its unmodified scripts (cvs_erosion.py:338-372, "
                              'float32 masks in, per-car dicts out)',
                 'port': "the oracle's numpy port (reference scripts not staged)"}


def bench_config(workload, frames, n_pts, world, step_bytes=None, exchange=None):
    """The `config` object both arms print (same workload, same keys)."""
    if step_bytes is None:
        step_bytes = frames * algorithmic_bytes_per_frame(n_pts, NCARS, NCARS)
    return {'workload': '%s: %d frames x %d pts x %d masks (%dx%d uint8) x <=%d boxes per GPU, '
                        '%dx%d ellipse erosion, 50 m gate' % (workload, frames, n_pts, NCARS, W_IMG, H_IMG,
                                                             NCARS, KSIZE, KSIZE),
            'frames_per_gpu': frames, 'points_per_frame': n_pts, 'masks_per_frame': NCARS,
            'l2_policy': 'inputs (%.1f GB per GPU) larger than the 126 MB L2; no flush' % (step_bytes / 1e9),
            'parallelism': 'frames sharded, dp%d' % world,
            'exchange': exchange}


# --------------------------------------------------------------------------- clocks

class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t_begin, t_end):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [l for (t, l) in self.lines if t_begin - 0.05 <= t <= t_end + 0.15] or [l for (_, l) in self.lines]
        mhz, mx, reasons = [], None, set()
        for l in rows:
            p = [x.strip() for x in l.split(',')]
            try:
                mhz.append(float(p[0])); mx = float(p[1])
            except Exception:
                continue
            for name, val in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), p[2:6]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': statistics.median(mhz) if mhz else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(mhz)}


# --------------------------------------------------------------------------- GPU arm

def run_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    frames = FRAMES

    # CPU baseline first (rank 0, N = 1 only): spawned host processes, GPU idle meanwhile
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference_throughput(steps=args.cpu_frames, warm=1, workload=args.workload)
        cpu = {'value': c['value'], 'unit': 'points/s', 'cores': c['cores'], 'kind': c['kind'],
               'sample': '%d synthetic frames (%d per host process, %d processes, OPENBLAS/cv2 threads = 1), '
                         '%.1f s wall; %s' % (c['frames'], args.cpu_frames, c['cores'], c['seconds'], CPU_KIND_NOTE[c['kind']])}

    from lidar_object_detection_b200 import distributed as D, kitti360 as K, synthetic as S
    cam, TVC, TVR = K.calibration_chain(CALIB_ROOT)
    sb = S.SyntheticBatch(frames, cam, TVC, TVR, ncars=NCARS, seed=SEED, rings=RINGS, azimuths=AZ, device=local,
                          first_frame=rank * frames)
    batch = sb.batch
    n_pts = sb.n
    exch = None
    if world > 1:
        nccl_exch = D.CounterExchange(world * frames, NCARS, rank, world, dev)
        nccl_exch.bind(batch.h_mask_off, sb.n_boxes)
        exch, exch_kind = nccl_exch, 'one NCCL all_reduce(SUM) over the zero-filled int64 table per batch'
        can = torch.tensor([int(args.exchange == 'peer' and D.PeerCounterExchange.available())], device=dev)
        dist.all_reduce(can, op=dist.ReduceOp.MIN)       # every rank must be able to, or none uses it
        if int(can.item()):
            try:
                exch = D.PeerCounterExchange(world * frames, NCARS, rank, world, dev)
                exch_kind = 'lf_publish_table: rows stored into every rank\'s table over NVLink peer memory + one ' \
                            'signal-pad barrier per batch'
            except Exception as e:                       # no peer access on this box: the NCCL exchange stays
                sys.stderr.write('peer exchange unavailable (%r), using NCCL\n' % (e,))
        if exch is not nccl_exch:                        # both exchanges must give the same table
            batch.run(KSIZE, 1, 50, 10, True)
            t1, a1 = nccl_exch.exchange_batch(batch)
            t2, _ = exch.exchange_batch(batch)
            torch.cuda.synchronize(dev)
            if not (torch.equal(t1.to(torch.int32), t2) and torch.equal(a1, exch.agg)):
                raise SystemExit('PARITY FAILURE: peer-store exchange differs from the NCCL all-reduce')

    def step(mark=None):
        batch.run(KSIZE, 1, 50, 10, True, mark=mark)
        if exch is not None:
            exch.exchange_batch(batch)

    stream = torch.cuda.current_stream(dev)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()

    # parity spot check against the oracle (checker only): two frames of this very batch
    parity_frames = 0
    if rank == 0 and not args.no_parity:
        from oracle import pipeline as P
        res = batch.results()
        for f in (0, frames - 1):
            pts, masks, boxes = sb.host_frame(f)
            want = P.process_frame(pts, boxes, masks.astype(np.float32), cam, TVC, TVR, erosion_kernel_size=KSIZE)
            got_t = res['total'][batch.h_mask_off[f]:batch.h_mask_off[f + 1]]
            got_i = res['inside'][batch.h_cnt_off[f]:batch.h_cnt_off[f + 1]].reshape(want['inside'].shape)
            if not (np.array_equal(got_t, want['total']) and np.array_equal(got_i, want['inside'])
                    and res['n_valid'][f] == want['valid'].sum()):
                raise SystemExit('PARITY FAILURE on synthetic frame %d: refusing to report a number' % f)
            parity_frames += 1

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    K_steps = args.steps
    marks = []

    def make_mark(lst):
        def mark(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            lst.append((name, e))
        return mark

    sync_all()
    t_begin = time.time()
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(K_steps):
        lst = []
        step(make_mark(lst))
        marks.append(lst)
    ev1.record(stream)
    sync_all()
    t_end = time.time()
    ms_total = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / K_steps
    value = world * frames * n_pts / (ms_step * 1e-3)

    # per-kernel device time from the marks (same stream, same timed region)
    kern = {}
    for lst in marks:
        d = dict(lst)
        pairs = (('zero', 'erode_pack'), ('erode_pack', 'fuse'), ('fuse', 'car_stats'))
        for a, b in pairs:
            kern.setdefault(b, []).append(d[a].elapsed_time(d[b]))
    kern_ms = {k: sum(v) / len(v) for k, v in kern.items()}

    # ---- end to end: host buffers in, host counters out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        h_pts = torch.empty(batch.points.shape, dtype=batch.points.dtype, pin_memory=True)
        h_masks = torch.empty(batch.masks.shape, dtype=batch.masks.dtype, pin_memory=True)
        h_out = torch.empty(batch.counters.shape, dtype=batch.counters.dtype, pin_memory=True)
        h_pts.copy_(batch.points)
        h_masks.copy_(batch.masks)
        torch.cuda.synchronize(dev)

        def e2e_step():
            batch.upload(h_pts, h_masks)
            step()
            batch.download(h_out)
            stream.synchronize()

        e2e_step()
        sync_all()
        n_e2e = max(1, min(args.e2e_steps, K_steps))
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        sync_all()
        dt = (time.perf_counter() - t0) / n_e2e
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {'value': world * frames * n_pts / dt, 'unit': 'points/s',
               'h2d_bytes_per_step': int(h_pts.numel() * 4 + h_masks.numel()),
               'd2h_bytes_per_step': int(h_out.numel() * 4), 'steps': n_e2e, 'ms_per_step': dt * 1e3,
               'api': 'FrameBatch.upload (pinned host points + uint8 masks as they are) / run / download (counters)'}
        t_end = time.time()

    # ---- secondary accounting (SURVEY 8(d) "packed-masks-in"): the masks cross PCIe and enter HBM at one
    # bit per pixel (lf_pack_masks_host on the sending side), erosion runs on the device from the words
    packed_in = None
    if not args.no_packed:
        from lidar_object_detection_b200 import io as IO, lidarfuse as LF
        words = LF.erode_pack(batch.masks, 0, 0, device=local).words
        pb = LF.FrameBatch(batch.points, [n_pts] * frames, LF.PackedMasks(words, H_IMG, W_IMG, raw=True), [NCARS] * frames,
                           batch.boxes, sb.n_boxes, cam, TVR, device=local)

        def pstep(mark=None):
            pb.run(KSIZE, 1, 50, 10, True, mark=mark)
            if exch is not None:
                exch.exchange_batch(pb)

        for _ in range(3):
            pstep()
        sync_all()
        same = all(torch.equal(a, b) for a, b in ((pb.total, batch.total), (pb.inside, batch.inside),
                                                  (pb.best_idx, batch.best_idx), (pb.packed.words, batch.packed.words)))
        if not same:
            raise SystemExit('PARITY FAILURE: packed-masks-in step differs from the uint8 step')
        pm = []
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record(stream)
        for _ in range(K_steps):
            lst = []
            pstep(make_mark(lst))
            pm.append(dict(lst))
        p1.record(stream)
        sync_all()
        pms = p0.elapsed_time(p1) / K_steps
        if world > 1:
            t = torch.tensor([pms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            pms = float(t.item())
        wpr = (W_IMG + 31) // 32
        pbytes = sum(16 * n_pts + NCARS * H_IMG * wpr * 4 + 192 * b + 8 * NCARS * (b + 3) for b in sb.n_boxes)
        er_ms = sum(d['zero'].elapsed_time(d['erode_pack']) for d in pm) / len(pm)
        packed_in = {'value': world * frames * n_pts / (pms * 1e-3), 'unit': 'points/s', 'ms_per_step': pms,
                     'kernel_ms': {'erode_pack': er_ms,
                                   'fuse': sum(d['erode_pack'].elapsed_time(d['fuse']) for d in pm) / len(pm)},
                     'algorithmic_bytes_per_step': pbytes, 'achieved_gbs': pbytes / (pms * 1e-3) / 1e9,
                     'erode_gbs_in_plus_out': 2 * frames * NCARS * H_IMG * wpr * 4 / (er_ms * 1e-3) / 1e9,
                     'note': 'lf_erode_pack mode 3 (masks arrive as uint32 words); results identical with the uint8 step'}
        if e2e is not None:
            hm = h_masks.numpy()
            h_words = torch.empty(words.shape, dtype=words.dtype, pin_memory=True)
            hw = h_words.numpy()
            threads = max(1, (os.cpu_count() or 1) // world)       # the ranks share the host's cores

            def e2e_packed_step():
                pb.upload(h_pts, h_masks, threads=threads)         # packs a slice while the previous one is on the wire
                pstep()
                pb.download(h_out)
                stream.synchronize()

            e2e_packed_step()
            if not torch.equal(h_out, batch.counters.cpu()):
                raise SystemExit('PARITY FAILURE: host-packed end-to-end step differs')
            sync_all()
            tp0 = time.perf_counter()
            for _ in range(n_e2e):
                e2e_packed_step()
            sync_all()
            dtp = (time.perf_counter() - tp0) / n_e2e
            if world > 1:
                t = torch.tensor([dtp], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dtp = float(t.item())
            tq0 = time.perf_counter()
            IO.pack_masks(hm, threads=threads, out=hw)
            tq = time.perf_counter() - tq0
            packed_in['e2e'] = {'value': world * frames * n_pts / dtp, 'unit': 'points/s', 'ms_per_step': dtp * 1e3,
                                'h2d_bytes_per_step': int(h_pts.numel() * 4 + words.numel() * 4),
                                'd2h_bytes_per_step': int(h_out.numel() * 4), 'host_pack_threads': threads,
                                'host_pack_gbs_alone': hm.nbytes / tq / 1e9,
                                'api': 'FrameBatch.upload on PackedMasks(raw=True): lf_pack_masks_host in 16 slices '
                                       'overlapped with their H2D / run / download; pinned uint8 masks in, counters out'}
        t_end = time.time()
        del pb
        # the end-to-end headline is the faster of the two host-to-host calls; the other stays beside it
        if e2e is not None and 'e2e' in packed_in and packed_in['e2e']['value'] > e2e['value']:
            plain = e2e
            e2e = dict(packed_in['e2e'], variant='masks bit-packed on the host for the transfer (erosion on the device)',
                       plain_transfer=plain)
        elif e2e is not None:
            e2e['variant'] = 'masks sent as uint8 arrays'

    clocks = sampler.stop(t_begin, t_end) if rank == 0 else None
    if rank == 0:
        peaks, peak_src = None, 'fallback'
        try:
            with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as fh:
                peaks = json.load(fh)
            peak, peak_src = float(peaks['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
        except Exception:
            peak = 6650.0
        step_bytes = sum(algorithmic_bytes_per_frame(n_pts, NCARS, b) for b in sb.n_boxes)
        se_name = 'SeCross3' if KSIZE == 3 else 'SeEllipse%d' % KSIZE
        # dominant kernel: the erode launch, F*M*H*W bytes in
        kname, kkey, tkey = 'lf_erode_fast_kernel<%s> (lf_erode_pack)' % se_name, 'erode_pack', 'lf_erode_fast_kernel'
        mask_bytes = frames * NCARS * H_IMG * W_IMG
        ach = mask_bytes / (kern_ms[kkey] * 1e-3) / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, 'profiles', 'dram_traffic.json' if args.workload == 'cfg3' else 'none')) as fh:
                traffic = json.load(fh).get(tkey + '_dram_bytes_per_launch')
        except Exception:
            pass
        line = {
            'metric': METRIC, 'value': value, 'unit': 'points/s', 'n_gpus': world, 'steps': K_steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f64 projection/containment, u32 bit-masks, i32 counters',
            'data': 'synthetic',
            'config': bench_config(args.workload, frames, n_pts, world, step_bytes, exch_kind if world > 1 else None),
            'gpu_launches': int(batch.launches_per_run - 1) * K_steps,
            'kernel_ms': kern_ms,
            'step_hbm': {'algorithmic_bytes_per_step': step_bytes, 'achieved_gbs': step_bytes / (ms_step * 1e-3) / 1e9,
                         'frac_of_measured_peak': step_bytes / (ms_step * 1e-3) / 1e9 / peak,
                         'frac_of_8000_nominal': step_bytes / (ms_step * 1e-3) / 1e9 / 8000.0},
            'roofline': {'kernel': kname, 'bound': 'hbm', 'achieved': ach, 'peak': peak,
                         'unit': 'GB/s', 'frac': ach / peak, 'traffic': traffic, 'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': mask_bytes},
            'e2e': e2e, 'packed_masks_in': packed_in, 'cpu_baseline': cpu, 'clocks': clocks, 'parity_checked_frames': parity_frames,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------- reference arm

def run_reference(args):
    """The reference's CPU implementation of the path on the host cores: every step is one frame per
    host process (a bounded sample of the workload: `cores` frames of the same shape), `--steps` timed
    steps after `--warmup` untimed ones, exactly as asked.  Rank 0 alone works."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    world = int(os.environ.get('WORLD_SIZE', args.gpus))
    steps, warm = max(1, args.steps), max(0, args.warmup)
    c = cpu_reference_throughput(steps=steps, warm=warm, workload=args.workload)
    n_pts = RINGS * AZ
    sample = '%d frames of the workload\'s shape per step (one per host process), %d steps, %.1f s wall; %s' % (
        c['cores'], steps, c['seconds'], CPU_KIND_NOTE[c['kind']])
    line = {'impl': 'reference', 'metric': METRIC, 'value': c['value'], 'unit': 'points/s',
            'n_gpus': world, 'steps': steps, 'warmup': warm,
            'ms_per_step': c['seconds'] / steps * 1e3, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f64 (numpy) + u8 (cv2)', 'data': 'synthetic',
            'config': bench_config(args.workload, FRAMES, n_pts, world),
            'cpu_baseline': {'value': c['value'], 'unit': 'points/s', 'cores': c['cores'], 'kind': c['kind'],
                             'sample': sample},
            'e2e': {'value': c['value'], 'unit': 'points/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--frames', type=int, default=None, help='frames per GPU (default: the workload\'s, cfg3 1024)')
    ap.add_argument('--workload', default='cfg3', choices=sorted(WORKLOADS),
                    help='cfg3 = the metric\'s configuration (default); cfg5 = dense stress, 2M points x 64 masks, 7x7')
    ap.add_argument('--cpu-frames', type=int, default=20,
                    help='frames each host process times for cpu_baseline (~70 ms each: ~1.5 s wall, ~20 core-seconds)')
    ap.add_argument('--e2e-steps', type=int, default=3)
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-parity', action='store_true')
    ap.add_argument('--no-packed', action='store_true', help='skip the packed-masks-in secondary measurement')
    ap.add_argument('--exchange', default='peer', choices=['peer', 'nccl'],
                    help='N > 1: per-car table exchanged by stores into peer memory (default) or by the NCCL all-reduce')
    args = ap.parse_args()
    set_workload(args.workload, args.frames)
    # stdout carries the ONE JSON line and nothing else: whatever libraries print while the job runs
    # (NCCL's version banner goes to stdout when NCCL_DEBUG is set) is sent to stderr at the fd level.
    sys.stdout.flush()
    keep = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(keep, 'w')
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_gpu(args)
    sys.stdout.flush()


if __name__ == '__main__':
    main()
