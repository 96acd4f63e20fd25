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
SPARSE_ROWS = True       # timed steps write only the rows of the packed masks that have a set pixel (--dense-rows: every row)
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
            'packed_rows': ('only the rows of the (internal) packed masks that have a set pixel are written and read '
                            '(lf_erode_pack_rows / lf_fuse_batch_rows)') if SPARSE_ROWS else 'every row written (lf_erode_pack / lf_fuse_batch)',
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


# --------------------------------------------------------------------------- host cores of a rank

def _gpu_numa_node(torch, index):
    """NUMA node of GPU `index` from sysfs (-1 when the platform does not say)."""
    try:
        p = torch.cuda.get_device_properties(index)
        bdf = '%04x:%02x:%02x.0' % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open('/sys/bus/pci/devices/%s/numa_node' % bdf) as fh:
            return int(fh.read().strip())
    except Exception:
        return -1


def _node_cpus(node):
    try:
        with open('/sys/devices/system/node/node%d/cpulist' % node) as fh:
            out = []
            for part in fh.read().strip().split(','):
                a, _, b = part.partition('-')
                out += list(range(int(a), int(b or a) + 1))
            return out
    except Exception:
        return []


def bind_host(torch, local, world):
    """At N > 1 every rank gets a disjoint set of host cores on its GPU's NUMA node (its pack threads
    and the pages of its pinned buffers then live there); at N = 1 the process keeps every core."""
    allowed = sorted(os.sched_getaffinity(0))
    node = _gpu_numa_node(torch, local)
    mine = allowed
    if world > 1:
        nodes = [_gpu_numa_node(torch, r) for r in range(world)]
        peers = [r for r in range(world) if nodes[r] == node]          # ranks that share this node's cores
        pool = [c for c in _node_cpus(node) if c in allowed] if node >= 0 else []
        if len(pool) < len(peers):                                      # no NUMA information: split what we may use
            pool, peers = allowed, list(range(world))
        k = peers.index(local)
        per = max(1, len(pool) // len(peers))
        mine = pool[k * per:(k + 1) * per] or pool
        try:
            os.sched_setaffinity(0, mine)
        except Exception:
            mine = allowed
    return {'threads': max(1, len(mine)), 'shared_host': world > 1,
            'report': {'cores': len(mine), 'first_core': mine[0], 'numa_node': node, 'bound': world > 1}}


# --------------------------------------------------------------------------- side measurements

def _peak_gbs():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as fh:
            return float(json.load(fh)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


def side_cfg4(torch, dist, D, S, world, rank, local, dev, cam, TVC, TVR, sb0, exch, total_frames=65536, pool=8,
              repeats=3):
    """BASELINE.json configs[3]: 65 536 HDL-64 frames sharded over the ranks, 1024 per batch, a pool of
    distinct resident batches per rank walked round-robin (uint8 masks of a whole shard do not fit HBM below
    N = 8), the per-car table of every batch exchanged as in the headline step.  Strong scaling: the job is
    fixed.  Checks: every exchanged table equals the one recorded for that pool batch, row totals add up."""
    FB = 1024
    per_rank = total_frames // world
    nb = per_rank // FB
    if nb * FB * world != total_frames:
        return {'skipped': 'world size does not divide the job into 1024-frame batches'}
    pool = min(pool, nb)
    stream = torch.cuda.current_stream(dev)
    sbs = [sb0] + [S.SyntheticBatch(FB, cam, TVC, TVR, ncars=NCARS, seed=SEED, rings=RINGS, azimuths=AZ, device=local,
                                    first_frame=(world + rank) * FB + (i - 1) * world * FB + 7 * total_frames)
                   for i in range(1, pool)]
    if exch is not None and hasattr(exch, 'bind'):
        for b in sbs:
            exch.bind(b.batch.h_mask_off, b.n_boxes)
    want, rows = [], []
    for b in sbs:
        b.batch.run(KSIZE, 1, 50, 10, True)
        if exch is not None:
            t, _ = exch.exchange_batch(b.batch)
            want.append(t.clone())
            rows.append(exch.agg.clone())
        else:
            want.append(torch.cat([b.batch.total, b.batch.best_idx, b.batch.best_cnt]).clone())
            rows.append(b.batch.agg.clone())
    torch.cuda.synchronize(dev)
    ok = torch.ones((), dtype=torch.int32, device=dev)
    agg_sum = torch.zeros_like(rows[0])

    def job(check):
        for i in range(nb):
            b = sbs[i % pool].batch
            b.run(KSIZE, 1, 50, 10, True, sparse_rows=SPARSE_ROWS)
            if exch is not None:
                t, _ = exch.exchange_batch(b)
                got, agg = t, exch.agg
            else:
                got, agg = None, b.agg
            if check:
                if got is None:
                    got = torch.cat([b.total, b.best_idx, b.best_cnt])
                ok.mul_((got == want[i % pool]).all().to(torch.int32))
                agg_sum.add_(agg)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    job(True)                                              # warm-up pass, with the checks
    sync()
    expect = sum(rows[i % pool] for i in range(nb))
    good = bool(ok.item()) and torch.equal(agg_sum, expect)
    times = []
    for _ in range(repeats):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync()
        e0.record(stream)
        job(False)
        e1.record(stream)
        sync()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    flag = torch.tensor([int(good)], device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if not bool(flag.item()):
        raise SystemExit('PARITY FAILURE: cfg4 job tables / row totals differ')
    ms = sorted(times)[len(times) // 2]
    n_pts = sb0.n
    out = {'value': total_frames * n_pts / (ms * 1e-3), 'unit': 'points/s', 'scaling': 'strong', 'job_ms': ms,
           'job_ms_all': times, 'frames': total_frames, 'batches_per_rank': nb, 'frames_per_batch': FB,
           'distinct_resident_frames_per_rank': pool * FB,
           'workload': 'cfg4: %d frames x %d pts x %d masks (%dx%d uint8) x <=%d boxes over %d GPU(s)' % (
               total_frames, n_pts, NCARS, W_IMG, H_IMG, NCARS, world),
           'checks': {'tables_equal_recorded_and_row_totals_add_up': True, 'rows': int(expect[1].item()),
                      'matched_rows': int(expect[2].item())}}
    del sbs[1:]
    torch.cuda.empty_cache()
    return out


def side_cfg5(torch, dist, D, S, world, rank, local, dev, cam, TVC, TVR, steps):
    """BASELINE.json configs[4]: 64 frames x 2 097 152 points x 64 masks, 7x7 ellipse erosion, per rank."""
    w = WORKLOADS['cfg5']
    sb5 = S.SyntheticBatch(w['frames'], cam, TVC, TVR, ncars=w['ncars'], seed=SEED + 5, rings=w['rings'],
                           azimuths=w['az'], device=local, first_frame=rank * w['frames'])
    b = sb5.batch
    ex = None
    if world > 1:
        ex = D.CounterExchange(world * w['frames'], w['ncars'], rank, world, dev)
        ex.bind(b.h_mask_off, sb5.n_boxes)
    stream = torch.cuda.current_stream(dev)

    def step(mark=None):
        b.run(w['ksize'], 1, 50, 10, True, mark=mark, sparse_rows=SPARSE_ROWS)
        if ex is not None:
            ex.exchange_batch(b)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(3):
        step()
    sync()
    marks = []

    def mk(lst):
        def mark(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            lst.append((name, e))
        return mark
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        lst = []
        step(mk(lst))
        marks.append(dict(lst))
    e1.record(stream)
    sync()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    km = {k2: sum(d[k1].elapsed_time(d[k2]) for d in marks) / len(marks)
          for k1, k2 in (('zero', 'erode_pack'), ('erode_pack', 'fuse'), ('fuse', 'car_stats'))}
    peak, _ = _peak_gbs()
    sbytes = sum(16 * sb5.n + w['ncars'] * H_IMG * W_IMG + 192 * nb + 8 * w['ncars'] * (nb + 3) for nb in sb5.n_boxes)
    fuse_gbs = 16.0 * sb5.n * w['frames'] / (km['fuse'] * 1e-3) / 1e9
    erode_gbs = float(w['ncars']) * H_IMG * W_IMG * w['frames'] / (km['erode_pack'] * 1e-3) / 1e9
    return {'value': world * w['frames'] * sb5.n / (ms * 1e-3), 'unit': 'points/s', 'ms_per_step': ms, 'steps': steps,
            'workload': 'cfg5: %d frames x %d pts x %d masks (%dx%d uint8) x <=%d boxes per GPU, 7x7 ellipse erosion'
                        % (w['frames'], sb5.n, w['ncars'], W_IMG, H_IMG, w['ncars']),
            'kernel_ms': km, 'algorithmic_bytes_per_step': sbytes,
            'step_frac_of_measured_peak': sbytes / (ms * 1e-3) / 1e9 / peak,
            'roofline': {'kernel': 'lf_fuse_kernel (dominant on this workload)', 'bound': 'hbm', 'achieved': fuse_gbs,
                         'peak': peak, 'unit': 'GB/s', 'frac': fuse_gbs / peak,
                         'algorithmic_bytes_per_launch': 16 * sb5.n * w['frames']},
            'erode_frac_of_measured_peak': erode_gbs / peak}


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
    host_bind = None

    # CPU baseline first (rank 0, N = 1 only): spawned host processes, GPU idle meanwhile
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference_throughput(steps=args.cpu_frames, warm=1, workload=args.workload)
        cpu = {'value': c['value'], 'unit': 'points/s', 'cores': c['cores'], 'kind': c['kind'],
               'sample': '%d synthetic frames (%d per host process, %d processes, OPENBLAS/cv2 threads = 1), '
                         '%.1f s wall; %s' % (c['frames'], args.cpu_frames, c['cores'], c['seconds'], CPU_KIND_NOTE[c['kind']])}

    host_bind = bind_host(torch, local, world)
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
        batch.run(KSIZE, 1, 50, 10, True, mark=mark, sparse_rows=SPARSE_ROWS)
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

    # parity spot check against the oracle (checker only): frames of this very batch, spread over it --
    # 32 for HDL-64 frames, 2 for the multi-sweep workload (the oracle needs seconds per such frame);
    # counters, the decision (best box, best count) and the per-car rows
    parity_frames = 0
    if rank == 0 and not args.no_parity:
        from oracle import pipeline as P
        res = batch.results()
        n_check = min(frames, args.parity_frames if n_pts <= (1 << 18) else 2)
        for f in sorted({int(x) for x in np.linspace(0, frames - 1, n_check)}):
            pts, masks, boxes = sb.host_frame(f)
            want = P.process_frame(pts, boxes, masks.astype(np.float32), cam, TVC, TVR, erosion_kernel_size=KSIZE)
            m0, m1 = batch.h_mask_off[f], batch.h_mask_off[f + 1]
            got_t = res['total'][m0:m1]
            got_i = res['inside'][batch.h_cnt_off[f]:batch.h_cnt_off[f + 1]].reshape(want['inside'].shape)
            ok = np.array_equal(got_t, want['total']) and np.array_equal(got_i, want['inside']) \
                and res['n_valid'][f] == want['valid'].sum()
            wi = want['inside']
            w_best = wi.max(axis=1) if wi.size else np.zeros(len(wi), np.int64)
            w_idx = np.where(w_best > 0, wi.argmax(axis=1), -1) if wi.size else -np.ones(len(wi), np.int64)
            ok = ok and np.array_equal(res['best_idx'][m0:m1], w_idx) and np.array_equal(res['best_cnt'][m0:m1], w_best)
            got_rows = [(s['car_id'], s['matched_bbox_id'], s['total_points'], s['points_inside_bbox'])
                        for s in batch.frame_stats(res, f)]
            want_rows = [(s['car_id'], s['matched_bbox_id'], s['total_points'], int(s['points_inside_bbox']))
                         for s in want['stats']]
            if not (ok and got_rows == want_rows):
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

    launches_per_step = int(batch.launches_per_run - 1) + (0 if exch is None else 2)   # kernels of ours (memsets not counted)

    # ---- end to end through the public API: pinned HOST points + masks in, per-car table on the HOST out;
    # the H2D copies, the kernels, the D2H of the counter block and the vectorised table are all inside the
    # timed region.  Three input encodings of the same masks:
    #   f32_packed  float32 host masks (the reference's boundary format, cvs_erosion.py:87, 103 -- what the
    #               CPU arm is fed), binarised + bit-packed on the host (lf_pack_masks_host) for the transfer
    #   u8_packed   uint8 host masks, packed the same way
    #   u8_plain    uint8 host masks copied as they are
    # The headline `e2e` is f32_packed (like for like with the reference arm); the others stay beside it.
    e2e, packed_in = None, None
    from lidar_object_detection_b200 import io as IO, lidarfuse as LF
    words = LF.erode_pack(batch.masks, 0, 0, device=local).words
    pb = LF.FrameBatch(batch.points, [n_pts] * frames, LF.PackedMasks(words, H_IMG, W_IMG, raw=True), [NCARS] * frames,
                       batch.boxes, sb.n_boxes, cam, TVR, device=local)

    def pstep(mark=None):
        pb.run(KSIZE, 1, 50, 10, True, mark=mark)
        if exch is not None:
            exch.exchange_batch(pb)

    def time_e2e(b, stepfn, h_pts, h_masks, n, **kw):
        h_out = b.download()                                  # pinned, allocated once

        def one():
            b.upload(h_pts, h_masks, **kw)
            stepfn()
            b.download(h_out)
            stream.synchronize()
            return b.table(h_out, 10)                         # the rows the caller gets (cvs_erosion.py:197-225)

        tab = one()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(n):
            tab = one()
        sync_all()
        dt = (time.perf_counter() - t0) / n
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt, tab, h_out

    if not args.no_packed:
        for _ in range(3):
            pstep()
        batch.run(KSIZE, 1, 50, 10, True)                     # every row of the packed masks, for the comparison
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
        t_end = time.time()

    if not args.no_e2e:
        import psutil
        n_e2e = max(1, min(args.e2e_steps, K_steps))
        threads = host_bind['threads']
        h_pts = torch.empty(batch.points.shape, dtype=batch.points.dtype, pin_memory=True)
        h_pts.copy_(batch.points)
        want = batch.table(None, 10)
        variants = {}
        d2h = int(batch.counters.numel() * 4)

        def record(name, dt, tab, h2d, note, extra=None):
            for k in want:
                if not np.array_equal(tab[k], want[k]):
                    raise SystemExit('PARITY FAILURE: end-to-end variant %s returns another per-car table (%s)' % (name, k))
            variants[name] = dict({'value': world * frames * n_pts / dt, 'unit': 'points/s', 'ms_per_step': dt * 1e3,
                                   'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': d2h, 'steps': n_e2e,
                                   'rows_per_step': int(len(tab['frame'])), 'api': note}, **(extra or {}))

        # uint8 host masks
        h_u8 = torch.empty(batch.masks.shape, dtype=torch.uint8, pin_memory=True)
        h_u8.copy_(batch.masks)
        torch.cuda.synchronize(dev)
        dt, tab, _ = time_e2e(batch, step, h_pts, h_u8, n_e2e)
        record('u8_plain', dt, tab, h_pts.numel() * 4 + h_u8.numel(),
               'FrameBatch.upload (pinned uint8 masks copied as they are) / run / download / table')
        tmp_words = np.empty(tuple(words.shape), np.uint32)
        pack_u8_gbs = 0.0
        sync_all()
        for _ in range(2):
            tq0 = time.perf_counter()
            IO.pack_masks(h_u8.numpy(), threads=threads, out=tmp_words)
            pack_u8_gbs = max(pack_u8_gbs, h_u8.numel() / (time.perf_counter() - tq0) / 1e9)
        del tmp_words
        dt, tab, _ = time_e2e(pb, pstep, h_pts, h_u8, n_e2e, threads=threads)
        record('u8_packed', dt, tab, h_pts.numel() * 4 + words.numel() * 4,
               'FrameBatch.upload on PackedMasks(raw=True): lf_pack_masks_host in 16 slices overlapped with their H2D / '
               'run / download / table', {'host_pack_threads': threads, 'host_pack_gbs_alone': pack_u8_gbs})
        del h_u8
        # float32 host masks: 4 bytes per pixel of pinned host memory -- only when the host has room for it
        f32_bytes = batch.masks.numel() * 4
        free = psutil.virtual_memory().available
        if not args.no_f32 and free > 2.5 * f32_bytes * max(1, world if host_bind['shared_host'] else 1):
            h_f32 = torch.empty(batch.masks.shape, dtype=torch.float32, pin_memory=True)
            sl = max(1, len(batch.masks) // 64)
            for a0 in range(0, len(batch.masks), sl):          # converted on the device slice by slice, copied back
                h_f32[a0:a0 + sl].copy_(batch.masks[a0:a0 + sl].to(torch.float32))
            torch.cuda.synchronize(dev)
            pack_f32_gbs = 0.0
            sync_all()                                        # every rank packs at the same time: the host is shared
            for _ in range(2):                                # best of two: the first pass touches the pages
                tq0 = time.perf_counter()
                IO.pack_masks(h_f32.numpy(), threads=threads, out=pb._h_words.numpy())
                pack_f32_gbs = max(pack_f32_gbs, f32_bytes / (time.perf_counter() - tq0) / 1e9)
            dt, tab, _ = time_e2e(pb, pstep, h_pts, h_f32, n_e2e, threads=threads)
            host_bytes = f32_bytes + 2 * words.numel() * 4 + h_pts.numel() * 4     # mask read + packed write + DMA reads
            record('f32_packed', dt, tab, h_pts.numel() * 4 + words.numel() * 4,
                   'FrameBatch.upload on PackedMasks(raw=True) from pinned FLOAT32 host masks (the reference\'s boundary '
                   'format): lf_pack_masks_host mode 1 in 16 slices overlapped with their H2D / run / download / table',
                   {'host_pack_threads': threads, 'host_pack_gbs_alone': pack_f32_gbs,
                    'host_mask_bytes_per_step': int(f32_bytes),
                    'host_dram_bytes_per_step': int(host_bytes),
                    'host_dram_floor_ms': host_bytes / (pack_f32_gbs * 1e9) * 1e3,
                    'host_floor_note': 'bytes through host DRAM per step / the read rate the packer reaches alone on '
                                       'this rank\'s cores = the time the host memory system needs whatever the GPU does'})
            del h_f32
        head = 'f32_packed' if 'f32_packed' in variants else 'u8_packed'
        e2e = dict(variants[head], variant=head, input_dtype='float32 masks' if head == 'f32_packed' else 'uint8 masks',
                   host=host_bind['report'], other_variants={k: v for k, v in variants.items() if k != head})
        if packed_in is not None:
            packed_in['e2e'] = variants['u8_packed']
        t_end = time.time()
    del pb

    # ---- BASELINE.json configs[3] and configs[4] beside the headline, measured by this very run with the
    # shipped kernels: cfg4 = the 65 536-frame job over the N ranks (STRONG scaling), cfg5 = dense stress
    cfg4 = None
    if not args.no_cfg4 and args.workload == 'cfg3' and frames == 1024:
        cfg4 = side_cfg4(torch, dist, D, S, world, rank, local, dev, cam, TVC, TVR, sb, exch)
    cfg5 = None
    if not args.no_cfg5 and args.workload == 'cfg3':
        del batch
        sb.batch = None
        torch.cuda.empty_cache()
        cfg5 = side_cfg5(torch, dist, D, S, world, rank, local, dev, cam, TVC, TVR, max(5, min(K_steps, 20)))

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
        # DRAM bytes of one launch of the dominant kernel from the committed ncu capture (not measured by this
        # run: `traffic_source` names the file and the capture it summarises; null when it does not cover the kernel)
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, 'profiles', 'dram_traffic.json')) as fh:
                tj = json.load(fh)
            if args.workload == 'cfg3' and tj.get(tkey + '_dram_bytes_per_launch'):
                traffic = tj[tkey + '_dram_bytes_per_launch']
                traffic_src = 'profiles/dram_traffic.json <- ' + str(tj.get('source'))
        except Exception:
            pass
        line = {
            'metric': METRIC, 'value': value, 'unit': 'points/s', 'n_gpus': world, 'steps': K_steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f64 projection/containment, u32 bit-masks, i32 counters',
            'data': 'synthetic',
            'config': bench_config(args.workload, frames, n_pts, world, step_bytes, exch_kind if world > 1 else None),
            'gpu_launches': launches_per_step * K_steps,
            'kernel_ms': kern_ms,
            'step_hbm': {'algorithmic_bytes_per_step': step_bytes, 'achieved_gbs': step_bytes / (ms_step * 1e-3) / 1e9,
                         'frac_of_measured_peak': step_bytes / (ms_step * 1e-3) / 1e9 / peak,
                         'frac_of_8000_nominal': step_bytes / (ms_step * 1e-3) / 1e9 / 8000.0},
            'roofline': {'kernel': kname, 'bound': 'hbm', 'achieved': ach, 'peak': peak,
                         'unit': 'GB/s', 'frac': ach / peak, 'traffic': traffic, 'traffic_source': traffic_src, 'peak_source': peak_src,
                         'note': ('the denominator is the measured COPY bandwidth (equal reads and writes); with the row map this '
                                  'kernel is an almost pure read stream, which HBM serves faster than a copy, so frac can exceed 1') if SPARSE_ROWS else None,
                         'algorithmic_bytes_per_launch': mask_bytes},
            'e2e': e2e, 'cfg4_job': cfg4, 'cfg5': cfg5, 'packed_masks_in': packed_in, 'cpu_baseline': cpu, 'clocks': clocks, 'parity_checked_frames': parity_frames,
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
    ap.add_argument('--no-cfg4', action='store_true', help='skip the 65 536-frame strong-scaling job (cfg4_job object)')
    ap.add_argument('--no-cfg5', action='store_true', help='skip the dense-stress side measurement (cfg5 object)')
    ap.add_argument('--no-f32', action='store_true', help='skip the float32-host-mask end-to-end variant (34.7 GB pinned)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-parity', action='store_true')
    ap.add_argument('--parity-frames', type=int, default=32, help='frames of the batch checked against the oracle')
    ap.add_argument('--no-packed', action='store_true', help='skip the packed-masks-in secondary measurement')
    ap.add_argument('--dense-rows', action='store_true',
                    help='write every row of the packed masks (lf_erode_pack / lf_fuse_batch) instead of only the rows '
                         'with a set pixel (lf_erode_pack_rows / lf_fuse_batch_rows, the default of the timed steps)')
    ap.add_argument('--exchange', default='peer', choices=['peer', 'nccl'],
                    help='N > 1: per-car table exchanged by stores into peer memory (default) or by the NCCL all-reduce')
    args = ap.parse_args()
    global SPARSE_ROWS
    SPARSE_ROWS = not args.dense_rows
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
