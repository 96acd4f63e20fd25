#!/usr/bin/env python
"""BASELINE.json configs[3] (cfg4) as one job: 65 536 synthetic HDL-64 frames (131 072 points, 16 masks
1408x376, <= 16 boxes each) sharded over the N GPUs of one box, 1024 frames per batch, the per-car table of
every batch exchanged between the ranks (peer stores, or --exchange nccl).  STRONG scaling: the job is fixed.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/cfg4_job.py [--frames 65536] [--pool 8] [--exchange peer|nccl]

A rank owns 65 536 / N frames = that many / 1024 batches.  uint8 masks for 8 192 frames are 69 GB, so a rank
keeps a pool of `--pool` distinct resident batches (8 = 90 GB of the 180 GB; at N = 8 every frame of the
shard is distinct and resident) and walks its batches over the pool; nothing is generated inside the timed
region.  One JSON line on stdout (rank 0): job time on the device (max over ranks), points/s, and the
size-independent checks: every exchanged table equals the one the NCCL all-reduce gives for the same batch,
and the job's row / matched-row totals equal the per-batch totals times the number of visits."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=65536)
    ap.add_argument('--pool', type=int, default=8)
    ap.add_argument('--exchange', default='peer', choices=['peer', 'nccl'])
    ap.add_argument('--repeats', type=int, default=3)
    args = ap.parse_args()
    keep = os.dup(1)
    os.dup2(2, 1)                                          # library banners go to stderr
    out = os.fdopen(keep, 'w')
    import torch
    import torch.distributed as dist
    from lidar_object_detection_b200 import distributed as D, kitti360 as K, synthetic as S
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', 0), ('WORLD_SIZE', 1), ('LOCAL_RANK', 0)))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    FB, NCARS = 1024, 16
    per_rank = args.frames // world
    nb = per_rank // FB
    assert nb * FB * world == args.frames, 'frames must be a multiple of 1024 * world'
    pool = min(args.pool, nb)
    cam, TVC, TVR = K.calibration_chain(os.path.join(ROOT, 'tests', 'golden', 'kitti360_sample'))
    first = rank * per_rank
    sbs = [S.SyntheticBatch(FB, cam, TVC, TVR, ncars=NCARS, device=local, first_frame=first + i * FB) for i in range(pool)]
    nccl = D.CounterExchange(world * FB, NCARS, rank, world, dev)
    exch = D.PeerCounterExchange(world * FB, NCARS, rank, world, dev) if args.exchange == 'peer' else nccl
    for sb in sbs:
        nccl.bind(sb.batch.h_mask_off, sb.n_boxes)         # same layout for every batch (16 masks per frame)
    stream = torch.cuda.current_stream(dev)
    # reference tables (NCCL all-reduce) and row totals of every pool batch, outside the timed region
    want, rows = [], []
    for sb in sbs:
        sb.batch.run(3, 1, 50, 10, True)
        t, a = nccl.exchange_batch(sb.batch)
        want.append(t.to(torch.int32).clone())
        rows.append(a.clone())
    torch.cuda.synchronize(dev)
    ok = torch.ones((), dtype=torch.int32, device=dev)
    agg_sum = torch.zeros(8, dtype=torch.int64, device=dev)

    def job(check):
        for i in range(nb):
            sb = sbs[i % pool]
            sb.batch.run(3, 1, 50, 10, True, sparse_rows=True)
            table, _ = exch.exchange_batch(sb.batch)
            if check:
                ok.mul_((table.to(torch.int32) == want[i % pool]).all().to(torch.int32))
                agg_sum.add_(exch.agg)

    job(True)                                              # warm-up pass, with the checks
    dist.barrier()
    torch.cuda.synchronize(dev)
    expect = sum(rows[i % pool] for i in range(nb))
    good = bool(ok.item()) and torch.equal(agg_sum, expect)
    times = []
    for _ in range(args.repeats):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize(dev)
        e0.record(stream)
        job(False)
        e1.record(stream)
        dist.barrier()
        torch.cuda.synchronize(dev)
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    flag = torch.tensor([int(good)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        ms = sorted(times)[len(times) // 2]
        n_pts = sbs[0].n
        out.write(json.dumps({
            'metric': 'points/sec (project+mask+bbox eval)', 'value': args.frames * n_pts / (ms * 1e-3), 'unit': 'points/s',
            'n_gpus': world, 'scaling': 'strong', 'job_ms': ms, 'job_ms_all': times, 'higher_is_better': True,
            'config': {'workload': 'cfg4: %d frames x %d pts x %d masks (1408x376 uint8) x <=16 boxes over %d GPU(s), '
                                   '%d batches of %d frames per rank over a pool of %d resident batches'
                                   % (args.frames, n_pts, NCARS, world, nb, FB, pool),
                       'exchange': args.exchange, 'distinct_frames_per_rank': pool * FB},
            'checks': {'tables_equal_nccl_all_reduce_and_row_totals_add_up': bool(flag.item()),
                       'rows': int(expect[1].item()), 'matched_rows': int(expect[2].item())},
            'data': 'synthetic'}) + '\n')
        out.flush()
    dist.destroy_process_group()
    if not bool(flag.item()):
        raise SystemExit('cfg4: exchanged tables / totals differ')


if __name__ == '__main__':
    main()
