"""world_size-2 gloo test of the frame sharding + the single counter all-reduce.
Runs on CPU: the counters are synthetic, the exchange logic is the one bench.py uses at N > 1."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lidar_object_detection_b200 import distributed as D


class _FakeBatch:
    """The attributes CounterExchange.exchange_batch reads from a lidarfuse.FrameBatch."""

    def __init__(self, total, best_cnt, best_idx, agg):
        self.total, self.best_cnt, self.best_idx, self.agg = total, best_cnt, best_idx, agg


def _make(rank, world, n_global, max_cars, seed=5):
    rng = np.random.default_rng(seed)                      # every rank draws the same global truth
    n_masks = rng.integers(0, max_cars + 1, n_global)
    n_boxes = rng.integers(0, 4, n_global)
    total = [rng.integers(0, 500, m) for m in n_masks]
    best = [np.minimum(t, rng.integers(0, 500, len(t))) for t in total]
    idx = [np.where(b > 0, rng.integers(0, 3, len(b)), -1) for b in best]
    lo, hi = D.shard_range(n_global, rank, world)
    return n_masks, n_boxes, total, best, idx, lo, hi


def _worker(rank, world, port, n_global, max_cars, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        n_masks, n_boxes, total, best, idx, lo, hi = _make(rank, world, n_global, max_cars)
        cat = lambda xs: torch.from_numpy(np.concatenate([np.asarray(x, np.int32) for x in xs] or [np.zeros(0, np.int32)]))
        mask_off = np.concatenate([[0], np.cumsum(n_masks[lo:hi])])
        rows_local = sum(int(((t > 0) & (nb > 0)).sum()) for t, nb in zip(total[lo:hi], n_boxes[lo:hi]))
        agg = torch.tensor([hi - lo, rows_local, 0, 0, 0, 0, 0, 0], dtype=torch.int64)
        ex = D.CounterExchange(n_global, max_cars, rank, world, torch.device('cpu'))
        ex.bind(mask_off, n_boxes[lo:hi])
        table, agg_sum = ex.exchange_batch(_FakeBatch(cat(total[lo:hi]), cat(best[lo:hi]), cat(idx[lo:hi]), agg))
        # the older two-step interface must agree
        t2 = D.local_table(cat(total[lo:hi]), cat(best[lo:hi]), cat(idx[lo:hi]), n_boxes[lo:hi], mask_off, max_cars)
        ex2 = D.CounterExchange(n_global, max_cars, rank, world, torch.device('cpu'))
        table2, _ = ex2.exchange(t2, agg)
        assert torch.equal(table, table2)
        if rank == 0:
            torch.save({'table': table.clone(), 'agg': agg_sum.clone()}, out)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


@pytest.mark.parametrize('n_global', [7, 16])
def test_counter_exchange_world2(tmp_path, n_global):
    world, max_cars = 2, 6
    out = str(tmp_path / 'r0.pt')
    mp.spawn(_worker, args=(world, _free_port(), n_global, max_cars, out), nprocs=world, join=True)
    got = torch.load(out)
    n_masks, n_boxes, total, best, idx, _, _ = _make(0, 1, n_global, max_cars)
    want = np.zeros((n_global, max_cars, 4), np.int64)
    for f in range(n_global):
        m = n_masks[f]
        want[f, :m, 0] = total[f]; want[f, :m, 1] = best[f]; want[f, :m, 2] = idx[f] + 1
        want[f, :m, 3] = int(n_boxes[f] > 0)
    assert np.array_equal(got['table'].numpy(), want)
    assert int(got['agg'][0]) == n_global                     # every frame counted exactly once (uneven shards too)
    rows = D.rows_from_table(got['table'], list(range(n_global)), min_points=10)
    n_rows = sum(int(((t > 0) & (nb > 0)).sum()) for t, nb in zip(total, n_boxes))
    assert len(rows) == n_rows == int(got['agg'][1])
    for fr, c, bi, tot, ins, outs in rows:
        assert tot == total[fr][c] and ins + outs == tot
        assert (bi >= 0) == (idx[fr][c] >= 0 and best[fr][c] >= 10)


def test_shard_ranges_cover_everything():
    for n in (1, 7, 1024, 65536):
        for w in (1, 2, 4, 8):
            r = [D.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))
