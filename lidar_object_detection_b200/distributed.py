"""Frame sharding over one process per GPU.

Frames are independent, so the data path has no collective: rank r of G owns
the contiguous frame block [r*F/G, (r+1)*F/G).  Only the small per-car counter
tables are combined, with ONE ``all_reduce(SUM)`` per batch over a zero-filled
integer table in which each rank writes only its own rows (a sum over disjoint
supports is an all-gather that tolerates uneven shards), plus the int64
``analyze_master_csv`` aggregates.  Backend: NCCL over NVLink on GPUs, gloo in
the CPU tests.
"""
import numpy as np
import torch
import torch.distributed as dist

TABLE_COLS = 4          # total_points, best_count, best_idx + 1, frame has boxes (a CSV row exists)


def shard_range(n_frames, rank, world):
    """Contiguous block of frames owned by ``rank``."""
    lo = (n_frames * rank) // world
    hi = (n_frames * (rank + 1)) // world
    return lo, hi


def local_table(total, best_cnt, best_idx, box_counts, mask_off, max_cars):
    """Per-car table of the local frames: int32 [F_local, max_cars, 4] (torch ops
    on whatever device the counters live on)."""
    F = len(mask_off) - 1
    dev = total.device
    table = torch.zeros((F, max_cars, TABLE_COLS), dtype=torch.int32, device=dev)
    if total.numel() == 0:
        return table
    mo = torch.as_tensor(mask_off, device=dev, dtype=torch.int64)
    counts = mo[1:] - mo[:-1]
    frame_of = torch.repeat_interleave(torch.arange(F, device=dev), counts)
    car_of = torch.arange(total.numel(), device=dev) - mo[frame_of]
    table[frame_of, car_of, 0] = total.to(torch.int32)
    table[frame_of, car_of, 1] = best_cnt.to(torch.int32)
    table[frame_of, car_of, 2] = best_idx.to(torch.int32) + 1
    has_boxes = (torch.as_tensor(box_counts, device=dev) > 0).to(torch.int32)
    table[frame_of, car_of, 3] = has_boxes[frame_of]
    return table


class CounterExchange:
    """Pre-allocated global table + aggregate vector and the single all-reduce."""

    def __init__(self, n_frames_global, max_cars, rank, world, device):
        self.rank, self.world = rank, world
        self.lo, self.hi = shard_range(n_frames_global, rank, world)
        self.max_cars = max_cars
        n = n_frames_global * max_cars * TABLE_COLS
        # one flat int64 buffer: [table as int64 | agg(8)] -> a single collective
        self.flat = torch.zeros(n + 8, dtype=torch.int64, device=device)
        self.table = self.flat[:n].view(n_frames_global, max_cars, TABLE_COLS)
        self.agg = self.flat[n:]

    def bind(self, mask_off, box_counts):
        """Pre-compute where the local cars land in the global table (once per batch layout)."""
        dev = self.flat.device
        mo = torch.as_tensor(np.asarray(mask_off), device=dev, dtype=torch.int64)
        F = mo.numel() - 1
        counts = mo[1:] - mo[:-1]
        if F != self.hi - self.lo:
            raise ValueError('layout has %d frames, this rank\'s shard of the table has %d' % (F, self.hi - self.lo))
        if F and int(counts.max()) > self.max_cars:
            raise ValueError('a frame has %d masks, the table has %d car slots per frame' % (int(counts.max()), self.max_cars))
        frame_of = torch.repeat_interleave(torch.arange(F, device=dev), counts)
        car_of = torch.arange(int(mo[-1]), device=dev) - mo[frame_of]
        self._dest = ((self.lo + frame_of) * self.max_cars + car_of) * TABLE_COLS
        self._has_boxes = (torch.as_tensor(np.asarray(box_counts), device=dev) > 0).to(torch.int64)[frame_of]
        self._n_cars = int(mo[-1])

    def exchange_batch(self, batch):
        """Scatter the local counters of a lidarfuse.FrameBatch into the table and
        all-reduce it: five small device ops and ONE collective per batch."""
        n = self._n_cars
        self.flat.zero_()
        self.flat[self._dest] = batch.total[:n].to(torch.int64)
        self.flat[self._dest + 1] = batch.best_cnt[:n].to(torch.int64)
        self.flat[self._dest + 2] = batch.best_idx[:n].to(torch.int64) + 1
        self.flat[self._dest + 3] = self._has_boxes
        self.agg.copy_(batch.agg)
        if self.world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        return self.table, self.agg

    def exchange(self, table_local, agg_local):
        """Fill this rank's rows, sum across ranks (async on the current stream
        for NCCL).  Afterwards every rank holds the full table."""
        self.flat.zero_()
        self.table[self.lo:self.hi] = table_local.to(torch.int64)
        self.agg.copy_(agg_local.to(torch.int64))
        if self.world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        return self.table, self.agg


def _check_batch(batch, n_frames, max_cars):
    """A batch must be this rank's shard and fit the table's car slots (rows would be dropped or mixed)."""
    if batch.F != n_frames:
        raise ValueError('batch holds %d frames, this rank\'s shard of the table has %d' % (batch.F, n_frames))
    if batch.max_m > max_cars:
        raise ValueError('a frame has %d masks, the table has %d car slots per frame' % (batch.max_m, max_cars))


class PeerCounterExchange:
    """The same exchange without a reduction: ``lf_publish_table`` stores this rank's rows and
    aggregates straight into every rank's copy of the table (NVLink peer stores into torch
    symmetric memory), then ONE signal-pad barrier orders the ranks -- two launches per batch
    instead of zero-fill + scatter + all-reduce.  Tables are int32 [F_global, max_cars, 4] and
    double-buffered: the table ``exchange_batch`` returns is valid for work ENQUEUED (on the same stream)
    before the next ``exchange_batch`` -- a peer that has passed barrier i+1 may already be storing step i+2
    into the half that holds step i, so clone it if it must outlive the next call.
    Needs one node with peer access (NVLink / NVSwitch); ``available()`` tells."""

    @staticmethod
    def available():
        try:
            import torch.distributed._symmetric_memory  # noqa: F401
        except Exception:
            return False
        return torch.cuda.is_available() and dist.is_initialized()

    def __init__(self, n_frames_global, max_cars, rank, world, device, group=None):
        import torch.distributed._symmetric_memory as symm
        self.rank, self.world, self.max_cars = rank, world, max_cars
        self.lo, self.hi = shard_range(n_frames_global, rank, world)
        self.n_frames = n_frames_global
        self._tbytes = n_frames_global * max_cars * TABLE_COLS * 4
        self._tbytes = (self._tbytes + 15) // 16 * 16
        self._half = self._tbytes + world * 8 * 8                      # table + [world][8] int64 aggregates
        self._half = (self._half + 15) // 16 * 16
        self.buf = symm.empty(2 * self._half // 8, dtype=torch.int64, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self._peers = int(self.hdl.buffer_ptrs_dev)
        self._step = 0
        torch.cuda.synchronize(device)
        self.hdl.barrier(channel=0)                                    # every copy is zeroed before the first store
        self.table = self.agg_slots = None

    def _views(self, half):
        base = half * self._half // 8
        t = self.buf[base:base + self._tbytes // 8].view(torch.int32)[:self.n_frames * self.max_cars * TABLE_COLS]
        a = self.buf[base + self._tbytes // 8:base + self._tbytes // 8 + self.world * 8]
        return t.view(self.n_frames, self.max_cars, TABLE_COLS), a.view(self.world, 8)

    def exchange_batch(self, batch):
        """Publish the counters of a lidarfuse.FrameBatch (after ``run``) to every rank; returns
        (table, agg_slots): the full int32 table and the aggregates per rank, int64 [world, 8]
        (``.agg`` sums them)."""
        from ._abi import check, ptr
        ctx = batch.ctx
        _check_batch(batch, self.hi - self.lo, self.max_cars)
        half = self._step & 1
        self._step += 1
        off = half * self._half
        check(ctx.lib.lf_publish_table(ctx.h, ptr(batch.total), ptr(batch.best_idx), ptr(batch.best_cnt),
                                       ptr(batch.mask_off), ptr(batch.box_off), ptr(batch.agg), batch.F, self.max_cars,
                                       self.lo, self._peers, self.world, self.rank, off, off + self._tbytes, ctx.stream),
              ctx.h)
        self.hdl.barrier(channel=0)
        self.table, self.agg_slots = self._views(half)
        return self.table, self.agg_slots

    @property
    def agg(self):
        """The eight aggregates summed over the ranks (int64 [8]); ``agg_slots`` holds them per rank."""
        return self.agg_slots.sum(dim=0)

    def bind(self, mask_off, box_counts):
        """Nothing to pre-compute (kept for interface parity with CounterExchange)."""


def rows_from_table(table, frame_ids, min_points=10):
    """Global per-car rows (frame, car_id, matched_bbox_id, total, inside, outside)
    from the reduced table -- what rank 0 turns into CSV lines."""
    t = table.cpu().numpy() if isinstance(table, torch.Tensor) else np.asarray(table)
    rows = []
    for fi, frame in enumerate(frame_ids):
        for c in range(t.shape[1]):
            total, best, idx1 = int(t[fi, c, 0]), int(t[fi, c, 1]), int(t[fi, c, 2])
            if total == 0 or t[fi, c, 3] == 0:          # no points, or no boxes (cvs_erosion.py:169-170, 178-179)
                continue
            if idx1 > 0 and best >= min_points:
                rows.append((frame, c, idx1 - 1, total, best, total - best))
            else:
                rows.append((frame, c, -1, total, 0, total))
    return rows
