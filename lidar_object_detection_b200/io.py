"""On-disk / wire formats either side of the hot path.

* ``SweepStreamer``: Velodyne ``%010d.bin`` sweeps (cvs_erosion.py:24-28:
  ``np.fromfile(float32).reshape(-1, 4)``) read by a background thread into pinned,
  double-buffered host memory and copied to the GPU on a side stream while the previous
  batch is being processed;
* ``pack_masks``: host-side binarise + bit-pack (``lf_pack_masks_host``, multi-threaded C++) for
  sending masks to the device at one bit per pixel;
* ``save_packed_masks`` / ``load_packed_masks``: instance masks as bit-packed words
  (``uint32 [M, H, ceil(W/32)]``, LSB = lowest x: the layout ``lf_erode_pack`` writes and
  ``lf_fuse_batch`` reads) with their confidence order in one ``.npz``: 1/8 byte per pixel;
* ``format_csv_rows`` / ``append_rows_to_master_csv``: the rows of
  ``master_car_statistics.csv`` for a whole batch formatted by one C call
  (``lf_format_csv_rows``), byte-identical with ``append_to_master_csv``
  (cvs_erosion.py:232-265).
"""
import ctypes as C
import os
import threading

import numpy as np

from . import _abi


# --------------------------------------------------------------------------
# sweeps
# --------------------------------------------------------------------------

def sweep_points(path):
    """Number of points in a ``.bin`` sweep (16 bytes per point)."""
    size = os.path.getsize(path)
    if size % 16:
        raise ValueError('%s: size %d is not a multiple of 16 bytes' % (path, size))
    return size // 16


def read_sweeps_into(paths, out, threads=1):
    """Read the sweeps ``paths`` back to back into ``out`` (float32 [>= sum N, 4], e.g. a
    view of a pinned tensor).  Returns the per-sweep point counts.  ``threads`` > 1 reads
    the files concurrently (``readinto`` releases the GIL; one thread copies ~8 GB/s out of
    the page cache)."""
    counts = [sweep_points(p) for p in paths]
    flat = out.reshape(-1)
    if sum(counts) * 4 > flat.shape[0]:
        raise ValueError('staging buffer too small')
    starts = np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.int64) if counts else []

    def one(i):
        o, n = int(starts[i]), counts[i]
        with open(paths[i], 'rb') as fh:
            got = fh.readinto(memoryview(flat[o * 4:(o + n) * 4]).cast('B'))
        if got != n * 16:
            raise IOError('%s: short read' % paths[i])

    if threads > 1 and len(paths) > 1:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(int(threads), len(paths))) as ex:
            list(ex.map(one, range(len(paths))))
    else:
        for i in range(len(paths)):
            one(i)
    return counts


class SweepStreamer:
    """Iterate over batches of sweeps as device tensors, overlapping disk reads, the
    host-to-device copy and the caller's kernels.

        for points, counts, paths in SweepStreamer(files, frames_per_batch=64):
            ...   # points: cuda float32 [sum N, 4]; valid until the next-but-one iteration

    Two pinned staging buffers and two device buffers: while batch i is being consumed,
    batch i+1 is read from disk by a worker thread and copied on a side stream."""

    def __init__(self, paths, frames_per_batch=64, device=None, max_points_per_frame=None, reader_threads=4):
        import torch
        if not torch.cuda.is_available():
            raise _abi.LidarFuseError('SweepStreamer needs a CUDA device (no CPU fallback)')
        self.torch = torch
        self.paths = list(paths)
        self.fpb = int(frames_per_batch)
        self.reader_threads = int(reader_threads)
        self.device = torch.device('cuda', torch.cuda.current_device() if device is None else device)
        counts = [sweep_points(p) for p in self.paths]
        self.counts = counts
        cap = max([sum(counts[i:i + self.fpb]) for i in range(0, len(counts), self.fpb)] or [1])
        if max_points_per_frame:
            cap = max(cap, int(max_points_per_frame) * self.fpb)
        self.host = [torch.empty((cap, 4), dtype=torch.float32, pin_memory=True) for _ in range(2)]
        self.dev = [torch.empty((cap, 4), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(self.device)
        self.done = [torch.cuda.Event(), torch.cuda.Event()]          # H2D of slot finished
        self.free = [torch.cuda.Event(), torch.cuda.Event()]          # consumer finished with slot

    def _stage(self, b, slot, result):
        try:
            paths = self.paths[b * self.fpb:(b + 1) * self.fpb]
            result['counts'] = read_sweeps_into(paths, self.host[slot].numpy(), self.reader_threads)
            result['paths'] = paths
        except Exception as e:                                        # surfaced in the consumer thread
            result['error'] = e

    def __len__(self):
        return (len(self.paths) + self.fpb - 1) // self.fpb

    def __iter__(self):
        torch = self.torch
        nb = len(self)
        if nb == 0:
            return
        pending = {}
        res = {}
        t = threading.Thread(target=self._stage, args=(0, 0, res))
        t.start()
        pending[0] = (t, res)
        used = [False, False]
        for b in range(nb):
            slot = b & 1
            t, res = pending.pop(b)
            t.join()
            if 'error' in res:
                raise res['error']
            n = sum(res['counts'])
            with torch.cuda.stream(self.copy_stream):
                if used[slot]:
                    self.copy_stream.wait_event(self.free[slot])      # the consumer of batch b-2 is done
                self.dev[slot][:n].copy_(self.host[slot][:n], non_blocking=True)
                self.done[slot].record(self.copy_stream)
            if b + 1 < nb:                                            # read the next batch meanwhile
                if b >= 1:
                    self.done[1 - slot].synchronize()                 # its staging buffer has been copied out
                res2 = {}
                t2 = threading.Thread(target=self._stage, args=(b + 1, 1 - slot, res2))
                t2.start()
                pending[b + 1] = (t2, res2)
            torch.cuda.current_stream(self.device).wait_event(self.done[slot])
            used[slot] = True
            yield self.dev[slot][:n], res['counts'], res['paths']
            self.free[slot].record(torch.cuda.current_stream(self.device))


# --------------------------------------------------------------------------
# packed mask container
# --------------------------------------------------------------------------

def save_packed_masks(path, words, H, W, conf=None, bbox=None):
    """``words``: uint32/int32 [M, H, ceil(W/32)] (numpy or tensor; ``PackedMasks.words``)."""
    w = words.cpu().numpy() if hasattr(words, 'cpu') else np.asarray(words)
    w = np.ascontiguousarray(w).view(np.uint32)
    M = w.shape[0]
    if w.ndim != 3 or w.shape[1:] != (H, (W + 31) // 32):
        raise ValueError('words must be [M, %d, %d]' % (H, (W + 31) // 32))
    conf = np.zeros(M, np.float32) if conf is None else np.asarray(conf, np.float32)
    if conf.shape != (M,):
        raise ValueError('conf must have one entry per mask')
    extra = {}
    if bbox is not None:
        extra['bbox'] = (bbox.cpu().numpy() if hasattr(bbox, 'cpu') else np.asarray(bbox)).astype(np.int32)
    np.savez_compressed(path, words=w, H=np.int32(H), W=np.int32(W), conf=conf,
                        order=np.argsort(-conf, kind='stable').astype(np.int32), **extra)


def load_packed_masks(path, device=None):
    """-> (lidarfuse.PackedMasks on the device, conf float32 [M])."""
    import torch
    from . import lidarfuse as LF
    with np.load(path) as z:
        w, H, W, conf = z['words'], int(z['H']), int(z['W']), z['conf']
        bbox = z['bbox'] if 'bbox' in z.files else None
    ctx = LF.get_ctx(device)
    words = torch.from_numpy(np.ascontiguousarray(w).view(np.int32)).to(ctx.device)
    tb = torch.from_numpy(np.ascontiguousarray(bbox)).to(ctx.device) if bbox is not None else None
    return LF.PackedMasks(words, H, W, tb), conf


def pack_masks(masks, float_threshold='ge_one', threads=0, out=None):
    """Binarise + bit-pack masks on the host with ``lf_pack_masks_host`` (multi-threaded C++; a
    transfer encoding: 1/8 of the uint8 bytes, 1/32 of the float32 bytes cross PCIe).  masks:
    numpy or CPU tensor [M,H,W], uint8 / bool (non-zero = set) or float32 (``>= 1.0``, or
    ``> 0.5`` with ``float_threshold='gt_half'``).  ``out``: optional uint32/int32 [M,H,ceil(W/32)]
    destination (e.g. ``pinned_tensor.numpy()``).  Returns the packed array; erosion is not done
    here -- hand the words to ``lidarfuse.PackedMasks(..., raw=True)``."""
    m = masks.numpy() if hasattr(masks, 'numpy') and not isinstance(masks, np.ndarray) else np.asarray(masks)
    if m.ndim != 3:
        raise ValueError('masks must be [M,H,W]')
    if m.dtype == np.bool_:
        m = m.view(np.uint8)
    if m.dtype == np.uint8:
        mode = _abi.MODE_U8
    elif m.dtype == np.float32:
        mode = _abi.MODE_F32_GT_HALF if float_threshold == 'gt_half' else _abi.MODE_F32_GE1
    else:
        raise ValueError('masks must be uint8, bool or float32, got %s' % m.dtype)
    m = np.ascontiguousarray(m)
    M, H, W = m.shape
    shape = (M, H, (W + 31) // 32)
    if out is None:
        out = np.empty(shape, dtype=np.uint32)
    elif tuple(out.shape) != shape or out.dtype.itemsize != 4 or not out.flags['C_CONTIGUOUS']:
        raise ValueError('out must be a C-contiguous 32-bit array of shape %s' % (shape,))
    _abi.check(_abi.load().lf_pack_masks_host(m.ctypes.data, mode, M, H, W, out.ctypes.data, int(threads)))
    return out


def pack_masks_host(masks):
    """numpy reference packer (uint8/bool/float [M,H,W], non-zero = set) -> uint32 [M,H,ceil(W/32)];
    the same layout the device writes (used by tests and offline converters)."""
    m = np.asarray(masks) != 0
    M, H, W = m.shape
    wpr = (W + 31) // 32
    pad = np.zeros((M, H, wpr * 32), dtype=bool)
    pad[:, :, :W] = m
    bits = pad.reshape(M, H, wpr, 32).astype(np.uint64)
    return (bits << np.arange(32, dtype=np.uint64)).sum(axis=-1).astype(np.uint32)


# --------------------------------------------------------------------------
# CSV
# --------------------------------------------------------------------------

def format_csv_rows(rows, timestamp, with_header=False):
    """``rows``: iterable of (frame, car_id, matched_bbox_id, total, inside[, outside]) as
    ``distributed.rows_from_table`` returns them -> bytes of the CSV lines."""
    if isinstance(rows, np.ndarray):
        a = np.ascontiguousarray(rows[:, :5], dtype=np.int64)
    else:
        rows = list(rows)
        a = np.array([r[:5] for r in rows], dtype=np.int64).reshape(len(rows), 5)
    n = len(a)
    frame = np.ascontiguousarray(a[:, 0])
    car = np.ascontiguousarray(a[:, 1].astype(np.int32))
    bid = np.ascontiguousarray(a[:, 2].astype(np.int32))
    total = np.ascontiguousarray(a[:, 3])
    inside = np.ascontiguousarray(a[:, 4])
    lib = _abi.load()
    ts = str(timestamp).encode()
    need = lib.lf_format_csv_rows(frame.ctypes.data, car.ctypes.data, bid.ctypes.data, total.ctypes.data,
                                  inside.ctypes.data, n, ts, int(with_header), None, 0)
    if need < 0:
        _abi.check(int(need))
    buf = C.create_string_buffer(int(need) + 1)
    lib.lf_format_csv_rows(frame.ctypes.data, car.ctypes.data, bid.ctypes.data, total.ctypes.data,
                           inside.ctypes.data, n, ts, int(with_header), buf, int(need))
    return buf.raw[:need]


def append_rows_to_master_csv(rows, master_csv_path, timestamp=None):
    """Batched ``append_to_master_csv`` (cvs_erosion.py:232-265): header only when the file
    is created, one write for all rows."""
    import datetime
    rows = list(rows)
    if not rows:
        return
    d = os.path.dirname(master_csv_path)
    if d:
        os.makedirs(d, exist_ok=True)
    new = not os.path.exists(master_csv_path)
    ts = timestamp if timestamp is not None else datetime.datetime.now().isoformat()
    with open(master_csv_path, 'ab') as fh:
        fh.write(format_csv_rows(rows, ts, with_header=new))
