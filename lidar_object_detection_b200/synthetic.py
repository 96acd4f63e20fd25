"""Synthetic HDL-64-shaped frames for the benchmark.

A frame = 64 rings x 2048 azimuth steps = 131 072 rays cast from the sensor
against a ground plane and ``ncars`` car-sized oriented boxes placed inside the
cam0 frustum; rays that hit nothing become 5-80 m clutter.  Masks are the
image-plane silhouettes of the (slightly inflated) cars, i.e. what a perfect
segmenter with a little bleed-out would return; boxes are the cars' 8 corners
in cam0 coordinates, pushed through the reference's own box preparation
(``filter_visible_bboxes`` + ``transform_bboxes_to_velodyne``).

Two generators with the same geometry: CUDA kernels (lf_synth_points /
lf_synth_masks) for device-resident 10 GB batches, and a numpy one for the
CPU reference arm.  They are inputs, not results; they need not agree bit for
bit with each other (parity tests copy the device batch to the host).
"""
import numpy as np

RINGS, AZIMUTHS = 64, 2048
GROUND_Z = -1.73
INFLATE = 1.06

# local corner order of KITTI360_sample/ReadMe.md:9-21 (bottom 0-1-2-3, top 4-5-6-7);
# the containment test uses the edges 0->1, 0->3, 0->4 (cvs_erosion.py:121-123)
_SIGNS = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1],
                   [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=np.float64)


def make_cars(F, ncars=16, seed=0x5EED0000):
    """[F,ncars,8] f64 = (cx, cy, cz, yaw, l, w, h, 0) in the velodyne frame."""
    rng = np.random.default_rng(seed)
    cars = np.zeros((F, ncars, 8))
    cars[..., 0] = rng.uniform(4.0, 45.0, (F, ncars))
    cars[..., 1] = rng.uniform(-12.0, 12.0, (F, ncars))
    cars[..., 3] = rng.uniform(0.0, 2 * np.pi, (F, ncars))
    scale = rng.uniform(0.8, 1.2, (F, ncars, 3))
    cars[..., 4:7] = np.array([4.2, 1.8, 1.5]) * scale
    cars[..., 2] = GROUND_Z + 0.5 * cars[..., 6]
    return cars


def car_corners_velo(cars):
    """[...,8] car rows -> [...,8,3] corners in the velodyne frame."""
    half = 0.5 * cars[..., None, 4:7] * _SIGNS                      # [...,8,3] local
    c, s = np.cos(cars[..., 3])[..., None], np.sin(cars[..., 3])[..., None]
    x = c * half[..., 0] - s * half[..., 1] + cars[..., None, 0]
    y = s * half[..., 0] + c * half[..., 1] + cars[..., None, 1]
    z = half[..., 2] + cars[..., None, 2]
    return np.stack([x, y, z], axis=-1)


def sort_cars_by_image_area(cars, camera, TrVeloToRect):
    """Order each frame's cars by the area of their projected bounding rectangle,
    largest first -- the stand-in for YOLO's confidence order (cvs_erosion.py:91-94)."""
    corners = car_corners_velo(cars)                                 # [F,C,8,3]
    homo = np.concatenate([corners, np.ones(corners.shape[:-1] + (1,))], axis=-1)
    rect = homo @ TrVeloToRect.T
    z = np.maximum(rect[..., 2], 0.1)
    u = camera.K[0, 0] * rect[..., 0] / z + camera.K[0, 2]
    v = camera.K[1, 1] * rect[..., 1] / z + camera.K[1, 2]
    u = np.clip(u, 0, camera.width)
    v = np.clip(v, 0, camera.height)
    area = (u.max(-1) - u.min(-1)) * (v.max(-1) - v.min(-1))
    order = np.argsort(-area, axis=1, kind='stable')
    return np.take_along_axis(cars, order[..., None], axis=1)


def reference_boxes(cars, camera, TrVeloToCam):
    """Per-frame lists of box dicts, prepared the way the reference prepares
    them: corners_cam0 -> visibility filter (cvs_erosion.py:55-74) -> corners_velo
    (cvs_erosion.py:41-52).  Vectorised over all frames."""
    F, C = cars.shape[:2]
    cv = car_corners_velo(cars).reshape(-1, 3)
    homo = np.hstack([cv, np.ones((len(cv), 1))])
    cam0 = np.matmul(TrVeloToCam, homo.T).T[:, :3]                   # [F*C*8,3]
    u, v, depth = camera.cam2image(np.ascontiguousarray(cam0.T))
    ok = ((depth > 0.1) & (u >= 0) & (u < camera.width) & (v >= 0) & (v < camera.height)).reshape(F, C, 8)
    keep = ok.sum(-1) >= 2
    back = np.matmul(np.linalg.inv(TrVeloToCam), np.hstack([cam0, np.ones((len(cam0), 1))]).T).T[:, :3]
    cam0 = cam0.reshape(F, C, 8, 3)
    back = back.reshape(F, C, 8, 3)
    out = []
    for f in range(F):
        out.append([{'index': int(c), 'corners_cam0': cam0[f, c].tolist(), 'corners_velo': back[f, c].tolist()}
                    for c in range(C) if keep[f, c]])
    return out


# ------------------------------------------------------------------ numpy generator (CPU arm)

def _ray_cars(cars_f, o, d, inflate):
    """Entry distance of rays (o [3] or [n,3], d [n,3]) into each car: [n,C], inf = miss."""
    c, s = np.cos(cars_f[:, 3]), np.sin(cars_f[:, 3])
    rel = o[..., None, :] - cars_f[None, :, :3] if o.ndim == 2 else (o[None, None, :] - cars_f[None, :, :3])
    lo = np.stack([c * rel[..., 0] + s * rel[..., 1], -s * rel[..., 0] + c * rel[..., 1], rel[..., 2]], -1)
    dd = d[:, None, :]
    ld = np.stack([c * dd[..., 0] + s * dd[..., 1], -s * dd[..., 0] + c * dd[..., 1],
                   np.broadcast_to(dd[..., 2], (len(d), len(cars_f)))], -1)
    hx = 0.5 * cars_f[None, :, 4:7] * inflate
    with np.errstate(divide='ignore', invalid='ignore'):
        inv = 1.0 / ld
        ta, tb = (-hx - lo) * inv, (hx - lo) * inv
    t_near, t_far = np.minimum(ta, tb), np.maximum(ta, tb)
    par = np.abs(ld) < 1e-9
    t_near = np.where(par, np.where(np.abs(lo) > hx, np.inf, -np.inf), t_near)
    t_far = np.where(par, np.where(np.abs(lo) > hx, -np.inf, np.inf), t_far)
    t0 = np.maximum(t_near.max(-1), 0.0)
    t1 = t_far.min(-1)
    return np.where(t0 <= t1, t0, np.inf)


def numpy_frame(cars_f, seed, rings=RINGS, azimuths=AZIMUTHS):
    """One frame of points [rings*azimuths,4] f32 (ring-major, like KITTI)."""
    rng = np.random.default_rng(seed)
    n = rings * azimuths
    ring, az = np.divmod(np.arange(n), azimuths)
    el = np.deg2rad(2.0 - 26.8 * ring / max(rings - 1, 1))
    a = np.deg2rad(az * (360.0 / azimuths) + (rng.random(n) - 0.5) * 0.1 - 180.0)
    d = np.stack([np.cos(el) * np.cos(a), np.cos(el) * np.sin(a), np.sin(el)], -1)
    t = np.full(n, np.inf)
    down = d[:, 2] < -1e-4
    t[down] = GROUND_Z / d[down, 2]
    if len(cars_f):
        for i0 in range(0, n, 16384):                        # bounded temporaries
            tc = _ray_cars(cars_f, np.zeros(3), d[i0:i0 + 16384], 1.0)
            t[i0:i0 + 16384] = np.minimum(t[i0:i0 + 16384], np.where(tc > 0, tc, np.inf).min(-1))
    clutter = 5.0 + 75.0 * rng.random(n)
    t = np.where(t > 120.0, clutter, t) + rng.normal(0.0, 0.02, n)
    pts = np.empty((n, 4), np.float32)
    pts[:, :3] = (t[:, None] * d).astype(np.float32)
    pts[:, 3] = rng.random(n).astype(np.float32)
    return pts


def numpy_masks(cars_f, camera, TrVeloToRect, inflate=INFLATE):
    """uint8 [C,H,W] silhouettes of the inflated cars."""
    H, W = camera.height, camera.width
    Tinv = np.linalg.inv(TrVeloToRect)
    out = np.zeros((len(cars_f), H, W), np.uint8)
    xs = (np.arange(W) - camera.K[0, 2]) / camera.K[0, 0]
    for y0 in range(0, H, 16):                               # bounded temporaries
        y1 = min(H, y0 + 16)
        ys = (np.arange(y0, y1) - camera.K[1, 2]) / camera.K[1, 1]
        r = np.stack([np.tile(xs, y1 - y0), np.repeat(ys, W), np.ones((y1 - y0) * W)], -1)
        t = _ray_cars(cars_f, Tinv[:3, 3], r @ Tinv[:3, :3].T, inflate)
        out[:, y0:y1, :] = np.isfinite(t).T.reshape(len(cars_f), y1 - y0, W)
    return out


# ------------------------------------------------------------------ device generator

class SyntheticBatch:
    """F synthetic frames resident on one GPU, as a ``lidarfuse.FrameBatch``."""

    def __init__(self, F, camera, TrVeloToCam, TrVeloToRect, ncars=16, seed=0x5EED0000, rings=RINGS,
                 azimuths=AZIMUTHS, device=None, first_frame=0):
        import torch
        from . import lidarfuse as LF
        from ._abi import check, ptr
        ctx = LF.get_ctx(device)
        ctx.set_camera(TrVeloToRect, camera.K, camera.width, camera.height)
        self.F, self.ncars, self.n = F, ncars, rings * azimuths
        self.seed = seed + first_frame
        cars = sort_cars_by_image_area(make_cars(F, ncars, self.seed), camera, TrVeloToRect)
        self.cars = cars
        self.bboxes = reference_boxes(cars, camera, TrVeloToCam)
        d_cars = torch.from_numpy(np.ascontiguousarray(cars)).to(ctx.device)
        self.points = torch.empty((F * self.n, 4), dtype=torch.float32, device=ctx.device)
        self.masks = torch.empty((F * ncars, camera.height, camera.width), dtype=torch.uint8, device=ctx.device)
        CH = 4096                                   # frames per launch (gridDim.y limit 65535)
        Tinv = np.ascontiguousarray(np.linalg.inv(TrVeloToRect))
        for f0 in range(0, F, CH):
            f1 = min(F, f0 + CH)
            check(ctx.lib.lf_synth_points(ctx.h, ptr(d_cars[f0:f1]), ncars, f1 - f0, rings, azimuths,
                                          int(self.seed + f0), ptr(self.points[f0 * self.n:]), ctx.stream), ctx.h)
            check(ctx.lib.lf_synth_masks(ctx.h, ptr(d_cars[f0:f1]), ncars, f1 - f0, Tinv.ctypes.data, INFLATE,
                                         ptr(self.masks[f0 * ncars:]), ctx.stream), ctx.h)
        torch.cuda.synchronize(ctx.device)
        corners = np.array([b['corners_velo'] for fr in self.bboxes for b in fr], dtype=np.float64).reshape(-1, 8, 3)
        self.n_boxes = [len(fr) for fr in self.bboxes]
        self.batch = LF.FrameBatch(self.points, [self.n] * F, self.masks, [ncars] * F, corners, self.n_boxes,
                                   camera, TrVeloToRect, device)

    def host_frame(self, f):
        """(points [n,4] f32, masks uint8 [C,H,W], box dicts) of frame f on the host."""
        pts = self.points[f * self.n:(f + 1) * self.n].cpu().numpy()
        masks = self.masks[f * self.ncars:(f + 1) * self.ncars].cpu().numpy()
        return pts, masks, self.bboxes[f]
