"""Deterministic synthetic indoor scenes: ScanNet / SUN RGB-D / S3DIS shaped point clouds.

Surfaces = floor + 4 walls + `n_boxes` furniture boxes (5 visible faces each); points are sampled
area-proportionally with 3 mm Gaussian noise; RGB ~ U[0, 255].  Returns fp32 [n, 6] points and GT boxes.
"""
import numpy as np

PRESETS = {
    # name: (room L, W, H, n_boxes, n_points, single_view, rotated)
    "S0": (6.0, 5.0, 2.7, 15, 100000, False, False),   # ScanNet
    "S1": (9.0, 7.0, 2.7, 30, 100000, False, False),   # ScanNet-large
    "S2": (14.0, 10.0, 3.0, 60, 1000000, False, False),  # S3DIS 1M stress shape
    "S3": (5.0, 5.0, 2.7, 10, 100000, True, True),     # SUN RGB-D single view, rotated GT
    "tiny": (2.0, 1.6, 1.2, 3, 3000, False, False),    # unit tests
}


def _sample_rect(rng, n, origin, u, v):
    a = rng.random((n, 1))
    b = rng.random((n, 1))
    return origin[None, :] + a * u[None, :] + b * v[None, :]


def make_scene(preset="S0", seed=0, n_points=None, n_classes=18):
    L, W, H, n_boxes, npts, single_view, rotated = PRESETS[preset]
    npts = int(n_points or npts)
    rng = np.random.default_rng(seed)
    rects = []  # (origin, u, v)
    e = np.eye(3)
    rects.append((np.array([0., 0., 0.]), L * e[0], W * e[1]))          # floor
    rects.append((np.array([0., 0., 0.]), L * e[0], H * e[2]))          # wall y=0
    rects.append((np.array([0., W, 0.]), L * e[0], H * e[2]))           # wall y=W
    rects.append((np.array([0., 0., 0.]), W * e[1], H * e[2]))          # wall x=0
    rects.append((np.array([L, 0., 0.]), W * e[1], H * e[2]))           # wall x=L
    gt = []
    for i in range(n_boxes):
        size = rng.uniform([.4, .4, .4], [2., 1.2, 1.6])
        size = np.minimum(size, [L * .45, W * .45, H * .8])
        c = np.array([rng.uniform(size[0] / 2, L - size[0] / 2), rng.uniform(size[1] / 2, W - size[1] / 2), size[2] / 2])
        yaw = rng.uniform(-np.pi / 2, np.pi / 2) if rotated else 0.0
        R = np.array([[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]])
        hx, hy, hz = size / 2
        ax, ay, az = R[:, 0], R[:, 1], R[:, 2]
        o = c - hx * ax - hy * ay - hz * az
        rects.append((o + 2 * hz * az, 2 * hx * ax, 2 * hy * ay))       # top
        rects.append((o, 2 * hx * ax, 2 * hz * az))
        rects.append((o + 2 * hy * ay, 2 * hx * ax, 2 * hz * az))
        rects.append((o, 2 * hy * ay, 2 * hz * az))
        rects.append((o + 2 * hx * ax, 2 * hy * ay, 2 * hz * az))
        gt.append(np.concatenate([c, size, [yaw]]))
    areas = np.array([np.linalg.norm(np.cross(u, v)) for _, u, v in rects])
    counts = rng.multinomial(npts, areas / areas.sum())
    pts = np.concatenate([_sample_rect(rng, k, o, u, v) for (o, u, v), k in zip(rects, counts) if k > 0])
    pts = pts + rng.normal(0, 0.003, pts.shape)
    if single_view:
        keep = pts[:, 0] > L * 0.02
        pts = pts[keep]
        extra = npts - len(pts)
        if extra > 0:  # IndoorPointSample semantics: sample with replacement up to n_points (duplicates reach the hash)
            pts = np.concatenate([pts, pts[rng.integers(0, len(pts), extra)]])
    pts[:, 0] -= L / 2
    pts[:, 1] -= W / 2
    gt = np.stack(gt).astype(np.float32)
    gt[:, 0] -= L / 2
    gt[:, 1] -= W / 2
    rgb = rng.uniform(0, 255, (len(pts), 3))
    perm = rng.permutation(len(pts))
    points = np.concatenate([pts, rgb], axis=1)[perm].astype(np.float32)
    labels = (np.arange(n_boxes) % n_classes).astype(np.int64)
    return points, gt, labels


def make_batch(preset="S0", batch=8, seed=0, n_points=None, n_classes=18):
    """Scenes with seeds seed, seed+100, ...  -> list of (points, gt_boxes, gt_labels)."""
    return [make_scene(preset, seed + 100 * i, n_points, n_classes) for i in range(batch)]


def random_boxes(n, seed=0, rotated=True, room=(6.0, 5.0, 2.7)):
    """Standalone NMS workload: centres U(room), sizes U(.3, 2), yaw U(-pi, pi) or 0, scores U(0, 1)."""
    rng = np.random.default_rng(seed)
    ctr = rng.uniform([-room[0] / 2, -room[1] / 2, 0], [room[0] / 2, room[1] / 2, room[2]], (n, 3))
    size = rng.uniform(.3, 2., (n, 3))
    yaw = rng.uniform(-np.pi, np.pi, (n, 1)) if rotated else np.zeros((n, 1))
    boxes = np.concatenate([ctr, size, yaw], 1).astype(np.float32)
    scores = rng.permutation(n).astype(np.float32) / n + rng.uniform(0, .5 / n, n).astype(np.float32)
    return boxes, scores.astype(np.float32)
