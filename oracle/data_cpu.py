"""ORACLE (test infrastructure only): numpy restatement of the point-cloud part of the reference's training pipeline
(configs/fcaf3d/fcaf3d_scannet-3d-18class.py:16-41):

  IndoorPointSample      mmdet3d/datasets/pipelines/transforms_3d.py:821-891
  RandomFlip3D           transforms_3d.py:59-162  -> DepthInstance3DBoxes.flip  core/bbox/structures/depth_box3d.py:176-208,
                                                     DepthPoints.flip           core/points/depth_points.py:28-33
  GlobalRotScaleTrans    transforms_3d.py:497-648 -> DepthInstance3DBoxes.rotate depth_box3d.py:119-174,
                                                     BasePoints.rotate / scale / translate  core/points/base_points.py:139-269,
                                                     BaseInstance3DBoxes.scale / translate  base_box3d.py:149-222

Boxes are (cx, cy, cz, dx, dy, dz, yaw) rows with the GRAVITY centre (the format the head consumes through
`gt_bboxes.gravity_center` / `.tensor[:, 3:]`, fcaf3d_neck_with_head.py:404-416); every transform here is linear about
the origin, so acting on the gravity centre equals acting on the reference's bottom-centre tensor and converting.
Pinned against the reference's own containers by tests/golden/augment_ref.npz (tests/golden/make_golden.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import numpy as np


def indoor_point_sample(points, num_points, rng):
    """transforms_3d.py:838-863: np.random.choice(N, num_points, replace=N < num_points); returns (points, choices)."""
    n = points.shape[0]
    choices = rng.choice(n, num_points, replace=n < num_points)
    return points[choices], choices


def random_flip(points, boxes, flip_h, flip_v, with_yaw=True):
    """horizontal: x -> -x, yaw -> pi - yaw; vertical: y -> -y, yaw -> -yaw (depth_box3d.py:190-197)."""
    points, boxes = points.copy(), boxes.copy()
    if flip_h:
        points[:, 0] = -points[:, 0]
        boxes[:, 0] = -boxes[:, 0]
        if with_yaw:
            boxes[:, 6] = -boxes[:, 6] + np.float32(np.pi)
    if flip_v:
        points[:, 1] = -points[:, 1]
        boxes[:, 1] = -boxes[:, 1]
        if with_yaw:
            boxes[:, 6] = -boxes[:, 6]
    return points, boxes


def global_rot_scale_trans(points, boxes, angle, scale, trans, with_yaw=True):
    """rotate about z by `angle` (xyz @ rot_mat_T, rot_mat_T = [[c, -s, 0], [s, c, 0], [0, 0, 1]]^T, yaw -= angle; without
    yaw the x / y sizes become the extents of the rotated footprint, depth_box3d.py:148-160), then scale, then translate
    (transforms_3d.py:624-648 order R, S, T)."""
    points, boxes = points.copy(), boxes.copy()
    c, s = np.float32(np.cos(np.float32(angle))), np.float32(np.sin(np.float32(angle)))
    rot_t = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float32).T
    if len(boxes):  # transforms_3d.py:582-586: points are rotated together with a non-empty box field
        centre = boxes[:, :3].copy()
        boxes[:, :3] = centre @ rot_t
        if with_yaw:
            boxes[:, 6] -= np.float32(angle)
        else:
            dx, dy = boxes[:, 3].copy(), boxes[:, 4].copy()
            boxes[:, 3] = np.abs(dx * c) + np.abs(dy * s)
            boxes[:, 4] = np.abs(dx * s) + np.abs(dy * c)
    points[:, :3] = points[:, :3] @ rot_t
    points[:, :3] *= np.float32(scale)
    boxes[:, :6] *= np.float32(scale)
    t = np.asarray(trans, dtype=np.float32)
    points[:, :3] += t
    boxes[:, :3] += t
    return points, boxes


def draw_params(rng, n_in, num_points=100000, flip_ratio_h=0.5, flip_ratio_v=0.5, rot_range=(-0.087266, 0.087266),
                scale_ratio_range=(0.9, 1.1), translation_std=(0.1, 0.1, 0.1)):
    """One scene's random numbers in the order the reference's pipeline consumes them: IndoorPointSample's choice
    (:859), RandomFlip's 2-D direction draw (mmdet 2.14 RandomFlip.__call__, not vendored: one choice over
    [horizontal, None]; its result is not used for the points), the two BEV flip draws (:144-150), the rotation angle
    (:572), the scale (:620) and the translation (:553)."""
    choices = rng.choice(n_in, num_points, replace=n_in < num_points)
    rng.choice(2, p=[flip_ratio_h, 1 - flip_ratio_h])
    flip_h = bool(rng.random() < flip_ratio_h)
    flip_v = bool(rng.random() < flip_ratio_v)
    angle = rng.uniform(rot_range[0], rot_range[1])
    scale = rng.uniform(scale_ratio_range[0], scale_ratio_range[1])
    trans = rng.normal(scale=np.asarray(translation_std, dtype=np.float32), size=3)
    return dict(choices=choices, flip_h=flip_h, flip_v=flip_v, angle=float(angle), scale=float(scale),
                trans=[float(v) for v in trans])
