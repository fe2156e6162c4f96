"""Runs the reference's OWN FCAF3D model files, unmodified, on the fcaf3d_b200 operator surfaces.

The three files
    mmdet3d/models/backbones/me_resnet.py
    mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py
    mmdet3d/models/detectors/single_stage_sparse.py
cannot be imported as `mmdet3d.*` without mmcv / mmdet / mmseg / MinkowskiEngine and every compiled op of mmdet3d
(mmdet3d/__init__.py:1-4, mmdet3d/ops/__init__.py:1-24).  What they actually use is small:

    MinkowskiEngine                       -> fcaf3d_b200.me                                   (operator surface 1)
    mmdet3d.ops.pcdet_nms                 -> fcaf3d_b200.pcdet_nms                            (operator surface 2)
    build_loss(dict(type=...))            -> fcaf3d_b200.losses.{FocalLoss, CrossEntropyLoss, IoU3DLoss}   (surface 3)
    mmcv.cnn.Scale / bias_init_with_prob, mmdet registries, BaseAssigner, reduce_mean, build_assigner,
    rotation_3d_in_axis (core/bbox/structures/utils.py:21-61), bbox3d2result (core/bbox/transforms.py:49-74),
    Base3DDetector                        -> the stand-ins below (pinned upstream behaviour: mmcv-full 1.3.8, mmdet 2.14.0)

`load_reference_models(root)` imports the three files BY PATH from a reference checkout (`root`/mmdet3d/...) after
registering those stand-ins in `sys.modules`; nothing of the reference is copied into this package.  Everything below runs
on CUDA tensors only - there is no CPU path behind it.
"""
import importlib.util
import math
import os
import sys
import types

import torch
import torch.nn as nn

from . import losses as L
from . import pcdet_nms as NMS

MODEL_FILES = (
    ("mmdet3d.models.backbones.me_resnet", "mmdet3d/models/backbones/me_resnet.py"),
    ("mmdet3d.models.dense_heads.fcaf3d_neck_with_head", "mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py"),
    ("mmdet3d.models.detectors.single_stage_sparse", "mmdet3d/models/detectors/single_stage_sparse.py"),
)


class Registry:
    """mmcv.utils.Registry, the part the three files touch: `@X.register_module()` and build-by-`type`."""

    def __init__(self, name):
        self.name, self.module_dict = name, {}

    def register_module(self, name=None, force=False, module=None):
        def deco(cls):
            self.module_dict[name or cls.__name__] = cls
            return cls
        return deco(module) if module is not None else deco

    def build(self, cfg, **extra):
        cfg = dict(cfg)
        cls = self.module_dict[cfg.pop("type")]
        cfg.update(extra)
        return cls(**cfg)


class Scale(nn.Module):
    """mmcv.cnn.Scale"""

    def __init__(self, scale=1.0):
        super().__init__()
        self.scale = nn.Parameter(torch.tensor(scale, dtype=torch.float))

    def forward(self, x):
        return x * self.scale


def bias_init_with_prob(prior_prob):
    """mmcv.cnn.bias_init_with_prob"""
    return float(-math.log((1 - prior_prob) / prior_prob))


def reduce_mean(tensor):
    """mmdet.core.reduce_mean: all-reduce mean over the data-parallel group, identity on one process"""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tensor
    tensor = tensor.clone()
    dist.all_reduce(tensor.div_(dist.get_world_size()), op=dist.ReduceOp.SUM)
    return tensor


class BaseAssigner:
    """mmdet.core.BaseAssigner (an interface only)"""


def rotation_3d_in_axis(points, angles, axis=0):
    """points [N, M, 3] rotated by angles [N] about `axis` (core/bbox/structures/utils.py:21-61): row-vector convention,
    out = points @ R_T with R_T = [[c, -s, 0], [s, c, 0], [0, 0, 1]] for axis 2."""
    s, c = torch.sin(angles), torch.cos(angles)
    o, z = torch.ones_like(c), torch.zeros_like(c)
    if axis in (2, -1):
        rows = ((c, -s, z), (s, c, z), (z, z, o))
    elif axis == 1:
        rows = ((c, z, -s), (z, o, z), (s, z, c))
    elif axis == 0:
        rows = ((z, c, -s), (z, s, c), (o, z, z))
    else:
        raise ValueError(f"axis should in range [0, 1, 2], got {axis}")
    rot_t = torch.stack([torch.stack(r) for r in rows])                # [3, 3, N]
    return torch.einsum("aij,jka->aik", (points, rot_t))


class DepthInstance3DBoxes:
    """The slice of mmdet3d's DepthInstance3DBoxes the FCAF3D head touches (base_box3d.py:36-70, depth_box3d.py:41-48):
    `tensor` [n, 7] with a bottom-centre origin, `volume`, `gravity_center`, `to`, `len`."""

    def __init__(self, tensor, box_dim=7, with_yaw=True, origin=(0.5, 0.5, 0)):
        tensor = torch.as_tensor(tensor, dtype=torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape((0, box_dim))
        if tensor.shape[-1] == 6:
            tensor = torch.cat((tensor, tensor.new_zeros(tensor.shape[0], 1)), dim=-1)
            with_yaw = False
        self.with_yaw, self.box_dim = with_yaw, 7
        self.tensor = tensor.clone()
        if tuple(origin) != (0.5, 0.5, 0):
            shift = self.tensor.new_tensor((0.5, 0.5, 0)) - self.tensor.new_tensor(origin)
            self.tensor[:, :3] += self.tensor[:, 3:6] * shift

    @property
    def volume(self):
        return self.tensor[:, 3] * self.tensor[:, 4] * self.tensor[:, 5]

    @property
    def gravity_center(self):
        gc = self.tensor[:, :3].clone()
        gc[:, 2] += self.tensor[:, 5] * 0.5
        return gc

    def to(self, device):
        out = DepthInstance3DBoxes(self.tensor.to(device), with_yaw=self.with_yaw)
        return out

    def __len__(self):
        return self.tensor.shape[0]


def bbox3d2result(bboxes, scores, labels, attrs=None):
    """core/bbox/transforms.py:49-74"""
    out = dict(boxes_3d=bboxes.to("cpu"), scores_3d=scores.cpu(), labels_3d=labels.cpu())
    if attrs is not None:
        out["attrs_3d"] = attrs.cpu()
    return out


class _IoU3DLoss(L.IoU3DLoss):
    """registry flavour: the reference hands 7-wide targets to a 6-wide axis-aligned prediction (iou3d_loss.py:21-35 reads
    columns 0..5 only)"""

    def forward(self, pred, target, weight=None, avg_factor=None, reduction_override=None, **kw):
        if target.shape[-1] != pred.shape[-1]:
            target = target[..., :pred.shape[-1]]
        return super().forward(pred, target, weight=weight, avg_factor=avg_factor,
                               reduction_override=reduction_override, **kw)


BACKBONES, HEADS, DETECTORS, LOSSES, NECKS = (Registry(n) for n in ("backbone", "head", "detector", "loss", "neck"))
BBOX_ASSIGNERS, IOU_CALCULATORS = Registry("bbox_assigner"), Registry("iou_calculator")
LOSSES.register_module(module=L.FocalLoss)
LOSSES.register_module(module=L.CrossEntropyLoss)
LOSSES.register_module(name="IoU3DLoss", module=_IoU3DLoss)


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []
    sys.modules[name] = m
    return m


_STAND_INS = ("mmcv", "mmcv.cnn", "mmdet", "mmdet.models", "mmdet.models.builder", "mmdet.core", "mmdet.core.bbox",
              "mmdet.core.bbox.builder", "mmdet3d", "mmdet3d.core", "mmdet3d.core.bbox", "mmdet3d.core.bbox.structures",
              "mmdet3d.ops", "mmdet3d.ops.pcdet_nms", "mmdet3d.models", "mmdet3d.models.backbones",
              "mmdet3d.models.dense_heads", "mmdet3d.models.detectors", "mmdet3d.models.detectors.base")


def install():
    """Registers `MinkowskiEngine` and the stand-in modules in sys.modules.  Returns the names it registered."""
    from . import install_as_minkowski_engine
    install_as_minkowski_engine()
    _module("mmcv")
    _module("mmcv.cnn", Scale=Scale, bias_init_with_prob=bias_init_with_prob)
    _module("mmdet")
    _module("mmdet.models", BACKBONES=BACKBONES, DETECTORS=DETECTORS, NECKS=NECKS, HEADS=HEADS, LOSSES=LOSSES,
            build_loss=LOSSES.build)
    _module("mmdet.models.builder", HEADS=HEADS, LOSSES=LOSSES, BACKBONES=BACKBONES, DETECTORS=DETECTORS,
            build_loss=LOSSES.build)
    _module("mmdet.core", BaseAssigner=BaseAssigner, reduce_mean=reduce_mean, build_assigner=BBOX_ASSIGNERS.build)
    _module("mmdet.core.bbox")
    _module("mmdet.core.bbox.builder", BBOX_ASSIGNERS=BBOX_ASSIGNERS)
    _module("mmdet3d")
    _module("mmdet3d.core", bbox3d2result=bbox3d2result)
    _module("mmdet3d.core.bbox")
    _module("mmdet3d.core.bbox.structures", rotation_3d_in_axis=rotation_3d_in_axis,
            DepthInstance3DBoxes=DepthInstance3DBoxes)
    _module("mmdet3d.ops")
    _module("mmdet3d.ops.pcdet_nms", pcdet_nms_gpu=NMS.pcdet_nms_gpu, pcdet_nms_normal_gpu=NMS.pcdet_nms_normal_gpu)
    _module("mmdet3d.models", build_backbone=BACKBONES.build, build_head=HEADS.build)
    _module("mmdet3d.models.backbones")
    _module("mmdet3d.models.dense_heads")
    _module("mmdet3d.models.detectors")
    _module("mmdet3d.models.detectors.base", Base3DDetector=nn.Module)
    return _STAND_INS


def uninstall():
    for name in _STAND_INS + tuple(n for n, _ in MODEL_FILES) + (
            "MinkowskiEngine", "MinkowskiEngine.utils", "MinkowskiEngine.modules", "MinkowskiEngine.modules.resnet_block"):
        sys.modules.pop(name, None)


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def reference_files_present(root):
    return all(os.path.isfile(os.path.join(root, rel)) for _, rel in MODEL_FILES)


def load_reference_models(root):
    """Imports the three model files from the reference checkout at `root`, unmodified, on this package's surfaces.
    Returns a namespace: MEResNet3D, Fcaf3DNeckWithHead, Fcaf3DAssigner, SingleStageSparse3DDetector."""
    if not reference_files_present(root):
        raise FileNotFoundError("reference model files not found under %s" % root)
    install()
    mods = [_load(name, os.path.join(root, rel)) for name, rel in MODEL_FILES]
    backbone, head, det = mods
    return types.SimpleNamespace(MEResNet3D=backbone.MEResNet3D, Fcaf3DNeckWithHead=head.Fcaf3DNeckWithHead,
                                 Fcaf3DAssigner=head.Fcaf3DAssigner,
                                 SingleStageSparse3DDetector=det.SingleStageSparse3DDetector)


def reference_detector(ns, dataset="scannet", **overrides):
    """The reference's detector class built with the hyper-parameters of configs/fcaf3d/fcaf3d.py:3-28 and the per-dataset
    overrides (same numbers as detector.build_fcaf3d, which tests/test_host_logic.py checks against the config files)."""
    from .detector import fcaf3d_config
    cfg = fcaf3d_config(dataset, **overrides)
    cfg["backbone"]["type"] = "MEResNet3D"
    cfg["neck_with_head"]["type"] = "Fcaf3DNeckWithHead"
    cfg["test_cfg"] = types.SimpleNamespace(**cfg["test_cfg"])        # the head reads test_cfg.nms_pre etc. as attributes
    return ns.SingleStageSparse3DDetector(**cfg)
