"""ORACLE (test infrastructure only; works only where /root/reference exists, i.e. in the build container).

Loads the reference's OWN model / loss files by path, unmodified, on top of the CPU oracle:
  mmdet3d/models/backbones/me_resnet.py, dense_heads/fcaf3d_neck_with_head.py, detectors/single_stage_sparse.py,
  models/losses/iou3d_loss.py, core/bbox/iou_calculators/iou3d_calculator.py, core/bbox/structures/utils.py,
  ops/rotated_iou/{oriented_iou_loss,box_intersection_2d}.py
`import mmdet3d` itself is impossible here (mmcv / mmdet / mmseg / MinkowskiEngine are absent), so the handful of
third-party symbols those files import are provided as stand-ins that restate the pinned upstream behaviour
(mmcv-full 1.3.8, mmdet 2.14.0, Rotated_IoU@3bdca6b).  tests/golden/make_golden.py uses
this to produce the committed fixtures; nothing on the GPU box imports it.
"""
import importlib.util
import math
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import me_cpu, nms_cpu

REF = os.environ.get("FCAF3D_REFERENCE", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF, "mmdet3d"))


class Registry:
    def __init__(self, name):
        self.name, self.modules = name, {}

    def register_module(self, name=None, force=False, module=None):
        def deco(cls):
            self.modules[name or cls.__name__] = cls
            return cls
        return deco(module) if module is not None else deco

    def build(self, cfg, **extra):
        cfg = dict(cfg)
        cls = self.modules[cfg.pop("type")]
        cfg.update(extra)
        return cls(**cfg)


BACKBONES, HEADS, DETECTORS, LOSSES = Registry("backbone"), Registry("head"), Registry("detector"), Registry("loss")
BBOX_ASSIGNERS, IOU_CALCULATORS, NECKS = Registry("assigner"), Registry("iou"), Registry("neck")


class Scale(nn.Module):  # mmcv.cnn.Scale
    def __init__(self, scale=1.0):
        super().__init__()
        self.scale = nn.Parameter(torch.tensor(scale, dtype=torch.float))

    def forward(self, x):
        return x * self.scale


def bias_init_with_prob(p):  # mmcv.cnn.bias_init_with_prob
    return float(-math.log((1 - p) / p))


def reduce_mean(t):  # mmdet.core.reduce_mean, single process
    return t


class BaseAssigner:
    pass


def weighted_loss(fn):  # mmdet.models.losses.utils.weighted_loss (2.14.0)
    def wrapper(pred, target, weight=None, reduction="mean", avg_factor=None, **kw):
        loss = fn(pred, target, **kw)
        if weight is not None:
            loss = loss * weight
        if avg_factor is None:
            return loss.mean() if reduction == "mean" else loss.sum() if reduction == "sum" else loss
        if reduction == "mean":
            return loss.sum() / avg_factor
        if reduction == "none":
            return loss
        raise ValueError('avg_factor can not be used with reduction="sum"')
    return wrapper


def sigmoid_focal_loss_ref(pred, labels, gamma=2.0, alpha=0.25):
    """mmcv.ops.sigmoid_focal_loss forward (1.3.8 CUDA kernel semantics, App. C): label -1 matches no class."""
    p = torch.sigmoid(pred)
    onehot = torch.zeros_like(pred)
    valid = labels >= 0
    onehot[valid.nonzero().squeeze(1), labels[valid]] = 1
    flt_min = torch.finfo(torch.float32).tiny
    pos = -alpha * (1 - p) ** gamma * torch.log(p.clamp(min=flt_min))
    neg = -(1 - alpha) * p ** gamma * torch.log((1 - p).clamp(min=flt_min))
    return onehot * pos + (1 - onehot) * neg


@LOSSES.register_module()
class FocalLoss(nn.Module):  # mmdet.models.losses.FocalLoss(use_sigmoid=True)
    def __init__(self, use_sigmoid=True, gamma=2.0, alpha=0.25, reduction="mean", loss_weight=1.0):
        super().__init__()
        self.gamma, self.alpha, self.loss_weight = gamma, alpha, loss_weight

    def forward(self, pred, target, weight=None, avg_factor=None, reduction_override=None):
        loss = sigmoid_focal_loss_ref(pred, target, self.gamma, self.alpha)
        return self.loss_weight * (loss.sum() / avg_factor if avg_factor is not None else loss.mean())


@LOSSES.register_module()
class CrossEntropyLoss(nn.Module):  # mmdet 2.14 CrossEntropyLoss(use_sigmoid=True) -> binary_cross_entropy + weight_reduce_loss
    def __init__(self, use_sigmoid=False, use_mask=False, reduction="mean", class_weight=None, loss_weight=1.0):
        super().__init__()
        assert use_sigmoid
        self.loss_weight, self.reduction = loss_weight, reduction

    def forward(self, cls_score, label, weight=None, avg_factor=None, reduction_override=None, **kw):
        reduction = reduction_override if reduction_override else self.reduction
        loss = F.binary_cross_entropy_with_logits(cls_score, label.float(), reduction="none")
        if weight is not None:
            loss = loss * weight.float()
        if avg_factor is None:
            loss = loss.mean() if reduction == "mean" else loss.sum() if reduction == "sum" else loss
        elif reduction == "mean":
            loss = loss.sum() / avg_factor
        elif reduction != "none":
            raise ValueError('avg_factor can not be used with reduction="sum"')
        return self.loss_weight * loss


def sort_v(vertices, mask, num_valid):
    """torch restatement of Rotated_IoU@3bdca6b cuda_op.sort_v (App. C): indices of the valid vertices in
    counter-clockwise angular order, first index repeated, padded with an invalid intersection slot."""
    B, N, M, _ = vertices.shape
    ang = torch.atan2(vertices[..., 1], vertices[..., 0])
    ang = torch.where(mask, ang, torch.full_like(ang, 1e9))
    order = torch.argsort(ang, dim=2, stable=True)  # valid first, ascending angle
    # padding index: first invalid slot among the 16 intersections (index >= 8)
    inter_invalid = (~mask[..., 8:]).float()
    pad = 8 + torch.argmax(inter_invalid, dim=2)
    out = pad.unsqueeze(-1).repeat(1, 1, 9)
    nv = num_valid.long()
    ar = torch.arange(9).view(1, 1, 9)
    take = ar < nv.unsqueeze(-1)
    out = torch.where(take & (nv.unsqueeze(-1) >= 3), order[..., :9], out)
    close = (ar == nv.unsqueeze(-1)) & (nv.unsqueeze(-1) >= 3)
    out = torch.where(close, order[..., :1].expand(-1, -1, 9), out)
    return out.int()


class DepthBoxes:
    """Minimal DepthInstance3DBoxes: core/bbox/structures/base_box3d.py:36-70, depth_box3d.py:41-48."""

    def __init__(self, tensor, box_dim=7, with_yaw=True, origin=(0.5, 0.5, 0)):
        tensor = torch.as_tensor(tensor, dtype=torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape((0, box_dim))
        if tensor.shape[-1] == 6:
            tensor = torch.cat((tensor, tensor.new_zeros(tensor.shape[0], 1)), dim=-1)
            with_yaw = False
        self.with_yaw = with_yaw
        self.tensor = tensor.clone()
        if origin != (0.5, 0.5, 0):
            dst = self.tensor.new_tensor((0.5, 0.5, 0))
            self.tensor[:, :3] += self.tensor[:, 3:6] * (dst - self.tensor.new_tensor(origin))

    @property
    def volume(self):
        return self.tensor[:, 3] * self.tensor[:, 4] * self.tensor[:, 5]

    @property
    def gravity_center(self):
        gc = self.tensor[:, :3].clone()
        gc[:, 2] = gc[:, 2] + self.tensor[:, 5] * 0.5
        return gc

    def to(self, device):
        return self

    def __len__(self):
        return self.tensor.shape[0]


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _load(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


_loaded = None


def load():
    """Returns a namespace with the reference classes: MEResNet3D, Fcaf3DNeckWithHead, Fcaf3DAssigner,
    SingleStageSparse3DDetector, IoU3DLoss, cal_iou_3d, axis_aligned_bbox_overlaps_3d, rotation_3d_in_axis."""
    global _loaded
    if _loaded is not None:
        return _loaded
    assert available(), "reference tree not found at %s" % REF
    me_cpu.install_as_minkowski_engine()
    build_loss = LOSSES.build
    build_assigner = BBOX_ASSIGNERS.build
    _mod("mmcv")
    _mod("mmcv.cnn", Scale=Scale, bias_init_with_prob=bias_init_with_prob)
    _mod("mmdet")
    _mod("mmdet.models", BACKBONES=BACKBONES, DETECTORS=DETECTORS, NECKS=NECKS, build_loss=build_loss)
    _mod("mmdet.models.builder", HEADS=HEADS, LOSSES=LOSSES, build_loss=build_loss)
    _mod("mmdet.models.losses")
    _mod("mmdet.models.losses.utils", weighted_loss=weighted_loss)
    _mod("mmdet.core", BaseAssigner=BaseAssigner, reduce_mean=reduce_mean, build_assigner=build_assigner)
    _mod("mmdet.core.bbox", bbox_overlaps=None)
    _mod("mmdet.core.bbox.builder", BBOX_ASSIGNERS=BBOX_ASSIGNERS)
    _mod("mmdet.core.bbox.iou_calculators")
    _mod("mmdet.core.bbox.iou_calculators.builder", IOU_CALCULATORS=IOU_CALCULATORS)
    _mod("mmdet3d")
    _mod("mmdet3d.core", bbox3d2result=lambda b, s, l: dict(boxes_3d=b, scores_3d=s, labels_3d=l))
    _mod("mmdet3d.core.bbox")
    utils = _load("mmdet3d.core.bbox.structures.utils", "mmdet3d/core/bbox/structures/utils.py")
    _mod("mmdet3d.core.bbox.structures", rotation_3d_in_axis=utils.rotation_3d_in_axis, get_box_type=None,
         utils=utils)
    _mod("mmdet3d.core.bbox.iou_calculators")
    iouc = _load("mmdet3d.core.bbox.iou_calculators.iou3d_calculator",
                 "mmdet3d/core/bbox/iou_calculators/iou3d_calculator.py")
    sys.modules["mmdet3d.core.bbox"].AxisAlignedBboxOverlaps3D = iouc.AxisAlignedBboxOverlaps3D
    _mod("mmdet3d.ops")
    _mod("mmdet3d.ops.pcdet_nms", pcdet_nms_gpu=nms_cpu.pcdet_nms_gpu, pcdet_nms_normal_gpu=nms_cpu.pcdet_nms_normal_gpu)
    _mod("mmdet3d.ops.rotated_iou")
    _mod("mmdet3d.ops.rotated_iou.cuda_op")
    _mod("mmdet3d.ops.rotated_iou.cuda_op.cuda_ext", sort_v=sort_v)
    _load("mmdet3d.ops.rotated_iou.box_intersection_2d", "mmdet3d/ops/rotated_iou/box_intersection_2d.py")
    # min_enclosing_box.py serves GIoU/DIoU only (out of scope) and needs numpy<1.24 (np.int): stand-in
    _mod("mmdet3d.ops.rotated_iou.min_enclosing_box", smallest_bounding_box=None)
    riou = _load("mmdet3d.ops.rotated_iou.oriented_iou_loss", "mmdet3d/ops/rotated_iou/oriented_iou_loss.py")
    sys.modules["mmdet3d.ops.rotated_iou"].cal_iou_3d = riou.cal_iou_3d
    sys.modules["mmdet3d.ops.rotated_iou"].cal_giou_3d = riou.cal_giou_3d
    losses = _load("mmdet3d.models.losses.iou3d_loss", "mmdet3d/models/losses/iou3d_loss.py")
    backbone = _load("mmdet3d.models.backbones.me_resnet", "mmdet3d/models/backbones/me_resnet.py")
    head = _load("mmdet3d.models.dense_heads.fcaf3d_neck_with_head",
                 "mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py")
    _mod("mmdet3d.models", build_backbone=BACKBONES.build, build_head=HEADS.build)
    _mod("mmdet3d.models.detectors")
    _mod("mmdet3d.models.detectors.base", Base3DDetector=nn.Module)
    det = _load("mmdet3d.models.detectors.single_stage_sparse", "mmdet3d/models/detectors/single_stage_sparse.py")
    fpn = _load("mmdet3d.models.necks.me_fpn", "mmdet3d/models/necks/me_fpn.py")
    ns = types.SimpleNamespace(
        MEFPN3D=fpn.MEFPN3D, MEResNet3D=backbone.MEResNet3D, Fcaf3DNeckWithHead=head.Fcaf3DNeckWithHead, Fcaf3DAssigner=head.Fcaf3DAssigner,
        compute_centerness=head.compute_centerness, SingleStageSparse3DDetector=det.SingleStageSparse3DDetector,
        IoU3DLoss=losses.IoU3DLoss, cal_iou_3d=riou.cal_iou_3d,
        axis_aligned_bbox_overlaps_3d=iouc.axis_aligned_bbox_overlaps_3d,
        rotation_3d_in_axis=utils.rotation_3d_in_axis, DepthBoxes=DepthBoxes, FocalLoss=FocalLoss,
        CrossEntropyLoss=CrossEntropyLoss)
    _loaded = ns
    return ns


def load_config(name):
    """exec()s configs/fcaf3d/<name> with `_base_` merging (dict-of-dicts deep merge like mmcv.Config)."""
    def merge(a, b):
        out = dict(a)
        for k, v in b.items():
            out[k] = merge(a[k], v) if isinstance(v, dict) and isinstance(a.get(k), dict) else v
        return out
    path = os.path.join(REF, "configs", "fcaf3d", name)
    scope = {}
    exec(compile(open(path).read(), path, "exec"), scope)
    cfg = {k: v for k, v in scope.items() if not k.startswith("__")}
    base = {}
    for b in cfg.pop("_base_", []):
        base = merge(base, load_config(b))
    return merge(base, cfg)


def build_reference_detector(config="fcaf3d_scannet-3d-18class.py", **head_overrides):
    ns = load()
    cfg = load_config(config)["model"]
    cfg = {k: (dict(v) if isinstance(v, dict) else v) for k, v in cfg.items()}
    cfg["neck_with_head"].update(head_overrides)
    cfg.pop("type")
    tc = types.SimpleNamespace(**cfg["test_cfg"])
    cfg["test_cfg"] = tc
    return ns.SingleStageSparse3DDetector(**cfg)


def load_structures():
    """The reference's own box / point containers, loaded by path (core/bbox/structures/{base_box3d,depth_box3d}.py,
    core/points/{base_points,depth_points}.py): their flip / rotate / scale / translate methods are what the training
    pipeline's RandomFlip3D and GlobalRotScaleTrans call (datasets/pipelines/transforms_3d.py:100-160,541-636).
    The CUDA-extension imports those files make (iou3d_cuda, points_in_boxes_batch) are not touched by these methods."""
    load()
    if "mmdet3d.core.points.depth_points" not in sys.modules:
        _mod("mmdet3d.ops.iou3d", iou3d_cuda=None)
        sys.modules["mmdet3d.ops"].points_in_boxes_batch = None
        bp = _load("mmdet3d.core.points.base_points", "mmdet3d/core/points/base_points.py")
        _mod("mmdet3d.core.points", BasePoints=bp.BasePoints)
        _load("mmdet3d.core.points.depth_points", "mmdet3d/core/points/depth_points.py")
        _load("mmdet3d.core.bbox.structures.base_box3d", "mmdet3d/core/bbox/structures/base_box3d.py")
        _load("mmdet3d.core.bbox.structures.depth_box3d", "mmdet3d/core/bbox/structures/depth_box3d.py")
    return (sys.modules["mmdet3d.core.bbox.structures.depth_box3d"].DepthInstance3DBoxes,
            sys.modules["mmdet3d.core.points.depth_points"].DepthPoints)


def load_indoor_eval():
    """The reference's own mmdet3d/core/evaluation/indoor_eval.py, loaded by path on top of its own box containers.
    Stand-ins: mmcv.utils.print_log / terminaltables.AsciiTable (printing only) and iou3d_cuda.boxes_overlap_bev_gpu
    (mmdet3d/ops/iou3d/src/iou3d_kernel.cu:111-119,127-245: XYXYR boxes, corners rotated by MINUS the angle) computed
    with the C restatement of the BEV overlap (oracle/nms_oracle.c) on the mirrored angle.  `BaseInstance3DBoxes.overlaps`
    calls `.cuda()` on its temporaries (base_box3d.py:422-426); call indoor_eval inside `cpu_tensors_as_cuda()` here."""
    Boxes, _ = load_structures()

    def boxes_overlap_bev_gpu(a, b, out):
        def conv(x):
            x = x.detach().cpu().float()
            r = torch.zeros((x.shape[0], 7))
            r[:, 0], r[:, 1] = (x[:, 0] + x[:, 2]) / 2, (x[:, 1] + x[:, 3]) / 2
            r[:, 3], r[:, 4], r[:, 5], r[:, 6] = x[:, 2] - x[:, 0], x[:, 3] - x[:, 1], 1.0, -x[:, 4]
            return r.numpy()
        out.copy_(torch.from_numpy(nms_cpu.pair_matrix(conv(a), conv(b), 0)))

    sys.modules["mmdet3d.ops.iou3d"].iou3d_cuda = types.SimpleNamespace(boxes_overlap_bev_gpu=boxes_overlap_bev_gpu)
    sys.modules["mmdet3d.core.bbox.structures.base_box3d"].iou3d_cuda = sys.modules["mmdet3d.ops.iou3d"].iou3d_cuda
    _mod("mmcv.utils", print_log=lambda *a, **k: None)
    _mod("terminaltables", AsciiTable=lambda data: types.SimpleNamespace(table="", inner_footing_row_border=False))
    _mod("mmdet3d.ops.roiaware_pool3d", points_in_boxes_gpu=None)
    _load("mmdet3d.core.bbox.structures.cam_box3d", "mmdet3d/core/bbox/structures/cam_box3d.py")
    _load("mmdet3d.core.bbox.structures.lidar_box3d", "mmdet3d/core/bbox/structures/lidar_box3d.py")
    mode = _load("mmdet3d.core.bbox.structures.box_3d_mode", "mmdet3d/core/bbox/structures/box_3d_mode.py")
    ev = _load("mmdet3d.core.evaluation.indoor_eval", "mmdet3d/core/evaluation/indoor_eval.py")
    return ev.indoor_eval, Boxes, mode.Box3DMode


class cpu_tensors_as_cuda:
    """Context manager: Tensor.cuda() returns the tensor itself (golden generation on a GPU-less host only)."""

    def __enter__(self):
        self._orig = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self

    def __exit__(self, *exc):
        torch.Tensor.cuda = self._orig
