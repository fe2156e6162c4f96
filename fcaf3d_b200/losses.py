"""Losses and target assignment of the FCAF3D head on the fcaf3d_b200 kernels.

Mirrors the callables the reference builds by registry name (fcaf3d_neck_with_head.py:40-43) and invokes as
`loss(pred, target, weight=None, avg_factor=None)` (:180,191-199):
  FocalLoss(use_sigmoid, gamma, alpha)            mmdet 2.14 -> mmcv.ops.sigmoid_focal_loss (label -1 = background)
  CrossEntropyLoss(use_sigmoid=True)              mmdet 2.14 -> binary_cross_entropy_with_logits
  IoU3DLoss(with_yaw)                             mmdet3d/models/losses/iou3d_loss.py:38-75
  Fcaf3DAssigner(limit, topk, n_scales)           fcaf3d_neck_with_head.py:387-466
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from ._lib import call, ptr


# ------------------------------------------------------------------------------------------- focal
class _FocalFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, gamma, alpha):
        logits = logits.contiguous().float()
        labels = labels.contiguous().long()
        loss = torch.empty_like(logits)
        grad = torch.empty_like(logits)
        call("b2_sigmoid_focal", ptr(logits), ptr(labels), logits.shape[0], logits.shape[1], float(gamma), float(alpha),
             ptr(loss), ptr(grad))
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return g * grad, None, None, None


def sigmoid_focal_loss(pred, labels, gamma=2.0, alpha=0.25):
    """element-wise [N, C] loss; rows whose label is outside [0, C) are all-negative"""
    return _FocalFn.apply(pred, labels, gamma, alpha)


class FocalLoss(nn.Module):
    def __init__(self, use_sigmoid=True, gamma=2.0, alpha=0.25, reduction="mean", loss_weight=1.0):
        super().__init__()
        assert use_sigmoid
        self.gamma, self.alpha, self.loss_weight = gamma, alpha, loss_weight

    def forward(self, pred, target, weight=None, avg_factor=None, reduction_override=None):
        loss = sigmoid_focal_loss(pred, target, self.gamma, self.alpha)
        if weight is not None:
            loss = loss * weight.reshape(-1, 1)
        return self.loss_weight * (loss.sum() / avg_factor if avg_factor is not None else loss.mean())


class CrossEntropyLoss(nn.Module):
    def __init__(self, use_sigmoid=False, use_mask=False, reduction="mean", class_weight=None, loss_weight=1.0):
        super().__init__()
        assert use_sigmoid, "FCAF3D uses the sigmoid (binary) flavour for centerness"
        self.loss_weight, self.reduction = loss_weight, reduction

    def forward(self, cls_score, label, weight=None, avg_factor=None, reduction_override=None, **kw):
        reduction = reduction_override if reduction_override else self.reduction     # mmdet weight_reduce_loss rules
        loss = F.binary_cross_entropy_with_logits(cls_score, label.float(), reduction="none")
        if weight is not None:
            loss = loss * weight.reshape(loss.shape[0], -1)
        if avg_factor is None:
            loss = loss.mean() if reduction == "mean" else loss.sum() if reduction == "sum" else loss
        elif reduction == "mean":
            loss = loss.sum() / avg_factor
        elif reduction != "none":
            raise ValueError('avg_factor can not be used with reduction="sum"')
        return self.loss_weight * loss


# ------------------------------------------------------------------------------------------- IoU losses
class _AxisAlignedIoUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        pred = pred.contiguous().float()
        target = target.contiguous().float()
        n, stride = pred.shape
        iou = torch.empty((n,), dtype=torch.float32, device=pred.device)
        dpred = torch.empty_like(pred)
        call("b2_aa_iou3d", ptr(pred), ptr(target), n, stride, ptr(iou), ptr(dpred))
        ctx.save_for_backward(dpred)
        return iou

    @staticmethod
    def backward(ctx, g):
        (dpred,) = ctx.saved_tensors
        return g.unsqueeze(1) * dpred, None


def axis_aligned_iou_3d(pred, target):
    """aligned pairs of (cx, cy, cz, dx, dy, dz[, ...]) boxes -> IoU [n]"""
    return _AxisAlignedIoUFn.apply(pred, target)


def _box2corners(box):
    """(n, 5) x, y, w, h, alpha -> (n, 4, 2) corners (oriented_iou_loss.py:6-35)"""
    x, y, w, h, alpha = box[:, 0:1], box[:, 1:2], box[:, 2:3], box[:, 3:4], box[:, 4:5]
    x4 = box.new_tensor([0.5, -0.5, -0.5, 0.5]) * w
    y4 = box.new_tensor([0.5, 0.5, -0.5, -0.5]) * h
    corners = torch.stack([x4, y4], dim=-1)
    sin, cos = torch.sin(alpha), torch.cos(alpha)
    rot_t = torch.stack([torch.cat([cos, sin], -1), torch.cat([-sin, cos], -1)], dim=-2)
    rotated = torch.bmm(corners, rot_t)
    return rotated + torch.stack([x, y], dim=-1)


def _sort_vertices(vertices, mask):
    """the un-vendored cuda_op.sort_v (box_intersection_2d.py:131-147) on our kernel"""
    num_valid = mask.sum(dim=1).int()
    mean = (vertices * mask.unsqueeze(-1).float()).sum(dim=1, keepdim=True) / num_valid.view(-1, 1, 1)
    vn = (vertices - mean).detach().contiguous().float()
    idx = torch.empty((vertices.shape[0], 9), dtype=torch.int32, device=vertices.device)
    m8 = mask.to(torch.uint8).contiguous()
    call("b2_sort_vertices", ptr(vn), ptr(m8), ptr(num_valid.contiguous()), vertices.shape[0], ptr(idx))
    return idx.long()


def rotated_iou_2d(box1, box2):
    """differentiable IoU of aligned pairs of rotated rectangles (n, 5); follows box_intersection_2d.py:13-183 and
    oriented_iou_loss.py:38-58 (torch composition + angular sort kernel).  Returns (iou, union)."""
    eps = 1e-8
    c1, c2 = _box2corners(box1), _box2corners(box2)
    # edge intersections (box_intersection_th)
    l1 = torch.cat([c1, c1[:, [1, 2, 3, 0], :]], dim=2)
    l2 = torch.cat([c2, c2[:, [1, 2, 3, 0], :]], dim=2)
    e1 = l1.unsqueeze(2).expand(-1, -1, 4, -1)
    e2 = l2.unsqueeze(1).expand(-1, 4, -1, -1)
    x1, y1, x2, y2 = e1[..., 0], e1[..., 1], e1[..., 2], e1[..., 3]
    x3, y3, x4, y4 = e2[..., 0], e2[..., 1], e2[..., 2], e2[..., 3]
    num = (x1 - x2) * (y3 - y4) - (y1 - y2) * (x3 - x4)
    den_t = (x1 - x3) * (y3 - y4) - (y1 - y3) * (x3 - x4)
    t = torch.where(num == 0, torch.full_like(num, -1.), den_t / num)
    mask_t = (t > 0) & (t < 1)
    den_u = (x1 - x2) * (y1 - y3) - (y1 - y2) * (x1 - x3)
    u = torch.where(num == 0, torch.full_like(num, -1.), -den_u / num)
    mask_u = (u > 0) & (u < 1)
    mask_inter = mask_t & mask_u
    t = den_t / (num + eps)
    inters = torch.stack([x1 + t * (x2 - x1), y1 + t * (y2 - y1)], dim=-1) * mask_inter.float().unsqueeze(-1)

    def in_box(ca, cb):  # corners of a inside b (box1_in_box2)
        a, b, d = cb[:, 0:1, :], cb[:, 1:2, :], cb[:, 3:4, :]
        ab, am, ad = b - a, ca - a, d - a
        p_ab, n_ab = (ab * am).sum(-1), (ab * ab).sum(-1)
        p_ad, n_ad = (ad * am).sum(-1), (ad * ad).sum(-1)
        return (p_ab / n_ab > -1e-6) & (p_ab / n_ab < 1 + 1e-6) & (p_ad / n_ad > -1e-6) & (p_ad / n_ad < 1 + 1e-6)

    c12, c21 = in_box(c1, c2), in_box(c2, c1)
    n = box1.shape[0]
    vertices = torch.cat([c1, c2, inters.reshape(n, 16, 2)], dim=1)
    mask = torch.cat([c12, c21, mask_inter.reshape(n, 16)], dim=1)
    idx = _sort_vertices(vertices, mask)
    sel = torch.gather(vertices, 1, idx.unsqueeze(-1).expand(-1, -1, 2))
    total = (sel[:, :-1, 0] * sel[:, 1:, 1] - sel[:, :-1, 1] * sel[:, 1:, 0]).sum(dim=1)
    inter_area = total.abs() / 2
    union = box1[:, 2] * box1[:, 3] + box2[:, 2] * box2[:, 3] - inter_area
    return inter_area / union, union


def rotated_iou_3d(b1, b2):
    """cal_iou_3d (oriented_iou_loss.py:86-109) for aligned pairs (n, 7) x, y, z, w, h, l, alpha"""
    zmax1, zmin1 = b1[:, 2] + b1[:, 5] * 0.5, b1[:, 2] - b1[:, 5] * 0.5
    zmax2, zmin2 = b2[:, 2] + b2[:, 5] * 0.5, b2[:, 2] - b2[:, 5] * 0.5
    z_overlap = (torch.min(zmax1, zmax2) - torch.max(zmin1, zmin2)).clamp_min(0.)
    iou_2d, u = rotated_iou_2d(b1[:, [0, 1, 3, 4, 6]], b2[:, [0, 1, 3, 4, 6]])
    inter3d = iou_2d * u * z_overlap
    v1 = b1[:, 3] * b1[:, 4] * b1[:, 5]
    v2 = b2[:, 3] * b2[:, 4] * b2[:, 5]
    return inter3d / (v1 + v2 - inter3d)


class IoU3DLoss(nn.Module):
    def __init__(self, with_yaw=True, reduction="mean", loss_weight=1.0):
        super().__init__()
        self.with_yaw, self.reduction, self.loss_weight = with_yaw, reduction, loss_weight

    def forward(self, pred, target, weight=None, avg_factor=None, reduction_override=None, **kw):
        # (the reference's check reads the device; callers that know their weights are positive pass check_weight=False)
        if weight is not None and kw.get("check_weight", True) and not torch.any(weight > 0):
            return pred.sum() * weight.sum()
        if weight is not None and weight.dim() > 1:
            weight = weight.mean(-1)
        iou = rotated_iou_3d(pred, target) if self.with_yaw else axis_aligned_iou_3d(pred, target)
        loss = 1 - iou
        if weight is not None:
            loss = loss * weight
        if avg_factor is None:
            return self.loss_weight * loss.mean()
        return self.loss_weight * loss.sum() / avg_factor


# ------------------------------------------------------------------------------------------- assigner
class Fcaf3DAssigner:
    def __init__(self, limit=27, topk=18, n_scales=4, type=None):
        self.limit, self.topk, self.n_scales = limit, topk, n_scales

    @torch.no_grad()
    def assign(self, points, gt_boxes, gt_labels):
        """points: list over levels of [P_l, 3]; gt_boxes [B, 7] (gravity centre, size, yaw); gt_labels int64 [B].
        Returns (centerness_targets [P], box_targets [P, 7], labels [P]; -1 = background)."""
        dev = points[0].device
        sizes = [int(p.shape[0]) for p in points]
        pts = torch.cat(points, 0).contiguous().float()
        P, B, L = int(pts.shape[0]), int(gt_boxes.shape[0]), len(points)
        off = [0]
        for s in sizes:
            off.append(off[-1] + s)
        level_off = torch.tensor(off, dtype=torch.int32, device=dev)
        boxes = gt_boxes.to(dev).contiguous().float()
        labels_in = gt_labels.to(dev).contiguous().long()
        counts = torch.empty((max(L * B, 1),), dtype=torch.int32, device=dev)
        best = torch.empty((max(B, 1),), dtype=torch.int32, device=dev)
        thr = torch.empty((max(B, 1),), dtype=torch.float32, device=dev)
        ct = torch.empty((P,), dtype=torch.float32, device=dev)
        bt = torch.empty((P, 7), dtype=torch.float32, device=dev)
        lb = torch.empty((P,), dtype=torch.int64, device=dev)
        call("b2_fcaf3d_assign", ptr(pts), P, ptr(level_off), L, ptr(boxes), B, ptr(labels_in), self.limit, self.topk,
             ptr(counts), ptr(best), ptr(thr), ptr(ct), ptr(bt), ptr(lb))
        return ct, bt, lb


    @torch.no_grad()
    def assign_batched(self, points, seg, gt_boxes, gt_labels):
        """All scenes at once (b2_fcaf3d_assign_batched).  points [P, 3] level-major; seg: host list over levels of lists
        over scenes of (begin, end) row ranges; gt_boxes / gt_labels: lists over scenes.
        Returns (centerness_targets [P], box_targets [P, 7], labels [P], scene_of int32 [P], n_pos int32 [S])."""
        dev = points.device
        L, S = len(seg), len(seg[0])
        pts = points.contiguous().float()
        P = int(pts.shape[0])
        box_off = [0]
        for g in gt_boxes:
            box_off.append(box_off[-1] + int(g.shape[0]))
        Btot = box_off[-1]
        meta = torch.tensor([v for lvl in seg for be in lvl for v in be] + box_off, dtype=torch.int32).to(dev)
        seg_dev, box_off_dev = meta[:2 * L * S], meta[2 * L * S:]
        boxes = (torch.cat([g.to(dev).float().reshape(-1, 7) for g in gt_boxes]) if Btot else
                 torch.zeros((1, 7), device=dev)).contiguous()
        labels_in = (torch.cat([l.to(dev).long().reshape(-1) for l in gt_labels]) if Btot else
                     torch.zeros((1,), dtype=torch.long, device=dev)).contiguous()
        counts = torch.empty((max(L * Btot, 1),), dtype=torch.int32, device=dev)
        best = torch.empty((max(Btot, 1),), dtype=torch.int32, device=dev)
        thr = torch.empty((max(Btot, 1),), dtype=torch.float32, device=dev)
        ct = torch.empty((P,), dtype=torch.float32, device=dev)
        bt = torch.empty((P, 7), dtype=torch.float32, device=dev)
        lb = torch.empty((P,), dtype=torch.int64, device=dev)
        scene_of = torch.empty((P,), dtype=torch.int32, device=dev)
        n_pos = torch.empty((S,), dtype=torch.int32, device=dev)
        call("b2_fcaf3d_assign_batched", ptr(pts), P, ptr(seg_dev), S, L, ptr(boxes), ptr(box_off_dev), Btot,
             ptr(labels_in), self.limit, self.topk, ptr(counts), ptr(best), ptr(thr), ptr(ct), ptr(bt), ptr(lb),
             ptr(scene_of), ptr(n_pos))
        return ct, bt, lb, scene_of, n_pos


LOSSES = {"FocalLoss": FocalLoss, "CrossEntropyLoss": CrossEntropyLoss, "IoU3DLoss": IoU3DLoss}


def build_loss(cfg):
    cfg = dict(cfg)
    return LOSSES[cfg.pop("type")](**cfg)
