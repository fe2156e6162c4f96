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

from ._lib import B2Error, call, ptr


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


class _BCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        logits = logits.contiguous().float()
        target = target.detach().contiguous().float()
        loss, grad = torch.empty_like(logits), torch.empty_like(logits)
        call("b2_bce_logits", ptr(logits), ptr(target), logits.numel(), ptr(loss), ptr(grad))
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return g * grad, None


def bce_with_logits(logits, target):
    """element-wise binary cross entropy with logits (value + gradient from one launch of b2_bce_logits)"""
    if logits.shape != target.shape:
        raise B2Error("bce_with_logits: logits %s and target %s differ in shape" % (tuple(logits.shape), tuple(target.shape)))
    return _BCEFn.apply(logits, target)


class CrossEntropyLoss(nn.Module):
    def __init__(self, use_sigmoid=False, use_mask=False, reduction="mean", class_weight=None, loss_weight=1.0):
        super().__init__()
        assert use_sigmoid, "FCAF3D uses the sigmoid (binary) flavour for centerness"
        self.loss_weight, self.reduction = loss_weight, reduction

    def forward(self, cls_score, label, weight=None, avg_factor=None, reduction_override=None, **kw):
        reduction = reduction_override if reduction_override else self.reduction     # mmdet weight_reduce_loss rules
        loss = bce_with_logits(cls_score, label.float())
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


class _RotatedIoUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        pred = pred.contiguous().float()
        target = target.detach().contiguous().float()
        n = pred.shape[0]
        iou = torch.empty((n,), dtype=torch.float32, device=pred.device)
        dpred = torch.empty_like(pred)
        call("b2_rot_iou3d", ptr(pred), ptr(target), n, ptr(iou), ptr(dpred))
        ctx.save_for_backward(dpred)
        return iou

    @staticmethod
    def backward(ctx, g):
        (dpred,) = ctx.saved_tensors
        return g.unsqueeze(1) * dpred, None


def rotated_iou_3d(pred, target):
    """cal_iou_3d (oriented_iou_loss.py:86-109) for aligned pairs (n, 7) x, y, z, w, l, h, alpha: value and the gradient
    w.r.t. `pred` from one kernel launch (b2_rot_iou3d); `target` is treated as constant, as the head uses it."""
    if pred.shape[-1] != 7 or target.shape[-1] != 7:
        raise B2Error("rotated_iou_3d takes (n, 7) boxes")
    return _RotatedIoUFn.apply(pred, target)


class _DecodedIoUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, bbox_pred, target, yaw_mode, rotated):
        points = points.detach().contiguous().float()
        bbox_pred = bbox_pred.contiguous().float()
        target = target.detach().contiguous().float()
        n, n_reg = bbox_pred.shape
        iou = torch.empty((n,), dtype=torch.float32, device=bbox_pred.device)
        dbbox = torch.empty_like(bbox_pred)
        call("b2_box_iou_decoded", ptr(points), ptr(bbox_pred), n_reg, int(yaw_mode), ptr(target), int(bool(rotated)), n,
             ptr(iou), ptr(dbbox))
        ctx.save_for_backward(dbbox)
        return iou

    @staticmethod
    def backward(ctx, g):
        (dbbox,) = ctx.saved_tensors
        return None, g.unsqueeze(1) * dbbox, None, None, None


def decoded_box_iou(points, bbox_pred, target, yaw_mode, rotated):
    """IoU [n] between the boxes `bbox_pred` [n, 6|7|8] decodes to at `points` [n, 3] (_bbox_pred_to_bbox,
    fcaf3d_neck_with_head.py:281-330) and `target` [n, 7]; decode, IoU and the gradient w.r.t. bbox_pred in one launch."""
    return _DecodedIoUFn.apply(points, bbox_pred, target, yaw_mode, rotated)


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
