"""Indoor detection metrics (mAP / mAR at IoU thresholds) with the 3-D IoU computed on the GPU.

Same protocol and result keys as the reference's `indoor_eval` (mmdet3d/core/evaluation/indoor_eval.py:202-309, called
by `Custom3DDataset.evaluate`, datasets/custom_3d.py:225-268):
  * per class, all detections of the dataset are ranked by score; a detection is a true positive at threshold t when its
    best-overlapping ground-truth box of the same scene and class has IoU > t and was not matched before (:129-154);
  * AP = area under the monotone precision envelope (:37-45), recall = TP / number of ground-truth boxes (:160-167);
  * classes are those that occur in the ground truth or in the detections (:236-247); a class without ground truth gives
    nan like the reference (division by zero positives).
3-D IoU (`BaseInstance3DBoxes.overlaps`, base_box3d.py:387-441): rotated BEV overlap x height overlap / union volume with
the clamp 1e-8.  The reference's BEV kernel for this op (mmdet3d/ops/iou3d/src/iou3d_kernel.cu:111-119) rotates the
corners by MINUS the stored angle, the pcdet kernel behind `b2_boxes_pair_matrix` by plus the angle, so the yaws are
negated before the call.

Boxes here are plain tensors `(cx, cy, cz, dx, dy, dz, yaw)` with the gravity centre (what the detector returns; the
reference's `gt_boxes_upright_depth` has the same layout, 6 columns for axis-aligned datasets).
"""
import numpy as np
import torch

from . import ops


def boxes_iou3d_eval(a, b):
    """[n,7] x [m,7] CUDA boxes -> IoU matrix [n,m] with the evaluation op's conventions."""
    a, b = a.float().contiguous().clone(), b.float().contiguous().clone()
    a[:, 6], b[:, 6] = -a[:, 6], -b[:, 6]
    overlap_bev = ops.boxes_pair_matrix(a, b, 0)
    top = torch.min((a[:, 2] + a[:, 5] / 2).view(-1, 1), (b[:, 2] + b[:, 5] / 2).view(1, -1))
    bottom = torch.max((a[:, 2] - a[:, 5] / 2).view(-1, 1), (b[:, 2] - b[:, 5] / 2).view(1, -1))
    inter = overlap_bev * torch.clamp(top - bottom, min=0)
    vol_a = (a[:, 3] * a[:, 4] * a[:, 5]).view(-1, 1)
    vol_b = (b[:, 3] * b[:, 4] * b[:, 5]).view(1, -1)
    return inter / torch.clamp(vol_a + vol_b - inter, min=1e-8)


def average_precision(recalls, precisions):
    """'area' mode of indoor_eval.py:7-52 for one curve."""
    mrec = np.concatenate(([0.0], recalls, [1.0]))
    mpre = np.concatenate(([0.0], precisions, [0.0]))
    for i in range(len(mpre) - 1, 0, -1):
        mpre[i - 1] = max(mpre[i - 1], mpre[i])
    ind = np.where(mrec[1:] != mrec[:-1])[0]
    return np.float32(np.sum((mrec[ind + 1] - mrec[ind]) * mpre[ind + 1]))


def _pad7(x):
    """[n, 6|7] -> float32 [n, 7] (axis-aligned boxes get yaw 0); an empty input becomes [0, 7]."""
    x = torch.as_tensor(x, dtype=torch.float32)
    if x.numel() == 0:
        return torch.zeros((0, 7))
    x = x.reshape(-1, x.shape[-1])
    if x.shape[1] == 6:
        x = torch.cat([x, x.new_zeros(x.shape[0], 1)], 1)
    return x


def _eval_class(dets, gts, thresholds, device):
    """dets: {scene: (boxes [n,7], scores [n])}, gts: {scene: boxes [m,7]} of ONE class."""
    npos = sum(int(g.shape[0]) for g in gts.values())
    scene_ids, scores, best_iou, best_gt = [], [], [], []
    for sid, (boxes, sc) in dets.items():
        n = int(boxes.shape[0])
        if n == 0:
            continue
        g = gts.get(sid)
        if g is not None and g.shape[0] > 0:
            iou = boxes_iou3d_eval(boxes.to(device), g.to(device)).cpu().numpy()
            j = iou.argmax(1)                      # first maximum, like the reference's strict '>' scan (:139-143)
            best_iou.append(iou[np.arange(n), j])
            best_gt.append(j)
        else:
            best_iou.append(np.zeros(n))           # the reference appends np.zeros(1) rows: IoU 0 with nothing
            best_gt.append(np.full(n, -1))
        scene_ids += [sid] * n
        scores.append(np.asarray(sc, dtype=np.float64).reshape(-1))
    nd = len(scene_ids)
    out = []
    if nd == 0:
        for _ in thresholds:
            rec = np.zeros(0) / float(npos) if npos else np.zeros(0)
            out.append((rec, np.zeros(0), average_precision(np.zeros(0), np.zeros(0))))
        return out
    scores = np.concatenate(scores)
    best_iou, best_gt = np.concatenate(best_iou), np.concatenate(best_gt)
    order = np.argsort(-scores)
    for thr in thresholds:
        taken = {sid: np.zeros(int(g.shape[0]), dtype=bool) for sid, g in gts.items()}
        tp, fp = np.zeros(nd), np.zeros(nd)
        for rank, d in enumerate(order):
            sid, j = scene_ids[d], best_gt[d]
            if j >= 0 and best_iou[d] > thr and not taken[sid][j]:
                tp[rank] = 1.0
                taken[sid][j] = True
            else:
                fp[rank] = 1.0
        ctp, cfp = np.cumsum(tp), np.cumsum(fp)
        with np.errstate(divide="ignore", invalid="ignore"):
            recall = ctp / float(npos)
        precision = ctp / np.maximum(ctp + cfp, np.finfo(np.float64).eps)
        out.append((recall, precision, average_precision(recall, precision)))
    return out


def indoor_eval(gt_annos, dt_annos, metric=(0.25, 0.5), label2cat=None, device=None):
    """gt_annos[i]: dict(gt_num, gt_boxes_upright_depth [n, 6|7], class [n]); dt_annos[i]: dict(boxes_3d [n, 6|7] gravity
    centre, scores_3d [n], labels_3d [n]).  Returns the reference's result dict: '<cat>_AP_<t>', '<cat>_rec_<t>',
    'mAP_<t>', 'mAR_<t>'."""
    assert len(gt_annos) == len(dt_annos)
    device = device or torch.device("cuda", torch.cuda.current_device())
    pred, gt = {}, {}
    for sid, det in enumerate(dt_annos):
        labels = np.asarray(torch.as_tensor(det["labels_3d"]).cpu()).reshape(-1)
        boxes = _pad7(torch.as_tensor(det["boxes_3d"]).cpu())
        scores = np.asarray(torch.as_tensor(det["scores_3d"]).cpu()).reshape(-1)
        for c in np.unique(labels):
            sel = labels == c
            pred.setdefault(int(c), {})[sid] = (boxes[torch.from_numpy(sel)], scores[sel])
            gt.setdefault(int(c), {}).setdefault(sid, torch.zeros((0, 7)))
    for sid, ann in enumerate(gt_annos):
        if ann["gt_num"] == 0:
            continue
        boxes = _pad7(torch.as_tensor(np.asarray(ann["gt_boxes_upright_depth"])))
        labels = np.asarray(ann["class"]).reshape(-1)
        for c in np.unique(labels):
            gt.setdefault(int(c), {})[sid] = boxes[torch.from_numpy(labels == c)]
    classes = list(gt.keys())
    label2cat = label2cat or {c: str(c) for c in classes}
    res = {c: _eval_class(pred[c], gt[c], metric, device) for c in classes if c in pred}
    ret = {}
    for ti, thr in enumerate(metric):
        aps, recs = [], []
        for c in classes:
            if c in res:
                recall, _, ap = res[c][ti]
                last = recall[-1] if len(recall) else 0.0
            else:
                ap, last = np.float32(0.0), 0.0
            ret[f"{label2cat[c]}_AP_{thr:.2f}"] = float(ap)
            ret[f"{label2cat[c]}_rec_{thr:.2f}"] = float(last)
            aps.append(float(ap))
            recs.append(float(last))
        ret[f"mAP_{thr:.2f}"] = float(np.mean(aps)) if aps else float("nan")
        ret[f"mAR_{thr:.2f}"] = float(np.mean(recs)) if recs else float("nan")
    return ret
