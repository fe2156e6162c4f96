"""The reference's `mmdet3d.ops.pcdet_nms` Python surface on the fcaf3d_b200 NMS kernels.

Mirrors mmdet3d/ops/pcdet_nms/pcdet_nms_utils.py:81-112 (`nms_gpu`, `nms_normal_gpu`, exported as
`pcdet_nms_gpu` / `pcdet_nms_normal_gpu` by mmdet3d/ops/pcdet_nms/__init__.py) and the pybind functions of
mmdet3d/ops/pcdet_nms/src/iou3d_nms_api.cpp:11-16.  Unlike the reference (`exit(-1)` on bad input,
iou3d_nms.cpp:14-38) errors raise exceptions.
"""
import torch

from . import ops
from ._lib import B2Error


def _check_boxes(boxes):
    if not boxes.is_cuda:
        raise B2Error("boxes must be a CUDA tensor")
    if boxes.dim() != 2 or boxes.shape[1] != 7:
        raise B2Error("boxes must be (N, 7) [x, y, z, dx, dy, dz, heading]")


def _nms_sorted(boxes_sorted, thresh, rotated):
    """keep flags (uint8, device) of greedy NMS over boxes already sorted by descending score."""
    n = int(boxes_sorted.shape[0])
    keep, _, _, _ = ops.nms_batched(boxes_sorted.contiguous().float(), None, [0, n], thresh, rotated, sort=False)
    return keep


def nms_gpu_raw(boxes_sorted, keep_out, thresh):
    """pybind-level `nms_gpu(boxes, keep, thresh) -> num_to_keep`: fills the caller's CPU int64 `keep`."""
    _check_boxes(boxes_sorted)
    idx = torch.nonzero(_nms_sorted(boxes_sorted, thresh, True)).squeeze(1).cpu()
    keep_out[: idx.numel()] = idx
    return int(idx.numel())


def nms_normal_gpu_raw(boxes_sorted, keep_out, thresh):
    _check_boxes(boxes_sorted)
    idx = torch.nonzero(_nms_sorted(boxes_sorted, thresh, False)).squeeze(1).cpu()
    keep_out[: idx.numel()] = idx
    return int(idx.numel())


def pcdet_nms_gpu(boxes, scores, thresh, pre_maxsize=None, **kwargs):
    """Rotated BEV NMS.  Returns (indices into the unsorted input of the kept boxes, None)."""
    _check_boxes(boxes)
    order = scores.sort(0, descending=True)[1]
    if pre_maxsize is not None:
        order = order[:pre_maxsize]
    keep = _nms_sorted(boxes[order], thresh, True)
    return order[torch.nonzero(keep).squeeze(1)].contiguous(), None


def pcdet_nms_normal_gpu(boxes, scores, thresh, **kwargs):
    """Axis-aligned BEV NMS (heading ignored)."""
    _check_boxes(boxes)
    order = scores.sort(0, descending=True)[1]
    keep = _nms_sorted(boxes[order], thresh, False)
    return order[torch.nonzero(keep).squeeze(1)].contiguous(), None


def boxes_iou_bev(boxes_a, boxes_b):
    _check_boxes(boxes_a), _check_boxes(boxes_b)
    return ops.boxes_pair_matrix(boxes_a.float(), boxes_b.float(), 1)


def boxes_overlap_bev(boxes_a, boxes_b):
    _check_boxes(boxes_a), _check_boxes(boxes_b)
    return ops.boxes_pair_matrix(boxes_a.float(), boxes_b.float(), 0)


def boxes_iou3d_gpu(boxes_a, boxes_b):
    """pcdet_nms_utils.py:47-78: BEV overlap x height overlap / union volume."""
    overlaps_bev = boxes_overlap_bev(boxes_a, boxes_b)
    a_max, a_min = (boxes_a[:, 2] + boxes_a[:, 5] / 2).view(-1, 1), (boxes_a[:, 2] - boxes_a[:, 5] / 2).view(-1, 1)
    b_max, b_min = (boxes_b[:, 2] + boxes_b[:, 5] / 2).view(1, -1), (boxes_b[:, 2] - boxes_b[:, 5] / 2).view(1, -1)
    overlaps_h = torch.clamp(torch.min(a_max, b_max) - torch.max(a_min, b_min), min=0)
    overlaps_3d = overlaps_bev * overlaps_h
    vol_a = (boxes_a[:, 3] * boxes_a[:, 4] * boxes_a[:, 5]).view(-1, 1)
    vol_b = (boxes_b[:, 3] * boxes_b[:, 4] * boxes_b[:, 5]).view(1, -1)
    return overlaps_3d / torch.clamp(vol_a + vol_b - overlaps_3d, min=1e-6)
