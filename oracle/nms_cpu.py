"""ORACLE (test infrastructure only): numpy/ctypes glue over oracle/nms_oracle.c and the compiled reference.

  nms_cpu(boxes_sorted, thr, rotated)   greedy BEV NMS exactly as pcdet_nms does it (mask 64x64 + host loop,
                                         mmdet3d/ops/pcdet_nms/src/iou3d_nms.cpp:90-136)
  pcdet_nms_gpu / pcdet_nms_normal_gpu  the Python surface of pcdet_nms_utils.py:81-112 on CPU tensors
  ref_iou_bev_cpu                       the reference's OWN iou3d_cpu.cpp (oracle/_ref), used to pin the restatement
  RefCuda                               the reference's OWN CUDA kernels (oracle/_ref, needs a GPU)
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_orc = None


def _lib():
    global _orc
    if _orc is None:
        path = os.path.join(HERE, "_build", "libnms_oracle.so")
        if not os.path.exists(path):
            from . import build_oracle
            build_oracle.build()
        _orc = ctypes.CDLL(path)
        _orc.orc_pair_matrix.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                         ctypes.c_void_p]
        _orc.orc_nms_mask.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_void_p]
        _orc.orc_nms_greedy.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        _orc.orc_nms_greedy.restype = ctypes.c_int
    return _orc


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def pair_matrix(a, b, mode):
    a, b = _f32(a), _f32(b)
    out = np.zeros((len(a), len(b)), dtype=np.float32)
    _lib().orc_pair_matrix(a.ctypes.data, len(a), b.ctypes.data, len(b), mode, out.ctypes.data)
    return out


def nms_mask(boxes_sorted, thr, rotated):
    b = _f32(boxes_sorted)
    n = len(b)
    cb = (n + 63) // 64
    mask = np.zeros((n, max(cb, 1)), dtype=np.uint64)
    if n:
        _lib().orc_nms_mask(b.ctypes.data, n, float(thr), int(rotated), mask.ctypes.data)
    return mask[:, :cb]


def greedy_keep(mask, n):
    keep = np.zeros(max(n, 1), dtype=np.int64)
    m = np.ascontiguousarray(mask)
    k = _lib().orc_nms_greedy(m.ctypes.data, n, keep.ctypes.data) if n else 0
    return keep[:k]


def nms_cpu(boxes_sorted, thr, rotated):
    return greedy_keep(nms_mask(boxes_sorted, thr, rotated), len(boxes_sorted))


def _nms_surface(boxes, scores, thr, rotated):
    import torch
    order = scores.sort(dim=0, descending=True, stable=True)[1]
    keep = nms_cpu(boxes[order].detach().cpu().numpy(), thr, rotated)
    return order[torch.from_numpy(keep)].contiguous(), None


def pcdet_nms_gpu(boxes, scores, thresh, pre_maxsize=None, **kw):
    return _nms_surface(boxes, scores, thresh, True)


def pcdet_nms_normal_gpu(boxes, scores, thresh, **kw):
    return _nms_surface(boxes, scores, thresh, False)


def ref_iou_bev_cpu(a, b):
    """IoU matrix from the reference's own iou3d_cpu.cpp; None when oracle/_ref was not built."""
    path = os.path.join(HERE, "_ref", "libpcdet_ref_cpu.so")
    if not os.path.exists(path):
        return None
    lib = ctypes.CDLL(path)
    a, b = _f32(a), _f32(b)
    out = np.zeros((len(a), len(b)), dtype=np.float32)
    lib.ref_iou_bev_matrix_cpu.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.ref_iou_bev_matrix_cpu(a.ctypes.data, len(a), b.ctypes.data, len(b), out.ctypes.data)
    return out


class RefCuda:
    """The reference's nms_kernel / nms_normal_kernel / boxes_iou_bev_kernel on torch CUDA tensors."""

    def __init__(self):
        path = os.path.join(HERE, "_ref", "libpcdet_ref_cuda.so")
        self.lib = ctypes.CDLL(path) if os.path.exists(path) else None

    @property
    def available(self):
        return self.lib is not None

    def nms_mask(self, boxes, thr, rotated):
        import torch
        n = boxes.shape[0]
        cb = (n + 63) // 64
        mask = torch.zeros((n, cb), dtype=torch.int64, device=boxes.device)
        torch.cuda.synchronize()
        self.lib.ref_nms_mask.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int]
        rc = self.lib.ref_nms_mask(boxes.data_ptr(), mask.data_ptr(), n, float(thr), int(rotated))
        assert rc == 0, rc
        return mask

    def iou_bev(self, a, b):
        import torch
        out = torch.zeros((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
        torch.cuda.synchronize()
        self.lib.ref_boxes_iou_bev.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                               ctypes.c_void_p]
        rc = self.lib.ref_boxes_iou_bev(a.shape[0], a.data_ptr(), b.shape[0], b.data_ptr(), out.data_ptr())
        assert rc == 0, rc
        return out
