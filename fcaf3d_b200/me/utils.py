"""ME.utils subset used by the reference (single_stage_sparse.py:34-36, me_resnet.py:46)."""
import math

import torch

from .. import ops
from .._lib import B2Error


def batch_sparse_collate(data, dtype=torch.int32, device=None):
    """list of (coords [N_i, 3] (float or int), feats [N_i, C]) -> (bcoords int32 [sum N, 4], feats [sum N, C]).

    bcoords[:, 0] is the position in the list, bcoords[:, 1:] = floor(coords) (ME A.1).  No dedup here.
    """
    coords_l, feats_l = [], []
    for b, (c, f) in enumerate(data):
        if device is not None:
            c, f = c.to(device), f.to(device)
        if not c.is_cuda:
            raise B2Error("fcaf3d_b200 has no CPU backend: collate CUDA tensors")
        ci = torch.floor(c).to(dtype) if c.is_floating_point() else c.to(dtype)
        bcol = torch.full((ci.shape[0], 1), b, dtype=dtype, device=ci.device)
        coords_l.append(torch.cat((bcol, ci), dim=1))
        feats_l.append(f)
    return torch.cat(coords_l, 0), torch.cat(feats_l, 0)


def sparse_collate(coords, feats, labels=None, dtype=torch.int32, device=None):
    out = batch_sparse_collate(list(zip(coords, feats)), dtype=dtype, device=device)
    return out if labels is None else out + (torch.cat(labels, 0),)


def batched_coordinates(coords, dtype=torch.int32, device=None):
    return batch_sparse_collate([(c, c[:, :0]) for c in coords], dtype=dtype, device=device)[0]


def _fans(tensor):
    if tensor.dim() < 2:
        raise ValueError("fan in / fan out need at least 2 dimensions")
    if tensor.dim() == 2:
        return tensor.size(1), tensor.size(0)
    rf = tensor.size(0)
    return tensor.size(1) * rf, tensor.size(2) * rf


def kaiming_normal_(tensor, a=0, mode="fan_in", nonlinearity="leaky_relu"):
    fan_in, fan_out = _fans(tensor)
    fan = fan_in if mode == "fan_in" else fan_out
    gain = torch.nn.init.calculate_gain(nonlinearity, a)
    std = gain / math.sqrt(fan)
    with torch.no_grad():
        return tensor.normal_(0, std)
