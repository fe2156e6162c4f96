"""Helpers shared by tests, smoke() and bench.py: deterministic weights that do not depend on RNG call order."""
import hashlib
import math

import torch


def _gen(name, seed):
    h = int.from_bytes(hashlib.sha256(("%d:%s" % (seed, name)).encode()).digest()[:4], "little")
    g = torch.Generator(device="cpu")
    g.manual_seed(h)
    return g


@torch.no_grad()
def deterministic_init(model, seed=0, cls_bias=None):
    """Fills every floating-point entry of the state dict from a per-tensor CPU generator keyed by its name, so two
    models with the same state-dict keys (ours and the reference's) get identical weights on any machine.
    Kernels ~ N(0, 2/fan_out); BN affine / running statistics are perturbed so that folding is exercised."""
    sd = model.state_dict()
    for name, t in sd.items():
        if not t.is_floating_point():
            continue
        g = _gen(name, seed)
        r = torch.randn(t.shape, generator=g, dtype=torch.float32)
        leaf = name.split(".")[-1]
        if leaf == "kernel":
            fan_out = t.shape[-1] * (t.shape[0] if t.dim() == 3 else 1)
            if "centerness_conv" in name or "reg_conv" in name or "cls_conv" in name:
                v = r * 0.01
            else:
                v = r * math.sqrt(2.0 / fan_out)
        elif leaf == "running_var":
            v = 1.0 + 0.1 * r.abs()
        elif leaf == "running_mean":
            v = 0.05 * r
        elif leaf == "weight":
            v = 1.0 + 0.1 * r
        elif leaf == "scale":
            v = 1.0 + 0.05 * r
        elif leaf == "bias":
            v = 0.05 * r
            if "cls_conv" in name and cls_bias is not None:
                v = torch.full_like(r, float(cls_bias))
        else:
            v = r * 0.1
        t.copy_(v.to(t.dtype))
    return model


def canonical_order(coords):
    """Row permutation that sorts int coordinate rows lexicographically (canonical form for parity checks)."""
    c = coords.long()
    if c.shape[1] == 3:
        c = torch.cat([torch.zeros_like(c[:, :1]), c], 1)
    key = (c[:, 0] << 54) | ((c[:, 1] + 131072) << 36) | ((c[:, 2] + 131072) << 18) | (c[:, 3] + 131072)
    return torch.argsort(key)
