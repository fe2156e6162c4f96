"""Checkpoints in the reference's on-disk format.

The released FCAF3D weights (README.md:81-93 of the reference) are mmcv checkpoints: a pickled dict
`{"meta": {...}, "state_dict": {...}[, "optimizer": {...}]}` written by `mmcv.runner.save_checkpoint`, whose state-dict
keys are those of `SingleStageSparse3DDetector` (`backbone.*`, `neck_with_head.*`), possibly prefixed with `module.` when
saved from a DataParallel wrapper, with MinkowskiEngine's parameter layout: convolution `kernel`
[K, Cin, Cout] ([Cin, Cout] for 1x1 stride-1), `MinkowskiBatchNorm` as `*.bn.{weight,bias,running_mean,running_var,
num_batches_tracked}`.  fcaf3d_b200.detector keeps exactly these keys and shapes (tests/test_host_logic.py), so loading
is a strict `load_state_dict` after unwrapping.
"""
import collections

import torch


def _unwrap(obj):
    sd = obj.get("state_dict", obj) if isinstance(obj, dict) else obj
    out = collections.OrderedDict()
    for k, v in sd.items():
        out[k[7:] if k.startswith("module.") else k] = v
    return out


def load_checkpoint(model, path, map_location="cpu", strict=True):
    """Loads an mmcv-format (or plain state-dict) checkpoint into the detector mirror; returns the checkpoint's meta."""
    obj = torch.load(path, map_location=map_location, weights_only=False)
    sd = _unwrap(obj)
    own = model.state_dict()
    bad = [k for k, v in sd.items() if k in own and tuple(own[k].shape) != tuple(v.shape)]
    if bad:
        raise RuntimeError("checkpoint tensors with another shape than the model: %s" % ", ".join(
            "%s %s vs %s" % (k, tuple(sd[k].shape), tuple(own[k].shape)) for k in bad[:5]))
    missing, unexpected = model.load_state_dict(sd, strict=False)
    if strict and (missing or unexpected):
        raise RuntimeError("checkpoint does not match the model: missing %s, unexpected %s" % (missing[:5], unexpected[:5]))
    return obj.get("meta", {}) if isinstance(obj, dict) else {}


def save_checkpoint(model, path, optimizer=None, meta=None):
    """Writes the same layout mmcv.runner.save_checkpoint produces (weights on the CPU)."""
    module = model.module if hasattr(model, "module") else model
    ckpt = {"meta": dict(meta or {}), "state_dict": collections.OrderedDict(
        (k, v.detach().cpu()) for k, v in module.state_dict().items())}
    if optimizer is not None:
        ckpt["optimizer"] = optimizer.state_dict()
    torch.save(ckpt, path)
