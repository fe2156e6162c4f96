"""Data-parallel training harness of the FCAF3D hot path (replaces the mmcv runner for benchmarking / tests).

Reference recipe (configs/fcaf3d/fcaf3d.py:30-31): AdamW(lr 1e-3, weight_decay 1e-4), gradient clipping at L2 norm 10;
one process per GPU, torch DistributedDataParallel over NCCL (dist_params backend 'nccl', fcaf3d.py:43);
BatchNorm statistics stay per GPU like in the reference (ME BatchNorm is never converted to SyncBN).
"""
import torch
import torch.distributed as dist


def shard_scenes(n_scenes, rank, world_size):
    """Round-robin scene indices of one rank (inference shards whole scenes; no collective is needed)."""
    return list(range(rank, n_scenes, world_size))


def make_optimizer(model, lr=1e-3, weight_decay=1e-4, max_norm=10.0):
    """AdamW + gradient clipping of the reference recipe on the library's kernels (fcaf3d_b200.optim.FusedAdamW: norm,
    clip and update of all parameters in three launches).  CPU parameters (host-logic tests) get torch's AdamW."""
    params = [p for p in model.parameters() if p.requires_grad]
    if params and all(p.is_cuda for p in params):
        from .optim import FusedAdamW
        return FusedAdamW(params, lr=lr, weight_decay=weight_decay, max_norm=max_norm)
    return torch.optim.AdamW(params, lr=lr, weight_decay=weight_decay)


def wrap_ddp(model, device):
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        import os
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[device.index], broadcast_buffers=False,
                                                        gradient_as_bucket_view=True,
                                                        bucket_cap_mb=int(os.environ.get("B2_DDP_BUCKET_MB", "100")))
        # measured at N=2: 100 MB buckets 480 scenes/s, 25 MB 460-474; the bf16 compression hook 320-360 (its cast kernels
        # cost more than NVLink saves), so it stays opt-in
        if os.environ.get("B2_DDP_BF16", "0") == "1":
            from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
            ddp.register_comm_hook(None, default_hooks.bf16_compress_hook)
        return ddp
    return model


class TrainStepModule(torch.nn.Module):
    """forward() = forward_train, so that DistributedDataParallel hooks the right call"""

    def __init__(self, detector):
        super().__init__()
        self.detector = detector

    def forward(self, points, gt_boxes, gt_labels):
        losses = self.detector.forward_train(points, gt_boxes, gt_labels)
        return sum(losses.values()), losses


def train_step(module, optimizer, points, gt_boxes, gt_labels, max_norm=10.0, next_points=None):
    """One optimisation step; returns (total loss tensor, dict of detached losses[, prepared next batch]).

    `points` is a list of point tensors or a batch prepared by `detector.prepare`.  With `next_points` the coordinate and
    kernel maps of the NEXT batch are built between this step's forward and backward (software pipelining: they depend
    on the points only), and the prepared batch is returned as third value to be passed as `points` of the next call -
    its forward then needs no host round trip and is enqueued while the GPU still runs this step's backward."""
    optimizer.zero_grad(set_to_none=True)
    total, losses = module(points, gt_boxes, gt_labels)
    nxt = None
    if next_points is not None:
        det = (module.module if hasattr(module, "module") else module).detector
        # (measured negative, round 2: building these maps on a side stream under the backward pass instead of between
        # forward and backward made the step SLOWER, 28.8 -> 32.1 ms - the small map kernels then compete with the
        # backward convolutions for the same SMs while the host still blocks on their row counts; building them in a host
        # thread of its own on a stream of its own, under forward + backward, changed nothing: 27.0 - 27.4 against 27.0 ms -
        # the step is bound by the python work of the enqueueing thread, not by these waits)
        with torch.no_grad():
            nxt = det.prepare(next_points)
    total.backward()
    if hasattr(optimizer, "max_norm"):
        optimizer.max_norm = max_norm                    # clipping happens inside the fused update
    else:
        torch.nn.utils.clip_grad_norm_(module.parameters(), max_norm)
    optimizer.step()
    out = (total.detach(), {k: v.detach() for k, v in losses.items()})
    return out + (nxt,) if next_points is not None else out
