"""MEFPN3D: the sparse feature-pyramid neck of the reference (mmdet3d/models/necks/me_fpn.py:8-143) on the fcaf3d_b200
MinkowskiEngine surface.  Same sub-module names and state-dict keys (`up_block_i`, `score_block_i`, `out_block_i`), same
forward contract: `forward(inputs)` -> list of output tensors fine -> coarse; with ground truth
`forward(inputs, gt_bboxes, gt_labels, img_metas)` -> (outs, dict(loss_pruning=...)).

The reference never instantiates it in a shipped config; it is mirrored for surface completeness.
Differences in execution only: the per-scene top-k masks of `_prune` are one `b2_topk_mask` launch over the batch
offsets instead of a python loop of `torch.topk` calls (rows kept: exactly min(pts_threshold, N_scene) per scene).
"""
import torch
from torch import nn

from . import me as ME
from . import ops
from .detector import BlockSequential
from .losses import build_loss


class MEFPN3D(nn.Module):
    def __init__(self, in_channels, out_channels, voxel_size, n_outs=4, pts_threshold=-1, gt_threshold=2.5,
                 loss=dict(type="CrossEntropyLoss", reduction="none", use_sigmoid=True, loss_weight=1.0)):
        super().__init__()
        self.voxel_size, self.n_outs = voxel_size, n_outs
        self.pts_threshold, self.gt_threshold = pts_threshold, gt_threshold
        self.loss = build_loss(loss)
        self.n_scales = len(in_channels)
        self.pruning = ME.MinkowskiPruning()
        for i in range(self.n_scales):
            if i > 0:
                setattr(self, f"up_block_{i}", self._make_up_block(in_channels[i], in_channels[i - 1]))
                if pts_threshold > 0:
                    setattr(self, f"score_block_{i}",
                            ME.MinkowskiConvolution(in_channels[i - 1], 1, kernel_size=1, bias=True, dimension=3))
            if i < n_outs:
                setattr(self, f"out_block_{i}", self._make_block(in_channels[i], out_channels))

    def init_weights(self):
        pass

    @staticmethod
    def _make_block(cin, cout):
        return BlockSequential(ME.MinkowskiConvolution(cin, cout, kernel_size=3, dimension=3),
                               ME.MinkowskiBatchNorm(cout), ME.MinkowskiELU())

    @staticmethod
    def _make_up_block(cin, cout):
        return BlockSequential(
            ME.MinkowskiGenerativeConvolutionTranspose(cin, cout, kernel_size=2, stride=2, dimension=3),
            ME.MinkowskiBatchNorm(cout), ME.MinkowskiELU(),
            ME.MinkowskiConvolution(cout, cout, kernel_size=3, dimension=3), ME.MinkowskiBatchNorm(cout),
            ME.MinkowskiELU())

    def forward(self, inputs, gt_bboxes=None, gt_labels=None, img_metas=None):
        outs, losses = [], []
        x = inputs[-1]
        for i in range(self.n_scales - 1, -1, -1):
            if i < self.n_scales - 1:
                x = inputs[i] + x
            if i < self.n_outs:
                outs.append(getattr(self, f"out_block_{i}")(x))
            if i > 0:
                x = getattr(self, f"up_block_{i}")(x)
                if self.pts_threshold > 0:
                    score = getattr(self, f"score_block_{i}")(x)
                    mask, loss = self._prune(score, gt_bboxes)
                    x = self.pruning(x, mask)
                    losses += loss
        outs = outs[::-1]
        if gt_bboxes is None:
            return outs
        return outs, (dict(loss_pruning=torch.mean(torch.cat(losses))) if self.pts_threshold > 0 else dict())

    def _prune(self, score, gt_bboxes):
        """Rows kept = per-scene top-`pts_threshold` scores, plus (training) every voxel within gt_threshold voxel sizes
        of a ground-truth centre; the BCE of the score against that ground-truth mask is the pruning loss."""
        m = score._map()
        s = score.F.detach().float().reshape(-1).contiguous()
        mask = ops.topk_mask(s, m.off_dev, m.num_batches, int(self.pts_threshold)).bool()
        losses = []
        if gt_bboxes is not None:
            radius = score.tensor_stride[0] * self.voxel_size * self.gt_threshold
            pts = m.coords[:, 1:] * self.voxel_size
            near = torch.zeros_like(mask)
            for b in range(m.num_batches):
                sl = slice(m.off_host[b], m.off_host[b + 1])
                boxes = gt_bboxes[b]
                centres = boxes.gravity_center if hasattr(boxes, "gravity_center") else boxes[:, :3]
                centres = centres.to(pts.device).float()
                if centres.shape[0]:
                    near[sl] = (torch.cdist(pts[sl], centres, compute_mode="donot_use_mm_for_euclid_dist") < radius).any(dim=1)
                losses.append(self.loss(score.F[sl], near[sl].unsqueeze(1)).squeeze(1))
            mask = mask | near
        return mask, losses
