"""ORACLE (test infrastructure / CPU baseline only): CPU restatement of the reference's FCAF3D forward pass on the
oracle ME surface, usable where /root/reference does not exist (the GPU box).

Follows mmdet3d/models/backbones/me_resnet.py:8-123 (ResNetBase / MEResNet3D),
mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py:47-126,256-279 (neck, pruning, heads) and
mmdet3d/models/detectors/single_stage_sparse.py:32-40 (extract_feat).  Same module tree and state-dict keys as the
reference, so name-keyed weights (fcaf3d_b200.testing.deterministic_init) are identical on both sides.
Pinned in tests/test_oracle_model.py against the golden vectors that the reference's own files produced.
"""
import torch
import torch.nn as nn

from . import me_cpu as ME

LAYERS = {14: (1, 1, 1, 1), 18: (2, 2, 2, 2), 34: (3, 4, 6, 3)}
PLANES = (64, 128, 256, 512)


class Scale(nn.Module):
    def __init__(self, scale=1.0):
        super().__init__()
        self.scale = nn.Parameter(torch.tensor(scale, dtype=torch.float))

    def forward(self, x):
        return x * self.scale


class Backbone(nn.Module):
    def __init__(self, in_channels=3, depth=34, n_outs=4):
        super().__init__()
        self.n_outs = n_outs
        self.conv1 = nn.Sequential(ME.MinkowskiConvolution(in_channels, 64, kernel_size=3, stride=2, dimension=3),
                                   ME.MinkowskiInstanceNorm(64), ME.MinkowskiReLU(inplace=True),
                                   ME.MinkowskiMaxPooling(kernel_size=2, stride=2, dimension=3))
        inplanes = 64
        for i in range(n_outs):
            planes = PLANES[i]
            down = nn.Sequential(ME.MinkowskiConvolution(inplanes, planes, kernel_size=1, stride=2, dimension=3),
                                 ME.MinkowskiBatchNorm(planes))
            blocks = [ME.BasicBlock(inplanes, planes, stride=2, downsample=down, dimension=3)]
            blocks += [ME.BasicBlock(planes, planes, dimension=3) for _ in range(1, LAYERS[depth][i])]
            setattr(self, "layer%d" % (i + 1), nn.Sequential(*blocks))
            inplanes = planes

    def forward(self, x):
        x = self.conv1(x)
        outs = []
        for i in range(self.n_outs):
            x = getattr(self, "layer%d" % (i + 1))(x)
            outs.append(x)
        return outs


class NeckWithHead(nn.Module):
    def __init__(self, n_classes, n_reg_outs, in_channels=(64, 128, 256, 512), out_channels=128, voxel_size=0.01,
                 pts_threshold=100000):
        super().__init__()
        self.voxel_size, self.pts_threshold = voxel_size, pts_threshold
        self.pruning = ME.MinkowskiPruning()
        for i, c in enumerate(in_channels):
            if i > 0:
                p = in_channels[i - 1]
                setattr(self, "up_block_%d" % i, nn.Sequential(
                    ME.MinkowskiGenerativeConvolutionTranspose(c, p, kernel_size=2, stride=2, dimension=3),
                    ME.MinkowskiBatchNorm(p), ME.MinkowskiELU(),
                    ME.MinkowskiConvolution(p, p, kernel_size=3, dimension=3), ME.MinkowskiBatchNorm(p),
                    ME.MinkowskiELU()))
            setattr(self, "out_block_%d" % i, nn.Sequential(
                ME.MinkowskiConvolution(c, out_channels, kernel_size=3, dimension=3),
                ME.MinkowskiBatchNorm(out_channels), ME.MinkowskiELU()))
        self.centerness_conv = ME.MinkowskiConvolution(out_channels, 1, kernel_size=1, dimension=3)
        self.reg_conv = ME.MinkowskiConvolution(out_channels, n_reg_outs, kernel_size=1, dimension=3)
        self.cls_conv = ME.MinkowskiConvolution(out_channels, n_classes, kernel_size=1, bias=True, dimension=3)
        self.scales = nn.ModuleList([Scale(1.) for _ in in_channels])

    def _prune(self, x, scores):
        """per-scene top-k of the interpolated coarser-level score; ties -> lowest row (documented rule)."""
        if self.pts_threshold < 0:
            return x
        with torch.no_grad():
            s = scores.features_at_coordinates(x.C.float()).squeeze(1)
            mask = torch.zeros(len(s), dtype=torch.bool)
            for perm in x.decomposition_permutations:
                k = min(len(perm), self.pts_threshold)
                mask[perm[torch.argsort(-s[perm], stable=True)[:k]]] = True
        return self.pruning(x, mask)

    def forward(self, inputs):
        """per level (fine -> coarse): dict(coords [n,4], centerness [n,1], bbox_pred [n,R], cls [n,C])"""
        levels = [None] * len(inputs)
        x, scores = inputs[-1], None
        for i in range(len(inputs) - 1, -1, -1):
            if i < len(inputs) - 1:
                x = getattr(self, "up_block_%d" % (i + 1))(x)
                x = inputs[i] + x
                x = self._prune(x, scores)
            out = getattr(self, "out_block_%d" % i)(x)
            centerness = self.centerness_conv(out).F
            cls = self.cls_conv(out)
            scores = ME.SparseTensor(cls.F.max(dim=1, keepdim=True).values, coordinate_map_key=cls.coordinate_map_key,
                                     coordinate_manager=cls.coordinate_manager)
            reg = self.reg_conv(out).F
            bbox = torch.cat((torch.exp(self.scales[i](reg[:, :6])), reg[:, 6:]), dim=1)
            levels[i] = dict(coords=out.C, centerness=centerness, bbox_pred=bbox, cls=cls.F)
        return levels


class Detector(nn.Module):
    def __init__(self, dataset="scannet", depth=34, voxel_size=0.01):
        super().__init__()
        n_classes, n_reg = {"scannet": (18, 6), "sunrgbd": (10, 8), "s3dis": (5, 6)}[dataset]
        self.voxel_size = voxel_size
        self.backbone = Backbone(3, depth)
        self.neck_with_head = NeckWithHead(n_classes, n_reg, voxel_size=voxel_size)

    @torch.no_grad()
    def forward(self, points):
        coords, feats = ME.utils.batch_sparse_collate([(p[:, :3] / self.voxel_size, p[:, 3:] / 255.) for p in points])
        x = ME.SparseTensor(coordinates=coords, features=feats)
        return self.neck_with_head(self.backbone(x))
