"""Host-side mirror of the reference's FCAF3D model code on the fcaf3d_b200 ME surface.

Mirrors (same module tree, attribute names and state-dict keys, same hyper-parameters):
  MEResNet3D                      mmdet3d/models/backbones/me_resnet.py:8-123
  Fcaf3DNeckWithHead              mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py:13-374
  SingleStageSparse3DDetector     mmdet3d/models/detectors/single_stage_sparse.py:9-62

Only the module tree (attribute names = state-dict keys, so released checkpoints load) and the hyper-parameters are
mirrored; the execution paths here are this package's own:
  * `simple_test_fused(...)`  – inference: conv + folded BN + activation + residual in one kernel, fused head, batched
                                top-k / decode / NMS with one host sync at the end;
  * `forward_train(...)`      – training: autograd through the ME-surface modules, whole-batch assigner and losses.
The reference's op-by-op execution (python loops over scenes / levels / classes) is not re-typed here: it is the
reference's own three files, loaded unmodified on this package's surfaces by `fcaf3d_b200.compat`.
"""
import math
import os
import queue
import threading

import torch
import torch.nn as nn

from . import autograd as AG
from . import me as ME
from . import ops
from .compat import Scale, bias_init_with_prob, reduce_mean
from .losses import Fcaf3DAssigner, build_loss, decoded_box_iou
from .me.modules.resnet_block import BasicBlock, Bottleneck


# =============================================================================================== backbone
class MEResNet3D(nn.Module):
    ARCH = {
        14: (BasicBlock, (1, 1, 1, 1)),
        18: (BasicBlock, (2, 2, 2, 2)),
        34: (BasicBlock, (3, 4, 6, 3)),
        50: (Bottleneck, (4, 3, 6, 3)),
        101: (Bottleneck, (3, 4, 23, 3)),
    }
    INIT_DIM = 64
    PLANES = (64, 128, 256, 512)

    def __init__(self, in_channels, depth, n_outs=4):
        super().__init__()
        if depth not in self.ARCH:
            raise ValueError(f"invalid depth={depth}")
        block, layers = self.ARCH[depth]
        self.n_outs = n_outs
        self.inplanes = self.INIT_DIM
        self.conv1 = nn.Sequential(
            ME.MinkowskiConvolution(in_channels, self.inplanes, kernel_size=3, stride=2, dimension=3),
            ME.MinkowskiInstanceNorm(self.inplanes),
            ME.MinkowskiReLU(inplace=True),
            ME.MinkowskiMaxPooling(kernel_size=2, stride=2, dimension=3),
        )
        for i in range(n_outs):
            setattr(self, f"layer{i + 1}", self._make_layer(block, self.PLANES[i], layers[i], stride=2))

    def _make_layer(self, block, planes, blocks, stride=1, dilation=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(
                ME.MinkowskiConvolution(self.inplanes, planes * block.expansion, kernel_size=1, stride=stride,
                                        dimension=3),
                ME.MinkowskiBatchNorm(planes * block.expansion),
            )
        mods = [block(self.inplanes, planes, stride=stride, dilation=dilation, downsample=downsample, dimension=3)]
        self.inplanes = planes * block.expansion
        mods += [block(self.inplanes, planes, stride=1, dilation=dilation, dimension=3) for _ in range(1, blocks)]
        return nn.Sequential(*mods)

    def init_weights(self):
        for m in self.modules():
            if isinstance(m, ME.MinkowskiConvolution):
                ME.utils.kaiming_normal_(m.kernel, mode="fan_out", nonlinearity="relu")
            if isinstance(m, ME.MinkowskiBatchNorm):
                nn.init.constant_(m.bn.weight, 1)
                nn.init.constant_(m.bn.bias, 0)

    def forward(self, x):
        """me_resnet.py:66-78, op by op.  The only additions are look-ahead hints to the coordinate manager: the next strided
        map is enqueued one level early so that its row count is read while the GPU still has work (stride_prefetch)."""
        outs = []
        cm = x.coordinate_manager
        stem, j = list(self.conv1), 0
        while j < len(stem):
            m = stem[j]
            nxt = stem[j + 1] if j + 1 < len(stem) else None
            if isinstance(m, ME.MinkowskiInstanceNorm) and isinstance(nxt, ME.MinkowskiReLU):
                x = m(x, act=ops.ACT_RELU)       # normalisation + ReLU as one pass
                j += 1
            else:
                x = m(x)
            if j == 0 and len(stem) > 3:
                cm.stride_prefetch(x.coordinate_map_key, stem[3].stride)            # the pooling map
            j += 1
        cm.stride_prefetch(x.coordinate_map_key, 2)
        for i in range(self.n_outs):
            for j, block in enumerate(getattr(self, f"layer{i + 1}")):
                x = block(x)
                if j == 0 and i + 1 < self.n_outs:
                    cm.stride_prefetch(x.coordinate_map_key, 2)
            outs.append(x)
        return outs

    def prebuild_maps(self, x, training=True):
        """Builds every coordinate map and kernel map the backbone will ask for on input tensor `x` - they depend on the
        coordinates only.  A trainer calls this for the NEXT batch between the forward and the backward of the current
        one: the host reads of the row counts then wait for little queued GPU work, and the next forward is enqueued
        without any host round trip while the GPU is still busy with this step's backward."""
        cm, key = x.coordinate_manager, x.coordinate_map_key

        def conv_maps(conv, in_key):
            out_key = in_key if conv.stride == 1 else cm.stride(in_key, conv.stride)
            if not (conv.kernel_size == 1 and conv.stride == 1):
                cm.kernel_map(in_key, out_key, conv.kernel_size, conv.dilation)
                if training:
                    cm.kernel_map_transposed(in_key, out_key, conv.kernel_size, conv.dilation)
            return out_key

        for m in self.conv1:
            if isinstance(m, ME.MinkowskiConvolution):
                key = conv_maps(m, key)
            elif isinstance(m, ME.MinkowskiMaxPooling):
                pkey = cm.stride(key, m.stride)
                cm.kernel_map(key, pkey, m.kernel_size, m.dilation)
                key = pkey
        for i in range(self.n_outs):
            for block in getattr(self, f"layer{i + 1}"):
                convs = [c for c in (getattr(block, n, None) for n in ("conv1", "conv2", "conv3")) if c is not None]
                in_key = key
                for c in convs:
                    key = conv_maps(c, key)
                if block.downsample is not None:
                    conv_maps(block.downsample[0], in_key)
        return x

    # ---- fused execution ---------------------------------------------------------------------
    def forward_fused(self, x):
        cm = x.coordinate_manager
        conv, inorm, _, pool = self.conv1
        dt = ME.get_compute_dtype()
        # Coordinate maps run one level ahead of the features: the build of the next strided map is enqueued before the
        # convolutions of the current level, so that when the host reads its row count the GPU still has work queued
        # (CoordinateManager.stride_prefetch) - the forward pass has no idle-GPU host round trips in the backbone.
        key, km = conv.maps(x)
        cm.stride_prefetch(key, pool.stride)
        f = ops.sparse_conv(x.F.contiguous(), conv._w3(), km, cm.size(key), out_dtype=dt)
        m = cm.get_map(key)
        scale, shift = inorm.scale_shift(f, m)
        pkey = cm.stride(key, pool.stride)
        first = getattr(self, "layer1")[0]
        if isinstance(first, BasicBlock):
            cm.stride_prefetch(pkey, first.conv1.stride)
        pkm = cm.kernel_map(key, pkey, pool.kernel_size, pool.dilation)
        if f.dtype == torch.bfloat16 and f.shape[1] % 8 == 0:
            # InstanceNorm + ReLU applied to the rows as the pooling gathers them (same bf16 values, one pass less)
            f = ops.maxpool_affine(f, pkm, cm.get_map(pkey).coords, scale, shift, ops.ACT_RELU)
        else:
            f = ops.affine_act(f, scale, shift, coords=m.coords, act=ops.ACT_RELU, out=f)
            f = ops.maxpool(f, pkm)
        key = pkey
        outs = []
        for i in range(self.n_outs):
            nxt = getattr(self, f"layer{i + 2}")[0] if i + 1 < self.n_outs else None
            for j, block in enumerate(getattr(self, f"layer{i + 1}")):
                f, key = _fused_block(block, f, key, cm)
                if j == 0 and isinstance(nxt, BasicBlock):
                    cm.stride_prefetch(key, nxt.conv1.stride)     # consumed by the first block of the next layer
            outs.append(ME.SparseTensor(f, coordinate_map_key=key, coordinate_manager=cm))
        return outs


class BlockSequential(nn.Sequential):
    """nn.Sequential (same children, same state-dict keys) whose forward hands a MinkowskiBatchNorm and the ReLU / ELU
    module that follows it to the normalisation as one call, so that the pair is one pass over the rows."""

    def forward(self, x):
        mods = list(self)
        i = 0
        while i < len(mods):
            m = mods[i]
            nxt = mods[i + 1] if i + 1 < len(mods) else None
            if isinstance(m, ME.MinkowskiBatchNorm) and isinstance(nxt, (ME.MinkowskiReLU, ME.MinkowskiELU)):
                x = m(x, act=ops.ACT_RELU if isinstance(nxt, ME.MinkowskiReLU) else ops.ACT_ELU)
                i += 2
            else:
                x = m(x)
                i += 1
        return x


def conv_bn_act(conv, bn, f, km, n_out, act, residual=None):
    """conv -> BatchNorm -> (+residual) -> activation.  Eval: one kernel (BN folded into the epilogue).
    Train: conv, batch statistics, then one fused normalise/residual/activation pass."""
    if bn.bn.training:
        y = ops.sparse_conv(f, conv._w3(), km, n_out, wt_bf16=conv._wt_bf16(), wt_tf32=conv._wt_tf32())
        scale, shift = bn.scale_shift(y)
        return ops.affine_act(y, scale, shift, residual=residual, act=act, out=y)
    scale, shift = bn.scale_shift(None)
    return ops.sparse_conv(f, conv._w3(), km, n_out, wt_bf16=conv._wt_bf16(), scale=scale.reshape(-1),
                           shift=shift.reshape(-1), residual=residual, act=act, wt_tf32=conv._wt_tf32())


def _fused_block(block, f, key, cm):
    if not isinstance(block, BasicBlock):  # Bottleneck: no fused schedule, run op by op
        out = block(ME.SparseTensor(f, coordinate_map_key=key, coordinate_manager=cm))
        return out.F, out.coordinate_map_key
    x = ME.SparseTensor(f, coordinate_map_key=key, coordinate_manager=cm)
    out_key, km1 = block.conv1.maps(x)
    n_out = cm.size(out_key)
    h = conv_bn_act(block.conv1, block.norm1, f, km1, n_out, ops.ACT_RELU)
    if block.downsample is not None:
        dconv, dbn = block.downsample
        kmd = cm.kernel_map(key, out_key, dconv.kernel_size, dconv.dilation) if dconv.stride != 1 else None
        res = conv_bn_act(dconv, dbn, f, kmd, n_out, ops.ACT_NONE)
    else:
        res = f
    km2 = cm.kernel_map(out_key, out_key, block.conv2.kernel_size, block.conv2.dilation)
    out = conv_bn_act(block.conv2, block.norm2, h, km2, n_out, ops.ACT_RELU, residual=res)
    return out, out_key


# =============================================================================================== neck + head
# forward_fused, eval: evaluate the up-block's 3^3 convolution at the rows that survive the prune only (B2_PRUNE_FIRST=0
# restores convolve-everything-then-prune for A/B runs; both give the same rows, tests/test_surface_gpu.py)
PRUNE_BEFORE_CONV = os.environ.get("B2_PRUNE_FIRST", "1") != "0"


class Fcaf3DNeckWithHead(nn.Module):
    def __init__(self, n_classes, in_channels, out_channels, n_reg_outs, voxel_size, pts_threshold, assigner=None,
                 yaw_parametrization="fcaf3d", loss_centerness=None, loss_bbox=None, loss_cls=None, train_cfg=None,
                 test_cfg=None):
        super().__init__()
        self.voxel_size = voxel_size
        self.yaw_parametrization = yaw_parametrization
        self.pts_threshold = pts_threshold
        self.train_cfg = train_cfg
        self.test_cfg = test_cfg or dict(nms_pre=1000, iou_thr=.5, score_thr=.01)
        self.n_classes, self.n_reg_outs = n_classes, n_reg_outs
        self.trace = None          # set to a dict to record every level's prune scores / mask in forward_fused
        assigner = dict(assigner or dict(limit=27, topk=18, n_scales=len(in_channels)))
        assigner.pop("type", None)
        self.assigner = Fcaf3DAssigner(**assigner)
        self.loss_centerness = build_loss(loss_centerness or dict(type="CrossEntropyLoss", use_sigmoid=True,
                                                                  loss_weight=1.0))
        self.loss_bbox = build_loss(loss_bbox or dict(type="IoU3DLoss", loss_weight=1.0))
        self.loss_cls = build_loss(loss_cls or dict(type="FocalLoss", use_sigmoid=True, gamma=2.0, alpha=0.25,
                                                    loss_weight=1.0))
        self._init_layers(in_channels, out_channels, n_reg_outs, n_classes)

    @staticmethod
    def _make_block(cin, cout):
        return BlockSequential(ME.MinkowskiConvolution(cin, cout, kernel_size=3, dimension=3),
                               ME.MinkowskiBatchNorm(cout), ME.MinkowskiELU())

    @staticmethod
    def _make_up_block(cin, cout):
        return BlockSequential(
            ME.MinkowskiGenerativeConvolutionTranspose(cin, cout, kernel_size=2, stride=2, dimension=3),
            ME.MinkowskiBatchNorm(cout), ME.MinkowskiELU(),
            ME.MinkowskiConvolution(cout, cout, kernel_size=3, dimension=3),
            ME.MinkowskiBatchNorm(cout), ME.MinkowskiELU())

    def _init_layers(self, in_channels, out_channels, n_reg_outs, n_classes):
        self.pruning = ME.MinkowskiPruning()
        for i in range(len(in_channels)):
            if i > 0:
                setattr(self, f"up_block_{i}", self._make_up_block(in_channels[i], in_channels[i - 1]))
            setattr(self, f"out_block_{i}", self._make_block(in_channels[i], out_channels))
        self.centerness_conv = ME.MinkowskiConvolution(out_channels, 1, kernel_size=1, dimension=3)
        self.reg_conv = ME.MinkowskiConvolution(out_channels, n_reg_outs, kernel_size=1, dimension=3)
        self.cls_conv = ME.MinkowskiConvolution(out_channels, n_classes, kernel_size=1, bias=True, dimension=3)
        self.scales = nn.ModuleList([Scale(1.) for _ in range(len(in_channels))])

    def init_weights(self):
        nn.init.normal_(self.centerness_conv.kernel, std=.01)
        nn.init.normal_(self.reg_conv.kernel, std=.01)
        nn.init.normal_(self.cls_conv.kernel, std=.01)
        nn.init.constant_(self.cls_conv.bias, bias_init_with_prob(.01))

    def _cfg(self, name):
        c = self.test_cfg
        return c[name] if isinstance(c, dict) else getattr(c, name)

    def _prune(self, x, scores):
        if self.pts_threshold < 0:
            return x
        with torch.no_grad():
            interpolated = scores.features_at_coordinates(x.C.float())
            m = x._map()
            mask = ops.topk_mask(interpolated.reshape(-1).contiguous(), m.off_dev, m.num_batches, self.pts_threshold)
        return self.pruning(x, mask.bool())

    # ---- batched training path: no python loop over scenes ------------------------------
    def forward_train_batched(self, inputs, gt_bboxes, gt_labels):
        """Same neck walk as forward(), but the heads keep every level as ONE tensor over all scenes and the losses are
        computed for the whole batch at once (loss_batched)."""
        levels = []
        x = inputs[-1]
        scores = None
        for i in range(len(inputs) - 1, -1, -1):
            if i < len(inputs) - 1:
                x = getattr(self, f"up_block_{i + 1}")(x)
                x = inputs[i] + x
                x = self._prune(x, scores)
            if i > 0:   # look-ahead: the next level's generated map and its union depend on coordinates only
                cm = x.coordinate_manager
                cm.union_prefetch(inputs[i - 1].coordinate_map_key, cm.generate_children(x.coordinate_map_key))
            out = getattr(self, f"out_block_{i}")(x)
            # the three 1x1 heads (ME: three F.mm calls, fcaf3d_neck_with_head.py:257-264) as ONE GEMM over the concatenated
            # kernels on the library's convolution kernels (identity map; tcgen05 forward / input gradient / weight
            # gradient with bf16 rows, CUDA-core fp32 otherwise), fp32 logits in both modes like forward_fused
            w = torch.cat((self.centerness_conv.kernel, self.reg_conv.kernel, self.cls_conv.kernel), dim=1)
            n_col = w.shape[1]
            wp = torch.nn.functional.pad(w, (0, (-n_col) % 64)).unsqueeze(0)
            h = AG.SparseConvFn.apply(out.F, wp, None, None, out.F.shape[0], out.F.dtype == torch.bfloat16,
                                      torch.float32)[:, :n_col]
            reg = h[:, 1:1 + self.n_reg_outs]
            bbox = torch.cat((torch.exp(self.scales[i](reg[:, :6])), reg[:, 6:]), dim=1)
            cls = h[:, 1 + self.n_reg_outs:] + self.cls_conv.bias
            scores = ME.SparseTensor(cls.max(dim=1, keepdim=True).values, coordinate_map_key=out.coordinate_map_key,
                                     coordinate_manager=out.coordinate_manager)
            levels.append((h[:, :1], bbox, cls, out))
        return self.loss_batched(levels[::-1], gt_bboxes, gt_labels)

    def loss_batched(self, levels, gt_bboxes, gt_labels):
        """levels: fine -> coarse list of (centerness [N_l, 1], bbox_pred [N_l, R], cls [N_l, C], SparseTensor).
        Per scene s the reference computes sum_p loss_p / normaliser_s and then averages over scenes
        (fcaf3d_neck_with_head.py:128-203); here every row carries the weight 1 / (S * normaliser_scene(row))."""
        S = len(gt_bboxes)
        maps = [lv[3]._map() for lv in levels]
        base, seg = 0, []
        for m in maps:
            seg.append([(base + m.off_host[b], base + m.off_host[b + 1]) for b in range(S)])
            base += m.n
        points = torch.cat([m.coords[:, 1:] * self.voxel_size for m in maps])
        centerness = torch.cat([lv[0] for lv in levels])
        bbox_preds = torch.cat([lv[1] for lv in levels])
        cls_scores = torch.cat([lv[2] for lv in levels])
        centerness_targets, bbox_targets, labels, scene_of, n_pos_i = self.assigner.assign_batched(
            points, seg, gt_bboxes, gt_labels)
        scene_of = scene_of.long()
        n_pos = torch.clamp(reduce_mean(n_pos_i.float()), min=1.)                      # [S]
        w_scene = 1.0 / (S * n_pos)
        loss_cls = self.loss_cls(cls_scores, labels, weight=w_scene[scene_of], avg_factor=1.0)
        pos_inds = torch.nonzero(labels >= 0).squeeze(1)                              # the one host sync of the loss
        pos_centerness = centerness[pos_inds]
        pos_bbox_preds = bbox_preds[pos_inds]
        pos_scene = scene_of[pos_inds]
        pos_ct = centerness_targets[pos_inds]
        # per-scene sum of the positives' centerness targets, in a FIXED order (index_add_ adds with atomics: its last bits
        # change from run to run, and through the bf16 activations that reaches every gradient of the step)
        per_scene = (torch.nn.functional.one_hot(pos_scene, S).to(pos_ct.dtype) * pos_ct.unsqueeze(1)).sum(0)
        denorm = torch.clamp(reduce_mean(per_scene), min=1e-6)
        if len(pos_inds) > 0:
            loss_centerness = self.loss_centerness(pos_centerness, pos_ct.unsqueeze(1),
                                                   weight=w_scene[pos_scene].unsqueeze(1), avg_factor=1.0)
            # decode + IoU (rotated or axis-aligned, as the configured IoU3DLoss) + their gradient: one launch
            iou = decoded_box_iou(points[pos_inds], pos_bbox_preds, bbox_targets[pos_inds],
                                  self._yaw_mode(pos_bbox_preds.shape[1]), self.loss_bbox.with_yaw)
            loss_bbox = self.loss_bbox.loss_weight * ((1 - iou) * (pos_ct / (S * denorm[pos_scene]))).sum()
        else:
            loss_centerness = pos_centerness.sum()
            loss_bbox = pos_bbox_preds.sum()
        return dict(loss_centerness=loss_centerness, loss_bbox=loss_bbox, loss_cls=loss_cls)

    def _yaw_mode(self, n_reg):
        if n_reg == 6:
            return ops.YAW_MODES["none"]
        return ops.YAW_MODES[self.yaw_parametrization]

    # ---- fused execution -----------------------------------------------------------------------
    def _head_weights(self):
        """One [1, C, 1 + n_reg + n_cls] kernel + bias for the three 1x1 head convolutions."""
        ks = (self.centerness_conv.kernel, self.reg_conv.kernel, self.cls_conv.kernel, self.cls_conv.bias)
        ver = tuple((k._version, k.data_ptr()) for k in ks)
        if getattr(self, "_hw_ver", None) != ver:
            w = torch.cat([k.detach() for k in ks[:3]], dim=1)
            b = torch.cat([w.new_zeros(1 + self.n_reg_outs), self.cls_conv.bias.detach().reshape(-1)])
            pad = (-w.shape[1]) % 32                      # columns padded to a multiple of 32 for the tcgen05 path
            w = torch.nn.functional.pad(w, (0, pad)).unsqueeze(0).contiguous()
            b = torch.nn.functional.pad(b, (0, pad)).contiguous()
            wt = w.transpose(1, 2).contiguous()
            self._hw, self._hw_ver = (w, b, wt.to(torch.bfloat16), wt), ver
        return self._hw[:3]

    def _head_weights_tf32(self):
        self._head_weights()
        return self._hw[3] if ME.get_tf32() else None

    def forward_fused(self, inputs, want_max_score=False):
        """Returns per level (fine -> coarse) dicts with centerness / bbox_pred / cls / prune score (+max score)
        tensors over the level's coordinate map."""
        cm = inputs[0].coordinate_manager
        levels = [None] * len(inputs)
        f, key = inputs[-1].F, inputs[-1].coordinate_map_key
        prev = pre_gen = None
        for i in range(len(inputs) - 1, -1, -1):
            if i < len(inputs) - 1:
                gen, bn1, _, conv, bn2, _ = getattr(self, f"up_block_{i + 1}")
                gkey = pre_gen
                if bn1.bn.training:
                    g = gen.generate(f)
                    s, b = bn1.scale_shift(g)
                    g = ops.affine_act(g, s, b, act=ops.ACT_ELU, out=g)
                else:
                    s, b = bn1.scale_shift(None)
                    g = gen.generate(f, scale=s.reshape(-1), shift=b.reshape(-1), act=ops.ACT_ELU)
                # inputs[i] + x  (union of the backbone map and the generated map)
                ukey, dst_a, dst_b = cm.union(inputs[i].coordinate_map_key, gkey)
                nu = cm.size(ukey)
                um = cm.get_map(ukey)
                need_prune = self.pts_threshold >= 0 and max(um.off_host[b + 1] - um.off_host[b]
                                                             for b in range(um.num_batches)) > self.pts_threshold
                if need_prune:
                    # prune by the coarser level's interpolated max-class score: the mask depends on the coordinates and
                    # on the previous level only, so it is ready before the up-block's 3^3 convolution runs
                    pm = cm.get_map(prev["key"])
                    table, cap = pm.ensure_table()
                    sc = ops.interp(table, cap, prev["prune"], pm.ts, um.coords.float())
                    mask = ops.topk_mask(sc.reshape(-1), um.off_dev, um.num_batches, self.pts_threshold)
                    if self.trace is not None:     # tests: the prune decision (scores, mask) of this level
                        self.trace[i] = dict(union_key=ukey, scores=sc, mask=mask)
                    keep = [min(um.off_host[b + 1] - um.off_host[b], self.pts_threshold) for b in range(um.num_batches)]
                    key, m8, pos, n_keep = cm.prune(ukey, mask, keep_counts=keep)   # top-k keeps exactly min(N_b, k) rows
                if (need_prune and PRUNE_BEFORE_CONV and not bn2.bn.training and nu == cm.size(gkey)
                        and int(conv.kernel_size) == 3 and int(conv.dilation) == 1):
                    # Eval: behind the convolution there is per-row work only (folded BatchNorm, ELU, + backbone row), so
                    # it is evaluated at the surviving rows alone; its table and the pruned map's own 3^3 table (the
                    # out_block's) both come from the parent's table by index arithmetic - no children-wide table, no
                    # hash of the pruned map.  Same values on the surviving rows as convolving everything and pruning.
                    self.prune_first_levels = getattr(self, "prune_first_levels", 0) + 1     # (tests look at this)
                    g2k, a2k, ksrc = ops.prune_plan(dst_b, dst_a, m8, pos, n_keep)
                    parent = cm.get_map(gkey).children_of
                    km_up, km_out = ops.kernel_map_children_pruned(cm.kernel_map(parent, parent, 3, 1), ksrc, g2k, n_keep)
                    cm.set_kernel_map(key, key, 3, 1, km_out)
                    cm.set_kernel_map(gkey, key, 3, 1, km_up)
                    f = conv_bn_act(conv, bn2, g, km_up, n_keep, ops.ACT_ELU)
                    ops.scatter_rows(inputs[i].F.contiguous(), a2k, f, accumulate=True)
                else:
                    km = cm.kernel_map(gkey, gkey, conv.kernel_size, conv.dilation)
                    g = conv_bn_act(conv, bn2, g, km, cm.size(gkey), ops.ACT_ELU)
                    f = torch.zeros((nu, g.shape[1]), dtype=g.dtype, device=g.device)
                    ops.scatter_rows(inputs[i].F.contiguous(), dst_a, f, accumulate=False)
                    ops.scatter_rows(g, dst_b, f, accumulate=True)
                    if need_prune:
                        f = ops.gather_rows(f, m8, pos, n_keep)
                    else:
                        key = ukey
            if i > 0:
                # the next (finer) level's generated map and its union with the backbone map depend on coordinates only:
                # enqueue them before this level's out_block / heads so that the union's row count is not waited for idle
                nkey = cm.generate_children(key)
                cm.union_prefetch(inputs[i - 1].coordinate_map_key, nkey)
                pre_gen = nkey
            conv, bn, _ = getattr(self, f"out_block_{i}")
            km = cm.kernel_map(key, key, conv.kernel_size, conv.dilation)
            o = conv_bn_act(conv, bn, f, km, cm.size(key), ops.ACT_ELU)
            w, b, wt = self._head_weights()
            lscale = self.scales[i].scale.detach().reshape(1)
            if ops.head_fused_ok(o, wt, self.n_reg_outs, self.n_classes):
                ctr, bbox, cls, prune, mx = ops.head_fused(o, wt, b, self.n_reg_outs, self.n_classes, lscale, want_max_score)
            else:
                h = ops.sparse_conv(o, w, None, o.shape[0], wt_bf16=wt, bias=b, out_dtype=torch.float32,
                                    wt_tf32=self._head_weights_tf32())
                ctr, bbox, cls, prune, mx = ops.head_post(h, self.n_reg_outs, self.n_classes, lscale, want_max_score)
            prev = dict(key=key, centerness=ctr, bbox_pred=bbox, cls=cls, prune=prune, max_score=mx)
            levels[i] = prev
        return levels

    def get_bboxes_fused(self, levels, cm):
        """Batched tail: per (scene, level) top-k -> decode -> per (scene, class) BEV NMS, all scenes at once.
        Returns a list over scenes of (boxes [n, 6|7], scores [n], labels [n]) on the device."""
        nms_pre, score_thr, iou_thr = self._cfg("nms_pre"), self._cfg("score_thr"), self._cfg("iou_thr")
        dev = levels[0]["cls"].device
        nb = cm.num_batches
        n_cls, n_reg = self.n_classes, self.n_reg_outs
        yaw_mode = self._yaw_mode(n_reg)
        rotated = n_reg != 6
        # candidate counts per (scene, level) are known on the host from the batch offsets
        maps = [cm.get_map(l["key"]) for l in levels]
        cnt = [[min(m.off_host[b + 1] - m.off_host[b], nms_pre) if nms_pre > 0 else m.off_host[b + 1] - m.off_host[b]
                for m in maps] for b in range(nb)]
        cand_off = [0]
        for b in range(nb):
            cand_off.append(cand_off[-1] + sum(cnt[b]))
        total = cand_off[-1]
        cand_boxes = torch.empty((total, 7), dtype=torch.float32, device=dev)
        cand_scores = torch.empty((total, n_cls), dtype=torch.float32, device=dev)
        bases = []
        for li in range(len(levels)):
            bases += [cand_off[b] + sum(cnt[b][:li]) for b in range(nb)]
        meta = ops.ints_to_device(cand_off + bases, dev)
        cand_off_dev = meta[: nb + 1]
        for li, (l, m) in enumerate(zip(levels, maps)):
            k = nms_pre if nms_pre > 0 else m.n
            mask = ops.topk_mask(l["max_score"], m.off_dev, nb, k)
            pos, _ = ops.scan_mask(mask)
            base = meta[nb + 1 + li * nb: nb + 1 + (li + 1) * nb]
            ops.call("b2_emit_candidates", ops.ptr(m.coords), ops.ptr(l["centerness"]), ops.ptr(l["bbox_pred"]),
                     ops.ptr(l["cls"]), ops.ptr(mask), ops.ptr(pos), ops.ptr(m.off_dev), ops.ptr(base), m.n, n_reg,
                     n_cls, yaw_mode, float(self.voxel_size), ops.ptr(cand_boxes), ops.ptr(cand_scores))
        # (scene, class) segments of candidates above the score threshold
        nf = total * n_cls
        nseg = nb * n_cls
        flags = torch.empty((max(nf, 1),), dtype=torch.uint8, device=dev)
        pos = torch.empty((max(nf, 1),), dtype=torch.int32, device=dev)
        seg_boxes = torch.empty((max(nf, 1), 7), dtype=torch.float32, device=dev)
        seg_scores = torch.empty((max(nf, 1),), dtype=torch.float32, device=dev)
        seg_src = torch.empty((max(nf, 1),), dtype=torch.int32, device=dev)
        seg_off = torch.zeros((nseg + 2,), dtype=torch.int32, device=dev)
        ws = ops.workspace(ops._lib.lib().b2_scan_ws_bytes(nf), dev)
        ops.call("b2_class_segments", ops.ptr(cand_boxes), ops.ptr(cand_scores), ops.ptr(cand_off_dev), nb, n_cls, total,
                 float(score_thr), ops.ptr(flags), ops.ptr(pos), ops.ptr(seg_off) + 4 * (nseg + 1), ops.ptr(seg_boxes),
                 ops.ptr(seg_scores), ops.ptr(seg_src), ops.ptr(seg_off), ops.ptr(ws), ws.numel())
        max_seg = max([sum(c) for c in cnt] + [0])       # a (scene, class) segment holds at most the scene's candidates
        if ops.NMS_GREEDY and max_seg <= ops.NMS_GREEDY_MAX_SEG:
            # no host read: the greedy kernels take the segment offsets from the device, buffers are sized by the bound
            n_in = nf
            keep, order, sorted_boxes, _ = ops.nms_batched(seg_boxes, seg_scores, None, iou_thr, rotated, sort=True,
                                                           seg_off_dev=seg_off[: nseg + 1], max_seg=max_seg)
        else:
            seg_off_host = ops.read_ints(seg_off)[: nseg + 1]
            n_in = seg_off_host[-1]
            keep, order, sorted_boxes, _ = ops.nms_batched(seg_boxes[:n_in], seg_scores[:n_in], seg_off_host, iou_thr,
                                                           rotated, sort=True)
        kpos, ktotal = ops.scan_mask(keep)
        out_boxes = torch.empty((n_in, 7), dtype=torch.float32, device=dev)
        out_scores = torch.empty((n_in,), dtype=torch.float32, device=dev)
        out_labels = torch.empty((n_in,), dtype=torch.int32, device=dev)
        out_scene = torch.empty((n_in,), dtype=torch.int32, device=dev)
        ops.call("b2_collect_dets", ops.ptr(sorted_boxes), ops.ptr(seg_scores), ops.ptr(order), ops.ptr(seg_off), nseg,
                 n_cls, ops.ptr(keep), ops.ptr(kpos), n_in, ops.ptr(out_boxes), ops.ptr(out_scores), ops.ptr(out_labels),
                 ops.ptr(out_scene))
        return dict(boxes=out_boxes, scores=out_scores, labels=out_labels, scene=out_scene, count=ktotal,
                    box_dim=7 if rotated else 6)


_PINNED_OUT = {}


def wait_ready(dets, stream=None):
    """Results of the native executor may still be in flight on its tail stream: makes `stream` (default: the current one)
    wait for them.  A no-op for results produced on the current stream."""
    ev = dets.get("ready")
    if ev is not None:
        (stream or torch.cuda.current_stream()).wait_event(ev)
    return dets


def split_detections(dets, num_scenes):
    """Device-side result dict -> host list of (boxes, scores, labels) per scene: one read of the count, then ONE packed
    device->host copy of (box, score, label, scene) rows into a reused pinned buffer; the rows come out of the NMS in
    (scene, class) order, so the per-scene split is a slice."""
    wait_ready(dets)
    n = int(dets["count"].item())
    d = dets["box_dim"]
    packed_dev = torch.cat((dets["boxes"][:n, :d], dets["scores"][:n, None], dets["labels"][:n, None].float(),
                            dets["scene"][:n, None].float()), dim=1)
    buf = _PINNED_OUT.get(d)
    if buf is None or buf.shape[0] < n:
        buf = torch.empty((max(n, 1) * 2, d + 3), dtype=torch.float32).pin_memory()
        _PINNED_OUT[d] = buf
    host = buf[:n]
    host.copy_(packed_dev, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return _split_packed(host.clone(), d, num_scenes)       # the pinned buffer is reused by the next call


def _split_packed(packed, d, num_scenes):
    n = packed.shape[0]
    boxes, scores, labels, scene = packed[:, :d], packed[:, d], packed[:, d + 1].long(), packed[:, d + 2].long()
    cnt = torch.bincount(scene, minlength=num_scenes).tolist() if n else [0] * num_scenes
    out, lo = [], 0
    for b in range(num_scenes):
        out.append((boxes[lo:lo + cnt[b]], scores[lo:lo + cnt[b]], labels[lo:lo + cnt[b]]))
        lo += cnt[b]
    return out


class DetectionRead:
    """Asynchronous read-back of one step's detections: the packed copy (count + the first `cap` rows) is enqueued on a side
    stream behind the step's last kernel, so the caller can enqueue the NEXT step before it looks at the result
    (`finish()`); the GPU never waits for the host between two steps.

    Staging: per (device, box width, count dtype) a pool of pinned buffers; a read takes a free one (or allocates one: the
    pool grows to the largest number of reads ever outstanding) and `finish()` hands out a copy, never the pinned rows,
    and puts the buffer back."""
    CAP = 16384
    _streams = {}      # device index -> side stream
    _pools = {}        # (device index, box width, count dtype) -> free (rows, count) pinned buffers
    _lock = threading.Lock()   # reads may be started from the detector's lane threads (submit)

    def __init__(self, dets, num_scenes):
        self.dets, self.num_scenes, self.d = dets, num_scenes, dets["box_dim"]
        dev = dets["boxes"].device
        di = dev.index if dev.index is not None else torch.cuda.current_device()
        with DetectionRead._lock:
            stream = DetectionRead._streams.get(di)
            if stream is None:
                stream = DetectionRead._streams[di] = torch.cuda.Stream(device=dev)
            self._pool = DetectionRead._pools.setdefault((di, self.d, str(dets["count"].dtype)), [])
            self._buf = self._pool.pop() if self._pool else None
        if self._buf is None:
            self._buf = (torch.empty((self.CAP, self.d + 3), dtype=torch.float32).pin_memory(),
                         torch.empty((1,), dtype=dets["count"].dtype).pin_memory())
        full, self.count_host = self._buf
        done = torch.cuda.Event()
        done.record()
        cap = min(self.CAP, int(dets["boxes"].shape[0]))
        self.cap = cap
        with torch.cuda.stream(stream):
            stream.wait_event(done)
            wait_ready(dets, stream)
            packed = torch.cat((dets["boxes"][:cap, :self.d], dets["scores"][:cap, None], dets["labels"][:cap, None].float(),
                                dets["scene"][:cap, None].float()), dim=1)
            self.host = full[:cap]
            if cap:
                self.host.copy_(packed, non_blocking=True)
            self.count_host.copy_(dets["count"].reshape(1), non_blocking=True)
            self.event = torch.cuda.Event()
            self.event.record()
            self._keep = packed

    def finish(self):
        if self._buf is None:                 # a second call: the staging buffer has gone back to the pool
            return self._out
        self.event.synchronize()
        n = int(self.count_host[0])
        out = None if n > self.cap else _split_packed(self.host[:n].clone(), self.d, self.num_scenes)
        if self._buf is not None:             # the staging buffer goes back to the pool (once)
            with DetectionRead._lock:
                self._pool.append(self._buf)
            self._buf = None
        if out is None:                       # more detections than the staged rows: fall back to the blocking read
            out = split_detections(self.dets, self.num_scenes)
        self._out = out
        return out


class PendingDetections:
    """Handle of a step handed to a lane thread (SingleStageSparse3DDetector.submit)."""

    def __init__(self, num_scenes, read):
        self.num_scenes, self.read = num_scenes, read
        self._done = threading.Event()
        self._dets = self._reader = self._error = None

    def dets(self):
        """device-side result dict (blocks until the lane thread has ENQUEUED the step; see wait_ready for its completion)"""
        self._done.wait()
        if self._error is not None:
            raise self._error
        return self._dets

    def result(self):
        """per-scene detections on the host (list of (boxes, scores, labels)), like simple_test_fused(to_host=True)"""
        d = self.dets()
        return self._reader.finish() if self._reader is not None else split_detections(d, self.num_scenes)


class _LaneWorker(threading.Thread):
    """One host thread per executor lane: the executor's call blocks on row-count reads from ITS coordinate stream (the GIL
    is released inside the library call), so with one thread for all lanes the host needed as long to enqueue a step as
    the GPU to run one; with a thread per lane the enqueues of consecutive steps overlap."""

    def __init__(self, engine, stream, dev):
        super().__init__(daemon=True)
        self.engine, self.stream, self.dev = engine, stream, dev
        self.jobs = queue.Queue(maxsize=1)     # one step being enqueued + one waiting per lane
        self.start()

    def run(self):
        torch.cuda.set_device(self.dev)
        while True:
            job = self.jobs.get()
            if job is None:
                return
            points, inputs_ready, after, pending = job
            try:
                with torch.no_grad(), torch.cuda.stream(self.stream):
                    self.stream.wait_event(after)       # whatever the caller queued before submit() comes first
                    dets = self.engine(points, inputs_ready=inputs_ready)
                    if "ready" not in dets:
                        dets["ready"] = torch.cuda.Event()
                        dets["ready"].record(self.stream)
                    pending._dets = dets
                    if pending.read:
                        pending._reader = DetectionRead(dets, pending.num_scenes)
            except BaseException as exc:            # handed to the caller of dets() / result()
                pending._error = exc
            pending._done.set()


# =============================================================================================== detector
class SingleStageSparse3DDetector(nn.Module):
    def __init__(self, backbone, neck_with_head, voxel_size, pretrained=False, train_cfg=None, test_cfg=None):
        super().__init__()
        backbone = dict(backbone)
        backbone.pop("type", None)
        neck_with_head = dict(neck_with_head)
        neck_with_head.pop("type", None)
        neck_with_head.update(train_cfg=train_cfg, test_cfg=test_cfg)
        self.backbone = MEResNet3D(**backbone)
        self.neck_with_head = Fcaf3DNeckWithHead(**neck_with_head)
        self.voxel_size = voxel_size
        self.train_cfg, self.test_cfg = train_cfg, test_cfg
        self.use_engine = os.environ.get("B2_ENGINE", "1") != "0"     # native executor for eval-mode inference
        self._engine = None
        # lanes > 1: consecutive simple_test_fused calls alternate between `lanes` executors, each with its own feature
        # stream, so two steps are in flight at once and the CTAs of one fill the SMs the other leaves idle (the deep
        # levels run 84 - 99 CTAs on 148 SMs, every layer ends in a partial wave).  Results carry a `ready` event.
        self.lanes = int(os.environ.get("B2_LANES", "2"))      # measured: 5.29 -> 4.88 ms per step with 2, no more with 3
        self._lane_engines, self._lane_streams, self._lane_next = [], [], 0
        self._lane_workers = []                                # host threads of submit(), one per lane
        self.init_weights()

    def init_weights(self, pretrained=None):
        self.backbone.init_weights()
        self.neck_with_head.init_weights()

    def __getstate__(self):
        """copy.deepcopy / pickle of the module: executors, their streams and lane threads are per-process run-time state
        (library handles, CUDA streams, threads) and are rebuilt on first use"""
        state = dict(self.__dict__)
        state.update(_engine=None, _lane_engines=[], _lane_streams=[], _lane_workers=[], _lane_next=0)
        return state

    def _input_tensor(self, points):
        coords, feats = ops.quantize_points([p.contiguous() for p in points], self.voxel_size, 255.0)
        stem = self.backbone.conv1[0]
        coords._b2_prefetch_stride = int(stem.stride) if isinstance(stem.stride, int) else int(stem.stride[0])
        return ME.SparseTensor(coordinates=coords, features=feats)

    def prepare(self, points, training=True):
        """Input tensor of a batch with all backbone coordinate / kernel maps already built (MEResNet3D.prebuild_maps);
        forward_train / extract_feat accept it in place of the point list."""
        return self.backbone.prebuild_maps(self._input_tensor(points), training=training)

    def forward_train(self, points, gt_bboxes_3d, gt_labels_3d, img_metas=None):
        """Losses of one batch (the call single_stage_sparse.py:42-49 makes), computed for all scenes at once.
        gt_bboxes_3d: list of [B_i, 7] (gravity centre, size, yaw) tensors.  `points`: list of [N_i, 6] tensors, or the
        tensor returned by prepare()."""
        x0 = points if isinstance(points, ME.SparseTensor) else self._input_tensor(points)
        return self.neck_with_head.forward_train_batched(self.backbone(x0), gt_bboxes_3d, gt_labels_3d)

    @torch.no_grad()
    def simple_test_fused(self, points, to_host=True, inputs_ready=None):
        """B200 inference path: fused backbone / neck / head and the batched tail.  In eval mode the whole step is handed to
        the native executor (fcaf3d_b200.engine: the same library calls issued from C++, bit-identical results);
        `use_engine = False`, per-kernel profiling, a prune trace or a Bottleneck backbone keep the python driver."""
        if self._engine_ok():
            if self.lanes > 1:
                dets = self._run_on_lane(points, inputs_ready)
            else:
                dets = self._engine(points, inputs_ready=inputs_ready)   # (see InferenceEngine.__call__ for inputs_ready)
            return split_detections(dets, len(points)) if to_host else dets
        x = self._input_tensor(points)
        feats = self.backbone.forward_fused(x)
        levels = self.neck_with_head.forward_fused(feats, want_max_score=True)
        dets = self.neck_with_head.get_bboxes_fused(levels, x.coordinate_manager)
        return split_detections(dets, len(points)) if to_host else dets

    def submit(self, points, read=True, inputs_ready=None):
        """Asynchronous form of simple_test_fused for serving loops: the step is handed to the next lane's host thread and a
        PendingDetections handle comes back at once; `handle.result()` gives the per-scene detections (read back through
        DetectionRead when `read`), `handle.dets()` the device-side dict.  Steps submitted from one thread are executed in
        order on alternating lanes; at most two steps per lane are outstanding (submit blocks beyond that).  Without the
        native executor (training mode, profiling, ...) the step runs synchronously in the caller's thread."""
        pending = PendingDetections(len(points), read)
        if not (self._engine_ok() and self.lanes > 1):
            pending._dets = self.simple_test_fused(points, to_host=False, inputs_ready=inputs_ready)
            if read:
                pending._reader = DetectionRead(pending._dets, len(points))
            pending._done.set()
            return pending
        dev = points[0].device
        self._make_lanes(dev)
        while len(self._lane_workers) < self.lanes:
            i = len(self._lane_workers)
            self._lane_workers.append(_LaneWorker(self._lane_engines[i], self._lane_streams[i], dev))
        lane = self._lane_next % self.lanes
        self._lane_next += 1
        after = torch.cuda.Event()
        after.record(torch.cuda.current_stream(dev))
        if inputs_ready is None:
            inputs_ready = getattr(points, "ready", None)
        self._lane_workers[lane].jobs.put((points, inputs_ready, after, pending))
        return pending

    def _make_lanes(self, dev):
        from . import engine as E
        while len(self._lane_engines) < self.lanes:
            self._lane_engines.append(self._engine if not self._lane_engines else E.InferenceEngine(self))
            self._lane_streams.append(torch.cuda.Stream(device=dev))

    def _run_on_lane(self, points, inputs_ready):
        dev = points[0].device
        self._make_lanes(dev)
        lane = self._lane_next % self.lanes
        self._lane_next += 1
        s = self._lane_streams[lane]
        s.wait_stream(torch.cuda.current_stream(dev))       # whatever the caller queued before this call comes first
        with torch.cuda.stream(s):
            dets = self._lane_engines[lane](points, inputs_ready=inputs_ready)
            if "ready" not in dets:
                dets["ready"] = torch.cuda.Event()
                dets["ready"].record(s)
        return dets

    def _engine_ok(self):
        from . import engine as E
        if not (self.use_engine and not self.training and ops.PROFILE is None and ops._TC_ENABLED and PRUNE_BEFORE_CONV
                and self.neck_with_head.trace is None and ME.STRIDED_FROM_INPUT == 1):
            return False
        if self._engine is None:
            if not E.supported(self):
                self.use_engine = False
                return False
            self._engine = E.InferenceEngine(self)
        return True

    @torch.no_grad()
    def forward_features_fused(self, points):
        x = self._input_tensor(points)
        feats = self.backbone.forward_fused(x)
        return self.neck_with_head.forward_fused(feats, want_max_score=True), x.coordinate_manager


def fcaf3d_config(dataset="scannet", **overrides):
    """Constructor arguments of the detector with the hyper-parameters of configs/fcaf3d/fcaf3d.py:3-28 and the per-dataset
    overrides (fcaf3d_scannet-3d-18class.py:4-8, fcaf3d_sunrgbd-3d-10class.py:4-7, fcaf3d_s3dis-3d-5class.py:4-8)."""
    heads = {"scannet": (18, 6), "sunrgbd": (10, 8), "s3dis": (5, 6)}
    n_classes, n_reg = heads[dataset]
    voxel_size = overrides.pop("voxel_size", 0.01)
    n_outs = overrides.pop("n_outs", 4)
    cfg = dict(
        backbone=dict(in_channels=3, depth=overrides.pop("depth", 34), n_outs=n_outs),
        neck_with_head=dict(n_classes=n_classes, in_channels=(64, 128, 256, 512)[:n_outs], out_channels=128,
                            n_reg_outs=n_reg, voxel_size=voxel_size, pts_threshold=overrides.pop("pts_threshold", 100000),
                            assigner=dict(type="Fcaf3DAssigner", limit=27, topk=18, n_scales=n_outs),
                            loss_bbox=dict(type="IoU3DLoss", loss_weight=1.0, with_yaw=(dataset == "sunrgbd"))),
        voxel_size=voxel_size, train_cfg=dict(), test_cfg=dict(nms_pre=1000, iou_thr=.5, score_thr=.01))
    assert not overrides, f"unknown overrides {overrides}"
    return cfg


def build_fcaf3d(dataset="scannet", **overrides):
    return SingleStageSparse3DDetector(**fcaf3d_config(dataset, **overrides))
