"""Generates the committed golden fixtures by running the REFERENCE's own python files
(/root/reference/mmdet3d/models/{backbones/me_resnet.py,dense_heads/fcaf3d_neck_with_head.py,
detectors/single_stage_sparse.py}, models/losses/iou3d_loss.py, ops/rotated_iou/*.py, loaded unmodified by
oracle/ref_loader.py) on the CPU oracle.  Run in the build container only:

    python tests/golden/make_golden.py

MinkowskiEngine itself is absent, so these vectors pin OUR restatement of ME + the reference's real model code; the
pcdet_nms arithmetic is pinned separately against the reference's compiled sources (oracle/_ref).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fcaf3d_b200.scenes import make_scene  # noqa: E402
from fcaf3d_b200.testing import canonical_order, deterministic_init  # noqa: E402
from oracle import ref_loader  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
SAMPLE = 256


def detector_fixture(config, seeds, name, n_points=3000):
    det = ref_loader.build_reference_detector(config)
    deterministic_init(det, seed=0, cls_bias=0.0)
    det.eval()
    pts = [make_scene("tiny", s, n_points)[0] for s in seeds]
    with torch.no_grad():
        x = det.extract_feat([torch.from_numpy(p) for p in pts], None)
        x = [list(a) for a in x]
        metas = [dict(box_type_3d=ref_loader.DepthBoxes)] * len(seeds)
        res = det.neck_with_head.get_bboxes(*x, metas)
    centernesses, bbox_preds, cls_scores, points = x
    out = {"seeds": np.array(seeds), "n_points": np.array(n_points), "voxel_size": np.array(det.voxel_size)}
    for i, p in enumerate(pts):
        out["points_%d" % i] = p
    for lvl in range(len(points)):
        for b in range(len(seeds)):
            vox = torch.round(points[lvl][b] / det.voxel_size).int()
            order = canonical_order(vox)
            feat = torch.cat([centernesses[lvl][b], bbox_preds[lvl][b], cls_scores[lvl][b]], 1)[order]
            vox = vox[order]
            step = max(1, len(vox) // SAMPLE)
            out["l%d_b%d_n" % (lvl, b)] = np.array(len(vox))
            out["l%d_b%d_vox" % (lvl, b)] = vox[::step].numpy()
            out["l%d_b%d_feat" % (lvl, b)] = feat[::step].numpy()
            out["l%d_b%d_colsum" % (lvl, b)] = feat.double().sum(0).numpy()
            out["l%d_b%d_abssum" % (lvl, b)] = feat.double().abs().sum(0).numpy()
    for b, (boxes, scores, labels) in enumerate(res):
        bt = boxes.tensor.clone()
        bt[:, 2] += bt[:, 5] * 0.5  # back to gravity centre, as the head produced them
        out["det%d_boxes" % b] = bt.numpy()
        out["det%d_scores" % b] = scores.numpy()
        out["det%d_labels" % b] = labels.numpy()
    np.savez_compressed(os.path.join(OUT, name), **out)
    print(name, {k: v.shape for k, v in out.items() if k.startswith("det")})


def ops_fixture():
    ns = ref_loader.load()
    g = torch.Generator().manual_seed(7)
    out = {}
    # docstring known-answer of axis_aligned_bbox_overlaps_3d (iou3d_calculator.py:228-248)
    b1 = torch.tensor([[0, 0, 0, 10, 10, 10], [10, 10, 10, 20, 20, 20], [32, 32, 32, 38, 40, 42.]])
    b2 = torch.tensor([[0, 0, 0, 10, 20, 20], [0, 10, 10, 10, 19, 20], [10, 10, 10, 20, 20, 20.]])
    out["aa_b1"], out["aa_b2"] = b1.numpy(), b2.numpy()
    out["aa_iou"] = ns.axis_aligned_bbox_overlaps_3d(b1, b2).numpy()
    # _bbox_pred_to_bbox for the three parametrisations the head supports
    head = ref_loader.build_reference_detector("fcaf3d_sunrgbd-3d-10class.py").neck_with_head
    pts = torch.rand(64, 3, generator=g) * 4
    bp8 = torch.cat([torch.rand(64, 6, generator=g) + .1, torch.randn(64, 2, generator=g)], 1)
    out["dec_points"], out["dec_pred8"] = pts.numpy(), bp8.numpy()
    out["dec_fcaf3d"] = head._bbox_pred_to_bbox(pts, bp8).numpy()
    head.yaw_parametrization = "sin-cos"
    out["dec_sincos"] = head._bbox_pred_to_bbox(pts, bp8).numpy()
    head.yaw_parametrization = "naive"
    out["dec_naive"] = head._bbox_pred_to_bbox(pts, bp8[:, :7]).numpy()
    out["dec_aligned"] = head._bbox_pred_to_bbox(pts, bp8[:, :6]).numpy()
    # IoU losses (value + gradient wrt prediction)
    n = 96
    tgt = torch.cat([torch.rand(n, 3, generator=g) * 3, torch.rand(n, 3, generator=g) + .3,
                     (torch.rand(n, 1, generator=g) - .5) * 3], 1)
    pred = (tgt + torch.randn(n, 7, generator=g) * torch.tensor([.2, .2, .2, .15, .15, .15, .3])).requires_grad_(True)
    w = torch.rand(n, generator=g)
    for with_yaw, tag in ((True, "rot"), (False, "aa")):
        loss_fn = ns.IoU3DLoss(with_yaw=with_yaw, loss_weight=1.0)
        p = pred if with_yaw else pred[:, :6]
        t = tgt if with_yaw else tgt[:, :6]
        loss = loss_fn(p, t, weight=w, avg_factor=w.sum())
        grad, = torch.autograd.grad(loss, pred, allow_unused=True)
        out["iou_%s_loss" % tag], out["iou_%s_grad" % tag] = loss.detach().numpy(), grad.numpy()
    out["iou_pred"], out["iou_tgt"], out["iou_w"] = pred.detach().numpy(), tgt.numpy(), w.numpy()
    out["iou_rot_values"] = ns.cal_iou_3d(pred.detach()[None], tgt[None])[0].numpy()
    # focal / centerness-BCE
    logits = torch.randn(200, 18, generator=g).requires_grad_(True)
    labels = torch.randint(-1, 18, (200,), generator=g)
    fl = ns.FocalLoss(use_sigmoid=True, gamma=2.0, alpha=0.25)(logits, labels, avg_factor=37.0)
    out["focal_logits"], out["focal_labels"] = logits.detach().numpy(), labels.numpy()
    out["focal_loss"] = fl.detach().numpy()
    out["focal_grad"] = torch.autograd.grad(fl, logits)[0].numpy()
    # assigner on a tiny scene (labels / centerness targets / box targets)
    pts_np, gt, gl = make_scene("tiny", 3, 3000)
    gtb = ns.DepthBoxes(torch.from_numpy(gt), origin=(.5, .5, .5))
    lv_pts = []
    for lvl, ts in enumerate((8, 16, 32, 64)):
        v = np.unique(np.floor(pts_np[:, :3] / 0.01 / ts).astype(np.int64), axis=0) * ts
        lv_pts.append(torch.from_numpy(v.astype(np.float32)) * 0.01)
    asg = ns.Fcaf3DAssigner(limit=27, topk=18, n_scales=4)
    ct, bt, lb = asg.assign(lv_pts, gtb, torch.from_numpy(gl))
    for i, p in enumerate(lv_pts):
        out["asg_points_%d" % i] = p.numpy()
    out["asg_gt"], out["asg_gt_labels"] = gt, gl
    out["asg_centerness"], out["asg_bbox_targets"], out["asg_labels"] = ct.numpy(), bt.numpy(), lb.numpy()
    np.savez_compressed(os.path.join(OUT, "ops_ref.npz"), **out)
    print("ops_ref", int((lb >= 0).sum()), "positives of", len(lb))


GRAD_KEYS = ["backbone.conv1.0.kernel", "backbone.conv1.1.weight", "backbone.layer1.0.conv1.kernel",
             "backbone.layer1.0.downsample.0.kernel", "backbone.layer1.0.norm1.bn.weight", "backbone.layer2.1.conv2.kernel",
             "backbone.layer4.2.conv2.kernel", "neck_with_head.up_block_1.0.kernel", "neck_with_head.up_block_3.3.kernel",
             "neck_with_head.out_block_0.0.kernel", "neck_with_head.out_block_2.1.bn.bias",
             "neck_with_head.centerness_conv.kernel", "neck_with_head.reg_conv.kernel", "neck_with_head.cls_conv.kernel",
             "neck_with_head.cls_conv.bias", "neck_with_head.scales.0.scale", "neck_with_head.scales.3.scale"]


def train_fixture(config, seeds, name, n_points=3000):
    """losses + parameter gradients of one training step of the reference's own model / loss / assigner code"""
    ns = ref_loader.load()
    det = ref_loader.build_reference_detector(config)
    deterministic_init(det, seed=1)
    det.train()
    scenes = [make_scene("tiny", s, n_points, n_classes=det.neck_with_head.cls_conv.out_channels) for s in seeds]
    if "sunrgbd" in config:  # rotated ground truth for the rotated-IoU loss
        rng = np.random.default_rng(11)
        for _, gt, _ in scenes:
            gt[:, 6] = rng.uniform(-0.6, 0.6, len(gt))
    pts = [torch.from_numpy(p) for p, _, _ in scenes]
    gtb = [ns.DepthBoxes(torch.from_numpy(g), origin=(.5, .5, .5)) for _, g, _ in scenes]
    gtl = [torch.from_numpy(l) for _, _, l in scenes]
    losses = det.forward_train(pts, gtb, gtl, [dict() for _ in seeds])
    total = sum(losses.values())
    total.backward()
    out = {"seeds": np.array(seeds), "n_points": np.array(n_points)}
    for i, (p, g, l) in enumerate(scenes):
        out["points_%d" % i], out["gt_%d" % i], out["gt_labels_%d" % i] = p, g, l
    for k, v in losses.items():
        out[k] = v.detach().numpy()
    params = dict(det.named_parameters())
    for k in GRAD_KEYS:
        g = params[k].grad
        flat = g.reshape(-1)
        step = max(1, flat.numel() // 4096)
        out["grad/" + k] = flat[::step].numpy()
        out["gradnorm/" + k] = np.array(float(g.double().norm()))
    np.savez_compressed(os.path.join(OUT, name), **out)
    print(name, {k: float(v) for k, v in losses.items()})


def mefpn_fixture(seeds=(3, 103), n_points=3000, name="mefpn_tiny.npz"):
    """The reference's own necks/me_fpn.py (MEFPN3D with score blocks and pruning) behind its own MEResNet3D, on the
    oracle: evaluation outputs per level and one training forward/backward (pruning loss + sampled gradients)."""
    ns = ref_loader.load()
    torch.manual_seed(0)
    model = torch.nn.Module()
    model.backbone = ns.MEResNet3D(in_channels=3, depth=34)
    model.neck = ns.MEFPN3D(in_channels=(64, 128, 256, 512), out_channels=128, voxel_size=.01, pts_threshold=400)
    deterministic_init(model, seed=2)
    import MinkowskiEngine as OME
    scenes = [make_scene("tiny", s, n_points, n_classes=18) for s in seeds]
    out = {"seeds": np.array(seeds), "n_points": np.array(n_points)}
    for i, (p, g, l) in enumerate(scenes):
        out["points_%d" % i], out["gt_%d" % i] = p, g

    def feats():
        coords, f = OME.utils.batch_sparse_collate(
            [(torch.from_numpy(p[:, :3] / .01), torch.from_numpy(p[:, 3:] / 255.)) for p, _, _ in scenes])
        return model.backbone(OME.SparseTensor(coordinates=coords, features=f))

    model.eval()
    with torch.no_grad():
        outs = model.neck(feats(), img_metas=[dict() for _ in seeds])
    for lvl, o in enumerate(outs):
        order = canonical_order(o.C)
        c, f = o.C[order], o.F[order]
        step = max(1, len(c) // SAMPLE)
        out["eval_l%d_n" % lvl] = np.array(len(c))
        out["eval_l%d_coords" % lvl] = c[::step].numpy()
        out["eval_l%d_feat" % lvl] = f[::step].numpy()
        out["eval_l%d_colsum" % lvl] = f.double().sum(0).numpy()
    model.train()
    gtb = [ns.DepthBoxes(torch.from_numpy(g), origin=(.5, .5, .5)) for _, g, _ in scenes]
    gtl = [torch.from_numpy(l) for _, _, l in scenes]
    outs, losses = model.neck(feats(), gtb, gtl, [dict() for _ in seeds])
    total = losses["loss_pruning"] + sum(o.F.square().mean() for o in outs)
    total.backward()
    out["loss_pruning"] = losses["loss_pruning"].detach().numpy()
    out["train_rows"] = np.array([len(o.F) for o in outs])
    params = dict(model.named_parameters())
    for k in ("neck.score_block_1.kernel", "neck.score_block_3.bias", "neck.up_block_2.0.kernel", "neck.out_block_0.0.kernel",
              "backbone.layer3.0.conv1.kernel"):
        g = params[k].grad.reshape(-1)
        step = max(1, g.numel() // 4096)
        out["grad/" + k] = g[::step].numpy()
    np.savez_compressed(os.path.join(OUT, name), **out)
    print(name, float(out["loss_pruning"]), out["train_rows"], [int(out["eval_l%d_n" % l]) for l in range(4)])


def augment_fixture(name="augment_ref.npz"):
    """RandomFlip3D / GlobalRotScaleTrans arithmetic through the reference's OWN containers (DepthInstance3DBoxes,
    DepthPoints loaded by path, oracle/ref_loader.load_structures) for a few parameter sets; boxes are stored with
    their gravity centre."""
    Boxes, Points = ref_loader.load_structures()
    rng = np.random.default_rng(7)
    out = {}
    cases = [dict(with_yaw=True, flip_h=True, flip_v=True, angle=0.05, scale=1.07, trans=(0.11, -0.07, 0.03)),
             dict(with_yaw=True, flip_h=False, flip_v=True, angle=-0.08, scale=0.93, trans=(-0.2, 0.05, 0.0)),
             dict(with_yaw=False, flip_h=True, flip_v=False, angle=0.0872, scale=1.1, trans=(0.0, 0.3, -0.1)),
             dict(with_yaw=False, flip_h=False, flip_v=False, angle=-0.03, scale=0.9, trans=(0.05, 0.05, 0.05))]
    for ci, c in enumerate(cases):
        pts = (rng.uniform(-3, 3, (500, 6))).astype(np.float32)
        pts[:, 3:] = rng.uniform(0, 255, (500, 3)).astype(np.float32)
        gt = np.concatenate([rng.uniform(-2, 2, (9, 3)), rng.uniform(0.3, 2, (9, 3)), rng.uniform(-1.5, 1.5, (9, 1))],
                            1).astype(np.float32)
        if not c["with_yaw"]:
            gt[:, 6] = 0
        b = Boxes(torch.from_numpy(gt[:, :7 if c["with_yaw"] else 6].copy()), box_dim=7 if c["with_yaw"] else 6,
                  with_yaw=c["with_yaw"], origin=(.5, .5, .5))
        p = Points(torch.from_numpy(pts.copy()), points_dim=6)
        if c["flip_h"]:
            p = b.flip("horizontal", points=p)
        if c["flip_v"]:
            p = b.flip("vertical", points=p)
        p, _ = b.rotate(c["angle"], p)
        p.scale(c["scale"])
        b.scale(c["scale"])
        t = np.array(c["trans"], dtype=np.float32)
        p.translate(t)
        b.translate(t)
        res = torch.cat([b.gravity_center, b.tensor[:, 3:]], 1).numpy()
        out["c%d_points_in" % ci], out["c%d_boxes_in" % ci] = pts, gt
        out["c%d_points_out" % ci], out["c%d_boxes_out" % ci] = p.tensor.numpy(), res
        out["c%d_params" % ci] = np.array([c["with_yaw"], c["flip_h"], c["flip_v"], c["angle"], c["scale"], *c["trans"]],
                                          dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, name), **out)
    print(name, len(cases), "cases")


def eval_fixture(name="indoor_eval_ref.npz"):
    """mAP / mAR of synthetic detections computed by the reference's own indoor_eval.py (see
    oracle/ref_loader.load_indoor_eval for the two stand-ins) for an axis-aligned (ScanNet-like) and a rotated
    (SUN RGB-D-like) case."""
    indoor_eval, Boxes, Box3DMode = ref_loader.load_indoor_eval()
    out = {}
    for case, with_yaw in (("aa", False), ("rot", True)):
        rng = np.random.default_rng(3 if with_yaw else 4)
        n_img, n_cls = 8, 5
        gt_annos, dt_annos = [], []
        for i in range(n_img):
            ng = int(rng.integers(4, 10)) if i else 0           # the first image has no ground truth
            gt = np.concatenate([rng.uniform(-3, 3, (ng, 3)), rng.uniform(0.4, 1.8, (ng, 3)),
                                 rng.uniform(-1.5, 1.5, (ng, 1)) * with_yaw], 1).astype(np.float32)
            gl = rng.integers(0, n_cls - 1, ng)                  # class n_cls-1 never appears in the ground truth
            gt_annos.append(dict(gt_num=ng, gt_boxes_upright_depth=gt if with_yaw else gt[:, :6], **{"class": gl}))
            # detections: jittered copies of the ground truth (some duplicated, some mislabelled) + random boxes
            k = int(rng.integers(3, 12))
            rnd = np.concatenate([rng.uniform(-3, 3, (k, 3)), rng.uniform(0.4, 1.8, (k, 3)),
                                  rng.uniform(-1.5, 1.5, (k, 1)) * with_yaw], 1).astype(np.float32)
            rl = rng.integers(0, n_cls if with_yaw else n_cls - 1, k)   # rot: a class without ground truth (nan AP)
            rep = np.repeat(gt, 2, 0) + rng.normal(0, 0.17, (2 * ng, 7)).astype(np.float32) * np.array([1, 1, 1, 1, 1, 1, with_yaw], np.float32)
            repl = np.repeat(gl, 2)
            if len(repl):
                repl[rng.random(len(repl)) < 0.15] = rng.integers(0, n_cls - 1)
            det = np.concatenate([rep, rnd], 0).astype(np.float32)
            lab = np.concatenate([repl, rl]).astype(np.int64)
            sc = rng.uniform(0.05, 1, len(det)).astype(np.float32)
            bottom = det.copy()
            bottom[:, 2] -= bottom[:, 5] / 2
            dt_annos.append(dict(boxes_3d=Boxes(torch.from_numpy(bottom), box_dim=7, with_yaw=with_yaw),
                                 scores_3d=torch.from_numpy(sc), labels_3d=torch.from_numpy(lab)))
            out["%s_gt%d" % (case, i)], out["%s_gtl%d" % (case, i)] = gt, gl
            out["%s_det%d" % (case, i)], out["%s_detl%d" % (case, i)], out["%s_dets%d" % (case, i)] = det, lab, sc
        label2cat = {c: "c%d" % c for c in range(n_cls)}
        with ref_loader.cpu_tensors_as_cuda(), np.errstate(all="ignore"):
            ret = indoor_eval(gt_annos, dt_annos, (0.25, 0.5), label2cat, box_type_3d=Boxes, box_mode_3d=Box3DMode.DEPTH)
        keys = sorted(ret)
        out["%s_keys" % case] = np.array(keys)
        out["%s_vals" % case] = np.array([ret[k] for k in keys], dtype=np.float64)
        print(case, {k: round(v, 4) for k, v in ret.items() if k.startswith("mA")})
    np.savez_compressed(os.path.join(OUT, name), **out)


if __name__ == "__main__":
    torch.set_num_threads(8)
    if "--eval-only" in sys.argv:
        eval_fixture()
        sys.exit(0)
    if "--augment-only" in sys.argv:
        augment_fixture()
        sys.exit(0)
    if "--mefpn-only" in sys.argv:
        mefpn_fixture()
        sys.exit(0)
    if "--train-only" in sys.argv:
        train_fixture("fcaf3d_scannet-3d-18class.py", (0, 100), "train_tiny_scannet.npz")
        train_fixture("fcaf3d_sunrgbd-3d-10class.py", (5,), "train_tiny_sunrgbd.npz")
        sys.exit(0)
    ops_fixture()
    detector_fixture("fcaf3d_scannet-3d-18class.py", (0, 100), "fcaf3d_tiny_scannet.npz")
    detector_fixture("fcaf3d_sunrgbd-3d-10class.py", (5,), "fcaf3d_tiny_sunrgbd.npz")
    train_fixture("fcaf3d_scannet-3d-18class.py", (0, 100), "train_tiny_scannet.npz")
    train_fixture("fcaf3d_sunrgbd-3d-10class.py", (5,), "train_tiny_sunrgbd.npz")
    mefpn_fixture()
    augment_fixture()
    eval_fixture()
