"""GPU parity of the training path: autograd of the sparse ops vs the CPU oracle's autograd, losses / assigner vs
golden vectors produced by the reference's own loss / assigner code, and one full training step (losses + parameter
gradients) vs the reference's own model code on the oracle.

Tolerances: fp32 conv gradients 1e-4 of max; bf16 tensor-core path 3e-2; focal / IoU loss values 1e-5 rel,
their gradients 1e-4 of max; full step: losses 2e-3 rel, gradients 1e-2 of max |grad| (45 layers of fp32 reassociation
+ batch statistics)."""
import os

import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from conftest import GOLDEN
from fcaf3d_b200 import losses as L
from fcaf3d_b200 import me as ME
from fcaf3d_b200.detector import build_fcaf3d
from fcaf3d_b200.testing import deterministic_init
from util import rel_err, surface_coords

pytestmark = pytest.mark.gpu


def _pair(dev, c_in, seed=0, n_points=2000, batch=2):
    c = torch.from_numpy(surface_coords(n_points, batch, seed))
    o0 = OME.SparseTensor(torch.zeros(len(c), 1), coordinates=c)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c.to(dev))
    torch.manual_seed(seed)
    f = torch.randn(len(o0), c_in)
    fo = f.clone().requires_grad_(True)
    fg = f.clone().to(dev).requires_grad_(True)
    o = OME.SparseTensor(fo, coordinate_map_key=o0.coordinate_map_key, coordinate_manager=o0.coordinate_manager)
    x = ME.SparseTensor(fg, coordinate_map_key=x0.coordinate_map_key, coordinate_manager=x0.coordinate_manager)
    return x, o, fg, fo


@pytest.mark.parametrize("cin,cout,k,stride", [(8, 16, 3, 1), (16, 8, 3, 2), (64, 64, 3, 1), (64, 128, 3, 2), (32, 64, 1, 2),
                                                (3, 64, 3, 2), (4, 32, 3, 1)])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 3e-2)])
def test_conv_autograd(dev, cin, cout, k, stride, dtype, tol):
    x, o, fg, fo = _pair(dev, cin, seed=cin + k)
    oc = OME.MinkowskiConvolution(cin, cout, kernel_size=k, stride=stride, dimension=3)
    gc = ME.MinkowskiConvolution(cin, cout, kernel_size=k, stride=stride, dimension=3)
    gc.load_state_dict(oc.state_dict())
    gc = gc.to(dev)
    ME.set_compute_dtype(dtype)
    try:
        y, ry = gc(x), oc(o)
        torch.manual_seed(1)
        w = torch.randn_like(ry.F)
        (y.F * w.to(dev)).sum().backward()
        (ry.F * w).sum().backward()
    finally:
        ME.set_compute_dtype(torch.float32)
    assert rel_err(y.F.detach(), ry.F.detach()) < tol
    assert rel_err(fg.grad, fo.grad) < tol
    assert rel_err(gc.kernel.grad, oc.kernel.grad) < tol


@pytest.mark.parametrize("cin,cout,k,stride,n_points", [(64, 64, 3, 1, 6000), (128, 128, 3, 1, 3000), (256, 64, 3, 2, 6000),
                                                        (512, 128, 3, 1, 1500), (64, 128, 1, 1, 3000), (128, 192, 3, 1, 700)])
def test_wgrad_tcgen05_matches_fp32_kernel(dev, cin, cout, k, stride, n_points):
    """b2_conv_wgrad, bf16 operands: the tcgen05 kernel (MN-major operands, split over row chunks) against the fp32
    CUDA-core kernel fed with the SAME bf16-rounded values; only the summation order differs: 2e-4 of max."""
    from fcaf3d_b200 import autograd as AG
    c = torch.from_numpy(surface_coords(n_points, 2, seed=cin + cout)).to(dev)
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c)
    cm, key = x0.coordinate_manager, x0.coordinate_map_key
    okey = key if stride == 1 else cm.stride(key, stride)
    km = cm.kernel_map(key, okey, k) if (k > 1 or stride > 1) else None
    n_in, n_out = cm.size(key), cm.size(okey)
    g = torch.Generator(device="cpu").manual_seed(3)
    a = torch.randn(n_in, cin, generator=g).to(dev).to(torch.bfloat16)
    d = torch.randn(n_out, cout, generator=g).to(dev).to(torch.bfloat16)
    dw = AG.conv_wgrad(a, d, km, k ** 3, cin, cout)
    ref = AG.conv_wgrad(a.float(), d.float(), km, k ** 3, cin, cout)
    assert dw.shape == (k ** 3, cin, cout) and torch.isfinite(dw).all()
    assert rel_err(dw, ref) < 2e-4


def test_generative_pool_norm_union_autograd(dev):
    x, o, fg, fo = _pair(dev, 16, seed=3)
    mods_o = [OME.MinkowskiConvolution(16, 32, kernel_size=3, stride=2, dimension=3), OME.MinkowskiInstanceNorm(32),
              OME.MinkowskiMaxPooling(2, 2, dimension=3), OME.MinkowskiBatchNorm(32),
              OME.MinkowskiGenerativeConvolutionTranspose(32, 32, kernel_size=2, stride=2, dimension=3)]
    mods_g = [ME.MinkowskiConvolution(16, 32, kernel_size=3, stride=2, dimension=3), ME.MinkowskiInstanceNorm(32),
              ME.MinkowskiMaxPooling(2, 2, dimension=3), ME.MinkowskiBatchNorm(32),
              ME.MinkowskiGenerativeConvolutionTranspose(32, 32, kernel_size=2, stride=2, dimension=3)]
    for g, r in zip(mods_g, mods_o):
        g.load_state_dict(r.state_dict())
        g.to(dev).train(), r.train()

    def run(mods, t, elu, relu):
        a = mods[0](t)
        b = relu(mods[1](a))
        c = mods[2](b)
        d = elu(mods[3](c))
        e = mods[4](d)           # children of the pooled map: stride 4 -> 2, same stride as `a`
        return a + e             # different maps -> union add

    yg = run(mods_g, x, ME.MinkowskiELU(), ME.MinkowskiReLU())
    yo = run(mods_o, o, OME.MinkowskiELU(), OME.MinkowskiReLU())
    assert torch.equal(yg.C.cpu(), yo.C)
    torch.manual_seed(2)
    w = torch.randn_like(yo.F)
    (yg.F * w.to(dev)).sum().backward()
    (yo.F * w).sum().backward()
    assert rel_err(yg.F.detach(), yo.F.detach()) < 1e-4
    assert rel_err(fg.grad, fo.grad) < 2e-4
    for g, r in zip(mods_g, mods_o):
        for (n1, p1), (n2, p2) in zip(g.named_parameters(), r.named_parameters()):
            # InstanceNorm.weight: ReLU and max-pool are positively homogeneous and the train-mode BatchNorm that follows
            # removes the scale again, so this gradient is ~0 (only BN's eps breaks the invariance): it is the small
            # residue of large cancelling terms and only reproducible to a few percent in fp32
            tol = 6e-2 if (isinstance(g, ME.MinkowskiInstanceNorm) and n1 == "weight") else 2e-4
            assert rel_err(p1.grad, p2.grad) < tol, n1


def test_focal_and_iou_losses_match_reference_golden(dev):
    g = np.load(os.path.join(GOLDEN, "ops_ref.npz"))
    logits = torch.from_numpy(g["focal_logits"]).to(dev).requires_grad_(True)
    labels = torch.from_numpy(g["focal_labels"]).to(dev)
    loss = L.FocalLoss(use_sigmoid=True, gamma=2.0, alpha=0.25)(logits, labels, avg_factor=37.0)
    loss.backward()
    assert abs(loss.item() - float(g["focal_loss"])) < 1e-5 * abs(float(g["focal_loss"]))
    assert rel_err(logits.grad, torch.from_numpy(g["focal_grad"])) < 1e-5
    tgt = torch.from_numpy(g["iou_tgt"]).to(dev)
    w = torch.from_numpy(g["iou_w"]).to(dev)
    # rotated: one fused kernel (value + analytic gradient); 1e-5 on the loss, 1e-4 of the largest entry on the gradient
    for with_yaw, tag, tol in ((False, "aa", 1e-4), (True, "rot", 1e-4)):
        pred = torch.from_numpy(g["iou_pred"]).to(dev).requires_grad_(True)
        p, t = (pred, tgt) if with_yaw else (pred[:, :6], tgt[:, :6])
        loss = L.IoU3DLoss(with_yaw=with_yaw)(p, t, weight=w, avg_factor=w.sum())
        loss.backward()
        assert abs(loss.item() - float(g["iou_%s_loss" % tag])) < 1e-5, tag
        assert rel_err(pred.grad, torch.from_numpy(g["iou_%s_grad" % tag])) < tol, tag
    iou = L.rotated_iou_3d(torch.from_numpy(g["iou_pred"]).to(dev), tgt)
    assert torch.allclose(iou.cpu(), torch.from_numpy(g["iou_rot_values"]), atol=1e-5)


@pytest.mark.parametrize("dataset,n_reg,param", [("scannet", 6, "fcaf3d"), ("sunrgbd", 8, "fcaf3d"), ("sunrgbd", 8, "sin-cos"),
                                                   ("sunrgbd", 7, "naive")])
def test_decoded_box_iou_equals_reference_decode_then_iou(dev, ref_models, dataset, n_reg, param):
    """b2_box_iou_decoded (decode + IoU + gradient w.r.t. the raw regression outputs, one launch) against the REFERENCE
    head's own `_bbox_pred_to_bbox` (fcaf3d_neck_with_head.py:281-330, torch autograd) followed by the IoU kernels that
    test_focal_and_iou_losses_match_reference_golden pins to the reference's golden values and gradients."""
    from fcaf3d_b200 import compat, ops
    head = compat.reference_detector(ref_models, dataset).neck_with_head
    head.yaw_parametrization = param
    g = torch.Generator().manual_seed(n_reg)
    n = 257
    pts = (torch.rand(n, 3, generator=g) * 4).to(dev)
    bp = torch.cat([torch.rand(n, 6, generator=g) + .1, torch.randn(n, n_reg - 6, generator=g)], 1).to(dev)
    with torch.no_grad():
        boxes = head._bbox_pred_to_bbox(pts, bp)
        tgt = torch.cat([boxes[:, :6], boxes[:, 6:7] if n_reg > 6 else boxes.new_zeros(n, 1)], 1)
        tgt = tgt + torch.randn(n, 7, generator=g).to(dev) * torch.tensor([.2, .2, .2, .1, .1, .1, .3], device=dev)
        tgt[:, 3:6].clamp_(min=0.05)
        tgt[::17] += 50.0                                  # some pairs do not overlap at all
    gy = torch.rand(n, generator=g).to(dev)
    rotated = n_reg > 6
    a = bp.clone().requires_grad_(True)
    iou_a = L.decoded_box_iou(pts, a, tgt, ops.YAW_MODES[param] if rotated else 0, rotated)
    (iou_a * gy).sum().backward()
    b = bp.clone().requires_grad_(True)
    dec = head._bbox_pred_to_bbox(pts, b)
    iou_b = L.rotated_iou_3d(dec, tgt) if rotated else L.axis_aligned_iou_3d(dec, tgt[:, :6].contiguous())
    (iou_b * gy).sum().backward()
    assert torch.allclose(iou_a, iou_b, atol=1e-5)
    assert float(iou_a[::17].abs().max()) == 0.0 and float(iou_a.max()) > 0.3
    assert rel_err(a.grad, b.grad) < 1e-4


def test_centerness_bce_kernel_matches_torch(dev):
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(5000, 1, generator=g) * 6).to(dev).requires_grad_(True)
    t = torch.rand(5000, 1, generator=g).to(dev)
    w = torch.rand(5000, 1, generator=g).to(dev)
    loss = L.CrossEntropyLoss(use_sigmoid=True)(x, t, weight=w, avg_factor=123.0)
    loss.backward()
    x2 = x.detach().clone().requires_grad_(True)
    want = (torch.nn.functional.binary_cross_entropy_with_logits(x2, t, reduction="none") * w).sum() / 123.0
    want.backward()
    assert abs(loss.item() - want.item()) < 1e-5 * abs(want.item())
    assert rel_err(x.grad, x2.grad) < 1e-5


def test_assigner_matches_reference_golden(dev):
    g = np.load(os.path.join(GOLDEN, "ops_ref.npz"))
    pts = [torch.from_numpy(g["asg_points_%d" % i]).to(dev) for i in range(4)]
    asg = L.Fcaf3DAssigner(limit=27, topk=18, n_scales=4)
    ct, bt, lb = asg.assign(pts, torch.from_numpy(g["asg_gt"]).to(dev), torch.from_numpy(g["asg_gt_labels"]).to(dev))
    want = torch.from_numpy(g["asg_labels"])
    assert int((want >= 0).sum()) > 20
    assert torch.equal(lb.cpu(), want)
    pos = want >= 0
    assert torch.allclose(ct.cpu()[pos], torch.from_numpy(g["asg_centerness"])[pos], atol=1e-5)
    assert torch.allclose(bt.cpu()[pos], torch.from_numpy(g["asg_bbox_targets"])[pos], atol=1e-6)
    # no ground truth -> everything background
    ct0, bt0, lb0 = asg.assign(pts, torch.zeros(0, 7, device=dev), torch.zeros(0, dtype=torch.long, device=dev))
    assert bool((lb0 == -1).all())


def test_batched_assigner_equals_per_scene(dev):
    """b2_fcaf3d_assign_batched over level-major rows == b2_fcaf3d_assign scene by scene (bit for bit), including a
    scene without ground truth."""
    asg = L.Fcaf3DAssigner(limit=27, topk=18, n_scales=4)
    S, Lv = 3, 4
    gen = torch.Generator().manual_seed(5)
    sizes = [[int(torch.randint(50, 4000 // (4 ** l) + 60, (1,), generator=gen)) for _ in range(S)] for l in range(Lv)]
    pts = [[(torch.rand(sizes[l][s], 3, generator=gen) * torch.tensor([6.0, 5.0, 2.7]) - torch.tensor([3.0, 2.5, 0.0])).to(dev)
            for s in range(S)] for l in range(Lv)]
    nbox = [6, 0, 9]
    gtb, gtl = [], []
    for s in range(S):
        c = torch.rand(nbox[s], 3, generator=gen) * torch.tensor([5.0, 4.0, 1.5]) - torch.tensor([2.5, 2.0, 0.0])
        sz = torch.rand(nbox[s], 3, generator=gen) * 1.5 + 0.4
        yaw = (torch.rand(nbox[s], 1, generator=gen) - 0.5) * 3.0
        gtb.append(torch.cat([c, sz, yaw], 1).to(dev))
        gtl.append(torch.randint(0, 10, (nbox[s],), generator=gen).to(dev))
    base, seg = 0, []
    for l in range(Lv):
        row = []
        for s in range(S):
            row.append((base, base + sizes[l][s]))
            base += sizes[l][s]
        seg.append(row)
    flat = torch.cat([pts[l][s] for l in range(Lv) for s in range(S)])
    ct, bt, lb, scene_of, n_pos = asg.assign_batched(flat, seg, gtb, gtl)
    for s in range(S):
        rct, rbt, rlb = asg.assign([pts[l][s] for l in range(Lv)], gtb[s], gtl[s])
        rows = torch.cat([torch.arange(*seg[l][s]) for l in range(Lv)]).to(dev)
        assert torch.equal(lb[rows], rlb)
        assert bool((scene_of[rows] == s).all())
        assert int(n_pos[s]) == int((rlb >= 0).sum())
        # centerness targets of background points are NaN in the reference as well (sqrt of a negative product)
        assert torch.allclose(ct[rows], rct, rtol=0, atol=0, equal_nan=True) and torch.equal(bt[rows], rbt)
    assert int(n_pos[1]) == 0 and int(n_pos.sum()) > 0


@pytest.mark.parametrize("mode", ["product_batched_loss", "reference_files"])
@pytest.mark.parametrize("name,dataset", [("train_tiny_scannet.npz", "scannet"), ("train_tiny_sunrgbd.npz", "sunrgbd")])
def test_training_step_matches_reference_golden(dev, ref_models, name, dataset, mode):
    """One training step (losses + sampled parameter gradients) against the golden the reference's own files produced on
    the CPU oracle: (1) the product's whole-batch path, (2) the reference's three files themselves (`forward_train`
    of single_stage_sparse.py:42-49 -> per-scene `_loss_single` + `Fcaf3DAssigner.assign`, unmodified) running on the GPU
    through the product's autograd-enabled ME surface and loss kernels."""
    from fcaf3d_b200 import compat
    from util import reference_twin
    g = np.load(os.path.join(GOLDEN, name))
    det = deterministic_init(build_fcaf3d(dataset), seed=1).to(dev).train()
    nb = len(g["seeds"])
    pts = [torch.from_numpy(g["points_%d" % i]).to(dev) for i in range(nb)]
    gtb = [torch.from_numpy(g["gt_%d" % i]).to(dev) for i in range(nb)]
    gtl = [torch.from_numpy(g["gt_labels_%d" % i]).to(dev) for i in range(nb)]
    ME.set_compute_dtype(torch.float32)
    if mode == "reference_files":
        det = reference_twin(ref_models, det, dataset)
        boxes = [compat.DepthInstance3DBoxes(b, origin=(.5, .5, .5)) for b in gtb]
        losses = det.forward_train(pts, boxes, gtl, [dict() for _ in pts])
    else:
        losses = det.forward_train(pts, gtb, gtl)
    for k, v in losses.items():
        assert abs(v.item() - float(g[k])) < 2e-3 * abs(float(g[k])), (k, v.item(), float(g[k]))
    sum(losses.values()).backward()
    params = dict(det.named_parameters())
    for key in [k[5:] for k in g.files if k.startswith("grad/")]:
        grad = params[key].grad
        flat = grad.reshape(-1)
        step = max(1, flat.numel() // 4096)
        want = torch.from_numpy(g["grad/" + key])
        err = rel_err(flat[::step], want)
        assert err < 1e-2, (key, err)
        assert abs(float(grad.double().norm()) - float(g["gradnorm/" + key])) <= 1e-2 * float(g["gradnorm/" + key]) + 1e-12, key


@pytest.mark.parametrize("bf16_act", [True, False], ids=["bf16_activations", "fp32_activations"])
def test_bf16_training_step_tracks_the_fp32_golden(dev, bf16_act):
    """bf16 mode (tcgen05 forward / input-gradient / weight-gradient convolutions, optionally bf16 activation storage
    between them): the losses of one step stay within 2 % of the reference's fp32 golden values; sampled gradients
    keep cosine similarity > 0.95 with the fp32 ones and deviate by < 35 % of the largest entry (measured on this tiny
    scene: neck / head kernels 3e-3 and cosine 0.99999; the deep backbone kernel 0.21-0.25 and 0.97-0.98 in BOTH
    flavours, i.e. the deviation comes from the bf16 operands of the convolutions, not from the activation storage)."""
    from fcaf3d_b200 import autograd as AG
    g = np.load(os.path.join(GOLDEN, "train_tiny_scannet.npz"))
    det = deterministic_init(build_fcaf3d("scannet"), seed=1).to(dev).train()
    nb = len(g["seeds"])
    pts = [torch.from_numpy(g["points_%d" % i]).to(dev) for i in range(nb)]
    gtb = [torch.from_numpy(g["gt_%d" % i]).to(dev) for i in range(nb)]
    gtl = [torch.from_numpy(g["gt_labels_%d" % i]).to(dev) for i in range(nb)]
    ME.set_compute_dtype(torch.bfloat16)
    old = AG.BF16_ACTIVATIONS
    AG.BF16_ACTIVATIONS = bf16_act
    try:
        prepared = det.prepare(pts)                      # the pipelined trainer's entry: maps prebuilt
        losses = det.forward_train(prepared, gtb, gtl)
        sum(losses.values()).backward()
    finally:
        ME.set_compute_dtype(torch.float32)
        AG.BF16_ACTIVATIONS = old
    for k, v in losses.items():
        assert abs(v.item() - float(g[k])) < 2e-2 * abs(float(g[k])), (k, v.item(), float(g[k]))
    params = dict(det.named_parameters())
    errs = {}
    for key in ("backbone.layer2.1.conv2.kernel", "neck_with_head.out_block_0.0.kernel", "neck_with_head.cls_conv.kernel"):
        flat = params[key].grad.reshape(-1)
        step = max(1, flat.numel() // 4096)
        want = torch.from_numpy(g["grad/" + key])
        got = flat[::step].float().cpu()
        cos = torch.nn.functional.cosine_similarity(got, want, dim=0).item()
        errs[key] = (rel_err(got, want), cos)
    print("bf16_act=%s" % bf16_act, errs)
    for key, (err, cos) in errs.items():
        assert cos > 0.95 and err < 0.35, (key, err, cos)


@pytest.mark.parametrize("act", [0, 1, 2], ids=["none", "relu", "elu"])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("n,c,with_res", [(50000, 64, True), (3001, 128, False), (700, 512, True)])
def test_fused_batchnorm_act_matches_torch(dev, act, dtype, tol, n, c, with_res):
    """BNActFn (bn_ops.cu) against nn.BatchNorm1d + residual add + ReLU / ELU in torch autograd: output, gradients of x,
    weight, bias and residual, running statistics and num_batches_tracked.  fp32: 2e-5 of max; bf16 rows: 2e-2."""
    from fcaf3d_b200 import autograd as AG
    from fcaf3d_b200 import ops
    g = torch.Generator().manual_seed(n + c + act)
    x0 = (torch.randn(n, c, generator=g) * 1.7 + 0.8).to(dev).to(dtype)
    r0 = torch.randn(n, c, generator=g).to(dev).to(dtype) if with_res else None
    gy = torch.randn(n, c, generator=g).to(dev).to(dtype)
    bn_a, bn_b = torch.nn.BatchNorm1d(c).to(dev), torch.nn.BatchNorm1d(c).to(dev)
    with torch.no_grad():
        bn_a.weight.copy_(torch.rand(c, generator=g) + 0.5)
        bn_a.bias.copy_(torch.randn(c, generator=g) * 0.1)
        bn_a.running_mean.copy_(torch.randn(c, generator=g) * 0.1)
        bn_b.load_state_dict(bn_a.state_dict())
    # ours
    xa = x0.clone().requires_grad_(True)
    ra = r0.clone().requires_grad_(True) if with_res else None
    assert AG.bn_act_supported(xa, bn_a)
    ya = AG.BNActFn.apply(xa, bn_a.weight, bn_a.bias, ra, bn_a, act)
    (ya.float() * gy.float()).sum().backward()
    # torch, fp32 maths on the same (rounded) inputs
    xb = x0.float().clone().requires_grad_(True)
    rb = r0.float().clone().requires_grad_(True) if with_res else None
    yb = bn_b(xb)
    if with_res:
        yb = yb + rb
    yb = torch.relu(yb) if act == 1 else torch.nn.functional.elu(yb) if act == 2 else yb
    (yb * gy.float()).sum().backward()
    assert ya.dtype == dtype and rel_err(ya, yb) < tol
    assert rel_err(xa.grad, xb.grad) < tol
    assert rel_err(bn_a.weight.grad, bn_b.weight.grad) < tol and rel_err(bn_a.bias.grad, bn_b.bias.grad) < tol
    if with_res:
        assert rel_err(ra.grad, rb.grad) < tol
    assert rel_err(bn_a.running_mean, bn_b.running_mean) < 1e-5 and rel_err(bn_a.running_var, bn_b.running_var) < 1e-5
    assert int(bn_a.num_batches_tracked) == int(bn_b.num_batches_tracked) == 1


@pytest.mark.parametrize("act", [0, 1], ids=["none", "relu"])
def test_fused_instance_norm_matches_torch_graph(dev, act):
    """InstanceNormActFn (per-scene BatchNorm kernels on row slices) against the per-scene torch graph it replaces
    (MinkowskiInstanceNorm with AG.FUSED_BN off): output and gradients of x, weight and bias, one scene empty."""
    from fcaf3d_b200 import autograd as AG
    from fcaf3d_b200 import ops
    c = torch.from_numpy(surface_coords(3000, 3, seed=2))
    c = c[c[:, 0] != 1].contiguous()                     # scene 1 has no rows
    x0 = ME.SparseTensor(torch.zeros(len(c), 1, device=dev), coordinates=c.to(dev))
    torch.manual_seed(3)
    f = (torch.randn(len(x0), 64, device=dev) * 2 + 0.5)
    gy = torch.randn(len(x0), 64, device=dev)
    outs = []
    for fused in (True, False):
        old = AG.FUSED_BN
        AG.FUSED_BN = fused
        try:
            inorm = ME.MinkowskiInstanceNorm(64).to(dev)
            with torch.no_grad():
                inorm.weight.copy_(torch.linspace(0.5, 1.5, 64).reshape(1, 64))
                inorm.bias.copy_(torch.linspace(-0.2, 0.2, 64).reshape(1, 64))
            fi = f.clone().requires_grad_(True)
            x = ME.SparseTensor(fi, coordinate_map_key=x0.coordinate_map_key, coordinate_manager=x0.coordinate_manager)
            y = inorm(x, act=act).F
            (y * gy).sum().backward()
            outs.append((y.detach(), fi.grad, inorm.weight.grad, inorm.bias.grad))
        finally:
            AG.FUSED_BN = old
    for a, b in zip(*outs):
        assert a.shape == b.shape and rel_err(a, b) < 5e-5


def test_fused_adamw_with_clipping_matches_torch(dev):
    """FusedAdamW (b2_grad_sqnorm + b2_adamw_step) against torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW over four
    steps on tensors from 1 element to ~1 M elements: parameters to 1e-6 of max, both moments to 1e-5 (summation order of
    the norm and FMA contraction differ from torch's kernels by a few ulp), gradient norm to 1e-6 relative; two steps
    clip (norm > max_norm), two do not."""
    from fcaf3d_b200.optim import FusedAdamW
    g = torch.Generator().manual_seed(0)
    shapes = [(27,), (3, 9), (64,), (27, 64, 64), (1, 128, 25), (27, 128, 256), (1,)]
    pa = [torch.nn.Parameter(torch.randn(s, generator=g).to(dev)) for s in shapes]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    oa = FusedAdamW(pa, lr=1e-3, weight_decay=1e-4, max_norm=10.0)
    ob = torch.optim.AdamW(pb, lr=1e-3, weight_decay=1e-4)
    for it, gscale in enumerate((1.0, 1e-3, 5.0, 5e-3)):      # gradient norm ~ 1050 x gscale
        grads = [torch.randn(s, generator=g).to(dev) * gscale for s in shapes]
        for p, q, gr in zip(pa, pb, grads):
            p.grad, q.grad = gr.clone(), gr.clone()
        sq = oa.step()
        norm = torch.nn.utils.clip_grad_norm_(pb, 10.0)
        ob.step()
        assert abs(float(sq.sqrt()) - float(norm)) < 1e-6 * float(norm)
        assert (float(norm) > 10.0) == (gscale >= 0.1)
        for i, (p, q) in enumerate(zip(pa, pb)):
            assert rel_err(p, q) < 1e-6, (it, i)
            st = ob.state[q]
            assert rel_err(oa.exp_avg[i], st["exp_avg"]) < 1e-5 and rel_err(oa.exp_avg_sq[i], st["exp_avg_sq"]) < 1e-5
    sd = oa.state_dict()
    assert int(sd["state"][0]["step"]) == 4 and len(sd["state"]) == len(shapes)
