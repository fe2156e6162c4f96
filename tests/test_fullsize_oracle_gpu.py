"""GPU vs the CPU oracle AT BENCHMARK SIZE (VERDICT r01 "weak" 1-2): the batches `bench.py` times - 8 ScanNet-shaped S0
scenes x 100 k points, 8 SUN RGB-D-shaped S3 scenes, one S3DIS-shaped S2 scene of 1 M points - go through the product's
fused inference path exactly as in the bench (same batch, same kernels, same tile variants), and single scenes of the batch
are checked against `oracle/model_cpu.py` (eval mode: scenes do not interact - InstanceNorm is per scene, BatchNorm uses
running statistics - so scene b of the batch equals that scene run alone on the oracle).

  * every coordinate map of the head levels: bit-exact (canonical order);
  * EVERY kernel map the pass built (stem strided-from-input, pooling 2^3, stride-2 3^3 / 1^3, hash-probed 3^3, derived
    children / children-pruned tables): bit-exact against the oracle's definition (kernel_offsets + lookup) on the scene's rows;
  * the prune decision (cluster radix-select top-k at >= 32 768 rows per scene): bit-exact against the oracle's rule
    (stable descending sort, ties -> lowest row) on the same scores; the interpolated scores themselves within 1e-4 of max;
  * head outputs per level: fp32 mode 2e-3 of the column's max |value| (same bound as the tiny goldens), bf16 mode 6e-2;
    the oracle prunes with the GPU's kept set so that a score tie-break of 1 ulp cannot move rows between the two sides
    (the decision itself is checked exactly, one bullet up);
  * bf16 against fp32 at the detection level: see test_bf16_detections_agree_with_fp32.
"""
import numpy as np
import pytest
import torch

import oracle.me_cpu as OME
from oracle import model_cpu
from fcaf3d_b200 import me as ME
from fcaf3d_b200.detector import build_fcaf3d
from fcaf3d_b200.scenes import make_scene
from fcaf3d_b200.testing import canonical_order, deterministic_init

pytestmark = pytest.mark.gpu

WORKLOADS = {
    # name: (dataset, preset, scenes in the batch, scenes checked against the oracle)
    "scannet_S0x8": ("scannet", "S0", 8, (0, 5)),
    "sunrgbd_S3x8": ("sunrgbd", "S3", 8, (3,)),
    "s3dis_S2x1": ("s3dis", "S2", 1, (0,)),
}


class _PinnedPruneHead(model_cpu.NeckWithHead):
    """oracle head that records its interpolated prune scores and, when told so, keeps a given set of voxels"""
    forced_keep = None      # level -> sorted int64 keys of the voxels to keep
    seen = None

    def _prune(self, x, scores):
        if self.pts_threshold < 0:
            return x
        with torch.no_grad():
            s = scores.features_at_coordinates(x.C.float()).squeeze(1)
            ts = x.tensor_stride[0]
            self.seen[ts] = (x.C.clone(), s.clone())
            if self.forced_keep is not None and ts in self.forced_keep:
                keys = OME.pack_keys(x.C.numpy())
                mask = torch.from_numpy(np.isin(keys, self.forced_keep[ts]))
                return self.pruning(x, mask)
        return super()._prune(x, scores)


def _oracle(dataset, seed=0):
    ref = model_cpu.Detector(dataset)
    head = ref.neck_with_head
    head.__class__ = _PinnedPruneHead
    head.seen, head.forced_keep = {}, {}
    return deterministic_init(ref, seed, cls_bias=0.0).eval()


def _scene_rows(m, b):
    return slice(m.off_host[b], m.off_host[b + 1])


def _check_kernel_maps(cm, b):
    """every table in the manager against the oracle's definition, restricted to scene b"""
    checked = {}
    for (in_key, out_key, ksize, dil, *flag), km in cm._kmaps.items():
        if flag:
            continue
        src, dst = cm.get_map(in_key), cm.get_map(out_key)
        si, so = _scene_rows(src, b), _scene_rows(dst, b)
        cin = src.coords[si].cpu().numpy()
        cout = dst.coords[so].cpu().numpy()
        omap = OME._Map(cin, src.ts)
        got = km.nbr[:, so].cpu().numpy().astype(np.int64)
        got = np.where(got >= 0, got - si.start, -1)
        for k, off in enumerate(OME.kernel_offsets(ksize, src.ts * dil)):
            q = cout.copy()
            q[:, 1] += off[0]
            q[:, 2] += off[1]
            q[:, 3] += off[2]
            want = omap.lookup(q)
            assert np.array_equal(got[k], want), (src.ts, dst.ts, ksize, k, int((got[k] != want).sum()))
        # rows of other scenes never appear
        assert got.max() < si.stop - si.start
        tag = "%d->%d k%d" % (src.ts, dst.ts, ksize) + (" children" if getattr(src, "children_of", None) is not None else "")
        checked[tag] = checked.get(tag, 0) + 1
    return checked


def _run_workload(dev, name, dtype):
    dataset, preset, nb, check = WORKLOADS[name]
    n_cls = {"scannet": 18, "sunrgbd": 10, "s3dis": 5}[dataset]
    scenes = [make_scene(preset, 100 * i, n_classes=n_cls)[0] for i in range(nb)]
    det = deterministic_init(build_fcaf3d(dataset), 0, cls_bias=0.0).to(dev).eval()
    det.neck_with_head.trace = {}
    ME.set_compute_dtype(dtype)
    try:
        with torch.no_grad():
            levels, cm = det.forward_features_fused([torch.from_numpy(p).to(dev) for p in scenes])
    finally:
        ME.set_compute_dtype(torch.float32)
    return dataset, scenes, det, levels, cm, check


@pytest.mark.parametrize("name", list(WORKLOADS))
def test_benchmark_batch_against_the_oracle_fp32(dev, name):
    dataset, scenes, det, levels, cm, check = _run_workload(dev, name, torch.float32)
    head = det.neck_with_head
    assert head.trace, "the benchmark shapes must prune at least one level"
    for b in check:
        kinds = _check_kernel_maps(cm, b)
        assert any("children" in k for k in kinds) and "1->2 k3" in kinds and "2->4 k2" in kinds, kinds
        # prune decision at this size: the kernel's mask == stable descending sort of the SAME scores, ties -> lowest row
        forced = {}
        for lvl, tr in head.trace.items():
            um = cm.get_map(tr["union_key"])
            sl = _scene_rows(um, b)
            sc = tr["scores"].reshape(-1)[sl].cpu()
            k = min(len(sc), head.pts_threshold)
            want = torch.zeros(len(sc), dtype=torch.bool)
            want[torch.argsort(-sc, stable=True)[:k]] = True
            assert torch.equal(tr["mask"][sl].cpu().bool(), want), (lvl, int((tr["mask"][sl].cpu().bool() != want).sum()))
            c = um.coords[sl].cpu().numpy().copy()
            c[:, 0] = 0
            forced[um.ts] = np.sort(OME.pack_keys(c[want.numpy()]))
            if len(sc) > 32768 and k < len(sc):
                kinds["cluster_topk"] = 1
        ref = _oracle(dataset)
        ref.neck_with_head.forced_keep = forced
        want_levels = ref([torch.from_numpy(scenes[b])])
        # interpolated prune scores of the oracle's own pass against the GPU's (same union rows, any order)
        for lvl, tr in head.trace.items():
            um = cm.get_map(tr["union_key"])
            sl = _scene_rows(um, b)
            oc, os_ = ref.neck_with_head.seen[um.ts]
            c = um.coords[sl].cpu().clone()
            c[:, 0] = 0
            o1, o2 = canonical_order(c), canonical_order(oc)
            assert torch.equal(c[o1], oc[o2].int())
            got, want = tr["scores"].reshape(-1)[sl].cpu()[o1], os_[o2]
            assert ((got - want).abs().max() / want.abs().max()).item() < 1e-4
        for lvl, (l, w) in enumerate(zip(levels, want_levels)):
            m = cm.get_map(l["key"])
            sl = _scene_rows(m, b)
            c = m.coords[sl].cpu().clone()
            c[:, 0] = 0
            assert len(c) == len(w["coords"]), (lvl, len(c), len(w["coords"]))
            o1, o2 = canonical_order(c), canonical_order(w["coords"])
            assert torch.equal(c[o1], w["coords"][o2].int())
            got = torch.cat([l["centerness"][sl], l["bbox_pred"][sl], l["cls"][sl]], 1).float().cpu()[o1]
            exp = torch.cat([w["centerness"], w["bbox_pred"], w["cls"]], 1)[o2]
            err = ((got - exp).abs() / exp.abs().max(0).values.clamp(min=1e-6)).max().item()
            assert err < 2e-3, (name, b, lvl, err)
        print(name, "scene", b, "kernel maps checked:", kinds)
    if name != "sunrgbd_S3x8":
        assert "cluster_topk" in kinds


def test_benchmark_batch_against_the_oracle_bf16(dev):
    """the benchmarked mode: same comparison of the head outputs with the bf16 bound (maps are dtype independent)"""
    dataset, scenes, det, levels, cm, check = _run_workload(dev, "scannet_S0x8", torch.bfloat16)
    head, b = det.neck_with_head, check[0]
    forced = {}
    for lvl, tr in head.trace.items():
        um = cm.get_map(tr["union_key"])
        sl = _scene_rows(um, b)
        c = um.coords[sl].cpu().numpy().copy()
        c[:, 0] = 0
        forced[um.ts] = np.sort(OME.pack_keys(c[tr["mask"][sl].cpu().numpy().astype(bool)]))
    ref = _oracle(dataset)
    ref.neck_with_head.forced_keep = forced
    want_levels = ref([torch.from_numpy(scenes[b])])
    for lvl, (l, w) in enumerate(zip(levels, want_levels)):
        m = cm.get_map(l["key"])
        sl = _scene_rows(m, b)
        c = m.coords[sl].cpu().clone()
        c[:, 0] = 0
        o1, o2 = canonical_order(c), canonical_order(w["coords"])
        assert torch.equal(c[o1], w["coords"][o2].int())
        got = torch.cat([l["centerness"][sl], l["bbox_pred"][sl], l["cls"][sl]], 1).float().cpu()[o1]
        exp = torch.cat([w["centerness"], w["bbox_pred"], w["cls"]], 1)[o2]
        err = ((got - exp).abs() / exp.abs().max(0).values.clamp(min=1e-6)).max().item()
        assert err < 6e-2, (lvl, err)


def _nearest(a, b):
    """per detection of `a` = (boxes, scores, labels): (max-abs box difference, |score difference|) to the same-class
    detection of `b` with the closest box (inf where the class is absent in b)"""
    ab, as_, al = [t.cpu().numpy() for t in a]
    bb, bs, bl = [t.cpu().numpy() for t in b]
    dbox, dsc = np.full(len(al), np.inf), np.full(len(al), np.inf)
    for cls in np.unique(al):
        ia, ib = np.nonzero(al == cls)[0], np.nonzero(bl == cls)[0]
        if len(ib) == 0:
            continue
        d = np.abs(ab[ia][:, None, :] - bb[ib][None, :, :]).max(-1)            # [na, nb]
        j = d.argmin(1)
        dbox[ia], dsc[ia] = d[np.arange(len(ia)), j], np.abs(as_[ia] - bs[ib][j])
    return dbox, dsc


@pytest.mark.parametrize("name", ["scannet_S0x8", "sunrgbd_S3x8"])
def test_bf16_detections_agree_with_fp32(dev, name):
    """Detection-level agreement of the benchmarked bf16 mode with the fp32 mode on the benchmark batch.  For every fp32
    detection the same-class bf16 detection with the closest box is looked up (and the other way round); asserted:
    counts within 3 %; >= 85 % have a partner within 5 cm (max-abs over the box parameters) and 0.05 in score; >= 70 %
    within 2 cm and 0.02 (measured on ScanNet-shape S0 x 8: 90 % and 80 %; the remaining ~10 % are NMS survivors that
    flipped).  These are bounds for RANDOM-INIT weights: ~3 600 of the 4 000 candidates of a scene survive NMS with scores
    in a narrow band, so a bf16-sized perturbation of the logits reorders survivors far more easily than with trained
    weights; the measured percentiles are printed (and recorded in DESIGN.md)."""
    dataset, preset, nb, _ = WORKLOADS[name]
    n_cls = {"scannet": 18, "sunrgbd": 10}[dataset]
    pts = [torch.from_numpy(make_scene(preset, 100 * i, n_classes=n_cls)[0]).to(dev) for i in range(nb)]
    det = deterministic_init(build_fcaf3d(dataset), 0, cls_bias=0.0).to(dev).eval()
    out = {}
    for dtype in (torch.float32, torch.bfloat16):
        ME.set_compute_dtype(dtype)
        try:
            out[dtype] = det.simple_test_fused(pts)
        finally:
            ME.set_compute_dtype(torch.float32)
    n32 = sum(len(a[1]) for a in out[torch.float32])
    n16 = sum(len(a[1]) for a in out[torch.bfloat16])
    stats = []
    for tag, (x, y) in (("fp32->bf16", (torch.float32, torch.bfloat16)), ("bf16->fp32", (torch.bfloat16, torch.float32))):
        dbox, dsc = [np.concatenate(v) for v in zip(*[_nearest(a, b) for a, b in zip(out[x], out[y])])]
        tight = float(((dbox <= 0.02) & (dsc <= 0.02)).mean())
        loose = float(((dbox <= 0.05) & (dsc <= 0.05)).mean())
        print(name, tag, "dets", n32, n16, "within 2cm/0.02: %.4f  within 5cm/0.05: %.4f" % (tight, loose),
              "box diff p50/p90/p99 [m]:", np.round(np.percentile(dbox, [50, 90, 99]), 4).tolist(),
              "score diff p50/p90/p99:", np.round(np.percentile(dsc, [50, 90, 99]), 4).tolist())
        stats.append((tag, tight, loose))
    for tag, tight, loose in stats:
        assert loose >= 0.85 and tight >= 0.70, (tag, tight, loose)
    assert abs(n32 - n16) < 0.03 * n32
