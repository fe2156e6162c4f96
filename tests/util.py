"""Shared helpers of the parity tests (seeded sparse inputs, canonical comparisons)."""
import numpy as np
import torch


def random_coords(n, batch=2, extent=40, seed=0, dup_frac=0.1, ts=1):
    """int32 (b, x, y, z) rows grouped by batch, negative coordinates included, some exact duplicates."""
    rng = np.random.default_rng(seed)
    rows = []
    for b in range(batch):
        c = rng.integers(-extent, extent, size=(n, 3)) * ts
        k = int(n * dup_frac)
        if k:
            c[rng.integers(0, n, k)] = c[rng.integers(0, n, k)]
        rows.append(np.concatenate([np.full((n, 1), b), c], 1))
    return np.concatenate(rows).astype(np.int32)


def surface_coords(n_points=3000, batch=2, seed=0):
    """Voxelised synthetic rooms (surface-like sparsity, realistic neighbour counts)."""
    from fcaf3d_b200.scenes import make_scene
    rows = []
    for b in range(batch):
        p = make_scene("tiny", seed + 100 * b, n_points)[0]
        c = np.floor(p[:, :3] / 0.01).astype(np.int32)
        rows.append(np.concatenate([np.full((len(c), 1), b, dtype=np.int32), c], 1))
    return np.concatenate(rows).astype(np.int32)


def sort_rows(coords, *others):
    """Canonical order of coordinate rows (and row-aligned companions)."""
    from fcaf3d_b200.testing import canonical_order
    order = canonical_order(torch.as_tensor(coords).cpu())
    return [torch.as_tensor(coords).cpu()[order]] + [torch.as_tensor(o).cpu()[order] for o in others]


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp(min=1e-12)).item()


def reference_twin(ref_models, det, dataset, **overrides):
    """The REFERENCE's detector class (its own three files, fcaf3d_b200.compat) with the same hyper-parameters, weights,
    device and train / eval mode as the product detector `det`."""
    from fcaf3d_b200 import compat
    ref = compat.reference_detector(ref_models, dataset, **overrides)
    missing = ref.load_state_dict(det.state_dict(), strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    ref = ref.to(next(det.parameters()).device)
    return ref.train() if det.training else ref.eval()


def reference_detections(ref, x, n_scenes):
    """get_bboxes of the reference head -> [(boxes [n, 6|7] gravity centre, scores, labels)] like the fused path returns"""
    from fcaf3d_b200 import compat
    metas = [dict(box_type_3d=compat.DepthInstance3DBoxes)] * n_scenes
    out = []
    for boxes, scores, labels in ref.neck_with_head.get_bboxes(*x, metas):
        t = boxes.tensor.clone()
        t[:, 2] += t[:, 5] * 0.5
        out.append((t if boxes.with_yaw else t[:, :6], scores, labels))
    return out
