"""Functional host wrappers over the C ABI (include/fcaf3d_b200.h).

Tensors are torch CUDA tensors used purely as device memory; every op enqueues on the current torch
stream.  No op has a CPU path.
"""
import ctypes
import os
import threading

import torch

from . import _lib
from ._lib import B2Error, DT, call, ptr, require_cuda, workspace

ACT_NONE, ACT_RELU, ACT_ELU = 0, 1, 2
I32 = torch.int32


def next_pow2(v):
    p = 1
    while p < v:
        p <<= 1
    return p


def table_capacity(n):
    return max(1024, next_pow2(2 * int(n)))


def alloc_table(n, device):
    cap = table_capacity(n)
    return torch.empty(cap * 16, dtype=torch.uint8, device=device), cap


_PINNED = threading.local()     # one ring per host thread (map building from a thread of its own must not share slots)


def _pinned_ints(n):
    """Small pinned int32 staging buffers, reused round-robin (64 outstanding reads per thread at most)."""
    if not hasattr(_PINNED, "bufs"):
        _PINNED.bufs = [torch.empty((256,), dtype=I32).pin_memory() for _ in range(64)]
        _PINNED.next = 0
    if n > 256:
        return torch.empty((n,), dtype=I32).pin_memory()
    i = _PINNED.next
    _PINNED.next = (i + 1) % len(_PINNED.bufs)
    return _PINNED.bufs[i][:n]


class HostRead:
    """Asynchronous device->host read of a small int32 tensor: the copy is enqueued now, `get()` waits for it only.
    Work enqueued on the stream between the two keeps the GPU busy while the host looks at the numbers."""

    def __init__(self, t):
        self.host = _pinned_ints(t.numel())
        self.host.copy_(t.reshape(-1), non_blocking=True)
        self.event = torch.cuda.Event()
        self.event.record()

    def get(self):
        self.event.synchronize()
        return self.host.tolist()


def ints_to_device(values, dev):
    """Small host list -> int32 device tensor through a pinned staging buffer (asynchronous H2D, no pageable staging)."""
    n = len(values)
    if n == 0:
        return torch.empty((0,), dtype=I32, device=dev)
    host = _pinned_ints(n)
    host.copy_(torch.tensor(values, dtype=I32))
    return host.to(dev, non_blocking=True)


def read_ints(t):
    """One device->host read of an int32 tensor (waits for the current stream up to the copy)."""
    if t.dtype != I32 or not t.is_cuda:
        return t.cpu().tolist()
    return HostRead(t).get()


# ----------------------------------------------------------------------------- coordinate maps
def quantize_points(points_list, voxel_size, feat_div=255.0):
    """[(N_i, 3+F) fp32 cuda] -> coords int32 [sum N, 4], feats fp32 [sum N, F]."""
    dev = points_list[0].device
    dim = points_list[0].shape[1]
    total = sum(int(p.shape[0]) for p in points_list)
    coords = torch.empty((total, 4), dtype=I32, device=dev)
    feats = torch.empty((total, dim - 3), dtype=torch.float32, device=dev)
    row0 = 0
    for b, p in enumerate(points_list):
        require_cuda(p, "points")
        call("b2_quantize_points", ptr(p), p.shape[0], dim, b, float(voxel_size), float(feat_div), row0, ptr(coords),
             ptr(feats))
        row0 += int(p.shape[0])
    coords._b2_num_batches = len(points_list)   # rows are batch-grouped by construction: insert_and_map skips two host reads
    return coords, feats


def map_build_start(coords, quant=0, num_batches=0, want_index=True):
    """Enqueues the dedup kernels and the asynchronous read of (n_out, err, batch offsets); see map_build_finish.
    want_index=False (strided maps inside a network): no unique_index / inverse map, two kernels instead of three."""
    require_cuda(coords, "coords")
    n = int(coords.shape[0])
    dev = coords.device
    table, cap = alloc_table(n, dev)
    out_coords = torch.empty((n, 4), dtype=I32, device=dev)
    unique_index = torch.empty((n,), dtype=I32, device=dev) if want_index else None
    inverse = torch.empty((n,), dtype=I32, device=dev) if want_index else None
    status = torch.zeros((2 + num_batches + 1,), dtype=I32, device=dev)  # [n_out, err, offsets...]
    ws_bytes = _lib.lib().b2_map_build_ws_bytes(n)
    ws = workspace(ws_bytes, dev)
    t0 = _prof_start()
    call("b2_map_build", ptr(coords), n, int(quant), ptr(table), cap, ptr(out_coords), ptr(unique_index), ptr(inverse),
         ptr(status), ptr(status) + 4, ptr(ws), ws.numel())
    if num_batches > 0:
        call("b2_batch_offsets_dev", ptr(out_coords), ptr(status), n, num_batches, ptr(status) + 8, ptr(status) + 4)
    _prof_stop(t0, "hash_build", 0.0, 40.0 * n)
    return dict(read=HostRead(status), status=status, out_coords=out_coords, unique_index=unique_index, inverse=inverse,
                table=table, cap=cap, num_batches=num_batches)


def map_build_finish(p):
    host = p["read"].get()
    n_out, err = host[0], host[1]
    if err == 1:
        raise _lib.B2Error("coordinate outside the packed key range (batch < 1023, |xyz| < 131072)")
    if err == 2:
        raise _lib.B2Error("coordinate rows must be grouped by ascending batch index")
    nb = p["num_batches"]
    uidx = p["unique_index"][:n_out] if p["unique_index"] is not None else None
    if nb > 0:
        return (p["out_coords"][:n_out], uidx, p["inverse"], p["table"], p["cap"], n_out,
                (p["status"][2:3 + nb], host[2:3 + nb]))
    return p["out_coords"][:n_out], uidx, p["inverse"], p["table"], p["cap"], n_out


def map_build(coords, quant=0, num_batches=0):
    """Dedup (first occurrence wins, first-occurrence order).  Returns
    (out_coords[n_out,4], unique_index[n_out], inverse_map[n], table, cap, n_out[, (off_dev, off_host)]).
    With num_batches > 0 the batch offsets of the new map come back from the same single host read."""
    return map_build_finish(map_build_start(coords, quant, num_batches))


def map_table(coords):
    n = int(coords.shape[0])
    table, cap = alloc_table(n, coords.device)
    err = torch.zeros((1,), dtype=I32, device=coords.device)
    call("b2_map_table", ptr(coords), n, ptr(table), cap, ptr(err))
    return table, cap


def map_lookup(table, cap, query):
    rows = torch.empty((query.shape[0],), dtype=I32, device=query.device)
    call("b2_map_lookup", ptr(table), cap, ptr(query), query.shape[0], ptr(rows))
    return rows


def map_children(coords, half):
    out = torch.empty((coords.shape[0] * 8, 4), dtype=I32, device=coords.device)
    call("b2_map_children", ptr(coords), coords.shape[0], int(half), ptr(out))
    return out


def batch_offsets(coords, num_batches):
    """Device int32 [B+1] + host list.  Raises if rows are not grouped by ascending batch index."""
    dev = coords.device
    buf = torch.zeros((num_batches + 2,), dtype=I32, device=dev)
    call("b2_batch_offsets", ptr(coords), coords.shape[0], num_batches, ptr(buf), ptr(buf) + 4 * (num_batches + 1))
    host = read_ints(buf)
    if host[-1]:
        raise _lib.B2Error("coordinate rows must be grouped by ascending batch index")
    return buf[: num_batches + 1], host[: num_batches + 1]


def map_union_start(a_table, a_cap, a_coords, a_off, b_coords, b_off, num_batches=0):
    dev = a_coords.device
    na, nb = int(a_coords.shape[0]), int(b_coords.shape[0])
    out_coords = torch.empty((na + nb, 4), dtype=I32, device=dev)
    dst_a = torch.empty((na,), dtype=I32, device=dev)
    dst_b = torch.empty((nb,), dtype=I32, device=dev)
    n_out = torch.zeros((2 + num_batches + 1,), dtype=I32, device=dev)  # [n, err, offsets...]
    ws_bytes = _lib.lib().b2_map_union_ws_bytes(nb)
    ws = workspace(ws_bytes, dev)
    call("b2_map_union", ptr(a_table), a_cap, ptr(a_coords), na, ptr(a_off), ptr(b_coords), nb, ptr(b_off),
         ptr(out_coords), ptr(dst_a), ptr(dst_b), ptr(n_out), ptr(ws), ws.numel())
    if num_batches > 0:
        call("b2_batch_offsets_dev", ptr(out_coords), ptr(n_out), na + nb, num_batches, ptr(n_out) + 8, ptr(n_out) + 4)
    return dict(read=HostRead(n_out), n_out=n_out, out_coords=out_coords, dst_a=dst_a, dst_b=dst_b, num_batches=num_batches)


def map_union_finish(p):
    host = p["read"].get()
    n, nb = host[0], p["num_batches"]
    if nb > 0:
        return p["out_coords"][:n], p["dst_a"], p["dst_b"], n, (p["n_out"][2:3 + nb], host[2:3 + nb])
    return p["out_coords"][:n], p["dst_a"], p["dst_b"], n


def map_union(a_table, a_cap, a_coords, a_off, b_coords, b_off, num_batches=0):
    return map_union_finish(map_union_start(a_table, a_cap, a_coords, a_off, b_coords, b_off, num_batches))


def scan_mask(mask):
    """mask uint8/bool [n] -> (pos int32 [n], total int)."""
    if mask.dtype == torch.bool:
        mask = mask.view(torch.uint8)
    n = int(mask.shape[0])
    dev = mask.device
    pos = torch.empty((n,), dtype=I32, device=dev)
    total = torch.zeros((1,), dtype=I32, device=dev)
    ws = workspace(_lib.lib().b2_scan_ws_bytes(n), dev)
    call("b2_scan_mask", ptr(mask), n, ptr(pos), ptr(total), ptr(ws), ws.numel())
    return pos, total


def gather_coords(coords, mask, pos, n_out):
    out = torch.empty((n_out, 4), dtype=I32, device=coords.device)
    call("b2_gather_coords", ptr(coords), ptr(mask), ptr(pos), coords.shape[0], ptr(out))
    return out


def gather_rows(x, mask, pos, n_out):
    x = x.contiguous()
    out = torch.empty((n_out,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    row_bytes = x.element_size() * (x.numel() // max(int(x.shape[0]), 1)) if x.shape[0] else x.element_size()
    call("b2_gather_rows", ptr(x), ptr(mask), ptr(pos), x.shape[0], int(row_bytes), ptr(out))
    return out


def index_rows(x, idx):
    """out[i] = x[idx[i]] (idx int32; negative -> zero row)."""
    x = x.contiguous()
    n = int(idx.shape[0])
    out = torch.empty((n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    row_bytes = x.element_size() * int(torch.Size(x.shape[1:]).numel())
    call("b2_index_rows", ptr(x), ptr(idx), n, int(row_bytes), ptr(out))
    return out


def scatter_rows(x, dst, out, accumulate):
    call("b2_scatter_rows", ptr(x), ptr(dst), x.shape[0], x.shape[1], DT[x.dtype], int(accumulate), ptr(out))
    return out


# ----------------------------------------------------------------------------- kernel maps / conv
class KernelMap:
    """Neighbour table of one (in_map, out_map, kernel) triple."""
    __slots__ = ("nbr", "ld", "tile_mask", "n_out", "kvol", "pairs", "sorted", "tc_uses")

    def __init__(self, nbr, ld, tile_mask, n_out, kvol):
        self.nbr, self.ld, self.tile_mask, self.n_out, self.kvol = nbr, ld, tile_mask, n_out, kvol
        self.pairs = None
        self.sorted = None      # (perm, nbr_sorted, tile_mask_sorted) once kmap_sorted() was asked for
        self.tc_uses = 0        # tcgen05 convolutions that walked this table so far


# Sparsity-grouped walk of the tcgen05 convolution (kmap_sort.cu): the rows of a 3^3 table are grouped by the kernel slabs
# they use (counting sort on a 6-bit key), a 128-row tile then walks ~16 instead of 27 offsets at tensor stride 8.  Round 1
# measured the walk itself (-25 % at 218 k rows of 64 channels, -14 % at 54 k x 128) but its full 27-bit radix sort cost
# 90-115 us per table and cancelled the gain; the counting sort is a fraction of that, so tables with >= SORT_MIN_ROWS rows
# are now grouped when their SORT_MIN_USES-th tcgen05 convolution arrives (the backbone's stride-1 tables serve 5-11
# convolutions each).  Measured (B200, 8 x S0): ts8 64->64 108 -> 78 us per convolution, ts16 128->128 60 -> 47 us, ts32 no gain
# (49 -> 48 us, hence the 32 k-row threshold); grouping costs ~35 us per table inside the step.  B2_SORT_MIN_ROWS=0 switches
# it off.  Results are bit-identical either way.
SORT_MIN_ROWS = int(os.environ.get("B2_SORT_MIN_ROWS", "32768"))
SORT_MIN_USES = int(os.environ.get("B2_SORT_MIN_USES", "2"))


def kmap_sorted(km):
    """(perm, nbr_sorted, tile_mask_sorted) of a neighbour table, built once per table."""
    if km.sorted is None:
        dev = km.nbr.device
        perm = torch.empty((km.ld,), dtype=I32, device=dev)
        nbr_s = torch.empty_like(km.nbr)
        tm_s = torch.empty_like(km.tile_mask)
        ws = workspace(_lib.lib().b2_kmap_sort_ws_bytes(km.n_out, km.kvol), dev, slot=2)
        t0 = _prof_start()
        call("b2_kmap_sort", ptr(km.nbr), km.ld, km.n_out, km.kvol, ptr(perm), ptr(nbr_s), ptr(tm_s), ptr(ws), ws.numel())
        _prof_stop(t0, "kmap_sort", 0.0, 4.0 * km.kvol * km.n_out * 3)
        km.sorted = (perm, nbr_s, tm_s)
    return km.sorted


def kernel_map(in_table, in_cap, out_coords, ksize, step):
    n_out = int(out_coords.shape[0])
    dev = out_coords.device
    ld = (n_out + 127) // 128 * 128
    kvol = ksize ** 3
    nbr = torch.empty((kvol, ld), dtype=I32, device=dev)
    tile_mask = torch.empty((max(ld // 128, 1),), dtype=I32, device=dev)
    t0 = _prof_start()
    call("b2_kernel_map", ptr(in_table), in_cap, ptr(out_coords), n_out, int(ksize), int(step), ptr(nbr), ld,
         ptr(tile_mask))
    _prof_stop(t0, "kernel_map", 0.0, 16.0 * n_out + 24.0 * kvol * n_out)
    return KernelMap(nbr, ld, tile_mask, n_out, kvol)


def kernel_map_strided(out_table, out_cap, in_coords, n_out, ksize, step):
    """Table of a stride-2 convolution / pooling built from the input rows (probes of the OUT map's hash table)."""
    n_in = int(in_coords.shape[0])
    dev = in_coords.device
    n_out = int(n_out)
    ld = (n_out + 127) // 128 * 128
    kvol = ksize ** 3
    nbr = torch.empty((kvol, ld), dtype=I32, device=dev)
    tile_mask = torch.empty((max(ld // 128, 1),), dtype=I32, device=dev)
    t0 = _prof_start()
    call("b2_kernel_map_strided", ptr(out_table), out_cap, ptr(in_coords), n_in, int(ksize), int(step), ptr(nbr), ld,
         ptr(tile_mask))
    probes = 3.375 if ksize == 3 else 1.0
    _prof_stop(t0, "kernel_map", 0.0, 16.0 * n_in + (20.0 + 4.0) * probes * n_in + 4.0 * kvol * ld)
    return KernelMap(nbr, ld, tile_mask, n_out, kvol)


def kernel_map_children(parent_km, n_parent):
    """3^3 table of the map of all children (rows 8 i + c) from the parent's 3^3 table, without hash probes."""
    dev = parent_km.nbr.device
    n_out = 8 * int(n_parent)
    ld = (n_out + 127) // 128 * 128
    nbr = torch.empty((27, ld), dtype=I32, device=dev)
    tile_mask = torch.empty((max(ld // 128, 1),), dtype=I32, device=dev)
    t0 = _prof_start()
    call("b2_kernel_map_children", ptr(parent_km.nbr), parent_km.ld, int(n_parent), ptr(nbr), ld, ptr(tile_mask))
    _prof_stop(t0, "kernel_map_children", 0.0, 4.0 * 27 * (n_parent + n_out))
    return KernelMap(nbr, ld, tile_mask, n_out, 27)


def prune_plan(dst_b, dst_a, mask, pos, n_keep):
    """Row maps of a prune that directly follows a children map + union: (children row -> pruned row or -1,
    backbone row -> pruned row or -1, pruned row -> children row)."""
    dev = dst_b.device
    gen_to_kept = torch.empty((dst_b.shape[0],), dtype=I32, device=dev)
    a_to_kept = torch.empty((dst_a.shape[0],), dtype=I32, device=dev)
    kept_src = torch.empty((max(int(n_keep), 1),), dtype=I32, device=dev)
    call("b2_prune_plan", ptr(dst_b), dst_b.shape[0], ptr(dst_a), dst_a.shape[0], ptr(mask), ptr(pos), ptr(gen_to_kept),
         ptr(a_to_kept), ptr(kept_src))
    return gen_to_kept, a_to_kept, kept_src


def kernel_map_children_pruned(parent_km, kept_src, gen_to_kept, n_keep):
    """(table from the pruned rows into all children rows, 3^3 table of the pruned map) from the parent's 3^3 table."""
    dev = parent_km.nbr.device
    n_keep = int(n_keep)
    ld = (n_keep + 127) // 128 * 128
    nbr_up = torch.empty((27, ld), dtype=I32, device=dev)
    nbr_out = torch.empty((27, ld), dtype=I32, device=dev)
    mask_up = torch.empty((max(ld // 128, 1),), dtype=I32, device=dev)
    mask_out = torch.empty((max(ld // 128, 1),), dtype=I32, device=dev)
    t0 = _prof_start()
    call("b2_kernel_map_children_pruned", ptr(parent_km.nbr), parent_km.ld, ptr(kept_src), ptr(gen_to_kept), n_keep, ld,
         ptr(nbr_up), ptr(mask_up), ptr(nbr_out), ptr(mask_out))
    _prof_stop(t0, "kernel_map_children", 0.0, 4.0 * n_keep * (1 + 27 + 27 + 2 * 27))
    return KernelMap(nbr_up, ld, mask_up, n_keep, 27), KernelMap(nbr_out, ld, mask_out, n_keep, 27)


_TC_ENABLED = True

# Optional per-launch profiling (bench.py): list of [name, start_event, end_event, algorithmic_flops, algorithmic_bytes]
PROFILE = None


def _prof_start():
    if PROFILE is None:
        return None
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def _prof_stop(start, name, flops=0.0, nbytes=0.0):
    if start is None:
        return
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    PROFILE.append([name, start, e, float(flops), float(nbytes)])


def kmap_pairs(kmap, n_out):
    """number of (in, out) pairs of a kernel map (one host sync, cached; profiling only)."""
    if kmap is None:
        return n_out
    if getattr(kmap, "pairs", None) is None:
        kmap.pairs = int((kmap.nbr >= 0).sum().item())
    return kmap.pairs


def set_tensor_cores(flag):
    global _TC_ENABLED
    _TC_ENABLED = bool(flag)


def tc_eligible(x, cin, cout):
    return _TC_ENABLED and x.dtype == torch.bfloat16 and cin % 64 == 0 and cout % 32 == 0


def sparse_conv(x, w, kmap, n_out, wt_bf16=None, bias=None, scale=None, shift=None, residual=None, act=ACT_NONE,
                ch_mod=0, out_dtype=None, wt_tf32=None):
    """out[o] = act((sum_k x[nbr[k][o]] @ w[k] + bias) * scale + shift + residual[o]).

    w: fp32 [K, Cin, Cout]; wt_bf16: optional bf16 [K, Cout, Cin] for the tcgen05 path; wt_tf32: optional fp32 [K, Cout, Cin]
    for fp32 rows on the tensor cores as TF32 operands; kmap None = identity (K = 1).
    """
    require_cuda(x, "features")
    wshape = w.shape if torch.is_tensor(w) else w          # a bare (K, Cin, Cout) is enough for the tcgen05 path
    kvol, cin, cout = int(wshape[0]), int(wshape[1]), int(wshape[2])
    out_dtype = out_dtype or x.dtype
    out = torch.empty((n_out, cout), dtype=out_dtype, device=x.device)
    nbr = kmap.nbr if kmap is not None else None
    ld = kmap.ld if kmap is not None else 0
    tmask = kmap.tile_mask if kmap is not None else None
    if residual is not None:
        assert residual.dtype == out_dtype and residual.shape == out.shape
    use_tc = wt_bf16 is not None and tc_eligible(x, cin, cout)
    use_tf32 = (not use_tc and wt_tf32 is not None and _TC_ENABLED and x.dtype == torch.float32
                and out_dtype == torch.float32 and cin % 32 == 0 and cout % 32 == 0)
    if not use_tc and not use_tf32 and not torch.is_tensor(w):
        raise B2Error("sparse_conv: fp32 weights are required off the tcgen05 path")
    if PROFILE is not None:
        pairs = kmap_pairs(kmap, n_out)
        esz = x.element_size()
        flops = 2.0 * pairs * cin * cout
        nbytes = esz * (x.shape[0] * cin + kvol * cin * cout) + out.element_size() * n_out * cout + 4.0 * kvol * ld
    t0 = _prof_start()
    if use_tc or use_tf32:
        perm = None
        if kmap is not None and kvol == 27:
            kmap.tc_uses += 1
            if kmap.sorted is not None or (0 < SORT_MIN_ROWS <= n_out and kmap.tc_uses >= SORT_MIN_USES):
                perm, nbr, tmask = kmap_sorted(kmap)
    if use_tf32:
        call("b2_conv_tc_tf32", ptr(x), x.shape[0], ptr(wt_tf32), ptr(nbr), ld, ptr(tmask), ptr(perm), n_out, cin, cout,
             kvol, ptr(bias), ptr(scale), ptr(shift), ptr(residual), act, ch_mod, ptr(out))
    elif use_tc:
        call("b2_conv_tc", ptr(x), x.shape[0], ptr(wt_bf16), ptr(nbr), ld, ptr(tmask), ptr(perm), n_out, cin, cout, kvol,
             ptr(bias), ptr(scale), ptr(shift), ptr(residual), act, ch_mod, ptr(out), DT[out_dtype])
    else:
        if cin <= 4 and x.shape[0] >= (1 << 26):
            raise B2Error("narrow-input convolution packs (offset, row) into 32 bits: at most 2^26 input rows")
        call("b2_conv_cc", ptr(x), DT[x.dtype], ptr(w), ptr(nbr), ld, ptr(tmask), n_out, cin, cout, kvol, ptr(bias),
             ptr(scale), ptr(shift), ptr(residual), act, ch_mod, ptr(out), DT[out_dtype])
    if t0 is not None:
        _prof_stop(t0, "conv_tc" if (use_tc or use_tf32) else "conv_cc", flops, nbytes)
    return out


def maxpool(x, kmap, want_argmax=False):
    out = torch.empty((kmap.n_out, x.shape[1]), dtype=x.dtype, device=x.device)
    arg = torch.empty((kmap.n_out, x.shape[1]), dtype=I32, device=x.device) if want_argmax else None
    call("b2_maxpool", ptr(x), DT[x.dtype], ptr(kmap.nbr), kmap.ld, kmap.kvol, kmap.n_out, x.shape[1], ptr(out),
         ptr(arg))
    return (out, arg) if want_argmax else out


def maxpool_affine(x, kmap, out_coords, scale, shift, act):
    """maxpool(affine_act(x, scale, shift, act)) in one pass (bf16 rows, per-scene scale / shift [B, C])."""
    if x.dtype != torch.bfloat16 or x.shape[1] % 8:
        raise B2Error("maxpool_affine: bf16 rows with a multiple of 8 channels")
    out = torch.empty((kmap.n_out, x.shape[1]), dtype=x.dtype, device=x.device)
    call("b2_maxpool_affine", ptr(x), ptr(kmap.nbr), kmap.ld, kmap.kvol, kmap.n_out, x.shape[1], ptr(out_coords),
         ptr(scale), ptr(shift), int(act), ptr(out))
    return out


def segment_stats(x, seg_off, nseg):
    c = int(x.shape[1])
    mean = torch.empty((nseg, c), dtype=torch.float32, device=x.device)
    var = torch.empty((nseg, c), dtype=torch.float32, device=x.device)
    ws = workspace(_lib.lib().b2_segment_stats_ws_bytes(nseg, c), x.device, slot=1)
    call("b2_segment_stats", ptr(x), DT[x.dtype], ptr(seg_off), nseg, c, ptr(mean), ptr(var), ptr(ws), ws.numel())
    return mean, var


def instnorm_supported(x):
    e = 16 // x.element_size()
    c = int(x.shape[1])
    return x.dtype in DT and c % e == 0 and c // e <= 256 and 256 % (c // e) == 0


def instnorm_scale_shift(x, seg_off, nseg, weight, bias, eps):
    """Per-scene folded InstanceNorm (scale, shift) [nseg, C] in one pass over the rows."""
    c = int(x.shape[1])
    scale = torch.empty((nseg, c), dtype=torch.float32, device=x.device)
    shift = torch.empty((nseg, c), dtype=torch.float32, device=x.device)
    ws = workspace(_lib.lib().b2_instnorm_ws_bytes(nseg, c), x.device, slot=1)
    call("b2_instnorm_scale_shift", ptr(x), DT[x.dtype], ptr(seg_off), nseg, c, ptr(weight), ptr(bias), float(eps),
         ptr(scale), ptr(shift), ptr(ws), ws.numel())
    return scale, shift


def affine_act(x, scale=None, shift=None, coords=None, residual=None, act=ACT_NONE, out=None):
    out = torch.empty_like(x) if out is None else out
    call("b2_affine_act", ptr(x), DT[x.dtype], ptr(coords), x.shape[0], x.shape[1], ptr(scale), ptr(shift),
         ptr(residual), act, ptr(out))
    return out


def pack_conv_weights(w, want_fwd=True, want_bwd=True):
    """bf16 operand layouts of w [K, Cin, Cout] fp32 in one launch: (w[k]^T as [K, Cout, Cin], w[K-1-k] as [K, Cin, Cout])."""
    kvol, cin, cout = (int(v) for v in w.shape)
    fwd = torch.empty((kvol, cout, cin), dtype=torch.bfloat16, device=w.device) if want_fwd else None
    bwd = torch.empty((kvol, cin, cout), dtype=torch.bfloat16, device=w.device) if want_bwd else None
    if want_fwd or want_bwd:
        call("b2_pack_conv_weights", ptr(w), kvol, cin, cout, ptr(fwd), ptr(bwd))
    return fwd, bwd


def cast(x, dtype):
    if x.dtype == dtype:
        return x
    out = torch.empty(x.shape, dtype=dtype, device=x.device)
    call("b2_cast", ptr(x), DT[x.dtype], x.numel(), ptr(out))
    return out


# ----------------------------------------------------------------------------- head tail
def interp(table, cap, feats, tensor_stride, query):
    feats = feats.contiguous()
    query = query.contiguous()
    out = torch.empty((query.shape[0], feats.shape[1]), dtype=torch.float32, device=feats.device)
    call("b2_interp", ptr(table), cap, ptr(feats), feats.shape[1], int(tensor_stride), ptr(query), query.shape[0],
         ptr(out))
    return out


def topk_mask(scores, seg_off, nseg, k):
    mask = torch.empty((scores.shape[0],), dtype=torch.uint8, device=scores.device)
    call("b2_topk_mask_n", ptr(scores), ptr(seg_off), nseg, int(k), int(scores.shape[0]), ptr(mask))
    return mask


def head_post(feat, n_reg, n_cls, scale, want_max_score=False):
    """feat [n, >= 1+n_reg+n_cls] fp32 (extra padding columns are ignored)."""
    n = int(feat.shape[0])
    dev = feat.device
    f32 = torch.float32
    centerness = torch.empty((n, 1), dtype=f32, device=dev)
    bbox_pred = torch.empty((n, n_reg), dtype=f32, device=dev)
    cls = torch.empty((n, n_cls), dtype=f32, device=dev)
    prune = torch.empty((n, 1), dtype=f32, device=dev)
    mx = torch.empty((n,), dtype=f32, device=dev) if want_max_score else None
    call("b2_head_post", ptr(feat), n, int(feat.shape[1]), n_reg, n_cls, ptr(scale), ptr(centerness), ptr(bbox_pred), ptr(cls), ptr(prune),
         ptr(mx))
    return centerness, bbox_pred, cls, prune, mx


def head_fused(x, wt_bf16, bias, n_reg, n_cls, scale, want_max_score=False):
    """x [n, cin] bf16, wt_bf16 [1, 32, cin], bias [32]: head GEMM + head_post in one launch (b2_head_fused); returns what
    head_post returns."""
    n = int(x.shape[0])
    dev = x.device
    f32 = torch.float32
    centerness = torch.empty((n, 1), dtype=f32, device=dev)
    bbox_pred = torch.empty((n, n_reg), dtype=f32, device=dev)
    cls = torch.empty((n, n_cls), dtype=f32, device=dev)
    prune = torch.empty((n, 1), dtype=f32, device=dev)
    mx = torch.empty((n,), dtype=f32, device=dev) if want_max_score else None
    t0 = _prof_start()
    call("b2_head_fused", ptr(x), n, ptr(wt_bf16), int(x.shape[1]), int(wt_bf16.shape[1]), ptr(bias), n_reg, n_cls, ptr(scale),
         ptr(centerness), ptr(bbox_pred), ptr(cls), ptr(prune), ptr(mx))
    if t0 is not None:
        cin = int(x.shape[1])
        _prof_stop(t0, "conv_tc", 2.0 * n * cin * 32, 2.0 * n * cin + 4.0 * n * (3 + n_reg + n_cls))
    return centerness, bbox_pred, cls, prune, mx


def head_fused_ok(x, wt_bf16, n_reg, n_cls):
    return (_TC_ENABLED and x.dtype == torch.bfloat16 and x.shape[1] % 64 == 0 and wt_bf16 is not None
            and wt_bf16.shape[1] == 32 and 1 + n_reg + n_cls <= 32)


YAW_MODES = {"none": 0, "naive": 1, "sin-cos": 2, "fcaf3d": 3}


def decode_boxes(points, bbox_pred, yaw_mode):
    n = int(points.shape[0])
    out = torch.empty((n, 7), dtype=torch.float32, device=points.device)
    call("b2_decode_boxes", ptr(points.contiguous()), ptr(bbox_pred.contiguous()), n, bbox_pred.shape[1], yaw_mode,
         ptr(out))
    return out


# ----------------------------------------------------------------------------- NMS
def nms_plan(seg_off_host):
    nseg = len(seg_off_host) - 1
    so = (ctypes.c_int * (nseg + 1))(*seg_off_host)
    to = (ctypes.c_int * (nseg + 1))()
    mo = (ctypes.c_longlong * (nseg + 1))()
    words = ctypes.c_longlong()
    mx = ctypes.c_int()
    rc = _lib.lib().b2_nms_plan(so, nseg, to, mo, ctypes.byref(words), ctypes.byref(mx))
    _lib.check(rc, "b2_nms_plan")
    return list(to), list(mo), words.value, mx.value


NMS_GREEDY = True  # blocked greedy kernel (rows of kept boxes only); False = full mask + device reduce
NMS_GREEDY_MAX_SEG = 4096  # in-kernel sort capacity (SORT_MAX, nms.cu)


def nms_batched(boxes, scores, seg_off_host, thresh, rotated, sort=True, greedy=None, seg_off_dev=None, max_seg=None):
    """boxes [n,7], scores [n] grouped into segments.  Returns (keep uint8 [n] in sorted order, order int32 [n]
    (segment-local source index of every sorted slot), sorted_boxes, counts int32 [nseg]).
    Without a host copy of the segment offsets (seg_off_host None) the greedy kernels run from the DEVICE offsets
    seg_off_dev int32 [nseg+1] and an upper bound max_seg of the segment length; boxes / scores may then be longer than
    seg_off[nseg] and keep is 0 past it - no host read anywhere."""
    dev = boxes.device
    n = int(boxes.shape[0])
    if seg_off_host is None:
        nseg = int(seg_off_dev.shape[0]) - 1
        keep = torch.zeros((n,), dtype=torch.uint8, device=dev)
        counts = torch.zeros((max(nseg, 1),), dtype=I32, device=dev)
        sorted_boxes = torch.empty_like(boxes) if sort else boxes
        order = torch.empty((n,), dtype=I32, device=dev) if sort else None
        if nseg > 0 and n > 0:
            ws = workspace(_lib.lib().b2_nms_greedy_ws_bytes(n, int(bool(rotated))), dev, slot=2)
            call("b2_nms_greedy_batched", ptr(boxes), ptr(scores), ptr(seg_off_dev), nseg, n, int(max_seg), float(thresh),
                 int(bool(rotated)), int(bool(sort)), ptr(sorted_boxes), ptr(order), ptr(ws), ws.numel(), ptr(keep),
                 ptr(counts))
        return keep, order, sorted_boxes, counts
    nseg = len(seg_off_host) - 1
    if (NMS_GREEDY if greedy is None else greedy) and nseg > 0:
        max_seg = max(seg_off_host[i + 1] - seg_off_host[i] for i in range(nseg))
        so = ints_to_device(list(seg_off_host), dev)
        keep = torch.empty((n,), dtype=torch.uint8, device=dev)
        counts = torch.zeros((nseg,), dtype=I32, device=dev)
        sorted_boxes = torch.empty_like(boxes) if sort else boxes
        order = torch.empty((n,), dtype=I32, device=dev) if sort else None
        ws = workspace(_lib.lib().b2_nms_greedy_ws_bytes(n, int(bool(rotated))), dev, slot=2)
        call("b2_nms_greedy_batched", ptr(boxes), ptr(scores), ptr(so), nseg, n, max_seg, float(thresh),
             int(bool(rotated)), int(bool(sort)), ptr(sorted_boxes), ptr(order), ptr(ws), ws.numel(), ptr(keep),
             ptr(counts))
        return keep, order, sorted_boxes, counts
    tile_off, mask_off, words, max_seg = nms_plan(seg_off_host)
    plan = torch.tensor(list(seg_off_host) + tile_off, dtype=I32, device=dev)
    moff = torch.tensor(mask_off, dtype=torch.int64, device=dev)
    mask = workspace(max(words, 1) * 8, dev, slot=2)
    keep = torch.empty((n,), dtype=torch.uint8, device=dev)
    counts = torch.zeros((max(nseg, 1),), dtype=I32, device=dev)
    sorted_boxes = torch.empty_like(boxes) if sort else boxes
    order = torch.empty((n,), dtype=I32, device=dev) if sort else None
    call("b2_nms_batched", ptr(boxes), ptr(scores), ptr(plan), ptr(plan) + 4 * (nseg + 1), ptr(moff), nseg,
         tile_off[-1], max_seg, float(thresh), int(bool(rotated)), int(bool(sort)), ptr(sorted_boxes), ptr(order),
         ptr(mask), ptr(keep), ptr(counts))
    return keep, order, sorted_boxes, counts


def boxes_pair_matrix(a, b, mode):
    out = torch.zeros((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
    call("b2_boxes_pair_matrix", ptr(a.contiguous()), a.shape[0], ptr(b.contiguous()), b.shape[0], mode, ptr(out))
    return out
