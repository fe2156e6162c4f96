"""The MinkowskiEngine Python surface that FCAF3D uses, on top of the fcaf3d_b200 CUDA library.

`fcaf3d_b200.install_as_minkowski_engine()` registers this package as `MinkowskiEngine`, so the reference's
model files (mmdet3d/models/backbones/me_resnet.py, dense_heads/fcaf3d_neck_with_head.py,
detectors/single_stage_sparse.py) import it unchanged.  Surface and semantics follow MinkowskiEngine v0.5.4
as used by those files; only CUDA tensors are accepted (no CPU backend).
"""
import math
import os

import torch
import torch.nn as nn

from .. import autograd as AG
from .. import ops
from .._lib import B2Error

__version__ = "0.5.4+b200"


class SparseTensorQuantizationMode:
    RANDOM_SUBSAMPLE = 0
    UNWEIGHTED_AVERAGE = 1
    UNWEIGHTED_SUM = 2
    NO_QUANTIZATION = 3


class CoordinateMapKey:
    """(tensor_stride, string_id) handle of a coordinate map (ME A.3)."""

    def __init__(self, tensor_stride, string_id=""):
        if isinstance(tensor_stride, int):
            tensor_stride = (tensor_stride,) * 3
        self._ts = tuple(int(t) for t in tensor_stride)
        self._id = str(string_id)

    def get_tensor_stride(self):
        return list(self._ts)

    def get_key(self):
        return (list(self._ts), self._id)

    def get_coordinate_size(self):
        return len(self._ts) + 1

    def __eq__(self, other):
        return isinstance(other, CoordinateMapKey) and self._ts == other._ts and self._id == other._id

    def __hash__(self):
        return hash((self._ts, self._id))

    def __repr__(self):
        return "coordinate map key:%s%s" % (list(self._ts), (":" + self._id) if self._id else "")


# stride-2 kernel maps from the input side (ops.kernel_map_strided): 0 = never (hash probes per output row), 1 = where
# there are few inputs per output row, 2 = always
STRIDED_FROM_INPUT = int(os.environ.get("B2_KMAP_STRIDED", "1"))


class _Map:
    """One coordinate map: int32 rows (b, x, y, z) grouped by batch, plus a lazily built hash table."""

    def __init__(self, coords, ts, num_batches, table=None, cap=0, off=None):
        self.coords = coords
        self.n = int(coords.shape[0])
        self.ts = int(ts)
        self.num_batches = num_batches
        self.table, self.cap = table, cap
        if off is None:
            off = ops.batch_offsets(coords, num_batches)
        self.off_dev, self.off_host = off

    def ensure_table(self):
        if self.table is None:
            self.table, self.cap = ops.map_table(self.coords)
        return self.table, self.cap


class CoordinateManager:
    """Owns every coordinate map and kernel map of one forward pass (ME CoordinateManager)."""

    def __init__(self, D=3, device=None):
        assert D == 3, "fcaf3d_b200 implements 3-D sparse tensors"
        self.D = D
        self.device = device
        self._maps = {}
        self._kmaps = {}
        self._stride_cache = {}
        self._pending = {}      # deferred map builds (stride / union) whose row count has not been read yet
        self._children_cache = {}
        self._auto = 0
        self.num_batches = 0

    # -- map creation ------------------------------------------------------------------------
    def _fresh_id(self, ts):
        while True:
            self._auto += 1
            key = CoordinateMapKey(ts, "m%d" % self._auto)
            if key not in self._maps:
                return key

    def insert_and_map(self, coordinates, tensor_stride=1, string_id=""):
        """Returns (key, (unique_index, inverse_mapping)) — int64 tensors like ME."""
        if not coordinates.is_cuda:
            raise B2Error("fcaf3d_b200 has no CPU coordinate manager: pass CUDA coordinates")
        ts = tensor_stride if isinstance(tensor_stride, int) else int(tensor_stride[0])
        coords = coordinates.to(torch.int32).contiguous()
        self.device = coords.device
        perm = None
        hint = getattr(coordinates, "_b2_num_batches", None)
        if hint is not None:        # produced by ops.quantize_points: batch-grouped by construction, batch count known
            nb = int(hint)
        elif coords.shape[0]:
            b = coords[:, 0]
            nb = int(b.max().item()) + 1
            if bool((b[1:] < b[:-1]).any().item()):  # not batch-grouped: stable regroup (rare path)
                perm = torch.sort(b, stable=True).indices
                coords = coords[perm].contiguous()
        else:
            nb = 1
        self.num_batches = max(self.num_batches, nb)
        p0 = ops.map_build_start(coords, 0, self.num_batches)
        # The first strided map can be built from the raw rows as well (dedup keeps first-occurrence order, so flooring the
        # raw rows gives the same map as flooring the deduplicated ones): callers that know the stride of their first
        # convolution ask for it here and both row counts come back behind one wait.
        pre = int(getattr(coordinates, "_b2_prefetch_stride", 0) or 0)
        p1 = (ops.map_build_start(coords, ts * pre, self.num_batches, want_index=False)
              if pre > 1 and perm is None else None)
        out_coords, unique_index, inverse, table, cap, n_out, off = ops.map_build_finish(p0)
        key = CoordinateMapKey(ts, string_id)
        if key in self._maps:
            key = self._fresh_id(ts)
        self._maps[key] = _Map(out_coords.contiguous(), ts, self.num_batches, table, cap, off=off)
        if p1 is not None:
            self._pending[("s", (key, pre))] = p1
        unique_index = unique_index.long()
        inverse = inverse.long()
        if perm is not None:
            unique_index = perm[unique_index]
            inv = torch.empty_like(inverse)
            inv[perm] = inverse
            inverse = inv
        return key, (unique_index, inverse)

    def stride_prefetch(self, in_key, stride):
        """Enqueues the build of the strided map now and leaves its row count in flight; a later stride() call picks it
        up.  Lets the caller put GPU work between the build and the host read (no idle GPU at the read)."""
        s = stride if isinstance(stride, int) else int(stride[0])
        ck = (in_key, s)
        if ck in self._stride_cache or ("s", ck) in self._pending:
            return
        src = self._maps[in_key]
        self._pending[("s", ck)] = ops.map_build_start(src.coords, src.ts * s, self.num_batches, want_index=False)

    def stride(self, in_key, stride):
        """Coordinates floored to multiples of ts*stride, deduplicated; cached per (key, stride) (ME A.3)."""
        s = stride if isinstance(stride, int) else int(stride[0])
        ck = (in_key, s)
        if ck in self._stride_cache:
            return self._stride_cache[ck]
        src = self._maps[in_key]
        ts_out = src.ts * s
        pending = self._pending.pop(("s", ck), None) or ops.map_build_start(src.coords, ts_out, self.num_batches,
                                                                             want_index=False)
        out_coords, _, _, table, cap, _, off = ops.map_build_finish(pending)
        key = CoordinateMapKey(ts_out, "")
        if key in self._maps:
            key = self._fresh_id(ts_out)
        m = _Map(out_coords.contiguous(), ts_out, self.num_batches, table, cap, off=off)
        self._maps[key] = m
        self._stride_cache[ck] = key
        return key

    def generate_children(self, in_key):
        """New map holding the 8 children of every voxel (generative transpose, ME A.5): row 8*i + k."""
        cached = self._children_cache.get(in_key)
        if cached is not None:      # one children map per parent map (lets callers create it ahead of the module call)
            return cached
        src = self._maps[in_key]
        if src.ts % 2:
            raise B2Error("generative transpose needs an even tensor stride")
        ts_out = src.ts // 2
        coords = ops.map_children(src.coords, ts_out)
        off_host = [8 * o for o in src.off_host]
        key = self._fresh_id(ts_out)
        m = _Map(coords, ts_out, self.num_batches, off=(src.off_dev * 8, off_host))
        m.children_of = in_key   # all 8 children of every parent, row 8 i + c: kernel maps can be derived, see kernel_map
        self._maps[key] = m
        self._children_cache[in_key] = key
        return key

    def _union_start(self, key_a, key_b):
        a, b = self._maps[key_a], self._maps[key_b]
        if a.ts != b.ts:
            raise B2Error("union of maps with different tensor strides")
        table, cap = a.ensure_table()
        return ops.map_union_start(table, cap, a.coords, a.off_dev, b.coords, b.off_dev, self.num_batches)

    def union_prefetch(self, key_a, key_b):
        """Enqueues the union now, row count left in flight (see stride_prefetch)."""
        if ("u", key_a, key_b) not in self._pending:
            self._pending[("u", key_a, key_b)] = self._union_start(key_a, key_b)

    def union(self, key_a, key_b):
        a = self._maps[key_a]
        pending = self._pending.pop(("u", key_a, key_b), None) or self._union_start(key_a, key_b)
        out_coords, dst_a, dst_b, n, off = ops.map_union_finish(pending)
        key = self._fresh_id(a.ts)
        self._maps[key] = _Map(out_coords.contiguous(), a.ts, self.num_batches, off=off)
        return key, dst_a, dst_b

    def prune(self, in_key, mask, keep_counts=None):
        """keep_counts: rows kept per scene when the caller knows them (a top-k mask) - saves the host read."""
        src = self._maps[in_key]
        m8 = mask.view(torch.uint8) if mask.dtype == torch.bool else mask
        pos, total = ops.scan_mask(m8)
        # batch offsets of the pruned map = scan position of every scene's first row; fetched with the row count
        nb = self.num_batches
        if src.n > 0:
            first = src.off_dev[:nb]
            off = torch.where(first < src.n, pos[first.clamp(max=src.n - 1).long()], total.expand(nb))
            off_dev = torch.cat((off, total)).to(torch.int32)
        else:
            off_dev = torch.zeros((nb + 1,), dtype=torch.int32, device=m8.device)
        if keep_counts is not None:
            off_host = [0]
            for c in keep_counts:
                off_host.append(off_host[-1] + int(c))
        else:
            off_host = ops.read_ints(off_dev)
        n_out = off_host[-1]
        coords = ops.gather_coords(src.coords, m8, pos, n_out)
        key = self._fresh_id(src.ts)
        self._maps[key] = _Map(coords, src.ts, self.num_batches, off=(off_dev, off_host))
        return key, m8, pos, n_out

    # -- queries -----------------------------------------------------------------------------
    def get_map(self, key):
        return self._maps[key]

    def get_coordinates(self, key):
        return self._maps[key].coords

    def size(self, key):
        return self._maps[key].n

    def number_of_unique_batch_indices(self):
        return self.num_batches

    def kernel_map(self, in_key, out_key, kernel_size, dilation=1):
        """Neighbour table from `out_key` rows into `in_key` rows; cached (ME A.11)."""
        ck = (in_key, out_key, int(kernel_size), int(dilation))
        km = self._kmaps.get(ck)
        if km is None:
            src, dst = self._maps[in_key], self._maps[out_key]
            parent = getattr(src, "children_of", None)
            if in_key == out_key and parent is not None and int(kernel_size) == 3 and int(dilation) == 1:
                # map of ALL children of `parent` (generative transpose): derive the table from the parent's, no hash
                pkm = self.kernel_map(parent, parent, 3, 1)
                km = ops.kernel_map_children(pkm, self._maps[parent].n)
            elif (STRIDED_FROM_INPUT and in_key != out_key and dst.ts == 2 * src.ts and int(dilation) == 1
                  and ((int(kernel_size) == 3 and (STRIDED_FROM_INPUT > 1 or 2 * src.n < 3 * dst.n))
                       or int(kernel_size) == 2)):
                # (kernel 2, the stem's pooling map: every input row has exactly ONE parent, so one probe per input row
                # replaces eight per output row)
                # stride 2, few inputs per output (the stem: 786 k -> 738 k rows): ~3.4 probes of the OUT table per input
                # row instead of 27 probes per output row (188 -> 95 us).  Deeper levels (548 k -> 218 k rows) have ~8
                # inputs per output, every 128-row tile's mask word is then hit ~3.7 k times (a hot spot in L2, atomics
                # or plain reads alike) and the per-output kernel wins (45 vs 148 us), so those stay on it.
                table, cap = dst.ensure_table()
                km = ops.kernel_map_strided(table, cap, src.coords, dst.n, int(kernel_size), src.ts)
            else:
                table, cap = src.ensure_table()
                km = ops.kernel_map(table, cap, dst.coords, int(kernel_size), src.ts * int(dilation))
            self._kmaps[ck] = km
        return km


def _record_stream(self, stream):
    """Marks every tensor of the manager as used on `stream` (torch's allocator then keeps their memory until that stream
    has passed the point of their release): for maps that were built on one stream and are consumed on another."""
    def rec(t):
        if torch.is_tensor(t) and t.is_cuda:
            t.record_stream(stream)
    for m in self._maps.values():
        for t in (m.coords, m.table, m.off_dev):
            rec(t)
    for km in self._kmaps.values():
        rec(km.nbr), rec(km.tile_mask)
        for t in (km.sorted or ()):
            rec(t)
    for p in self._pending.values():
        for t in p.values():
            rec(t)


CoordinateManager.record_stream = _record_stream


def _set_kernel_map(self, in_key, out_key, kernel_size, dilation, km):
    """Registers a neighbour table a caller derived by other means (e.g. ops.kernel_map_children_pruned)."""
    self._kmaps[(in_key, out_key, int(kernel_size), int(dilation))] = km


CoordinateManager.set_kernel_map = _set_kernel_map


def _kernel_map_transposed(self, in_key, out_key, kernel_size, dilation=1):
    """For every INPUT row the output rows at coord_in + offset_k (backward of a convolution in_key -> out_key)."""
    if in_key == out_key:
        return self.kernel_map(in_key, out_key, kernel_size, dilation)
    if int(kernel_size) % 2 == 0:
        raise B2Error("backward of even-sized convolution kernels is not implemented")
    ck = (in_key, out_key, int(kernel_size), int(dilation), "T")
    km = self._kmaps.get(ck)
    if km is None:
        src, dst = self._maps[in_key], self._maps[out_key]
        table, cap = dst.ensure_table()
        km = ops.kernel_map(table, cap, src.coords, int(kernel_size), src.ts * int(dilation))
        self._kmaps[ck] = km
    return km


CoordinateManager.kernel_map_transposed = _kernel_map_transposed


def _iso(v):
    if isinstance(v, (list, tuple)):
        assert all(x == v[0] for x in v), "only isotropic kernels/strides are supported"
        return int(v[0])
    return int(v)


class SparseTensor:
    """ME.SparseTensor (A.2): features [N, C] over a coordinate map."""

    def __init__(self, features, coordinates=None, tensor_stride=1, coordinate_map_key=None, coordinate_manager=None,
                 quantization_mode=SparseTensorQuantizationMode.RANDOM_SUBSAMPLE, device=None, **kwargs):
        if device is not None:
            features = features.to(device)
            if coordinates is not None:
                coordinates = coordinates.to(device)
        self.quantization_mode = quantization_mode
        self.unique_index = None
        self.inverse_mapping = None
        if coordinates is not None:
            assert coordinate_map_key is None
            if coordinate_manager is None:
                coordinate_manager = CoordinateManager(D=coordinates.shape[1] - 1, device=coordinates.device)
            key, (uidx, inv) = coordinate_manager.insert_and_map(coordinates, _iso(tensor_stride))
            self.unique_index, self.inverse_mapping = uidx, inv
            if quantization_mode == SparseTensorQuantizationMode.RANDOM_SUBSAMPLE:
                features = ops.index_rows(features.contiguous(), uidx.to(torch.int32))
            elif quantization_mode == SparseTensorQuantizationMode.NO_QUANTIZATION:
                if uidx.shape[0] != features.shape[0]:
                    raise B2Error("NO_QUANTIZATION requires unique coordinates")
            else:
                n = uidx.shape[0]
                acc = torch.zeros((n, features.shape[1]), dtype=features.dtype, device=features.device)
                acc.index_add_(0, inv, features)
                if quantization_mode == SparseTensorQuantizationMode.UNWEIGHTED_AVERAGE:
                    cnt = torch.zeros((n, 1), dtype=features.dtype, device=features.device)
                    cnt.index_add_(0, inv, torch.ones_like(features[:, :1]))
                    acc = acc / cnt
                features = acc
            coordinate_map_key = key
        else:
            assert coordinate_map_key is not None and coordinate_manager is not None
        if not features.is_cuda:
            raise B2Error("fcaf3d_b200 SparseTensor needs CUDA features (no CPU backend)")
        self._F = features
        self.coordinate_map_key = coordinate_map_key
        self.coordinate_manager = coordinate_manager
        if self.coordinate_manager.size(coordinate_map_key) != features.shape[0]:
            raise B2Error("feature rows (%d) != coordinate rows (%d)" %
                          (features.shape[0], self.coordinate_manager.size(coordinate_map_key)))

    # -- ME properties -------------------------------------------------------------------------
    @property
    def F(self):
        return self._F

    features = F

    @property
    def C(self):
        return self.coordinate_manager.get_coordinates(self.coordinate_map_key)

    coordinates = C

    @property
    def tensor_stride(self):
        return self.coordinate_map_key.get_tensor_stride()

    @property
    def D(self):
        return 3

    dimension = D

    @property
    def device(self):
        return self._F.device

    @property
    def dtype(self):
        return self._F.dtype

    @property
    def shape(self):
        return self._F.shape

    def size(self, *a):
        return self._F.size(*a)

    def __len__(self):
        return self._F.shape[0]

    def _map(self):
        return self.coordinate_manager.get_map(self.coordinate_map_key)

    @property
    def decomposition_permutations(self):
        m = self._map()
        dev = self._F.device
        return [torch.arange(m.off_host[b], m.off_host[b + 1], device=dev) for b in range(m.num_batches)]

    @property
    def decomposed_coordinates(self):
        m = self._map()
        return [m.coords[m.off_host[b]:m.off_host[b + 1], 1:] for b in range(m.num_batches)]

    @property
    def decomposed_features(self):
        m = self._map()
        return [self._F[m.off_host[b]:m.off_host[b + 1]] for b in range(m.num_batches)]

    @property
    def decomposed_coordinates_and_features(self):
        return self.decomposed_coordinates, self.decomposed_features

    def features_at_coordinates(self, query_coordinates):
        """Trilinear interpolation of the features at float (b, x, y, z) queries (ME A.9)."""
        m = self._map()
        table, cap = m.ensure_table()
        return ops.interp(table, cap, ops.cast(self._F, torch.float32), m.ts, query_coordinates.float())

    def _same_map(self, other):
        return (self.coordinate_manager is other.coordinate_manager
                and self.coordinate_map_key == other.coordinate_map_key)

    def __add__(self, other):
        if isinstance(other, (int, float)):
            return SparseTensor(self._F + other, coordinate_map_key=self.coordinate_map_key,
                                coordinate_manager=self.coordinate_manager)
        if self.coordinate_manager is not other.coordinate_manager:
            raise B2Error("SparseTensor addition needs a shared coordinate manager")
        grad = AG.wants_grad(self._F, other._F)
        if self._same_map(other):
            f = self._F + other._F if grad else ops.affine_act(self._F.contiguous(), residual=other._F.contiguous())
            return SparseTensor(f, coordinate_map_key=self.coordinate_map_key,
                                coordinate_manager=self.coordinate_manager)
        # different maps: union (ME A.8)
        key, dst_a, dst_b = self.coordinate_manager.union(self.coordinate_map_key, other.coordinate_map_key)
        n = self.coordinate_manager.size(key)
        if grad:  # autograd flows through the row scatter, as through ME's indexing
            rows_ok = (self._F.shape[1] * self._F.element_size()) % 16 == 0 and self._F.dtype in ops.DT
            if rows_ok:
                out = AG.UnionAddFn.apply(self._F, other._F, dst_a, dst_b, n)
            else:
                out = torch.zeros((n, self._F.shape[1]), dtype=self._F.dtype, device=self._F.device)
                out = out.index_add(0, dst_a.long(), self._F).index_add(0, dst_b.long(), other._F.to(self._F.dtype))
            return SparseTensor(out, coordinate_map_key=key, coordinate_manager=self.coordinate_manager)
        out = torch.zeros((n, self._F.shape[1]), dtype=self._F.dtype, device=self._F.device)
        ops.scatter_rows(self._F.contiguous(), dst_a, out, accumulate=False)
        ops.scatter_rows(other._F.contiguous().to(self._F.dtype), dst_b, out, accumulate=True)
        return SparseTensor(out, coordinate_map_key=key, coordinate_manager=self.coordinate_manager)

    def __iadd__(self, other):
        if isinstance(other, SparseTensor) and self._same_map(other) and not AG.wants_grad(self._F, other._F):
            ops.affine_act(self._F, residual=other._F.contiguous(), out=self._F)
            return self
        return self.__add__(other)

    def __repr__(self):
        return "SparseTensor(F=%s, key=%r)" % (tuple(self._F.shape), self.coordinate_map_key)


class TensorField:
    """ME.TensorField: features on the ORIGINAL (unquantised) points plus the map to their voxels.

    `sparse()` quantises (UNWEIGHTED_AVERAGE by default, as in ME), `SparseTensor.slice(field)` brings voxel features
    back to the points.  Not used by the reference's FCAF3D files; provided for ME-surface completeness."""

    def __init__(self, features, coordinates, tensor_stride=1,
                 quantization_mode=SparseTensorQuantizationMode.UNWEIGHTED_AVERAGE, coordinate_manager=None, device=None):
        if device is not None:
            features, coordinates = features.to(device), coordinates.to(device)
        if not coordinates.is_cuda:
            raise B2Error("fcaf3d_b200 TensorField needs CUDA tensors (no CPU backend)")
        self._F = features
        self._C = coordinates
        self.quantization_mode = quantization_mode
        self.tensor_stride = _iso(tensor_stride)
        self.coordinate_manager = coordinate_manager or CoordinateManager(D=coordinates.shape[1] - 1,
                                                                          device=coordinates.device)
        self.inverse_mapping = None
        self.coordinate_map_key = None

    @property
    def F(self):
        return self._F

    features = F

    @property
    def C(self):
        return self._C

    coordinates = C

    def sparse(self, quantization_mode=None):
        mode = self.quantization_mode if quantization_mode is None else quantization_mode
        c = torch.floor(self._C).to(torch.int32) if self._C.is_floating_point() else self._C.to(torch.int32)
        st = SparseTensor(self._F, coordinates=c, tensor_stride=self.tensor_stride, quantization_mode=mode,
                          coordinate_manager=self.coordinate_manager)
        self.inverse_mapping, self.coordinate_map_key = st.inverse_mapping, st.coordinate_map_key
        return st


def _slice(self, field):
    """SparseTensor.slice(TensorField): feature of every original point = feature of its voxel"""
    if field.coordinate_map_key != self.coordinate_map_key or field.inverse_mapping is None:
        raise B2Error("slice() needs the TensorField this tensor was created from (same coordinate map)")
    out = TensorField(ops.index_rows(self._F.contiguous(), field.inverse_mapping.to(torch.int32)), field.C,
                      tensor_stride=field.tensor_stride, quantization_mode=field.quantization_mode,
                      coordinate_manager=self.coordinate_manager)
    out.inverse_mapping, out.coordinate_map_key = field.inverse_mapping, field.coordinate_map_key
    return out


SparseTensor.slice = _slice


# ------------------------------------------------------------------------------------------- modules
_POLICY = {"dtype": torch.float32, "tf32": False}


def set_compute_dtype(dtype):
    """torch.float32 (CUDA-core fp32 path, or the TF32 tensor-core path with set_tf32) or torch.bfloat16 (tcgen05 path for
    wide layers)."""
    assert dtype in (torch.float32, torch.bfloat16)
    _POLICY["dtype"] = dtype


def set_tf32(flag):
    """fp32 feature storage with the wide convolutions on the tensor cores as TF32 operands (b2_conv_tc_tf32; fp32
    accumulation; ~1e-3 of max against the exact fp32 path, which stays the 2e-5 parity mode).  Inference / forward only."""
    _POLICY["tf32"] = bool(flag)


def get_tf32():
    return _POLICY["tf32"]


def get_compute_dtype():
    return _POLICY["dtype"]


class MinkowskiModuleBase(nn.Module):
    pass


class _ConvBase(MinkowskiModuleBase):
    def _wt_bf16(self):
        """bf16 [K, Cout, Cin] copy of the kernel for the tensor-core path (cached per parameter version)."""
        k = self.kernel
        ver = (k._version, k.data_ptr())
        if getattr(self, "_wt_cache_ver", None) != ver:
            w = k.detach()
            if w.dim() == 2:
                w = w.unsqueeze(0)
            self._wt_cache = w.transpose(1, 2).contiguous().to(torch.bfloat16)
            self._wt_cache_ver = ver
        return self._wt_cache

    def _w3(self):
        w = self.kernel.detach()
        return (w.unsqueeze(0) if w.dim() == 2 else w).contiguous()

    def _wt_tf32(self):
        """fp32 [K, Cout, Cin] copy of the kernel for the TF32 tensor-core path; None unless set_tf32(True)."""
        if not _POLICY["tf32"]:
            return None
        k = self.kernel
        ver = (k._version, k.data_ptr())
        if getattr(self, "_wt32_ver", None) != ver:
            self._wt32 = self._w3().transpose(1, 2).contiguous()
            self._wt32_ver = ver
        return self._wt32


class MinkowskiConvolution(_ConvBase):
    """ME.MinkowskiConvolution (A.4).  kernel: [K, in, out], or [in, out] when K == 1 and stride == 1."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, convolution_mode=None, dimension=None):
        super().__init__()
        assert dimension == 3, "dimension must be 3"
        assert not expand_coordinates
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.dilation = _iso(kernel_size), _iso(stride), _iso(dilation)
        self.dimension = dimension
        kvol = self.kernel_size ** 3
        self.kernel_volume = kvol
        self.use_mm = kvol == 1 and self.stride == 1
        shape = (in_channels, out_channels) if self.use_mm else (kvol, in_channels, out_channels)
        self.kernel = nn.Parameter(torch.empty(shape))
        self.bias = nn.Parameter(torch.empty(1, out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        with torch.no_grad():
            n = (self.out_channels if False else self.in_channels) * self.kernel_volume
            stdv = 1.0 / math.sqrt(n)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def maps(self, x):
        """(out_key, kernel map or None for the identity) for input tensor x."""
        cm = x.coordinate_manager
        in_key = x.coordinate_map_key
        out_key = in_key if self.stride == 1 else cm.stride(in_key, self.stride)
        if self.kernel_size == 1 and self.stride == 1:
            return out_key, None
        return out_key, cm.kernel_map(in_key, out_key, self.kernel_size, self.dilation)

    def forward(self, x, coordinates=None):
        out_key, km = self.maps(x)
        n_out = x.coordinate_manager.size(out_key)
        if AG.wants_grad(x.F, self.kernel, self.bias):
            cm = x.coordinate_manager
            if self.use_mm:
                f = x.F @ self.kernel.to(x.F.dtype)          # ME: F.mm(kernel) for 1x1 stride-1 convolutions
            else:
                km_t = None if km is None else cm.kernel_map_transposed(x.coordinate_map_key, out_key,
                                                                         self.kernel_size, self.dilation)
                f = AG.SparseConvFn.apply(x.F, self.kernel, km, km_t, n_out, get_compute_dtype() == torch.bfloat16)
            if self.bias is not None:
                f = f + self.bias.to(f.dtype)
            return SparseTensor(f, coordinate_map_key=out_key, coordinate_manager=cm)
        bias = self.bias.detach().reshape(-1).contiguous() if self.bias is not None else None
        f = ops.sparse_conv(x.F.contiguous(), self._w3(), km, n_out, wt_bf16=self._wt_bf16(), bias=bias,
                            wt_tf32=self._wt_tf32())
        return SparseTensor(f, coordinate_map_key=out_key, coordinate_manager=x.coordinate_manager)

    def __repr__(self):
        return "MinkowskiConvolution(in=%d, out=%d, kernel_size=%s, stride=%s, dilation=%s)" % (
            self.in_channels, self.out_channels, [self.kernel_size] * 3, [self.stride] * 3, [self.dilation] * 3)


class MinkowskiGenerativeConvolutionTranspose(_ConvBase):
    """ME.MinkowskiGenerativeConvolutionTranspose (A.5), kernel_size = stride = 2: every input voxel emits its
    8 children; out[8*i + k] = in[i] @ kernel[k] — a dense GEMM [N, Cin] x [Cin, 8*Cout]."""

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, convolution_mode=None, dimension=None):
        super().__init__()
        assert dimension == 3
        self.kernel_size, self.stride = _iso(kernel_size), _iso(stride)
        if self.kernel_size != 2 or self.stride != 2:
            raise B2Error("generative transpose is implemented for kernel_size=2, stride=2 (FCAF3D's use)")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_volume = 8
        self.kernel = nn.Parameter(torch.empty(8, in_channels, out_channels))
        self.bias = nn.Parameter(torch.empty(1, out_channels)) if bias else None
        with torch.no_grad():
            stdv = 1.0 / math.sqrt(in_channels * 8)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def _packed(self):
        """fp32 [1, Cin, 8*Cout] and bf16 [1, 8*Cout, Cin] views of the kernel (column k*Cout + co)."""
        k = self.kernel
        ver = (k._version, k.data_ptr())
        if getattr(self, "_pk_ver", None) != ver:
            w = k.detach()
            cin, cout = self.in_channels, self.out_channels
            wp = w.permute(1, 0, 2).reshape(1, cin, 8 * cout).contiguous()
            wpt = wp.transpose(1, 2).contiguous()
            self._pk = (wp, wpt.to(torch.bfloat16), wpt)
            self._pk_ver = ver
        return self._pk[:2]

    def _packed_tf32(self):
        self._packed()
        return self._pk[2] if _POLICY["tf32"] else None

    def forward(self, x):
        cm = x.coordinate_manager
        out_key = cm.generate_children(x.coordinate_map_key)
        if AG.wants_grad(x.F, self.kernel):
            packed = self.kernel.permute(1, 0, 2).reshape(1, self.in_channels, 8 * self.out_channels)
            f = AG.SparseConvFn.apply(x.F, packed, None, None, x.F.shape[0], get_compute_dtype() == torch.bfloat16)
            f = f.reshape(x.F.shape[0] * 8, self.out_channels)
        else:
            f = self.generate(x.F.contiguous())
        return SparseTensor(f, coordinate_map_key=out_key, coordinate_manager=cm)

    def generate(self, feats, scale=None, shift=None, act=ops.ACT_NONE):
        wp, wpt = self._packed()
        bias = self.bias.detach().reshape(-1).contiguous() if self.bias is not None else None
        out = ops.sparse_conv(feats, wp, None, feats.shape[0], wt_bf16=wpt, bias=bias, scale=scale, shift=shift, act=act,
                              ch_mod=self.out_channels, wt_tf32=self._packed_tf32())
        return out.view(feats.shape[0] * 8, self.out_channels)


class MinkowskiPruning(MinkowskiModuleBase):
    """ME.MinkowskiPruning (A.10): keeps rows where mask is True, order preserved."""

    def forward(self, x, mask):
        cm = x.coordinate_manager
        key, m8, pos, n_out = cm.prune(x.coordinate_map_key, mask.contiguous())
        f = AG.PruneFn.apply(x.F, m8, pos, n_out) if AG.wants_grad(x.F) else ops.gather_rows(x.F, m8, pos, n_out)
        return SparseTensor(f, coordinate_map_key=key, coordinate_manager=cm)


class MinkowskiMaxPooling(MinkowskiModuleBase):
    def __init__(self, kernel_size, stride=1, dilation=1, kernel_generator=None, dimension=None):
        super().__init__()
        self.kernel_size, self.stride, self.dilation = _iso(kernel_size), _iso(stride), _iso(dilation)

    def forward(self, x):
        cm = x.coordinate_manager
        out_key = x.coordinate_map_key if self.stride == 1 else cm.stride(x.coordinate_map_key, self.stride)
        km = cm.kernel_map(x.coordinate_map_key, out_key, self.kernel_size, self.dilation)
        if AG.wants_grad(x.F):
            f = AG.MaxPoolFn.apply(x.F, km, x.F.shape[0])
        else:
            f = ops.maxpool(x.F.contiguous(), km)
        return SparseTensor(f, coordinate_map_key=out_key, coordinate_manager=cm)


class MinkowskiInstanceNorm(MinkowskiModuleBase):
    """Per batch element, per channel: (x - mean) / sqrt(var + 1e-8) * weight + bias (ME A.6)."""

    def __init__(self, num_features):
        super().__init__()
        self.num_features = num_features
        self.eps = 1e-8
        self.weight = nn.Parameter(torch.ones(1, num_features))
        self.bias = nn.Parameter(torch.zeros(1, num_features))

    def scale_shift(self, feats, m):
        w, b = self.weight.detach(), self.bias.detach()
        if ops.instnorm_supported(feats) and w.dtype == torch.float32 and w.is_contiguous() and b.is_contiguous():
            return ops.instnorm_scale_shift(feats, m.off_dev, m.num_batches, w.reshape(-1), b.reshape(-1), self.eps)
        mean, var = ops.segment_stats(feats, m.off_dev, m.num_batches)
        scale = self.weight.detach() / torch.sqrt(var + self.eps)
        shift = self.bias.detach() - mean * scale
        return scale.contiguous(), shift.contiguous()

    def forward(self, x, act=ops.ACT_NONE):
        m = x._map()
        f = x.F.contiguous()
        if (AG.wants_grad(f, self.weight, self.bias) and AG.FUSED_BN and f.shape[1] % 8 == 0 and f.shape[1] <= 2048
                and f.dtype in ops.DT):
            out = AG.InstanceNormActFn.apply(f, self.weight, self.bias, m.off_host, self.eps, act)
            return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)
        if AG.wants_grad(f, self.weight, self.bias):
            # differentiable per-scene statistics (ME: global pooling + broadcast, both with autograd).  Rows of a
            # scene are contiguous, so plain slices are used: indexing the [B, C] statistics with the per-row batch
            # index would make autograd accumulate ~100k duplicate rows per scene in its backward
            outs = []
            for bi in range(m.num_batches):
                fs = f[m.off_host[bi]:m.off_host[bi + 1]]
                if fs.shape[0] == 0:
                    outs.append(fs)
                    continue
                mean = fs.mean(0, keepdim=True)
                xc = fs - mean
                var = (xc * xc).mean(0, keepdim=True)
                outs.append(xc / torch.sqrt(var + self.eps) * self.weight + self.bias)
            out = torch.cat(outs, 0)
            if act == ops.ACT_RELU:
                out = torch.relu(out)
            return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)
        scale, shift = self.scale_shift(f, m)
        out = ops.affine_act(f, scale, shift, coords=m.coords, act=act)
        return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)


class MinkowskiBatchNorm(MinkowskiModuleBase):
    """nn.BatchNorm1d over the feature rows (ME A.6); statistics are per GPU, never synchronised."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.bn = nn.BatchNorm1d(num_features, eps=eps, momentum=momentum, affine=affine,
                                 track_running_stats=track_running_stats)

    def scale_shift(self, feats):
        """Per-channel (scale, shift) [1, C]; in training mode also updates the running statistics."""
        bn = self.bn
        if bn.training or not bn.track_running_stats:
            n = feats.shape[0]
            off = torch.tensor([0, n], dtype=torch.int32, device=feats.device)
            mean, var = ops.segment_stats(feats, off, 1)
            if bn.training and bn.track_running_stats:
                with torch.no_grad():
                    mom = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked + 1)
                    bn.num_batches_tracked += 1
                    unbiased = var[0] * (n / max(n - 1, 1))
                    bn.running_mean.mul_(1 - mom).add_(mean[0], alpha=mom)
                    bn.running_var.mul_(1 - mom).add_(unbiased, alpha=mom)
        else:
            # eval: the folded (scale, shift) only changes when a parameter / buffer does -> cache it
            ts = [t for t in (bn.weight, bn.bias, bn.running_mean, bn.running_var) if t is not None]
            ver = tuple((t._version, t.data_ptr()) for t in ts)
            if getattr(self, "_fold_ver", None) == ver:
                return self._fold
            mean, var = bn.running_mean.unsqueeze(0), bn.running_var.unsqueeze(0)
        inv = torch.rsqrt(var + bn.eps)
        w = bn.weight.detach().unsqueeze(0) if bn.affine else torch.ones_like(inv)
        b = bn.bias.detach().unsqueeze(0) if bn.affine else torch.zeros_like(inv)
        scale = w * inv
        shift = b - mean * scale
        out = (scale.float().contiguous(), shift.float().contiguous())
        if not (bn.training or not bn.track_running_stats):
            self._fold, self._fold_ver = out, ver
        return out

    def forward(self, x, act=ops.ACT_NONE, residual=None):
        f = x.F.contiguous()
        if AG.wants_grad(f, self.bn.weight, self.bn.bias):
            if self.bn.training and AG.bn_act_supported(f, self.bn):
                # batch statistics, normalisation, residual and activation on the library's kernels (bn_ops.cu)
                res = None if residual is None else residual.to(f.dtype)
                out = AG.BNActFn.apply(f, self.bn.weight, self.bn.bias, res, self.bn, act)
                return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)
            # exactly ME: nn.BatchNorm1d on the feature rows (bf16 rows: torch's mixed-dtype kernels, fp32 statistics)
            out = self.bn(f) if f.dtype == torch.bfloat16 else self.bn(f.float()).to(f.dtype)
            if residual is not None:
                out = out + residual
            out = torch.relu(out) if act == ops.ACT_RELU else torch.nn.functional.elu(out) if act == ops.ACT_ELU else out
            return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)
        scale, shift = self.scale_shift(f)
        out = ops.affine_act(f, scale, shift, residual=residual, act=act)
        return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)


class MinkowskiReLU(MinkowskiModuleBase):
    def __init__(self, inplace=False):
        super().__init__()
        self.inplace = inplace

    def forward(self, x):
        f = x.F
        if AG.wants_grad(f):
            return SparseTensor(torch.relu(f), coordinate_map_key=x.coordinate_map_key,
                                coordinate_manager=x.coordinate_manager)
        out = ops.affine_act(f, act=ops.ACT_RELU, out=f if self.inplace else None)
        if self.inplace:
            return x
        return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)


class MinkowskiELU(MinkowskiModuleBase):
    def __init__(self, alpha=1.0, inplace=False):
        super().__init__()
        assert alpha == 1.0
        self.inplace = inplace

    def forward(self, x):
        if AG.wants_grad(x.F):
            return SparseTensor(torch.nn.functional.elu(x.F), coordinate_map_key=x.coordinate_map_key,
                                coordinate_manager=x.coordinate_manager)
        out = ops.affine_act(x.F.contiguous(), act=ops.ACT_ELU)
        return SparseTensor(out, coordinate_map_key=x.coordinate_map_key, coordinate_manager=x.coordinate_manager)


from . import utils  # noqa: E402
from . import modules  # noqa: E402
