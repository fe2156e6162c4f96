"""ORACLE (test infrastructure, never shipped or measured as the product): a CPU restatement of the
MinkowskiEngine v0.5.4 Python surface that FCAF3D uses, in numpy + torch-CPU.

PARITY UNPINNED: MinkowskiEngine is an un-vendored dependency of the reference (docker/Dockerfile:24-32),
absent from /root/reference and not installable here, and no reference test covers this path.
This file restates ME's published behaviour and is anchored on the reference call sites:
  single_stage_sparse.py:34-37 (batch_sparse_collate, SparseTensor), me_resnet.py:18-25,45-78 (conv, instance
  norm, relu, max pool, BasicBlock), fcaf3d_neck_with_head.py:49-72,83-85,101,110-126,256-279 (generative
  transpose, union add, features_at_coordinates, pruning, heads).
Self-checks that need no reference (tests/test_oracle.py): dense-grid conv == torch conv3d, generative
transpose == conv_transpose3d, max pool == max_pool3d.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
Everything is differentiable through torch autograd (index_select / mm / index_add), which is what the
training-path tests use as the gradient oracle.
"""
import math

import numpy as np
import torch
import torch.nn as nn

BIAS = 131072


def pack_keys(coords):
    """(b, x, y, z) int rows -> int64 keys whose order is the lexicographic order of the rows."""
    c = np.asarray(coords, dtype=np.int64)
    return (c[:, 0] << 54) | ((c[:, 1] + BIAS) << 36) | ((c[:, 2] + BIAS) << 18) | (c[:, 3] + BIAS)


class SparseTensorQuantizationMode:
    RANDOM_SUBSAMPLE = 0
    UNWEIGHTED_AVERAGE = 1
    UNWEIGHTED_SUM = 2
    NO_QUANTIZATION = 3


class CoordinateMapKey:
    def __init__(self, tensor_stride, string_id=""):
        if isinstance(tensor_stride, int):
            tensor_stride = (tensor_stride,) * 3
        self._ts = tuple(int(t) for t in tensor_stride)
        self._id = str(string_id)

    def get_tensor_stride(self):
        return list(self._ts)

    def get_key(self):
        return (list(self._ts), self._id)

    def __eq__(self, o):
        return isinstance(o, CoordinateMapKey) and self._ts == o._ts and self._id == o._id

    def __hash__(self):
        return hash((self._ts, self._id))

    def __repr__(self):
        return "oracle key:%s:%s" % (list(self._ts), self._id)


class _Map:
    def __init__(self, coords, ts):
        self.coords = np.ascontiguousarray(coords, dtype=np.int32)
        self.n = len(self.coords)
        self.ts = int(ts)
        keys = pack_keys(self.coords)
        self.order = np.argsort(keys, kind="stable")
        self.sorted_keys = keys[self.order]

    def lookup(self, coords):
        """rows of the given coordinates (-1 where absent)."""
        if self.n == 0:
            return np.full(len(coords), -1, dtype=np.int64)
        keys = pack_keys(coords)
        pos = np.searchsorted(self.sorted_keys, keys)
        pos_c = np.minimum(pos, self.n - 1)
        hit = self.sorted_keys[pos_c] == keys
        return np.where(hit, self.order[pos_c], -1)

    def batch_offsets(self, nb):
        b = self.coords[:, 0]
        assert np.all(b[1:] >= b[:-1]), "rows must be grouped by batch"
        return np.searchsorted(b, np.arange(nb + 1), side="left")


def dedup_first_occurrence(coords):
    """unique rows in first-occurrence order, representative = lowest row (ME CPU backend rule, App. A.2)."""
    keys = pack_keys(coords)
    _, first, inv = np.unique(keys, return_index=True, return_inverse=True)
    rank = np.argsort(np.argsort(first, kind="stable"), kind="stable")  # unique id -> output row
    unique_index = np.sort(first)
    return unique_index, rank[inv]


def kernel_offsets(ksize, step):
    """ME region order: x fastest; odd kernels centred, even kernels start at the voxel."""
    base = -(ksize // 2) if ksize % 2 else 0
    offs = []
    for iz in range(ksize):
        for iy in range(ksize):
            for ix in range(ksize):
                offs.append(((ix + base) * step, (iy + base) * step, (iz + base) * step))
    return offs


class CoordinateManager:
    def __init__(self, D=3):
        self.D = D
        self._maps = {}
        self._kmaps = {}
        self._stride_cache = {}
        self._auto = 0
        self.num_batches = 0

    def _fresh(self, ts):
        self._auto += 1
        return CoordinateMapKey(ts, "o%d" % self._auto)

    def insert_and_map(self, coordinates, tensor_stride=1, string_id=""):
        c = coordinates.detach().cpu().numpy().astype(np.int32)
        self.num_batches = max(self.num_batches, int(c[:, 0].max()) + 1 if len(c) else 1)
        uidx, inv = dedup_first_occurrence(c)
        key = CoordinateMapKey(tensor_stride, string_id)
        if key in self._maps:
            key = self._fresh(tensor_stride)
        self._maps[key] = _Map(c[uidx], tensor_stride if isinstance(tensor_stride, int) else tensor_stride[0])
        return key, (torch.from_numpy(uidx.astype(np.int64)), torch.from_numpy(inv.astype(np.int64)))

    def stride(self, in_key, s):
        ck = (in_key, s)
        if ck in self._stride_cache:
            return self._stride_cache[ck]
        src = self._maps[in_key]
        ts = src.ts * s
        c = src.coords.copy()
        c[:, 1:] = np.floor_divide(c[:, 1:], ts) * ts
        uidx, _ = dedup_first_occurrence(c)
        key = CoordinateMapKey(ts, "")
        if key in self._maps:
            key = self._fresh(ts)
        self._maps[key] = _Map(c[uidx], ts)
        self._stride_cache[ck] = key
        return key

    def generate_children(self, in_key):
        src = self._maps[in_key]
        half = src.ts // 2
        offs = np.array([(0,) + o for o in kernel_offsets(2, half)], dtype=np.int32)  # [8, 4]
        c = (src.coords[:, None, :] + offs[None, :, :]).reshape(-1, 4)
        key = self._fresh(half)
        self._maps[key] = _Map(c, half)
        return key

    def union(self, key_a, key_b):
        """scene s of the union = [a rows of s][b-only rows of s] (same documented order as the CUDA path)."""
        a, b = self._maps[key_a], self._maps[key_b]
        assert a.ts == b.ts
        found = a.lookup(b.coords)
        missing = found < 0
        nb = self.num_batches
        a_off, b_off = a.batch_offsets(nb), b.batch_offsets(nb)
        pos = np.concatenate([[0], np.cumsum(missing)])  # exclusive scan, pos[len] = total
        shift_of_scene = pos[b_off[:nb]]
        dst_a = np.arange(a.n) + shift_of_scene[a.coords[:, 0]]
        dst_b = np.where(missing, a_off[b.coords[:, 0] + 1] + pos[:-1], found + shift_of_scene[b.coords[:, 0]])
        n = a.n + int(missing.sum())
        out = np.zeros((n, 4), dtype=np.int32)
        out[dst_a] = a.coords
        out[dst_b[missing]] = b.coords[missing]
        key = self._fresh(a.ts)
        self._maps[key] = _Map(out, a.ts)
        return key, torch.from_numpy(dst_a.astype(np.int64)), torch.from_numpy(dst_b.astype(np.int64))

    def prune(self, in_key, mask):
        src = self._maps[in_key]
        key = self._fresh(src.ts)
        self._maps[key] = _Map(src.coords[mask.cpu().numpy().astype(bool)], src.ts)
        return key

    def get_map(self, key):
        return self._maps[key]

    def size(self, key):
        return self._maps[key].n

    def kernel_map(self, in_key, out_key, ksize, dilation=1):
        """list over kernel offsets of (in_rows, out_rows) int64 tensors."""
        ck = (in_key, out_key, ksize, dilation)
        if ck not in self._kmaps:
            src, dst = self._maps[in_key], self._maps[out_key]
            pairs = []
            for off in kernel_offsets(ksize, src.ts * dilation):
                q = dst.coords.copy()
                q[:, 1] += off[0]
                q[:, 2] += off[1]
                q[:, 3] += off[2]
                rows = src.lookup(q)
                o = np.nonzero(rows >= 0)[0]
                pairs.append((torch.from_numpy(rows[o].astype(np.int64)), torch.from_numpy(o.astype(np.int64))))
            self._kmaps[ck] = pairs
        return self._kmaps[ck]

    def neighbour_table(self, in_key, out_key, ksize, dilation=1):
        """[K, n_out] int64 table with -1 for absent neighbours (the CUDA path's kernel-map format)."""
        pairs = self.kernel_map(in_key, out_key, ksize, dilation)
        n_out = self._maps[out_key].n
        t = np.full((len(pairs), n_out), -1, dtype=np.int64)
        for k, (i, o) in enumerate(pairs):
            t[k, o.numpy()] = i.numpy()
        return t


def _iso(v):
    return int(v[0]) if isinstance(v, (list, tuple)) else int(v)


class SparseTensor:
    def __init__(self, features, coordinates=None, tensor_stride=1, coordinate_map_key=None, coordinate_manager=None,
                 quantization_mode=SparseTensorQuantizationMode.RANDOM_SUBSAMPLE, device=None, **kw):
        self.unique_index = self.inverse_mapping = None
        if coordinates is not None:
            if coordinate_manager is None:
                coordinate_manager = CoordinateManager()
            key, (uidx, inv) = coordinate_manager.insert_and_map(coordinates, _iso(tensor_stride))
            self.unique_index, self.inverse_mapping = uidx, inv
            features = features[uidx]
            coordinate_map_key = key
        self._F = features
        self.coordinate_map_key = coordinate_map_key
        self.coordinate_manager = coordinate_manager
        assert coordinate_manager.size(coordinate_map_key) == features.shape[0]

    @property
    def F(self):
        return self._F

    features = F

    @property
    def C(self):
        return torch.from_numpy(self.coordinate_manager.get_map(self.coordinate_map_key).coords)

    coordinates = C

    @property
    def tensor_stride(self):
        return self.coordinate_map_key.get_tensor_stride()

    @property
    def D(self):
        return 3

    @property
    def device(self):
        return self._F.device

    @property
    def dtype(self):
        return self._F.dtype

    @property
    def shape(self):
        return self._F.shape

    def __len__(self):
        return self._F.shape[0]

    def _map(self):
        return self.coordinate_manager.get_map(self.coordinate_map_key)

    def _offsets(self):
        return self._map().batch_offsets(self.coordinate_manager.num_batches)

    @property
    def decomposition_permutations(self):
        off = self._offsets()
        return [torch.arange(off[b], off[b + 1]) for b in range(len(off) - 1)]

    @property
    def decomposed_coordinates(self):
        off = self._offsets()
        c = self.C
        return [c[off[b]:off[b + 1], 1:] for b in range(len(off) - 1)]

    @property
    def decomposed_features(self):
        off = self._offsets()
        return [self._F[off[b]:off[b + 1]] for b in range(len(off) - 1)]

    def features_at_coordinates(self, q):
        """trilinear interpolation, absent corners contribute 0 (App. A.9)."""
        m = self._map()
        ts = m.ts
        qn = q.detach().cpu().numpy().astype(np.float32)
        b = qn[:, 0].astype(np.int32)
        base = (np.floor(qn[:, 1:] / np.float32(ts)).astype(np.int32)) * ts
        out = torch.zeros((len(qn), self._F.shape[1]), dtype=torch.float32)
        for k in range(8):
            off = np.array([(k & 1) * ts, ((k >> 1) & 1) * ts, ((k >> 2) & 1) * ts], dtype=np.int32)
            corner = base + off[None, :]
            w = np.prod(np.float32(1) - np.abs(qn[:, 1:] - corner.astype(np.float32)) / np.float32(ts), axis=1)
            rows = m.lookup(np.concatenate([b[:, None], corner], axis=1))
            ok = (rows >= 0) & (w != 0)
            idx = torch.from_numpy(np.nonzero(ok)[0])
            out[idx] += torch.from_numpy(w[ok].astype(np.float32))[:, None] * self._F[torch.from_numpy(rows[ok])].float()
        return out

    def __add__(self, other):
        cm = self.coordinate_manager
        if self.coordinate_map_key == other.coordinate_map_key:
            return SparseTensor(self._F + other._F, coordinate_map_key=self.coordinate_map_key, coordinate_manager=cm)
        key, dst_a, dst_b = cm.union(self.coordinate_map_key, other.coordinate_map_key)
        out = torch.zeros((cm.size(key), self._F.shape[1]), dtype=self._F.dtype)
        out = out.index_add(0, dst_a, self._F).index_add(0, dst_b, other._F)
        return SparseTensor(out, coordinate_map_key=key, coordinate_manager=cm)

    def __iadd__(self, other):
        return self.__add__(other)


# ------------------------------------------------------------------------------------------- modules
def sparse_conv_ref(feats, kernel, pairs, n_out):
    """out[o] = sum_k sum_{(i,o) in M_k} feats[i] @ kernel[k]  (App. A.4), fp32, differentiable."""
    out = torch.zeros((n_out, kernel.shape[-1]), dtype=feats.dtype)
    for k, (i, o) in enumerate(pairs):
        if len(i):
            out = out.index_add(0, o, feats.index_select(0, i) @ kernel[k])
    return out


class MinkowskiConvolution(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, convolution_mode=None, dimension=None):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.dilation = _iso(kernel_size), _iso(stride), _iso(dilation)
        kvol = self.kernel_size ** 3
        self.kernel_volume = kvol
        self.use_mm = kvol == 1 and self.stride == 1
        shape = (in_channels, out_channels) if self.use_mm else (kvol, in_channels, out_channels)
        self.kernel = nn.Parameter(torch.empty(shape))
        self.bias = nn.Parameter(torch.empty(1, out_channels)) if bias else None
        with torch.no_grad():
            stdv = 1.0 / math.sqrt(in_channels * kvol)
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def forward(self, x):
        cm = x.coordinate_manager
        in_key = x.coordinate_map_key
        if self.use_mm:
            f = x.F @ self.kernel
            out_key = in_key
        else:
            out_key = in_key if self.stride == 1 else cm.stride(in_key, self.stride)
            pairs = cm.kernel_map(in_key, out_key, self.kernel_size, self.dilation)
            f = sparse_conv_ref(x.F, self.kernel, pairs, cm.size(out_key))
        if self.bias is not None:
            f = f + self.bias
        return SparseTensor(f, coordinate_map_key=out_key, coordinate_manager=cm)


class MinkowskiGenerativeConvolutionTranspose(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, convolution_mode=None, dimension=None):
        super().__init__()
        assert _iso(kernel_size) == 2 and _iso(stride) == 2
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.kernel_volume = 2, 2, 8
        self.kernel = nn.Parameter(torch.empty(8, in_channels, out_channels))
        self.bias = None
        with torch.no_grad():
            stdv = 1.0 / math.sqrt(in_channels * 8)
            self.kernel.uniform_(-stdv, stdv)

    def forward(self, x):
        cm = x.coordinate_manager
        out_key = cm.generate_children(x.coordinate_map_key)
        # out[8 i + k] = in[i] @ kernel[k]
        f = torch.einsum("nc,kcd->nkd", x.F, self.kernel).reshape(-1, self.out_channels)
        return SparseTensor(f, coordinate_map_key=out_key, coordinate_manager=cm)


class MinkowskiPruning(nn.Module):
    def forward(self, x, mask):
        key = x.coordinate_manager.prune(x.coordinate_map_key, mask)
        return SparseTensor(x.F[mask], coordinate_map_key=key, coordinate_manager=x.coordinate_manager)


class MinkowskiMaxPooling(nn.Module):
    def __init__(self, kernel_size, stride=1, dilation=1, kernel_generator=None, dimension=None):
        super().__init__()
        self.kernel_size, self.stride, self.dilation = _iso(kernel_size), _iso(stride), _iso(dilation)

    def forward(self, x):
        cm = x.coordinate_manager
        out_key = x.coordinate_map_key if self.stride == 1 else cm.stride(x.coordinate_map_key, self.stride)
        # per output voxel the per-channel max over its present inputs; the gradient goes to the arg-max input only
        # (ME keeps an arg-max index per output element; torch.max(dim) returns the first maximal index)
        table = torch.from_numpy(cm.neighbour_table(x.coordinate_map_key, out_key, self.kernel_size, self.dilation))
        n_in, c = x.F.shape
        padded = torch.cat([x.F, torch.full((1, c), -float("inf"), dtype=x.F.dtype)], 0)
        cand = padded[torch.where(table >= 0, table, torch.full_like(table, n_in)).t()]      # [n_out, K, C]
        out = cand.max(dim=1).values
        out = torch.where(torch.isinf(out), torch.zeros_like(out), out)
        return SparseTensor(out, coordinate_map_key=out_key, coordinate_manager=cm)


class MinkowskiInstanceNorm(nn.Module):
    def __init__(self, num_features):
        super().__init__()
        self.num_features, self.eps = num_features, 1e-8
        self.weight = nn.Parameter(torch.ones(1, num_features))
        self.bias = nn.Parameter(torch.zeros(1, num_features))

    def forward(self, x):
        outs = []
        for f in x.decomposed_features:
            if len(f) == 0:
                outs.append(f)
                continue
            mean = f.mean(0, keepdim=True)
            var = ((f - mean) ** 2).mean(0, keepdim=True)
            outs.append((f - mean) / torch.sqrt(var + self.eps) * self.weight + self.bias)
        return SparseTensor(torch.cat(outs, 0), coordinate_map_key=x.coordinate_map_key,
                            coordinate_manager=x.coordinate_manager)


class MinkowskiBatchNorm(nn.Module):
    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.bn = nn.BatchNorm1d(num_features, eps=eps, momentum=momentum, affine=affine,
                                 track_running_stats=track_running_stats)

    def forward(self, x):
        return SparseTensor(self.bn(x.F), coordinate_map_key=x.coordinate_map_key,
                            coordinate_manager=x.coordinate_manager)


class MinkowskiReLU(nn.Module):
    def __init__(self, inplace=False):
        super().__init__()

    def forward(self, x):
        return SparseTensor(torch.relu(x.F), coordinate_map_key=x.coordinate_map_key,
                            coordinate_manager=x.coordinate_manager)


class MinkowskiELU(nn.Module):
    def __init__(self, alpha=1.0, inplace=False):
        super().__init__()

    def forward(self, x):
        return SparseTensor(torch.nn.functional.elu(x.F), coordinate_map_key=x.coordinate_map_key,
                            coordinate_manager=x.coordinate_manager)


class _Utils:
    """ME.utils subset (App. A.1, A.4)."""

    @staticmethod
    def batch_sparse_collate(data, dtype=torch.int32, device=None):
        cs, fs = [], []
        for b, (c, f) in enumerate(data):
            ci = torch.floor(c).to(dtype) if c.is_floating_point() else c.to(dtype)
            cs.append(torch.cat((torch.full((len(ci), 1), b, dtype=dtype), ci), 1))
            fs.append(f)
        return torch.cat(cs, 0), torch.cat(fs, 0)

    @staticmethod
    def kaiming_normal_(tensor, a=0, mode="fan_in", nonlinearity="leaky_relu"):
        if tensor.dim() == 2:
            fan_in, fan_out = tensor.size(1), tensor.size(0)
        else:
            fan_in, fan_out = tensor.size(1) * tensor.size(0), tensor.size(2) * tensor.size(0)
        fan = fan_in if mode == "fan_in" else fan_out
        std = torch.nn.init.calculate_gain(nonlinearity, a) / math.sqrt(fan)
        with torch.no_grad():
            return tensor.normal_(0, std)


utils = _Utils()


class BasicBlock(nn.Module):
    """MinkowskiEngine.modules.resnet_block.BasicBlock (App. A.7)."""
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, dilation=1, downsample=None, bn_momentum=0.1, dimension=-1):
        super().__init__()
        self.conv1 = MinkowskiConvolution(inplanes, planes, kernel_size=3, stride=stride, dilation=dilation,
                                          dimension=dimension)
        self.norm1 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.conv2 = MinkowskiConvolution(planes, planes, kernel_size=3, stride=1, dilation=dilation,
                                          dimension=dimension)
        self.norm2 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.relu = MinkowskiReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        residual = x
        out = self.relu(self.norm1(self.conv1(x)))
        out = self.norm2(self.conv2(out))
        if self.downsample is not None:
            residual = self.downsample(x)
        out = out + residual
        return self.relu(out)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, dilation=1, downsample=None, bn_momentum=0.1, dimension=-1):
        super().__init__()
        self.conv1 = MinkowskiConvolution(inplanes, planes, kernel_size=1, dimension=dimension)
        self.norm1 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.conv2 = MinkowskiConvolution(planes, planes, kernel_size=3, stride=stride, dilation=dilation,
                                          dimension=dimension)
        self.norm2 = MinkowskiBatchNorm(planes, momentum=bn_momentum)
        self.conv3 = MinkowskiConvolution(planes, planes * 4, kernel_size=1, dimension=dimension)
        self.norm3 = MinkowskiBatchNorm(planes * 4, momentum=bn_momentum)
        self.relu = MinkowskiReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        residual = x
        out = self.relu(self.norm1(self.conv1(x)))
        out = self.relu(self.norm2(self.conv2(out)))
        out = self.norm3(self.conv3(out))
        if self.downsample is not None:
            residual = self.downsample(x)
        return self.relu(out + residual)


def install_as_minkowski_engine():
    """Registers this oracle as `MinkowskiEngine` so the reference's model files run on it (CPU, this container)."""
    import sys
    import types
    me = sys.modules[__name__]
    mods = types.ModuleType("MinkowskiEngine.modules")
    rb = types.ModuleType("MinkowskiEngine.modules.resnet_block")
    rb.BasicBlock, rb.Bottleneck = BasicBlock, Bottleneck
    mods.resnet_block = rb
    sys.modules["MinkowskiEngine"] = me
    sys.modules["MinkowskiEngine.modules"] = mods
    sys.modules["MinkowskiEngine.modules.resnet_block"] = rb
    return me
