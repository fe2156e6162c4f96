"""Device-side data path of FCAF3D training.

The reference augments every scene on the host with numpy (configs/fcaf3d/fcaf3d_scannet-3d-18class.py:16-41 ->
mmdet3d/datasets/pipelines/transforms_3d.py: IndoorPointSample :821-891, RandomFlip3D :59-162, GlobalRotScaleTrans
:497-648) in 4 dataloader workers per GPU; at hundreds of scenes per second per B200 that pipeline starves the GPU.
Here the raw scene is copied once from pinned memory and sampled / flipped / rotated / scaled / translated by ONE kernel
per scene (`b2_augment_points`), the boxes by `b2_augment_boxes`:

  DeviceAugment      the three transforms with the reference's parameters; random numbers either from a host numpy
                     Generator in the reference's draw order (`draw`) or, for the sampling indices, from the device
  ScenePrefetcher    double-buffered pinned-host -> device staging on a side stream: the copy of batch i+1 overlaps the
                     training step of batch i

No CPU fallback: CUDA tensors only.
"""
import numpy as np
import torch

from ._lib import B2Error, call, ptr


class DeviceAugment:
    def __init__(self, num_points=100000, flip_ratio_bev_horizontal=0.5, flip_ratio_bev_vertical=0.5,
                 rot_range=(-0.087266, 0.087266), scale_ratio_range=(0.9, 1.1), translation_std=(0.1, 0.1, 0.1),
                 with_yaw=False, seed=None, device_sampling=True):
        self.num_points = int(num_points)
        self.flip_h, self.flip_v = float(flip_ratio_bev_horizontal), float(flip_ratio_bev_vertical)
        self.rot_range = tuple(rot_range) if not isinstance(rot_range, (int, float)) else (-rot_range, rot_range)
        self.scale_ratio_range = tuple(scale_ratio_range)
        std = translation_std if not isinstance(translation_std, (int, float)) else (translation_std,) * 3
        self.translation_std = tuple(float(s) for s in std)
        self.with_yaw = bool(with_yaw)
        self.device_sampling = bool(device_sampling)
        self.rng = np.random.default_rng(seed)

    def draw(self, n_in):
        """One scene's random numbers, in the order the reference's pipeline consumes them (sampling indices, the unused
        2-D flip draw of mmdet's RandomFlip, the two BEV flips, angle, scale, translation)."""
        r = self.rng
        choices = None
        if not self.device_sampling:
            choices = r.choice(n_in, self.num_points, replace=n_in < self.num_points)
        r.choice(2, p=[self.flip_h, 1 - self.flip_h])
        flip_h = bool(r.random() < self.flip_h)
        flip_v = bool(r.random() < self.flip_v)
        angle = float(r.uniform(self.rot_range[0], self.rot_range[1]))
        scale = float(r.uniform(self.scale_ratio_range[0], self.scale_ratio_range[1]))
        trans = [float(v) for v in r.normal(scale=np.asarray(self.translation_std, dtype=np.float32), size=3)]
        return dict(choices=choices, flip_h=flip_h, flip_v=flip_v, angle=angle, scale=scale, trans=trans)

    def apply(self, points, boxes, params):
        """points [n_in, 3 + attrs] fp32 cuda, boxes [B, 7] (gravity centre, size, yaw) fp32 cuda or None."""
        if not points.is_cuda:
            raise B2Error("DeviceAugment works on CUDA tensors only (no CPU path)")
        points = points.contiguous().float()
        dev, n_in, dim = points.device, int(points.shape[0]), int(points.shape[1])
        choices = params.get("choices")
        if choices is None and self.num_points > 0:
            if n_in >= self.num_points:          # without replacement, like np.random.choice(replace=False)
                choices = torch.randperm(n_in, device=dev)[: self.num_points]
            else:
                choices = torch.randint(0, n_in, (self.num_points,), device=dev)
        elif choices is not None and not torch.is_tensor(choices):
            choices = torch.from_numpy(np.ascontiguousarray(choices, dtype=np.int64)).to(dev, non_blocking=True)
        n_out = int(choices.shape[0]) if choices is not None else n_in
        out = torch.empty((n_out, dim), dtype=torch.float32, device=dev)
        err = torch.zeros((1,), dtype=torch.int32, device=dev)
        t = params["trans"]
        call("b2_augment_points", ptr(points), n_in, dim, ptr(choices.contiguous() if choices is not None else None),
             n_out, int(params["flip_h"]), int(params["flip_v"]), float(params["angle"]), float(params["scale"]),
             float(t[0]), float(t[1]), float(t[2]), ptr(out), ptr(err))
        out_boxes = None
        if boxes is not None:
            boxes = boxes.contiguous().float()
            out_boxes = torch.empty_like(boxes)
            call("b2_augment_boxes", ptr(boxes), int(boxes.shape[0]), int(self.with_yaw), int(params["flip_h"]),
                 int(params["flip_v"]), float(params["angle"]), float(params["scale"]), float(t[0]), float(t[1]),
                 float(t[2]), ptr(out_boxes))
        return out, out_boxes, err

    def __call__(self, points_list, boxes_list=None):
        """lists over scenes -> (augmented points list, augmented boxes list); raises on an out-of-range index."""
        outs, bouts, errs = [], [], []
        for i, p in enumerate(points_list):
            b = boxes_list[i] if boxes_list is not None else None
            o, bo, e = self.apply(p, b, self.draw(int(p.shape[0])))
            outs.append(o)
            bouts.append(bo)
            errs.append(e)
        if errs and int(torch.stack(errs).max().item()):
            raise B2Error("sampling index outside the scene")
        return outs, (bouts if boxes_list is not None else None)


class ReadyList(list):
    """list of device tensors + `ready`: the event after which their host->device copies are complete"""
    ready = None


class ScenePrefetcher:
    """Iterates over `batches` (an iterable of lists of CPU tensors / numpy arrays, e.g. the raw scenes of one training
    step) and yields lists of CUDA tensors; while the consumer works on batch i, batch i+1 is staged through pinned
    memory and copied on a side stream."""

    _streams = {}

    def __init__(self, batches, device):
        self.batches = iter(batches)
        self.device = torch.device(device)
        # one copy stream per device for the life of the process: torch's allocator keeps a pool per stream, so a fresh
        # stream per prefetcher would cudaMalloc its first batches again (inside whatever loop is being timed)
        key = self.device.index if self.device.index is not None else torch.cuda.current_device()
        if key not in ScenePrefetcher._streams:
            ScenePrefetcher._streams[key] = torch.cuda.Stream(device=self.device)
        self.stream = ScenePrefetcher._streams[key]
        self._next = None
        self._preload()

    def _preload(self):
        try:
            batch = next(self.batches)
        except StopIteration:
            self._next = None
            return
        host = []
        for t in batch:
            t = torch.from_numpy(np.ascontiguousarray(t)) if isinstance(t, np.ndarray) else t.contiguous()
            host.append(t if t.is_pinned() else t.pin_memory())
        with torch.cuda.stream(self.stream):
            dev = [h.to(self.device, non_blocking=True) for h in host]
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._next = (dev, ev, host)          # the pinned buffers stay alive until the copy has been consumed

    def __iter__(self):
        return self

    def __next__(self):
        if self._next is None:
            raise StopIteration
        dev, ev, _ = self._next
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for t in dev:
            t.record_stream(cur)
        self._preload()
        out = ReadyList(dev)
        out.ready = ev               # consumers with streams of their own (the native executor) wait for the copy directly
        return out
