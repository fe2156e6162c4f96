"""Native executor of the fused inference step (csrc/engine.cu) bound to a detector's weights.

`detector.simple_test_fused` drives one step as ~200 C-ABI calls from python; on a B200 the interpreter then needs as long
per batch as the GPU.  `InferenceEngine(det)` hands the SAME sequence of library calls to one C++ entry point
(`b2_engine_infer`): same kernels, same arguments, bit-identical detections, a host cost of the launches alone.
Eval mode only (BatchNorm folded), BasicBlock backbones (depth 14 / 18 / 34); anything else stays on the python path.
"""
import ctypes
import os

import torch

from . import me as ME
from . import ops
from ._lib import B2Error, last_error, lib, ptr, stream
from .me.modules.resnet_block import BasicBlock


def supported(det):
    bb = det.backbone
    return all(isinstance(b, BasicBlock) for i in range(bb.n_outs) for b in getattr(bb, "layer%d" % (i + 1))) and \
        det.neck_with_head._cfg("nms_pre") > 0 and bb.n_outs <= 4


class InferenceEngine:
    def __init__(self, det):
        if not supported(det):
            raise B2Error("InferenceEngine: BasicBlock backbones with nms_pre > 0 only")
        self.det = det
        self._handle = None
        self._sig = None
        self._arenas = [None, None]      # alternate: the tail of step i may still read arena i % 2 while step i + 1 runs
        self._arena_free = [None, None]  # event after which the arena may be overwritten (its last tail has drained)
        self._step = 0
        self._hold = [None, None]
        self._tail = None                # side stream of the top-k / NMS tail (B2_ENGINE_OVERLAP=0: none)
        self.overlap_tail = os.environ.get("B2_ENGINE_OVERLAP", "1") != "0"
        self.split_streams = os.environ.get("B2_ENGINE_SPLIT", "1") != "0"
        self._keep = []
        self._tensors = None

    # ---- weight table -------------------------------------------------------------------------------------------------
    def _signature(self):
        if self._tensors is None:
            self._tensors = list(self.det.parameters()) + list(self.det.buffers())
        return (ME.get_tf32(),) + tuple((t._version, t.data_ptr()) for t in self._tensors)

    def _rec(self, words, conv=None, bn=None, wt=None, w32=None, bias=None, kvol=None, cin=None, cout=None):
        """one 8-word record: wt bf16, w fp32, scale, shift, bias, kvol, cin, cout"""
        scale = shift = None
        if conv is not None:
            w32 = conv._w3()
            if ME.get_tf32():
                wt = conv._wt_tf32() if (w32.shape[1] % 32 == 0 and w32.shape[2] % 32 == 0) else None
            else:
                wt = conv._wt_bf16() if (w32.shape[1] % 64 == 0 and w32.shape[2] % 32 == 0) else None
            kvol, cin, cout = (int(v) for v in w32.shape)
            bias = conv.bias.detach().reshape(-1).contiguous() if conv.bias is not None else None
        if bn is not None:
            s, b = bn.scale_shift(None)
            scale, shift = s.reshape(-1).contiguous(), b.reshape(-1).contiguous()
        self._keep += [wt, w32, scale, shift, bias]
        words += [ptr(t) or 0 for t in (wt, w32, scale, shift, bias)] + [int(kvol or 0), int(cin or 0), int(cout or 0)]

    def _build(self):
        det, head, bb = self.det, self.det.neck_with_head, self.det.backbone
        if any(m.training for m in det.modules() if isinstance(m, torch.nn.BatchNorm1d)):
            raise B2Error("InferenceEngine runs the eval-mode network (call det.eval())")
        self._destroy()
        self._keep = []
        L = bb.n_outs
        conv, inorm, _, pool = bb.conv1
        blocks = [len(getattr(bb, "layer%d" % (i + 1))) for i in range(L)] + [0] * (4 - L)
        w, b, wt = head._head_weights()
        yaw_mode = head._yaw_mode(head.n_reg_outs)
        words = [0xB2E1, L, head.n_classes, head.n_reg_outs, yaw_mode, int(head.pts_threshold), int(head._cfg("nms_pre")),
                 3 + conv.in_channels, int(w.shape[2])] + blocks + [0, 0, 0]
        assert len(words) == 16
        self._rec(words, conv=conv)
        iw, ib = inorm.weight.detach().reshape(-1).contiguous(), inorm.bias.detach().reshape(-1).contiguous()
        self._rec(words, wt=iw, w32=ib)
        for i in range(L):
            layer = getattr(bb, "layer%d" % (i + 1))
            for blk in layer:
                self._rec(words, conv=blk.conv1, bn=blk.norm1)
                self._rec(words, conv=blk.conv2, bn=blk.norm2)
            if layer[0].downsample is None:
                raise B2Error("InferenceEngine: the first block of every stage must have a downsample branch")
            self._rec(words, conv=layer[0].downsample[0], bn=layer[0].downsample[1])
        for i in range(L):
            c, bn, _ = getattr(head, "out_block_%d" % i)
            self._rec(words, conv=c, bn=bn)
        for i in range(1, L):
            gen, bn1, _, c, bn2, _ = getattr(head, "up_block_%d" % i)
            wp, wpt = gen._packed()
            if ME.get_tf32():
                wpt = gen._packed_tf32()
            gb = gen.bias.detach().reshape(-1).contiguous() if gen.bias is not None else None
            s, sh = bn1.scale_shift(None)
            s, sh = s.reshape(-1).contiguous(), sh.reshape(-1).contiguous()
            self._keep += [wp, wpt, gb, s, sh]
            words += [ptr(wpt) or 0, ptr(wp), ptr(s), ptr(sh), ptr(gb) or 0, 1, gen.in_channels, 8 * gen.out_channels]
            self._rec(words, conv=c, bn=bn2)
        if ME.get_tf32():
            wt = head._head_weights_tf32()
        self._rec(words, wt=wt, w32=w, bias=b, kvol=1, cin=int(w.shape[1]), cout=int(w.shape[2]))
        for i in range(L):
            sc = head.scales[i].scale.detach().reshape(1)
            self._rec(words, wt=sc)
        self._yaw_rotated = head.n_reg_outs != 6
        desc = (ctypes.c_longlong * len(words))(*words)
        fl = (ctypes.c_double * 4)(float(head.voxel_size), float(head._cfg("score_thr")), float(head._cfg("iou_thr")),
                                   float(inorm.eps))
        h = lib().b2_engine_create(desc, len(words), fl)
        if not h:
            raise B2Error("b2_engine_create failed: %s" % last_error())
        self._handle = ctypes.c_void_p(h)
        self._sig = self._signature()

    def _destroy(self):
        if self._handle is not None:
            lib().b2_engine_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    # ---- one step -------------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def __call__(self, points, inputs_ready=None):
        """points: list of [N_i, 6] fp32 CUDA tensors.  Returns the device-side result dict of get_bboxes_fused
        (boxes [n,7], scores, labels, scene, count, box_dim[, ready]); detector.split_detections / DetectionRead bring it to
        the host.  inputs_ready: a torch.cuda.Event after which the points are complete (e.g. the end of their host->device
        copy), or True when they are complete already; the executor's coordinate stream then starts on them right away
        instead of waiting for everything queued on the current stream (default: it waits)."""
        if self._handle is None or self._sig != self._signature():
            self._build()
        dev = points[0].device
        n = len(points)
        pts = [p if (p.is_contiguous() and p.dtype == torch.float32) else p.contiguous().float() for p in points]
        for p in pts:
            if not p.is_cuda:
                raise B2Error("InferenceEngine needs CUDA points (no CPU path)")
        ptrs = (ctypes.c_void_p * n)(*[p.data_ptr() for p in pts])
        counts = (ctypes.c_longlong * n)(*[int(p.shape[0]) for p in pts])
        total = sum(int(p.shape[0]) for p in pts)
        l = lib()
        cap = int(l.b2_engine_out_rows(self._handle, n))
        boxes = torch.empty((cap, 7), dtype=torch.float32, device=dev)
        scores = torch.empty((cap,), dtype=torch.float32, device=dev)
        labels = torch.empty((cap,), dtype=torch.int32, device=dev)
        scene = torch.empty((cap,), dtype=torch.int32, device=dev)
        count = torch.empty((1,), dtype=torch.int32, device=dev)
        slot = self._step % 2
        self._step += 1
        self._hold[slot] = pts           # the coordinate stream reads them: alive until this slot's next turn
        if inputs_ready is None:
            inputs_ready = getattr(points, "ready", None)
        in_ev = inputs_ready.cuda_event if isinstance(inputs_ready, torch.cuda.Event) else None
        in_flag = 1 if inputs_ready is True else 0
        free_ev = self._arena_free[slot].cuda_event if self._arena_free[slot] is not None else None
        main = torch.cuda.current_stream(dev)
        if self.overlap_tail and self._tail is None:
            self._tail = torch.cuda.Stream(device=dev)
        tail = self._tail if self.overlap_tail else None
        if self._arena_free[slot] is not None:
            main.wait_event(self._arena_free[slot])       # the tail that last read this arena has finished
        need = (4096 * total + (256 << 20)) if self._arenas[slot] is None else 0
        rows, used = ctypes.c_longlong(0), ctypes.c_size_t(0)
        dt = 1 if ME.get_compute_dtype() == torch.bfloat16 else (2 if ME.get_tf32() else 0)
        for _ in range(6):
            arena = self._arenas[slot]
            if arena is None or arena.numel() < need or arena.device != dev:
                self._arenas[slot] = arena = None
                self._arenas[slot] = arena = torch.empty(max(need, 1 << 20), dtype=torch.uint8, device=dev)
                other = self._arenas[1 - slot]            # both arenas at once: no allocation in a later (timed) call
                if other is None or other.numel() < arena.numel() or other.device != dev:
                    if self._arena_free[1 - slot] is not None:
                        self._arena_free[1 - slot].synchronize()
                    self._arenas[1 - slot] = None
                    self._arenas[1 - slot] = torch.empty(arena.numel(), dtype=torch.uint8, device=dev)
            rc = l.b2_engine_infer(self._handle, ptrs, counts, n, dt, ptr(arena), arena.numel(), ptr(boxes),
                                   ptr(scores), ptr(labels), ptr(scene), ptr(count), cap, ctypes.byref(rows),
                                   ctypes.byref(used), stream(), tail.cuda_stream if tail is not None else None, in_ev,
                                   in_flag, free_ev)
            if rc == -3:            # arena too small: the launches so far wrote only into the arena - grow it and redo the step
                torch.cuda.synchronize(dev)
                need = 2 * arena.numel()
                continue
            if rc != 0:
                raise B2Error("b2_engine_infer failed (%d): %s" % (rc, last_error()))
            break
        else:
            raise B2Error("b2_engine_infer: arena kept overflowing: %s" % last_error())
        ready = None
        if tail is None and self.split_streams:
            ready = torch.cuda.Event()           # (arena reuse is then ordered behind the feature stream)
            ready.record(main)
            self._arena_free[slot] = ready
            ready = None
        if tail is not None:
            for t in (boxes, scores, labels, scene, count):
                t.record_stream(tail)                     # allocated on the main stream, written on the tail stream
            ready = torch.cuda.Event()
            ready.record(tail)
            self._arena_free[slot] = ready
        self.arena_used = int(used.value)
        nr = int(rows.value)
        out = dict(boxes=boxes[:nr], scores=scores[:nr], labels=labels[:nr], scene=scene[:nr], count=count,
                   box_dim=7 if self._yaw_rotated else 6)
        if ready is not None:
            out["ready"] = ready     # the detections are complete once this event has fired (detector.wait_ready)
        return out
