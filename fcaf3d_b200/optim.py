"""AdamW with fused gradient clipping on the library's kernels (csrc/optim_ops.cu): the optimiser of the reference recipe
(configs/fcaf3d/fcaf3d.py:30-31: AdamW lr 1e-3, weight_decay 1e-4, grad_clip max_norm 10) as three launches per step over
all parameter tensors - squared gradient norm (deterministic), its reduction, and the update with the clip coefficient
applied on the fly - instead of torch's clip_grad_norm_ + multi-tensor AdamW kernel chains.  CUDA fp32 parameters only."""
import ctypes

import torch

from ._lib import B2Error, call, lib, ptr


class FusedAdamW:
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, max_norm=None):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise B2Error("FusedAdamW: no parameters")
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise B2Error("FusedAdamW: contiguous fp32 CUDA parameters only")
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = lr, betas, eps, weight_decay, max_norm
        self.step_count = 0
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        dev = self.params[0].device
        chunk = int(lib().b2_opt_chunk_elems())
        pairs = [(i, c) for i, p in enumerate(self.params) for c in range((p.numel() + chunk - 1) // chunk)]
        self.n_chunks = len(pairs)
        self._chunks = torch.tensor(pairs, dtype=torch.int32).reshape(-1, 2).to(dev)
        self._partial = torch.empty((max(self.n_chunks, 1),), dtype=torch.float32, device=dev)
        self.sqnorm = torch.zeros((1,), dtype=torch.float32, device=dev)
        self._table = torch.empty((len(self.params), 5), dtype=torch.int64, device=dev)
        self._table_host = torch.empty((len(self.params), 5), dtype=torch.int64).pin_memory()
        self._ptrs = None

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()

    def _refresh_table(self):
        grads = []
        for p in self.params:
            if p.grad is None:                       # parameter unused in this step: torch skips it; a zero gradient
                p.grad = torch.zeros_like(p)         # changes only weight decay and the moment decay, applied here
            g = p.grad
            if g.dtype != torch.float32 or not g.is_contiguous():
                raise B2Error("FusedAdamW: contiguous fp32 gradients only")
            grads.append(g)
        ptrs = tuple(g.data_ptr() for g in grads) + tuple(p.data_ptr() for p in self.params)
        if ptrs != self._ptrs:
            rows = [[p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel()]
                    for p, g, m, v in zip(self.params, grads, self.exp_avg, self.exp_avg_sq)]
            self._table_host.copy_(torch.tensor(rows, dtype=torch.int64))
            self._table.copy_(self._table_host, non_blocking=True)
            self._ptrs = ptrs

    @torch.no_grad()
    def step(self):
        """One update; returns the (device, 1-element) squared gradient norm BEFORE clipping when clipping is on."""
        self._refresh_table()
        self.step_count += 1
        clip = self.max_norm is not None and self.max_norm > 0
        if clip:
            call("b2_grad_sqnorm", ptr(self._table), ptr(self._chunks), self.n_chunks, ptr(self._partial), ptr(self.sqnorm))
        call("b2_adamw_step", ptr(self._table), ptr(self._chunks), self.n_chunks, float(self.lr), float(self.betas[0]),
             float(self.betas[1]), float(self.eps), float(self.weight_decay), int(self.step_count),
             float(self.max_norm) if clip else 0.0, ptr(self.sqnorm) if clip else None)
        return self.sqnorm if clip else None

    # checkpointing in torch.optim.AdamW's layout (state per parameter index: step, exp_avg, exp_avg_sq)
    def state_dict(self):
        return {"state": {i: {"step": torch.tensor(float(self.step_count)), "exp_avg": m, "exp_avg_sq": v}
                          for i, (m, v) in enumerate(zip(self.exp_avg, self.exp_avg_sq))},
                "param_groups": [{"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay,
                                  "params": list(range(len(self.params)))}]}

    def load_state_dict(self, sd):
        for i, st in sd["state"].items():
            self.exp_avg[int(i)].copy_(st["exp_avg"])
            self.exp_avg_sq[int(i)].copy_(st["exp_avg_sq"])
            self.step_count = int(st["step"])
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps, self.weight_decay = g["lr"], tuple(g["betas"]), g["eps"], g["weight_decay"]
