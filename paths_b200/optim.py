"""AdamW on the flat parameter buffer: `torch.optim.AdamW(model.parameters(), lr, weight_decay)` of the reference
(train.py:49,66) as ONE fused kernel (csrc/optim.cu) instead of a multi-tensor loop over 340 parameter tensors.

It subclasses torch.optim.Optimizer only so that the reference's `config.get_lr_scheduler(opt)` (ExponentialLR,
config.py:130-131), `opt.zero_grad()` and `param_groups[0]["lr"]` keep working; the arithmetic is the CUDA kernel.
Semantics match torch: parameters whose .grad is None (classifiers of levels 0..L-2, modules the forward pass never
evaluates such as procs.0.hctx_mlp of the RNN variant) are skipped, parameters with exact-zero gradients (the dead
encoder / cross-attention weights) are only weight-decayed.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


class FlatAdamW(torch.optim.Optimizer):
    def __init__(self, model, lr: float = 2e-5, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2):
        super().__init__(list(model.parameters()), dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.model = model
        self.rt = model.runtime()
        lay = self.rt.layout
        self.exp_avg = torch.zeros(lay.total_all, dtype=torch.float32, device=lay.device)
        self.exp_avg_sq = torch.zeros(lay.total_all, dtype=torch.float32, device=lay.device)
        self.t = 0
        self._params = [q for g in self.param_groups for q in g["params"]]

    def zero_grad(self, set_to_none: bool = True) -> None:
        """`opt.zero_grad()` of train.py:59.  The gradients live in one flat buffer: dropping the views costs a plain loop over a
        cached list (no module-tree walk, no per-tensor kernels); the buffer itself is cleared by the next backward pass."""
        if set_to_none:
            for q in self._params:
                q.grad = None
            return
        super().zero_grad(set_to_none=False)

    def _skip_ranges(self, lay):
        rng = []
        for n, prm, off, k in lay.entries + lay.dead_entries:
            if prm.grad is None:                     # torch skips parameters without a gradient (in the dead tail: modules the
                                                     # forward pass never evaluates, e.g. procs.0.hctx_mlp of the RNN variant)
                if rng and rng[-1][1] == off:
                    rng[-1][1] = off + k
                else:
                    rng.append([off, off + k])
        # merge ranges separated only by alignment padding
        merged = []
        for a, b in rng:
            if merged and a - merged[-1][1] < 64 and not any(merged[-1][1] <= o < a for _, _, o, _ in lay.entries + lay.dead_entries):
                merged[-1][1] = b
            else:
                merged.append([a, b])
        return merged

    @torch.no_grad()
    def step(self, closure=None):
        assert closure is None
        lay = self.rt.layout
        if not lay.valid():
            raise RuntimeError("FlatAdamW: the model's parameters were re-allocated; rebuild the optimizer")
        for n, prm, off, k in lay.entries:           # a gradient produced outside the engine: fold it into the flat buffer
            if prm.grad is not None and prm.grad.data_ptr() != lay.gviews[n].data_ptr():
                lay.gviews[n].copy_(prm.grad)
        skip = self._skip_ranges(lay)
        if len(skip) > 16:
            raise RuntimeError(f"FlatAdamW: {len(skip)} gradient-less parameter ranges (at most 16 supported)")
        g = self.param_groups[0]
        self.t += 1
        arr = (C.c_int64 * max(1, 2 * len(skip)))(*[x for ab in skip for x in ab])
        L.check(L.lib().paths_adamw_step(lay.flat.data_ptr(), lay.gflat.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                                         lay.total, lay.total_all, arr, len(skip), float(g["lr"]), float(g["betas"][0]),
                                         float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]), self.t,
                                         torch.cuda.current_stream().cuda_stream), "adamw_step")
        return None
