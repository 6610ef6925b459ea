"""Fused Adam for the training loop.

`FusedAdam` is a drop-in for the reference's `torch.optim.Adam(param_groups, lr=0.0, eps=1e-15)` (REF/train.py:119-135):
same param_groups ('name', 'lr' per group), same state keys ('step', 'exp_avg', 'exp_avg_sq') so that the densification
surgery (densify.replace_param, REF/external.py:143-208) works unchanged -- but optimizer.step() is ONE kernel over all
tensors instead of ~7 multi-tensor kernels per group.  The step counters are device tensors, so step() can be captured
into a CUDA graph.  No CPU path."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self._scratch = {}

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        batches = {}              # (device, betas, eps) -> list of GsrAdamTensor
        keep = []
        for group in self.param_groups:
            for p in group['params']:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise ValueError("FusedAdam needs contiguous float32 CUDA parameters (no CPU path)")
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                st = self.state[p]
                if len(st) == 0:
                    st['step'] = torch.zeros((), dtype=torch.float32, device=p.device)
                    st['exp_avg'] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                t = _lib.GsrAdamTensor(p.data_ptr(), g.data_ptr(), st['exp_avg'].data_ptr(), st['exp_avg_sq'].data_ptr(),
                                       st['step'].data_ptr(), p.numel(), float(group['lr']), 0)
                keep.append(g)
                key = (p.device, tuple(group['betas']), float(group['eps']))
                batches.setdefault(key, []).append(t)
        for (dev, betas, eps), ts in batches.items():
            scratch = self._scratch.get(dev)
            if scratch is None:
                scratch = self._scratch[dev] = torch.empty(_lib.ADAM_SCRATCH_BYTES, dtype=torch.uint8, device=dev)
            with torch.cuda.device(dev):
                stream = torch.cuda.current_stream(dev).cuda_stream
                for i in range(0, len(ts), _lib.ADAM_MAX_TENSORS):
                    chunk = ts[i:i + _lib.ADAM_MAX_TENSORS]
                    arr = (_lib.GsrAdamTensor * len(chunk))(*chunk)
                    _lib.check(lib.gsr_adam_step(len(chunk), arr, betas[0], betas[1], eps, scratch.data_ptr(), stream),
                               "adam_step")
        return loss
