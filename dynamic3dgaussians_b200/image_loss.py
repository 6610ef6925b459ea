"""Fused L1 + SSIM image loss: the step on the other side of the rasterizer in the training loop.

`l1_ssim_loss(im, target)` == 0.8 * l1_loss_v1(im, target) + 0.2 * (1 - calc_ssim(im, target)) of REF/train.py:183
(REF/helpers.py l1_loss_v1, REF/external.py:90-133), value and gradient w.r.t. `im`, in two CUDA kernels instead of five
grouped convolutions, ~20 elementwise kernels and their autograd twins.  The gradient is what the rasterizer backward
takes as dL/dcolor.  No CPU path."""
from __future__ import annotations

import torch

from . import _lib


def _check(t: torch.Tensor, name: str) -> torch.Tensor:
    if not torch.is_tensor(t) or not t.is_cuda or t.dtype != torch.float32:
        raise ValueError(f"{name} must be a float32 CUDA tensor (this loss has no CPU path)")
    if t.dim() not in (3, 4):
        raise ValueError(f"{name} must be [C,H,W] or [N,C,H,W]")
    return t.contiguous()


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, target, w_l1, w_ssim):
        lib = _lib.load()
        img = _check(img, "img")
        target = _check(target, "target")
        if img.shape != target.shape:
            raise ValueError(f"img {tuple(img.shape)} and target {tuple(target.shape)} differ in shape")
        H, W = int(img.shape[-2]), int(img.shape[-1])
        n = img.numel() // (H * W)
        dev = img.device
        nbytes = lib.gsr_image_loss_state_bytes(n, W, H)
        state = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        out = torch.empty(3, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(lib.gsr_image_loss_forward(n, W, H, img.data_ptr(), target.data_ptr(), w_l1, w_ssim,
                                                  state.data_ptr(), nbytes, out.data_ptr(),
                                                  torch.cuda.current_stream(dev).cuda_stream), "image_loss_forward")
        ctx.save_for_backward(img, target, state)
        ctx.w = (float(w_l1), float(w_ssim), n, W, H, nbytes)
        ctx.mark_non_differentiable(out)
        return out[0], out

    @staticmethod
    def backward(ctx, g_loss, _g_parts):
        img, target, state = ctx.saved_tensors
        w_l1, w_ssim, n, W, H, nbytes = ctx.w
        lib = _lib.load()
        dev = img.device
        g = g_loss.to(torch.float32).reshape(1).contiguous()
        grad = torch.empty_like(img)
        with torch.cuda.device(dev):
            _lib.check(lib.gsr_image_loss_backward(n, W, H, img.data_ptr(), target.data_ptr(), w_l1, w_ssim,
                                                   state.data_ptr(), nbytes, g.data_ptr(), grad.data_ptr(),
                                                   torch.cuda.current_stream(dev).cuda_stream), "image_loss_backward")
        return grad, None, None, None


def l1_ssim_loss(img: torch.Tensor, target: torch.Tensor, w_l1: float = 0.8, w_ssim: float = 0.2,
                 return_parts: bool = False):
    """w_l1 * mean|img - target| + w_ssim * (1 - mean SSIM).  With return_parts also the detached [total, l1, ssim]."""
    loss, parts = _L1SSIM.apply(img, target, float(w_l1), float(w_ssim))
    return (loss, parts) if return_parts else loss
