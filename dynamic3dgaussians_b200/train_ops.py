"""Fused versions of the small per-iteration ops around the rasterizer in the train.py-shaped loop.

At 300k Gaussians one optimizer iteration is ~1 ms of GPU time, so the ~40 eager elementwise / reduction kernels (and
their autograd twins) of `params2rendervar` (REF/helpers.py:98-107), the per-camera colour correction (REF/train.py:181)
and the densification statistics (REF/train.py:288-289, REF/external.py:136-140) are a fifth of it.  Each group is one
CUDA kernel here (`csrc/train_glue.cu`), same formulas as the torch ops they replace.  No CPU path."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not torch.is_tensor(t) or not t.is_cuda or t.dtype != torch.float32:
        raise ValueError(f"{name} must be a float32 CUDA tensor (no CPU path)")
    return t.contiguous()


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class _Activations(torch.autograd.Function):
    @staticmethod
    def forward(ctx, unnorm_rotations, logit_opacities, log_scales):
        q, lo, ls = _f32(unnorm_rotations, "unnorm_rotations"), _f32(logit_opacities, "logit_opacities"), \
            _f32(log_scales, "log_scales")
        P = q.shape[0]
        if q.shape != (P, 4) or lo.numel() != P or ls.shape != (P, 3):
            raise ValueError("expected unnorm_rotations [P,4], logit_opacities [P,1], log_scales [P,3]")
        rot, op, sc = torch.empty_like(q), torch.empty_like(lo), torch.empty_like(ls)
        with torch.cuda.device(q.device):
            _lib.check(_lib.load().gsr_activations_forward(P, q.data_ptr(), lo.data_ptr(), ls.data_ptr(), rot.data_ptr(),
                                                           op.data_ptr(), sc.data_ptr(), _stream(q.device)),
                       "activations_forward")
        ctx.save_for_backward(q, op, sc)
        return rot, op, sc

    @staticmethod
    def backward(ctx, g_rot, g_op, g_sc):
        q, op, sc = ctx.saved_tensors
        P = q.shape[0]
        g_rot = None if g_rot is None else _f32(g_rot, "dL/drotations")
        g_op = None if g_op is None else _f32(g_op, "dL/dopacities")
        g_sc = None if g_sc is None else _f32(g_sc, "dL/dscales")
        need = ctx.needs_input_grad
        d_q = torch.empty_like(q) if need[0] else None
        d_lo = torch.empty_like(op) if need[1] else None
        d_ls = torch.empty_like(sc) if need[2] else None
        with torch.cuda.device(q.device):
            _lib.check(_lib.load().gsr_activations_backward(P, q.data_ptr(), op.data_ptr(), sc.data_ptr(), _ptr(g_rot),
                                                            _ptr(g_op), _ptr(g_sc), _ptr(d_q), _ptr(d_lo), _ptr(d_ls),
                                                            _stream(q.device)), "activations_backward")
        return d_q, d_lo, d_ls


def activations(unnorm_rotations: torch.Tensor, logit_opacities: torch.Tensor, log_scales: torch.Tensor):
    """(F.normalize(unnorm_rotations), sigmoid(logit_opacities), exp(log_scales)) in one kernel, one more for the backward."""
    return _Activations.apply(unnorm_rotations, logit_opacities, log_scales)


class _CamAffine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, im, cam_m, cam_c, cam_id):
        im, cam_m, cam_c = _f32(im, "im"), _f32(cam_m, "cam_m"), _f32(cam_c, "cam_c")
        if im.dim() != 3 or im.shape[0] != 3 or cam_m.dim() != 2 or cam_m.shape[1] != 3 or cam_c.shape != cam_m.shape:
            raise ValueError("expected im [3,H,W], cam_m / cam_c [num_cams,3]")
        cam_id = int(cam_id)
        if not 0 <= cam_id < cam_m.shape[0]:
            raise IndexError(f"camera {cam_id} out of range")
        HW = im.shape[1] * im.shape[2]
        out = torch.empty_like(im)
        off = 12 * cam_id
        with torch.cuda.device(im.device):
            _lib.check(_lib.load().gsr_cam_affine_forward(HW, im.data_ptr(), cam_m.data_ptr() + off, cam_c.data_ptr() + off,
                                                          out.data_ptr(), _stream(im.device)), "cam_affine_forward")
        ctx.save_for_backward(im, cam_m)
        ctx.cam_id = cam_id
        ctx.n_cams = cam_m.shape[0]
        return out

    @staticmethod
    def backward(ctx, g):
        im, cam_m = ctx.saved_tensors
        g = _f32(g, "dL/dim")
        HW = im.shape[1] * im.shape[2]
        need = ctx.needs_input_grad
        d_im = torch.empty_like(im) if need[0] else None
        d_mc = torch.zeros((2, ctx.n_cams, 3), dtype=torch.float32, device=im.device) if (need[1] or need[2]) else None
        off = 12 * ctx.cam_id
        with torch.cuda.device(im.device):
            _lib.check(_lib.load().gsr_cam_affine_backward(
                HW, im.data_ptr(), cam_m.data_ptr() + off, g.data_ptr(), _ptr(d_im),
                None if d_mc is None else d_mc[0].data_ptr() + off, None if d_mc is None else d_mc[1].data_ptr() + off,
                _stream(im.device)), "cam_affine_backward")
        return d_im, (d_mc[0] if need[1] else None), (d_mc[1] if need[2] else None), None


def cam_affine(im: torch.Tensor, cam_m: torch.Tensor, cam_c: torch.Tensor, cam_id: int) -> torch.Tensor:
    """exp(cam_m[cam_id])[:, None, None] * im + cam_c[cam_id][:, None, None]  (REF/train.py:181)."""
    return _CamAffine.apply(im, cam_m, cam_c, cam_id)


@torch.no_grad()
def update_max_radius(radii: torch.Tensor, max_2D_radius: torch.Tensor) -> torch.Tensor:
    """REF/train.py:288-289 in place: max_2D_radius[seen] = max(radii[seen], max_2D_radius[seen]); returns seen (bool)."""
    if not radii.is_cuda or radii.dtype != torch.int32:
        raise ValueError("radii must be an int32 CUDA tensor")
    radii = radii.contiguous()
    P = radii.numel()
    seen = torch.empty(P, dtype=torch.bool, device=radii.device)
    with torch.cuda.device(radii.device):
        _lib.check(_lib.load().gsr_densify_stats(P, radii.data_ptr(), None, None, _f32(max_2D_radius, "max_2D_radius").data_ptr(),
                                                 None, None, seen.data_ptr(), _stream(radii.device)), "densify_stats")
    return seen


@torch.no_grad()
def accumulate_mean2d_gradient_(seen: torch.Tensor, means2D_grad: torch.Tensor, accum: torch.Tensor,
                                denom: torch.Tensor) -> None:
    """accumulate_mean2d_gradient (REF/external.py:136-140) with in-place updates, one kernel."""
    if seen.dtype != torch.bool or not seen.is_cuda:
        raise ValueError("seen must be a bool CUDA tensor")
    g = _f32(means2D_grad, "means2D.grad")
    P = seen.numel()
    if g.shape != (P, 3) or not accum.is_contiguous() or not denom.is_contiguous():
        raise ValueError("expected means2D.grad [P,3] and contiguous accumulators")
    with torch.cuda.device(seen.device):
        _lib.check(_lib.load().gsr_densify_stats(P, None, seen.contiguous().data_ptr(), g.data_ptr(), None,
                                                 _f32(accum, "accum").data_ptr(), _f32(denom, "denom").data_ptr(), None,
                                                 _stream(seen.device)), "densify_stats")


class _L1U8(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, target, row_weight):
        img = _f32(img, "img")
        if not target.is_cuda or target.dtype != torch.uint8 or target.shape != img.shape:
            raise ValueError("target must be a uint8 CUDA tensor of img's shape")
        if img.dim() != 4 or img.shape[1] != 3:
            raise ValueError("expected img [V,3,H,W]")
        V, _, H, W = img.shape
        target = target.contiguous()
        if row_weight is not None:
            row_weight = _f32(row_weight, "row_weight")
            if row_weight.numel() != V * H:
                raise ValueError("row_weight must be [V,H]")
        weight = 1.0 / max(img.numel(), 1)
        loss = torch.empty((), dtype=torch.float32, device=img.device)
        with torch.cuda.device(img.device):
            _lib.check(_lib.load().gsr_l1_u8_loss_forward(V, W, H, img.data_ptr(), target.data_ptr(), _ptr(row_weight),
                                                          weight, loss.data_ptr(), _stream(img.device)), "l1_u8_forward")
        ctx.save_for_backward(img, target, row_weight)
        ctx.weight = weight
        return loss

    @staticmethod
    def backward(ctx, g):
        img, target, row_weight = ctx.saved_tensors
        V, _, H, W = img.shape
        g = _f32(g, "dL/dloss").reshape(1)
        d_img = torch.empty_like(img)
        with torch.cuda.device(img.device):
            _lib.check(_lib.load().gsr_l1_u8_loss_backward(V, W, H, img.data_ptr(), target.data_ptr(), _ptr(row_weight),
                                                           ctx.weight, g.data_ptr(), d_img.data_ptr(),
                                                           _stream(img.device)), "l1_u8_backward")
        return d_img, None, None


def l1_loss_u8(img: torch.Tensor, target_u8: torch.Tensor, row_weight: Optional[torch.Tensor] = None) -> torch.Tensor:
    """mean |img - target / 255| over [V,3,H,W] against uint8 frames (l1_loss_v1, REF/helpers.py), or
    sum row_weight[v, y] * |.| when per-(view, row) weights are given -- one kernel forward, one backward, no float copy
    of the targets."""
    return _L1U8.apply(img, target_u8, row_weight)
