"""Densification and optimizer-state surgery -- REF/external.py:136-205,244-292.

The reference applies clone, split and two prune rounds as four separate passes, each of which re-allocates every
parameter and both Adam moments with torch.cat / boolean indexing and ends with torch.cuda.empty_cache().  Here the
same decisions (same thresholds, same order of the surviving rows: originals, clones, split children) are turned into
ONE gather list, and every parameter and Adam moment is rebuilt once with a single index_select."""
from __future__ import annotations

from typing import Dict, Optional

import torch

PER_GAUSSIAN = ('means3D', 'rgb_colors', 'seg_colors', 'unnorm_rotations', 'logit_opacities', 'log_scales')
GRAD_THRESH = 0.0002          # REF/external.py:247
DENSIFY_UNTIL = 5000          # :245
DENSIFY_FROM, DENSIFY_EVERY = 500, 100       # :248
OPACITY_RESET_EVERY = 3000    # :288
SPLIT_N = 2                   # :261


def accumulate_mean2d_gradient(variables: dict) -> dict:
    """REF/external.py:136-140: per-view norm of dL/dmean2D.xy for the Gaussians seen in this step."""
    seen = variables['seen']
    g = variables['means2D'].grad
    norm = torch.norm(g[:, :2], dim=-1)
    variables['means2D_gradient_accum'] = variables['means2D_gradient_accum'] + torch.where(seen, norm, torch.zeros_like(norm))
    variables['denom'] = variables['denom'] + seen.to(variables['denom'].dtype)
    return variables


def replace_param(params: Dict[str, torch.nn.Parameter], optimizer, name: str, value: torch.Tensor,
                  exp_avg: Optional[torch.Tensor] = None, exp_avg_sq: Optional[torch.Tensor] = None,
                  reset_state: bool = False):
    """Swaps the tensor of one Adam group, carrying the step count over (REF/external.py:143-156,159-183)."""
    group = next(g for g in optimizer.param_groups if g['name'] == name)
    old = group['params'][0]
    state = optimizer.state.pop(old, None)
    new = torch.nn.Parameter(value.requires_grad_(True))
    group['params'][0] = new
    if state is not None:
        if reset_state:
            state['exp_avg'] = torch.zeros_like(value)
            state['exp_avg_sq'] = torch.zeros_like(value)
        else:
            state['exp_avg'] = exp_avg
            state['exp_avg_sq'] = exp_avg_sq
        optimizer.state[new] = state
    params[name] = new
    return new


def build_rotation(q: torch.Tensor) -> torch.Tensor:
    """Rotation matrices of (un-normalised) quaternions (w, x, y, z) -- REF/helpers.py build_rotation."""
    q = torch.nn.functional.normalize(q)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=-1).reshape(-1, 3, 3)


def densify_plan(grads: torch.Tensor, max_scale: torch.Tensor, opacity: torch.Tensor, scene_radius: float, i: int):
    """The reference's clone / split / prune decisions (REF/external.py:249-285) as one gather list.

    Returns (src [M] long: the original row every surviving row is gathered from, kind [M] int8: 0 original, 1 clone,
    2 split child).  Row order == the reference's: surviving originals, surviving clones, surviving children.
    Pruning uses the row's own opacity and (for i >= 3000) its scale, which for a child is scale / (0.8 * n)."""
    P = grads.shape[0]
    dev = grads.device
    small = max_scale <= 0.01 * scene_radius
    hot = grads >= GRAD_THRESH
    to_clone = hot & small
    to_split = hot & ~small                     # clones carry a padded gradient of 0 and are never split (:256-258)
    remove_threshold = 0.25 if i == 5000 else 0.005
    low = opacity < remove_threshold
    child_scale = max_scale / (0.8 * SPLIT_N)
    big = max_scale > 0.1 * scene_radius if i >= 3000 else torch.zeros_like(low)
    big_child = child_scale > 0.1 * scene_radius if i >= 3000 else torch.zeros_like(low)
    idx = torch.arange(P, device=dev)
    keep_orig = idx[~to_split & ~low & ~big]
    keep_clone = idx[to_clone & ~low & ~big]
    child = idx[to_split & ~low & ~big_child]
    src = torch.cat([keep_orig, keep_clone, child.repeat(SPLIT_N)])
    kind = torch.cat([torch.zeros_like(keep_orig, dtype=torch.int8), torch.ones_like(keep_clone, dtype=torch.int8),
                      torch.full((child.numel() * SPLIT_N,), 2, dtype=torch.int8, device=dev)])
    return src, kind


def densify(params, variables, optimizer, i: int, generator: Optional[torch.Generator] = None):
    """REF/external.py:244-292, one gather per tensor."""
    if i > DENSIFY_UNTIL:
        return params, variables
    variables = accumulate_mean2d_gradient(variables)
    if i >= DENSIFY_FROM and i % DENSIFY_EVERY == 0:
        grads = variables['means2D_gradient_accum'] / variables['denom']
        grads = torch.nan_to_num(grads, nan=0.0)                                   # :250
        scales = torch.exp(params['log_scales'])
        max_scale = scales.max(dim=1).values
        opacity = torch.sigmoid(params['logit_opacities']).squeeze(-1)
        src, kind = densify_plan(grads, max_scale, opacity, variables['scene_radius'], i)
        is_child = kind == 2
        fresh = (kind != 0)[:, None]
        new_vals = {}
        for k in PER_GAUSSIAN:
            new_vals[k] = params[k].detach().index_select(0, src)
        # split children: sample around the parent in its own frame, shrink (:262-268)
        n_child = int(is_child.sum())
        if n_child:
            c_src = src[is_child]
            stds = scales.index_select(0, c_src)
            if generator is not None and generator.device != stds.device:      # e.g. a CPU generator shared with get_batch
                s_gen = stds.to(generator.device)
                samples = torch.normal(mean=torch.zeros_like(s_gen), std=s_gen, generator=generator).to(stds.device)
            else:
                samples = torch.normal(mean=torch.zeros_like(stds), std=stds, generator=generator)
            rots = build_rotation(params['unnorm_rotations'].detach().index_select(0, c_src))
            new_vals['means3D'][is_child] += torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1)
            new_vals['log_scales'][is_child] = torch.log(stds / (0.8 * SPLIT_N))
        for k in PER_GAUSSIAN:
            group = next(g for g in optimizer.param_groups if g['name'] == k)
            st = optimizer.state.get(group['params'][0], None)
            ea = es = None
            if st is not None and 'exp_avg' in st:
                zero = torch.zeros((), device=src.device, dtype=st['exp_avg'].dtype)
                ea = torch.where(fresh, zero, st['exp_avg'].index_select(0, src))   # new rows start with zero moments
                es = torch.where(fresh, zero, st['exp_avg_sq'].index_select(0, src))
            replace_param(params, optimizer, k, new_vals[k], ea, es)
        M = src.numel()
        dev = src.device
        variables['means2D_gradient_accum'] = torch.zeros(M, device=dev)             # :271-273
        variables['denom'] = torch.zeros(M, device=dev)
        variables['max_2D_radius'] = torch.zeros(M, device=dev)
    if i > 0 and i % OPACITY_RESET_EVERY == 0:                                       # :288-290
        v = torch.full_like(params['logit_opacities'].detach(), 0.01)
        replace_param(params, optimizer, 'logit_opacities', torch.log(v / (1 - v)), reset_state=True)
    return params, variables
