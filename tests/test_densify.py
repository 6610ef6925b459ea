"""Densification as one gather (dynamic3dgaussians_b200/densify.py) against the reference's four passes restated in
oracle/eager_densify.py: same surviving rows in the same order, same values, same Adam moments.  Runs on the CPU."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402,F401
from dynamic3dgaussians_b200 import densify as DZ  # noqa: E402
from oracle import eager_densify as OD  # noqa: E402


def _state(P, seed):
    g = torch.Generator().manual_seed(seed)
    t = {
        'means3D': torch.randn(P, 3, generator=g),
        'rgb_colors': torch.rand(P, 3, generator=g),
        'seg_colors': torch.rand(P, 3, generator=g).round(),
        'unnorm_rotations': torch.randn(P, 4, generator=g),
        'logit_opacities': torch.randn(P, 1, generator=g) * 3.0,
        'log_scales': torch.log(torch.rand(P, 3, generator=g) * 0.08 + 1e-3),
    }
    grads = torch.rand(P, generator=g) * 0.0005
    return t, grads, g


@pytest.mark.parametrize("i", [500, 3000, 5000])
def test_one_gather_equals_the_reference_passes(i):
    P, radius = 4000, 3.0
    t, grads, g = _state(P, i)
    params = {k: torch.nn.Parameter(v.clone()) for k, v in t.items()}
    params['cam_m'] = torch.nn.Parameter(torch.zeros(4, 3))
    params['cam_c'] = torch.nn.Parameter(torch.zeros(4, 3))
    opt = torch.optim.Adam([{'params': [v], 'name': k, 'lr': 1e-3} for k, v in params.items()], lr=0.0, eps=1e-15)
    for v in params.values():
        v.grad = torch.randn(v.shape, generator=g)
    opt.step()                                                            # populates exp_avg / exp_avg_sq / step
    moments = {k: (opt.state[params[k]]['exp_avg'].clone(), opt.state[params[k]]['exp_avg_sq'].clone()) for k in OD.KEYS}
    values = {k: params[k].detach().clone() for k in OD.KEYS}

    denom = torch.randint(0, 5, (P,), generator=g).float()
    variables = {'means2D_gradient_accum': grads * denom, 'denom': denom.clone(), 'scene_radius': radius,
                 'max_2D_radius': torch.zeros(P), 'seen': torch.zeros(P, dtype=torch.bool)}
    m2d = torch.zeros(P, 3, requires_grad=True)
    m2d.grad = torch.zeros(P, 3)
    variables['means2D'] = m2d
    g_eff = torch.nan_to_num(variables['means2D_gradient_accum'] / variables['denom'], nan=0.0)

    noise = {}

    def noise_fn(src, stds):                                              # deterministic stand-in for torch.normal
        z = torch.sin(src.float()[:, None] * torch.tensor([1.0, 2.0, 3.0]) + torch.arange(src.numel())[:, None] * 0.1)
        noise['z'] = z * stds
        return noise['z']

    ot, om, osrc, okind = OD.densify_step(values, moments, g_eff, radius, i, noise_fn)

    src, kind = DZ.densify_plan(g_eff, torch.exp(values['log_scales']).max(dim=1).values,
                                torch.sigmoid(values['logit_opacities']).squeeze(-1), radius, i)
    assert torch.equal(src, osrc) and torch.equal(kind, okind)
    assert (kind == 1).any() and (kind == 2).any() and src.numel() != P

    step_before = opt.state[params['means3D']]['step'].clone()
    torch_normal = torch.normal
    try:
        torch.normal = lambda mean, std, generator=None: (                # the same stand-in noise, restricted to survivors
            torch.sin(src[kind == 2].float()[:, None] * torch.tensor([1.0, 2.0, 3.0])
                      + _child_positions(osrc, okind, OD, values, g_eff, radius)[:, None] * 0.1) * std)
        params, variables = DZ.densify(params, variables, opt, i)
    finally:
        torch.normal = torch_normal
    for k in OD.KEYS:
        if k == 'logit_opacities' and i % 3000 == 0:
            continue                                                      # reset to logit(0.01) right after (:288-290)
        assert torch.allclose(params[k].detach(), ot[k], atol=1e-6), k
        st = opt.state[params[k]]
        assert torch.equal(st['exp_avg'], om[k][0]) and torch.equal(st['exp_avg_sq'], om[k][1]), k
        assert torch.equal(st['step'], step_before)
    if i % 3000 == 0:
        assert torch.allclose(torch.sigmoid(params['logit_opacities'].detach()), torch.full_like(params['logit_opacities'], 0.01))
        assert float(opt.state[params['logit_opacities']]['exp_avg'].abs().max()) == 0.0
    M = src.numel()
    assert variables['denom'].shape == (M,) and float(variables['means2D_gradient_accum'].abs().max()) == 0.0
    assert params['cam_m'].shape == (4, 3)


def _child_positions(osrc, okind, OD, values, g_eff, radius):
    """Index of every surviving child in the reference's (unpruned) child list: the stand-in noise is keyed on it."""
    P = g_eff.shape[0]
    scale = torch.exp(values['log_scales']).max(dim=1).values
    to_split = (g_eff >= 0.0002) & (scale > 0.01 * radius)
    order = torch.arange(P)[to_split].repeat(2)                           # the reference's child list (parents tiled)
    pos = torch.arange(order.numel())
    surv = osrc[okind == 2]
    # survivors keep the list order; match them up copy by copy
    out = []
    j = 0
    for p in range(order.numel()):
        if j < surv.numel() and order[p] == surv[j]:       # both copies of a parent survive or are pruned together
            out.append(pos[p])
            j += 1
    return torch.stack(out).float() if out else torch.zeros(0)


def test_accumulate_and_schedule():
    P = 10
    variables = {'means2D_gradient_accum': torch.zeros(P), 'denom': torch.zeros(P),
                 'seen': torch.tensor([True] * 5 + [False] * 5)}
    m = torch.zeros(P, 3, requires_grad=True)
    m.grad = torch.ones(P, 3) * 3.0
    variables['means2D'] = m
    DZ.accumulate_mean2d_gradient(variables)
    assert torch.allclose(variables['means2D_gradient_accum'][:5], torch.full((5,), 18.0 ** 0.5))
    assert float(variables['means2D_gradient_accum'][5:].abs().max()) == 0.0
    assert torch.equal(variables['denom'], torch.tensor([1.0] * 5 + [0.0] * 5))
