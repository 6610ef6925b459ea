"""The train.py-shaped loop on the GPU: the loss goes down on a small synthetic scene, densification changes the
point count without breaking the optimizer, the second timestep runs with the fused rigidity losses."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import settings_from_spec  # noqa: E402

pytestmark = pytest.mark.gpu


def _dataset(sc, cams, bg=(0.0, 0.0, 0.0)):
    from dynamic3dgaussians_b200 import rasterizer as RZ
    rv = {k: v.cuda() for k, v in sc.rendervar().items()}
    seg = sc.seg_colors.cuda()
    data = []
    with torch.no_grad():
        for i, c in enumerate(cams):
            st_, _ = settings_from_spec(c, bg=bg)
            color, extra, *_ = RZ.forward_views([st_], rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"],
                                                rv["rotations"], extra_colors=seg)
            data.append({'cam': st_, 'im': color[0].clone(), 'seg': extra[0].clone(), 'id': i})
    return data, rv, seg


def test_loss_decreases_densifies_and_second_timestep_runs():
    from dynamic3dgaussians_b200 import scenes, train_loop as TL, densify as DZ
    sc = scenes.make_scene_c2(6_000, seed=7)
    cams = scenes.cameras_c2(6, w=160, h=96, f=120.0)
    data, rv, seg = _dataset(sc, cams)
    g = torch.Generator().manual_seed(0)
    # start from a perturbed copy of the scene that produced the targets
    pts = rv["means3D"] + 0.01 * torch.randn(rv["means3D"].shape, generator=g).cuda()
    rgb = (rv["colors_precomp"] + 0.2 * torch.randn(rv["colors_precomp"].shape, generator=g).cuda()).clamp(0, 1)
    params, variables = TL.initialize_params(pts, rgb, seg[:, 0].contiguous(), num_cams=len(cams), scene_radius=3.3)
    opt = TL.initialize_optimizer(params, variables)
    hist = TL.train_timestep(params, variables, opt, data, 60, True, generator=g, do_densify=False)
    first = torch.stack(hist[:10]).mean().item()
    last = torch.stack(hist[-10:]).mean().item()
    assert last < first, (first, last)
    assert all(torch.isfinite(h) for h in hist)

    # one densification round in the loop (iteration 500 of the schedule), then more steps with the new point count
    P0 = params['means3D'].shape[0]
    hist = TL.train_timestep(params, variables, opt, data, 3, True, generator=g, do_densify=True, start_iter=498)
    P1 = params['means3D'].shape[0]
    assert P1 != P0 and variables['denom'].shape[0] == P1
    for k in DZ.PER_GAUSSIAN:
        assert params[k].shape[0] == P1
        grp = next(x for x in opt.param_groups if x['name'] == k)
        assert grp['params'][0] is params[k] and opt.state[params[k]]['exp_avg'].shape == params[k].shape
    assert all(torch.isfinite(h) for h in hist)

    # timestep 1: rigidity graph, constant-velocity init, fused dynamic losses
    variables = TL.initialize_post_first_timestep(params, variables, opt)
    params, variables = TL.initialize_per_timestep(params, variables, opt)
    loss, variables, losses = TL.get_loss(params, data[0], variables, False)
    assert set(losses) >= {'im', 'seg', 'rigid', 'rot', 'iso', 'floor', 'bg', 'soft_col_cons'}
    loss.backward()
    assert torch.isfinite(params['means3D'].grad).all() and float(params['means3D'].grad.abs().max()) > 0
    assert torch.isfinite(params['unnorm_rotations'].grad).all()
    hist = TL.train_timestep(params, variables, opt, data, 5, False, generator=g)
    assert all(torch.isfinite(h) for h in hist)
