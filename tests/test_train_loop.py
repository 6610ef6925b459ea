"""The train.py-shaped loop on the GPU: the loss goes down on a small synthetic scene, densification changes the
point count without breaking the optimizer, the second timestep runs with the fused rigidity losses."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import rel_err, settings_from_spec  # noqa: E402

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


def test_graphed_trainer_matches_eager_iterations():
    """A whole iteration replayed from a CUDA graph gives the same parameters as the eager loop on the same camera
    sequence (Adam is deterministic; the rasterizer's atomics reorder sums, hence the tolerance)."""
    from dynamic3dgaussians_b200 import scenes, train_loop as TL
    sc = scenes.make_scene_c2(5_000, seed=9)
    cams = scenes.cameras_c2(3, w=160, h=96, f=120.0)
    data, rv, seg = _dataset(sc, cams)
    results = []
    for graphed in (False, True):
        params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone() * 0.8,
                                                 seg[:, 0].contiguous(), num_cams=len(cams), scene_radius=3.3)
        opt = TL.initialize_optimizer(params, variables)
        g = torch.Generator().manual_seed(3)
        if graphed:
            tr = TL.GraphedTrainer(params, variables, opt, data, True, capacity=400_000, warmup=1)
            # the warm-up iteration above moved the parameters: start both arms from the same state
            with torch.no_grad():
                for k, v in results[0]["init"].items():
                    params[k].copy_(v)
                for st in opt.state.values():
                    st["exp_avg"].zero_(); st["exp_avg_sq"].zero_(); st["step"].zero_()
                variables["denom"].zero_(); variables["means2D_gradient_accum"].zero_(); variables["max_2D_radius"].zero_()
            losses = [tr.step(g).clone() for _ in range(12)]
        else:
            init = {k: v.detach().clone() for k, v in params.items()}
            losses = TL.train_timestep(params, variables, opt, data, 12, True, generator=g, do_densify=False,
                                       capacity=400_000)
            results.append({"init": init})
        results[-1 if not graphed else 0].update({("g" if graphed else "e"): (
            {k: v.detach().clone() for k, v in params.items()}, torch.stack(losses), variables["denom"].clone())})
    pe, le, de = results[0]["e"]
    pg, lg, dg = results[0]["g"]
    assert torch.allclose(le, lg, rtol=1e-4, atol=1e-6)
    init = results[0]["init"]
    for k in pe:
        # Adam normalises the gradient, so Gaussians whose gradient is at rounding-noise level may step differently
        # (atomics reorder the sums); the bulk of the update must agree
        moved = float((pe[k] - init[k]).abs().mean())
        assert float((pe[k] - pg[k]).abs().mean()) <= 0.05 * moved + 1e-7, k
    assert float(dg.sum()) > 0            # densification statistics accumulate under replay


def test_fused_adam_matches_torch_adam():
    from dynamic3dgaussians_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(4)
    shapes = [(5000, 3), (5000, 1), (5000, 4), (7, 3), (1,)]
    lrs = [1e-3, 0.05, 0.0, 1e-4, 0.01]
    pa = [torch.nn.Parameter(torch.randn(s, generator=g).cuda()) for s in shapes]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    oa = FusedAdam([{'params': [p], 'name': str(i), 'lr': lr} for i, (p, lr) in enumerate(zip(pa, lrs))], lr=0.0, eps=1e-15)
    ob = torch.optim.Adam([{'params': [p], 'name': str(i), 'lr': lr} for i, (p, lr) in enumerate(zip(pb, lrs))], lr=0.0,
                          eps=1e-15)
    for it in range(25):
        for a, b in zip(pa, pb):
            gr = torch.randn(a.shape, generator=g).cuda() * (10.0 ** (it % 5 - 3))
            a.grad, b.grad = gr.clone(), gr.clone()
        if it == 7:
            pa[3].grad = None                  # a tensor without a gradient is skipped, its step counter stays behind
            pb[3].grad = None
        oa.step()
        ob.step()
    for a, b, lr in zip(pa, pb, lrs):
        assert torch.allclose(a, b, rtol=2e-5, atol=1e-7), (a.shape, lr)
        # torch updates the first moment as lerp(m, g, 1 - b1), here fma(b1, m, (1 - b1) g): same value, other rounding
        assert rel_err(oa.state[a]['exp_avg'], ob.state[b]['exp_avg']) <= 1e-6
        assert rel_err(oa.state[a]['exp_avg_sq'], ob.state[b]['exp_avg_sq']) <= 1e-6
        assert float(oa.state[a]['step']) == float(ob.state[b]['step'])
    assert torch.equal(pa[2], pb[2])           # lr 0: untouched
    with pytest.raises(ValueError):
        cpu = torch.nn.Parameter(torch.zeros(3))
        o = FusedAdam([cpu], lr=1e-3)
        cpu.grad = torch.ones(3)
        o.step()
