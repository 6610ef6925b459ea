"""GPU tests (pytest -m gpu) of the boundary behaviour added with ABI version 4:

  * dL/dmeans2D of the COLOUR render alone out of the fused colour + segmentation pass (gsr_backward's
    dL_dmeans2D_color_only; REF/train.py:245, REF/external.py:136-140) -- what densification accumulates;
  * `prefiltered` (R/cuda_rasterizer/auxiliary.h:160-167) and `debug` (auxiliary.h:172-179) of the settings tuple;
  * gsr_set_option instead of environment reads; NVTX ranges;
  * GraphedTrainer: capacity overflow is visible, building it does not train, learning-rate changes are refused;
  * the camera cache sees in-place updates of the pose / background tensors."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import RZ, elem_violations, rel_err, settings_from_spec  # noqa: E402
from dynamic3dgaussians_b200 import _lib, scenes  # noqa: E402

pytestmark = pytest.mark.gpu
GRAD_TOL = 1e-4


def _rv(sc):
    return {k: v.cuda() for k, v in sc.rendervar().items()}


def _scene(P=20_000, views=2):
    sc = scenes.make_scene_c2(P, seed=12)
    cams = scenes.cameras_c2(views, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c, bg=(0.1, 0.2, 0.3))[0] for c in cams]
    return sc, sets, _rv(sc)


def test_color_only_means2D_equals_the_separate_colour_render():
    sc, sets, rv = _scene()
    V, H, W = len(sets), 180, 320
    seg = sc.seg_colors.cuda()
    gen = torch.Generator().manual_seed(3)
    gc = torch.randn(V, 3, H, W, generator=gen).cuda()
    ge = torch.randn(V, 3, H, W, generator=gen).cuda() * 3.0          # the seg loss weighs 3x in the loop
    gd = torch.randn(V, 1, H, W, generator=gen).cuda() * 0.1
    ga = torch.randn(V, 1, H, W, generator=gen).cuda()
    args = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    bg_extra = torch.tensor([0.3, 0.0, 0.6])
    *_, st = RZ.forward_views(sets, *args, extra_colors=seg, bg_extra=bg_extra)
    g = RZ.backward_views(st, *args, None, None, seg, gc, ge, gd, ga, want_means2D_color=True)
    # the reference's two calls: colour render (colour, depth, alpha cotangents), then the segmentation render
    *_, st_c = RZ.forward_views(sets, *args)
    g_c = RZ.backward_views(st_c, *args, None, None, None, gc, None, gd, ga)
    seg_sets = [s._replace(bg=bg_extra.cuda()) for s in sets]
    *_, st_s = RZ.forward_views(seg_sets, rv["means3D"], rv["opacities"], seg, rv["scales"], rv["rotations"])
    g_s = RZ.backward_views(st_s, rv["means3D"], rv["opacities"], seg, rv["scales"], rv["rotations"], None, None, None,
                            ge, None, None, None)
    assert rel_err(g["means2D_color"], g_c["means2D"]) <= GRAD_TOL
    assert elem_violations(g["means2D_color"], g_c["means2D"])[0] == 0
    assert rel_err(g["means2D"], g_c["means2D"] + g_s["means2D"]) <= GRAD_TOL
    assert rel_err(g["means2D_color"], g["means2D"]) > 0.05               # the two really differ on this input
    # asking for the colour-only share changes nothing else
    g0 = RZ.backward_views(st, *args, None, None, seg, gc, ge, gd, ga)
    for k in ("means3D", "opacity", "scales", "rotations", "colors", "extra", "means2D"):
        assert rel_err(g[k], g0[k]) <= 1e-5, k
    # without extra colours the colour render is the whole render
    g1 = RZ.backward_views(st_c, *args, None, None, None, gc, None, gd, ga, want_means2D_color=True)
    assert torch.equal(g1["means2D_color"], g1["means2D"])
    # Gaussians that touch no tile get exactly zero in both
    dead = st.radii <= 0
    assert float(g["means2D_color"][dead].abs().max()) == 0.0
    # through autograd: the handle's .grad is the colour-only share, summed over the views for a [P,3] handle
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in rv.items()}
    h_all = torch.zeros_like(rv["means3D"]).requires_grad_(True)
    h_col = torch.zeros(V, rv["means3D"].shape[0], 3, device="cuda", requires_grad=True)
    color, extra, depth, alpha, _ = RZ.rasterize_views(
        sets, leaves["means3D"], leaves["opacities"], means2D=h_all, colors_precomp=leaves["colors_precomp"],
        scales=leaves["scales"], rotations=leaves["rotations"], extra_colors=seg, bg_extra=bg_extra, means2D_color=h_col)
    ((color * gc).sum() + (extra * ge).sum() + (depth * gd).sum() + (alpha * ga).sum()).backward()
    assert rel_err(h_col.grad, g_c["means2D"]) <= GRAD_TOL
    assert rel_err(h_all.grad, (g_c["means2D"] + g_s["means2D"]).sum(0)) <= GRAD_TOL


def test_train_loop_feeds_densification_from_the_colour_render_only():
    """get_loss + accumulate: `variables['means2D'].grad` must be what the reference's colour render alone would give
    (REF/train.py:245), i.e. independent of the segmentation loss; and the selection grads >= 0.0002 follows it."""
    from dynamic3dgaussians_b200 import train_loop as TL
    sc = scenes.make_scene_c2(20_000, seed=4)
    cams = scenes.cameras_c2(3, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c)[0] for c in cams]
    dev = "cuda"
    rv = _rv(sc)
    seg = sc.seg_colors.cuda()[:, 0].contiguous()
    g = torch.Generator().manual_seed(2)
    data = [{'cam': s, 'id': i, 'im': torch.rand(3, 180, 320, generator=g).to(dev),
             'seg': torch.rand(3, 180, 320, generator=g).to(dev)} for i, s in enumerate(sets)]

    def stats(weights):
        params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, 3, 3.3)
        loss, variables, _ = TL.get_loss(params, data[1], variables, True, weights=weights)
        loss.backward()
        return variables['means2D'].grad.clone(), variables['seen'].clone()

    w = dict(TL.LOSS_WEIGHTS)
    g_with_seg, seen = stats(w)
    g_no_seg, seen2 = stats(dict(w, seg=0.0))
    assert torch.equal(seen, seen2)
    assert rel_err(g_with_seg, g_no_seg) <= GRAD_TOL
    norm_a = torch.norm(g_with_seg[:, :2], dim=-1)
    norm_b = torch.norm(g_no_seg[:, :2], dim=-1)
    margin = (norm_b - 0.0002).abs() > 1e-7                                 # away from the threshold's knife edge
    assert torch.equal((norm_a >= 0.0002)[margin], (norm_b >= 0.0002)[margin])
    # and it is NOT the gradient of the fused loss: with the segmentation term the total differs
    params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, 3, 3.3)
    rvv = TL.params2rendervar(params)
    color, segim, _, _, _ = RZ.rasterize_views([sets[1]], rvv['means3D'], rvv['opacities'], means2D=rvv['means2D'],
                                               colors_precomp=rvv['colors_precomp'], scales=rvv['scales'],
                                               rotations=rvv['rotations'], extra_colors=params['seg_colors'])
    from dynamic3dgaussians_b200.image_loss import l1_ssim_loss
    (l1_ssim_loss(color[0], data[1]['im']) + 3.0 * l1_ssim_loss(segim[0], data[1]['seg'])).backward()
    assert rel_err(rvv['means2D'].grad, g_with_seg) > 0.05


def test_prefiltered_flag():
    cs = scenes.camera_c1()
    cam, _ = settings_from_spec(cs)
    sc = scenes.make_scene_c1(3000, seed=5)
    rv = _rv(sc)
    args = lambda m: (m, rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    pf = cam._replace(prefiltered=True)
    # everything in front of the camera: the promise holds
    c1, *_ , st1 = RZ.forward_views([pf], *args(rv["means3D"]))
    c0, *_ = RZ.forward_views([cam], *args(rv["means3D"]))
    assert torch.equal(c1, c0) and not RZ.read_flags(st1)["prefilter_violation"]
    # a third of the cloud behind the camera although prefiltered=True: the reference traps the kernel
    # (auxiliary.h:160-167), the drop-in raises
    far = rv["means3D"].clone()
    far[::3] = far[::3] * 0.1 + torch.tensor([0.9, -0.75, -4.35], device="cuda") * 2.0     # behind the eye of camera_c1
    with pytest.raises(RuntimeError, match="filtered although prefiltered"):
        RZ.forward_views([pf], *args(far))
    # with a caller-chosen capacity nothing synchronises: the flag is there to be read afterwards
    *_, st2 = RZ.forward_views([pf], *args(far), capacity=400_000)
    f = RZ.read_flags(st2)
    assert f["prefilter_violation"] and not f["overflowed"]
    *_, st3 = RZ.forward_views([cam], *args(far), capacity=400_000)
    assert not RZ.read_flags(st3)["prefilter_violation"]


def test_debug_mode_syncs_every_stage_and_changes_no_result():
    sc, sets, rv = _scene(10_000, 2)
    args = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    dbg = [s._replace(debug=True) for s in sets]
    c0, _, d0, a0, r0, st0 = RZ.forward_views(sets, *args)
    c1, _, d1, a1, r1, st1 = RZ.forward_views(dbg, *args)
    assert st1.flags == _lib.FLAG_DEBUG and st0.flags == 0
    assert torch.equal(c0, c1) and torch.equal(d0, d1) and torch.equal(r0, r1)
    gc = torch.ones_like(c0)
    g0 = RZ.backward_views(st0, *args, None, None, None, gc, None, None, None)
    g1 = RZ.backward_views(st1, *args, None, None, None, gc, None, None, None)
    for k in ("means3D", "opacity", "scales"):
        assert rel_err(g1[k], g0[k]) <= 1e-5
    # the drop-in module with debug=True, as the reference's callers may set it (REF/helpers.py:91)
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in rv.items()}
    im, radius, depth = Renderer(raster_settings=dbg[0])(means2D=torch.zeros_like(rv["means3D"]), **leaves)
    im.sum().backward()
    assert torch.equal(im, c0[0]) and torch.isfinite(leaves["means3D"].grad).all()


def test_options_replace_environment_reads():
    lib = _lib.load()
    assert lib.gsr_abi_version() == 4
    assert lib.gsr_set_option(b"no_such_option", 1) == -1 and lib.gsr_set_option(b"blend_ilp", 7) == -1
    with pytest.raises(RuntimeError):
        _lib.set_option("no_such_option", 1)
    for key in ("blend_ilp", "bwd_cost_order", "nvtx"):
        old = _lib.get_option(key)
        _lib.set_option(key, 1)
        assert _lib.get_option(key) == 1
        _lib.set_option(key, old)
    # NVTX ranges on: the stages still run (the ranges resolve to no-ops without a profiler attached)
    sc, sets, rv = _scene(5_000, 1)
    args = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    c0, *_ = RZ.forward_views(sets, *args)
    _lib.set_option("nvtx", 1)
    try:
        c1, *_, st = RZ.forward_views(sets, *args)
        RZ.backward_views(st, *args, None, None, None, torch.ones_like(c1), None, None, None)
    finally:
        _lib.set_option("nvtx", 0)
    assert torch.equal(c0, c1)


def test_camera_cache_sees_in_place_updates():
    sc, sets, rv = _scene(5_000, 1)
    args = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    s = sets[0]
    c0, *_ = RZ.forward_views([s], *args)
    s.bg.copy_(torch.tensor([0.9, 0.8, 0.7]))                  # random backgrounds are drawn in place
    c1, *_ = RZ.forward_views([s], *args)
    assert not torch.equal(c0, c1)
    assert float((c1[0, :, 0, 0].cpu() - torch.tensor([0.9, 0.8, 0.7])).abs().max()) < 0.5
    vm = s.viewmatrix.clone()
    s.viewmatrix.mul_(1.0)                                     # bumps the version, same values: same image
    c2, *_ = RZ.forward_views([s], *args)
    assert torch.equal(c1, c2)
    s.viewmatrix[..., 3, 2] += 0.25                            # pose optimisation moves the camera in place
    c3, *_ = RZ.forward_views([s], *args)
    assert not torch.equal(c2, c3)
    s.viewmatrix.copy_(vm)
    # a temporary bg_extra tensor freed and re-created at the same address must not serve a stale background
    seg = sc.seg_colors.cuda()
    outs = []
    for val in (0.1, 0.7):
        be = torch.full((3,), val)
        _, e, *_ = RZ.forward_views([s], *args, extra_colors=seg, bg_extra=be)
        outs.append(e.clone())
        del be
    assert not torch.equal(outs[0], outs[1])


def test_graphed_trainer_capacity_warmup_and_learning_rates():
    from dynamic3dgaussians_b200 import train_loop as TL
    sc = scenes.make_scene_c2(15_000, seed=6)
    cams = scenes.cameras_c2(3, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c)[0] for c in cams]
    rv = _rv(sc)
    seg = sc.seg_colors.cuda()[:, 0].contiguous()
    g = torch.Generator().manual_seed(2)
    data = [{'cam': s, 'id': i, 'im': torch.rand(3, 180, 320, generator=g).cuda(),
             'seg': torch.rand(3, 180, 320, generator=g).cuda()} for i, s in enumerate(sets)]
    params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, 3, 3.3)
    opt = TL.initialize_optimizer(params, variables)
    before = {k: v.detach().clone() for k, v in params.items()}
    trainer = TL.GraphedTrainer(params, variables, opt, data, True, capacity=6_000_000)
    # building the trainer (3 warm-up iterations + one capture per camera) must not train
    for k, v in params.items():
        assert torch.equal(v.detach(), before[k]), k
    assert float(variables['denom'].abs().max()) == 0.0 and float(variables['means2D_gradient_accum'].abs().max()) == 0.0
    assert all(float(opt.state[v]['step']) == 0.0 for v in params.values() if v in opt.state)
    assert all(float(opt.state[v]['exp_avg'].abs().max()) == 0.0 for v in params.values() if v in opt.state)
    for _ in range(6):
        trainer.step(g)
    torch.cuda.synchronize()
    assert not trainer.overflowed() and 0 < trainer.max_num_rendered() <= 6_000_000
    assert float(variables['denom'].max()) >= 1.0 and float(opt.state[params['means3D']]['step']) == 6.0
    assert not torch.equal(params['means3D'].detach(), before['means3D'])
    # a capacity below what a view needs: the render drops instances -- and says so
    small = TL.GraphedTrainer(params, variables, opt, data, True, capacity=trainer.max_num_rendered() // 4)
    for _ in range(3):
        small.step(g)
    assert small.overflowed()
    # learning rates are baked into the captured iterations
    for group in opt.param_groups:
        if group['name'] == 'log_scales':
            group['lr'] = 0.0
    trainer.order = []
    with pytest.raises(RuntimeError, match="learning rate changed"):
        trainer.step(g)


def test_blend_ring_is_repeatable():
    """compute-sanitizer is closed on this GPU pool (profiles/r02_sanitize_closed_on_pool.txt), so the barrier-free TMA
    ring of the blend kernels is checked the way a race would show: deep tile lists (many ring wraps, slots re-issued
    while other warps still read their neighbours), the same launch repeated, every forward output bit-identical and the
    backward equal to summation-order noise.  (The same kernels are also bit-identical to the reference binary at 300k x
    27 views, test_gpu_branches.py.)"""
    sc = scenes.make_scene_c2(80_000, seed=7)
    cams = scenes.cameras_c2(27)[2:27:9]
    sets = [settings_from_spec(c, bg=(0.3, 0.2, 0.1))[0] for c in cams]
    rv = _rv(sc)
    seg = sc.seg_colors.cuda()
    args = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    gen = torch.Generator().manual_seed(8)
    gc = torch.randn(3, 3, 360, 640, generator=gen).cuda()
    first = None
    for ilp in (1, 2):
        _lib.set_option("blend_ilp", ilp)
        for rep in range(4):
            c, e, d, a, r, st = RZ.forward_views(sets, *args, extra_colors=seg)
            L = RZ.workspace_layout(st.P, st.V, st.W, st.H, st.channels, st.sh_coeffs, st.capacity)
            n = st.V * st.H * st.W
            ncon = st.image[L.n_contrib: L.n_contrib + 4 * n].view(torch.int32).clone()
            g = RZ.backward_views(st, *args, None, None, seg, gc, gc, None, None)
            cur = (c, e, d, a, ncon)
            if first is None:
                first, g_first = cur, g
                assert int(ncon.max()) > 3 * 128                       # the ring wrapped
                continue
            assert all(torch.equal(x, y) for x, y in zip(cur, first)), (ilp, rep)
            for k in ("means3D", "opacity", "scales", "rotations", "colors", "extra"):
                assert rel_err(g[k], g_first[k]) <= 2e-6, (k, ilp, rep)
    _lib.set_option("blend_ilp", 0)
