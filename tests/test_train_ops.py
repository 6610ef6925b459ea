"""Fused training-loop glue (train_ops.py / csrc/train_glue.cu) against the eager torch ops it replaces
(REF/helpers.py:98-107, REF/train.py:181,288-289, REF/external.py:136-140), values and gradients, on the GPU; the CPU
part checks that there is no CPU path."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import rel_err, settings_from_spec  # noqa: E402

TOL = 2e-6   # float32: same formulas, different association in the reductions


def test_no_cpu_path():
    from dynamic3dgaussians_b200 import train_ops as TO
    with pytest.raises(ValueError):
        TO.activations(torch.zeros(4, 4), torch.zeros(4, 1), torch.zeros(4, 3))
    with pytest.raises(ValueError):
        TO.cam_affine(torch.zeros(3, 8, 8), torch.zeros(2, 3), torch.zeros(2, 3), 0)
    with pytest.raises(ValueError):
        TO.update_max_radius(torch.zeros(4, dtype=torch.int32), torch.zeros(4))
    with pytest.raises(ValueError):
        TO.l1_loss_u8(torch.zeros(1, 3, 8, 8), torch.zeros(1, 3, 8, 8, dtype=torch.uint8))


@pytest.mark.gpu
def test_activations_match_torch():
    from dynamic3dgaussians_b200 import train_ops as TO
    g = torch.Generator().manual_seed(3)
    P = 10_007
    q = torch.randn(P, 4, generator=g).cuda()
    lo = (3 * torch.randn(P, 1, generator=g)).cuda()
    ls = torch.randn(P, 3, generator=g).cuda()
    w = [torch.randn(P, 4, generator=g).cuda(), torch.randn(P, 1, generator=g).cuda(), torch.randn(P, 3, generator=g).cuda()]
    res = []
    for fused in (True, False):
        a, b, c = (t.clone().requires_grad_(True) for t in (q, lo, ls))
        if fused:
            r, o, s = TO.activations(a, b, c)
        else:
            r, o, s = torch.nn.functional.normalize(a), torch.sigmoid(b), torch.exp(c)
        ((r * w[0]).sum() + (o * w[1]).sum() + (s * w[2]).sum()).backward()
        res.append((r, o, s, a.grad, b.grad, c.grad))
    for x, y in zip(*res):
        assert rel_err(x, y) <= TOL
    # F.normalize's eps clamp: a zero quaternion stays zero
    z = torch.zeros(3, 4).cuda()
    assert float(TO.activations(z, lo[:3], ls[:3])[0].abs().max()) == 0
    # a cotangent that never arrives is a zero gradient, not garbage
    a, b, c = (t.clone().requires_grad_(True) for t in (q, lo, ls))
    r, o, s = TO.activations(a, b, c)
    (o * w[1]).sum().backward()
    assert float(a.grad.abs().max()) == 0 and float(c.grad.abs().max()) == 0 and rel_err(b.grad, res[1][4]) <= TOL


@pytest.mark.gpu
def test_cam_affine_matches_torch():
    from dynamic3dgaussians_b200 import train_ops as TO
    g = torch.Generator().manual_seed(4)
    im = torch.rand(3, 90, 161, generator=g).cuda()
    m = (0.3 * torch.randn(5, 3, generator=g)).cuda()
    c = (0.1 * torch.randn(5, 3, generator=g)).cuda()
    w = torch.randn(3, 90, 161, generator=g).cuda()
    res = []
    for fused in (True, False):
        a, b, d = (t.clone().requires_grad_(True) for t in (im, m, c))
        out = TO.cam_affine(a, b, d, 3) if fused else torch.exp(b[3])[:, None, None] * a + d[3][:, None, None]
        (out * w).sum().backward()
        res.append((out, a.grad, b.grad, d.grad))
    for x, y in zip(*res):
        assert x.shape == y.shape and rel_err(x, y) <= 2e-5      # 14k-term float32 sums in a different order
    with pytest.raises(IndexError):
        TO.cam_affine(im, m, c, 5)


@pytest.mark.gpu
def test_densify_stats_match_torch():
    from dynamic3dgaussians_b200 import train_ops as TO
    g = torch.Generator().manual_seed(5)
    P = 5_003
    radii = torch.randint(-1, 40, (P,), generator=g).to(torch.int32).cuda()
    maxr = (30 * torch.rand(P, generator=g)).cuda()
    grad = torch.randn(P, 3, generator=g).cuda()
    accum, denom = torch.rand(P, generator=g).cuda(), torch.randint(0, 9, (P,), generator=g).float().cuda()
    seen_ref = radii > 0
    maxr_ref = torch.where(seen_ref, torch.maximum(radii.float(), maxr), maxr)
    norm = torch.norm(grad[:, :2], dim=-1)
    accum_ref = accum + torch.where(seen_ref, norm, torch.zeros_like(norm))
    denom_ref = denom + seen_ref.float()
    seen = TO.update_max_radius(radii, maxr)
    TO.accumulate_mean2d_gradient_(seen, grad, accum, denom)
    assert seen.dtype == torch.bool and torch.equal(seen, seen_ref)
    assert torch.equal(maxr, maxr_ref) and torch.equal(denom, denom_ref)
    assert rel_err(accum, accum_ref) <= TOL


@pytest.mark.gpu
def test_get_loss_fused_glue_matches_eager_glue():
    from dynamic3dgaussians_b200 import rasterizer as RZ, scenes, train_loop as TL
    sc = scenes.make_scene_c2(4_000, seed=11)
    cams = scenes.cameras_c2(3, w=160, h=96, f=120.0)
    rv = {k: v.cuda() for k, v in sc.rendervar().items()}
    seg = sc.seg_colors.cuda()
    st_, _ = settings_from_spec(cams[1], bg=(0.0, 0.0, 0.0))
    with torch.no_grad():
        color, extra, *_ = RZ.forward_views([st_], rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"],
                                            rv["rotations"], extra_colors=seg)
    data = {'cam': st_, 'im': color[0].clone(), 'seg': extra[0].clone(), 'id': 1}
    gen = torch.Generator().manual_seed(1)
    pts = rv["means3D"] + 0.01 * torch.randn(rv["means3D"].shape, generator=gen).cuda()
    out = []
    for fused in (True, False):
        params, variables = TL.initialize_params(pts.clone(), rv["colors_precomp"].clone(), seg[:, 0].contiguous(), 3, 3.3)
        with torch.no_grad():
            params['cam_m'].add_(0.05)
            params['cam_c'].sub_(0.02)
            params['unnorm_rotations'].mul_(1.7)
        loss, variables, _ = TL.get_loss(params, data, variables, True, fused=fused)
        loss.backward()
        TL._accumulate_inplace(variables, fused=fused)
        out.append((loss.detach(), variables['max_2D_radius'], variables['means2D_gradient_accum'], variables['denom'],
                    variables['seen'], *[params[k].grad for k in ('means3D', 'rgb_colors', 'unnorm_rotations',
                                                                  'logit_opacities', 'log_scales', 'cam_m', 'cam_c')]))
    for x, y in zip(*out):
        assert x.shape == y.shape
        if x.dtype == torch.bool:
            assert torch.equal(x, y)
        else:
            assert rel_err(x, y) <= 1e-4          # the rasterizer backward's atomics


@pytest.mark.gpu
@pytest.mark.parametrize("W", [64, 61])          # 4 pixels per thread, and the scalar path
def test_l1_loss_u8_matches_torch(W):
    from dynamic3dgaussians_b200 import train_ops as TO
    g = torch.Generator().manual_seed(6)
    V, H = 3, 37
    img = torch.rand(V, 3, H, W, generator=g).cuda()
    tgt = torch.randint(0, 256, (V, 3, H, W), dtype=torch.uint8, generator=g).cuda()
    img[0, 0, 0, :4] = tgt[0, 0, 0, :4].float() * (1.0 / 255.0)  # sign(0) = 0
    rw = torch.rand(V, H, generator=g).cuda()
    rw[1, 5:] = 0.0
    rw /= rw.sum() * 3 * W
    for weights in (None, rw):
        a = img.clone().requires_grad_(True)
        b = img.clone().requires_grad_(True)
        la = TO.l1_loss_u8(a, tgt, weights)
        d = (b - tgt.float() * (1.0 / 255.0)).abs()
        lb = d.mean() if weights is None else (d * weights[:, None, :, None]).sum()
        (2.5 * la).backward()
        (2.5 * lb).backward()
        assert abs(float(la.detach()) - float(lb.detach())) <= 2e-6 * abs(float(lb.detach()))
        assert rel_err(a.grad, b.grad) <= 1e-6
