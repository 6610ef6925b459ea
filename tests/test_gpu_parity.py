"""GPU parity tests (pytest -m gpu): the CUDA path, through the C ABI, against the CPU oracle, the golden fixture and the
reference CUDA build (oracle/_ref).  Tolerances are BASELINE.json's: bit-exact radii / sorted lists / tile ranges /
n_contrib; images and depth <= 1e-5 max-abs; gradients <= 1e-4 relative (max-abs error over max-abs value)."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import GOLDEN, O, RZ, export_view, rel_err, settings_from_spec  # noqa: E402
from dynamic3dgaussians_b200 import _lib, scenes  # noqa: E402
from oracle import ref_cuda  # noqa: E402

pytestmark = pytest.mark.gpu
IMG_TOL = 1e-5
GRAD_TOL = 1e-4


def _rv(sc):
    return {k: v.cuda() for k, v in sc.rendervar().items()}


def _fwd(sets, rv, **kw):
    return RZ.forward_views(sets, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], **kw)


@pytest.mark.parametrize("key_bytes,n,bits", [(4, 1, 32), (4, 4096, 8), (4, 100_003, 15), (4, 2_500_000, 10),
                                              (8, 5000, 37), (8, 3073, 64), (8, 1_500_000, 42)])
def test_onesweep_sort_is_stable_and_exact(key_bytes, n, bits):
    lib = _lib.load()
    g = torch.Generator().manual_seed(n)
    hi = 2 ** 31 - 1 if key_bytes == 4 else 2 ** 62
    keys = torch.randint(0, hi, (n,), generator=g, dtype=torch.int64)
    if bits < key_bytes * 8:
        keys &= (1 << bits) - 1
    if n > 1000:
        keys[: n // 2] = keys[n // 2: 2 * (n // 2)]            # many duplicates: exercises stability
    kd = (keys.to(torch.int32) if key_bytes == 4 else keys).cuda()
    vd = torch.arange(n, dtype=torch.int32).cuda()
    ko, vo = torch.empty_like(kd), torch.empty_like(vd)
    tb = lib.gsr_sort_temp_bytes(n, key_bytes)
    tmp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    _lib.check(lib.gsr_sort_pairs(key_bytes, kd.data_ptr(), ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), n, 0, bits,
                                  tmp.data_ptr(), tb, torch.cuda.current_stream().cuda_stream), "sort")
    rk, ri = torch.sort(keys, stable=True)
    assert torch.equal(ko.cpu().to(torch.int64), rk)
    assert torch.equal(vo.cpu().to(torch.int64), ri)


def test_golden_fixture_forward_and_backward():
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden
    sc, cs, cot = make_golden.build_inputs()
    z = np.load(GOLDEN)
    st_, _ = settings_from_spec(cs, bg=(0.15, 0.25, 0.35))
    rv = _rv(sc)
    seg = sc.seg_colors.cuda()
    color, extra, depth, alpha, radii, st = _fwd([st_], rv, extra_colors=seg)
    keys, plist, ranges, ncon = export_view(st)
    assert np.array_equal(radii[0].cpu().numpy(), z["radii"])
    assert np.array_equal(plist.cpu().numpy(), z["point_list"])
    assert np.array_equal(keys.cpu().numpy().astype(np.uint64), z["keys_sorted"])
    assert np.array_equal(ranges.cpu().numpy(), z["ranges"])
    ok = ~z["fragile"]
    assert np.array_equal(ncon.cpu().numpy()[ok], z["n_contrib"][ok])
    img = torch.cat([color[0], extra[0]], 0).cpu().numpy()
    assert np.abs(img - z["color"])[:, ok].max() <= IMG_TOL
    assert np.abs(depth[0].cpu().numpy() - z["depth"])[:, ok].max() <= IMG_TOL
    assert np.abs(alpha[0].cpu().numpy() - z["alpha"])[:, ok].max() <= IMG_TOL
    okt = torch.from_numpy(ok).float()
    g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                          None, seg, (cot["color"] * okt).cuda()[None], (cot["extra"] * okt).cuda()[None],
                          (cot["depth"] * okt).cuda()[None], (cot["alpha"] * okt).cuda()[None])
    pairs = dict(means3D="means3D", opacity="opacities", scales="scales", rotations="rotations",
                 colors="colors_precomp", extra="seg")
    for mine, gold in pairs.items():
        assert rel_err(g[mine], torch.from_numpy(z["g_" + gold])) <= GRAD_TOL, mine
    assert rel_err(g["means2D"][0], torch.from_numpy(z["g_means2D"])) <= GRAD_TOL


def test_c1_against_oracle_on_the_same_inputs():
    sc = scenes.make_scene_c1(10_000)
    cs = scenes.camera_c1()
    st_, oc = settings_from_spec(cs, bg=(0.1, 0.2, 0.3))
    rv = _rv(sc)
    color, _, depth, alpha, radii, st = _fwd([st_], rv)
    keys, plist, ranges, ncon = export_view(st)
    leaves = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in rv.items()}
    out = O.rasterize(oc, leaves["means3D"], leaves["opacities"], colors_precomp=leaves["colors_precomp"],
                      scales=leaves["scales"], rotations=leaves["rotations"], collect_fragile=True)
    assert torch.equal(radii[0].cpu(), out["pre"]["radii"])
    assert st.num_rendered == out["lists"]["num_rendered"]
    assert np.array_equal(keys.cpu().numpy().astype(np.uint64), out["lists"]["keys_sorted"])
    assert np.array_equal(plist.cpu().numpy(), out["lists"]["point_list"])
    assert np.array_equal(ranges.cpu().numpy(), out["lists"]["ranges"])
    ok = ~out["fragile"]
    assert int(out["fragile"].sum()) < 0.01 * ok.numel()
    assert torch.equal(ncon.cpu().long()[ok], out["n_contrib"][ok])
    assert float((color[0].cpu() - out["color"]).abs()[:, ok].max()) <= IMG_TOL
    assert float((depth[0].cpu() - out["depth"]).abs()[:, ok].max()) <= IMG_TOL
    assert float((alpha[0].cpu() - out["alpha"]).abs()[:, ok].max()) <= IMG_TOL
    gen = torch.Generator().manual_seed(1)
    okf = ok.float()
    gc, gd, ga = [torch.randn(c, cs.h, cs.w, generator=gen) * okf for c in (3, 1, 1)]
    ((out["color"] * gc).sum() + (out["depth"] * gd).sum() + (out["alpha"] * ga).sum()).backward()
    g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                          None, None, gc.cuda()[None], None, gd.cuda()[None], ga.cuda()[None])
    assert rel_err(g["means3D"], leaves["means3D"].grad) <= GRAD_TOL
    assert rel_err(g["colors"], leaves["colors_precomp"].grad) <= GRAD_TOL
    assert rel_err(g["opacity"], leaves["opacities"].grad) <= GRAD_TOL
    assert rel_err(g["scales"], leaves["scales"].grad) <= GRAD_TOL
    assert rel_err(g["rotations"], leaves["rotations"].grad) <= GRAD_TOL
    assert rel_err(g["means2D"][0], O.means2D_grad_ndc(out["pre"])) <= GRAD_TOL


@pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref not built")
def test_vs_reference_cuda_bit_exact_c1_and_c2_views():
    """The reference's own CUDA, compiled from /root/reference into oracle/_ref: every integer artefact and the images
    must be IDENTICAL (the per-pair arithmetic is pinned to the reference binary's)."""
    cases = [(scenes.make_scene_c1(10_000), [scenes.camera_c1()]),
             (scenes.make_scene_c2(60_000, seed=2), scenes.cameras_c2(3))]
    for sc, cams in cases:
        rv = _rv(sc)
        sets = [settings_from_spec(c, bg=(0.2, 0.0, 0.1))[0] for c in cams]
        color, _, depth, alpha, radii, st = _fwd(sets, rv)
        for v, s in enumerate(sets):
            rr = ref_cuda.RefRasterizer(s)
            ro = rr.forward(rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
            rs = rr.export_state(ro["num_rendered"])
            keys, plist, ranges, ncon = export_view(st, v)
            assert torch.equal(radii[v], ro["radii"])
            assert keys.numel() == ro["num_rendered"]
            assert torch.equal(keys, rs["keys"]) and torch.equal(plist, rs["point_list"])
            assert torch.equal(ranges, rs["ranges"]) and torch.equal(ncon, rs["n_contrib"])
            assert torch.equal(color[v], ro["color"]) and torch.equal(depth[v], ro["depth"])


def test_batched_views_equal_single_views_and_fused_extra_colours():
    sc = scenes.make_scene_c2(30_000, seed=3)
    cams = scenes.cameras_c2(4, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c)[0] for c in cams]
    rv = _rv(sc)
    seg = sc.seg_colors.cuda()
    cb, eb, db, ab, rb, stb = _fwd(sets, rv, extra_colors=seg)
    gen = torch.Generator().manual_seed(2)
    gc = torch.randn(4, 3, 180, 320, generator=gen).cuda()
    ge = torch.randn(4, 3, 180, 320, generator=gen).cuda()
    gb = RZ.backward_views(stb, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"],
                           None, None, seg, gc, ge, None, None)
    acc = None
    for v, s in enumerate(sets):
        c1, _, d1, a1, r1, st1 = _fwd([s], rv)
        rv2 = dict(rv, colors_precomp=seg)
        c2, _, _, _, _, st2 = _fwd([s], rv2)                       # the reference's second (segmentation) render
        assert torch.equal(cb[v], c1[0]) and torch.equal(eb[v], c2[0]) and torch.equal(rb[v], r1[0])
        assert torch.equal(db[v], d1[0]) and torch.equal(ab[v], a1[0])
        g1 = RZ.backward_views(st1, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"],
                               rv["rotations"], None, None, None, gc[v:v + 1], None, None, None)
        g2 = RZ.backward_views(st2, rv["means3D"], rv["opacities"], seg, rv["scales"], rv["rotations"], None, None,
                               None, ge[v:v + 1], None, None, None)
        s_ = {k: g1[k] + g2[k] for k in ("means3D", "opacity", "scales", "rotations")}
        acc = s_ if acc is None else {k: acc[k] + s_[k] for k in acc}
    for k in acc:
        assert rel_err(gb[k], acc[k]) <= GRAD_TOL, k


def test_tile_row_bands_partition_a_view():
    """Band units (dist.partition_units): a view rendered as bands of tile rows gives, band by band, exactly the rows
    of the whole-view images, and the bands' gradients add up to the whole view's (what the all-reduce sums)."""
    from dynamic3dgaussians_b200 import dist as D
    sc = scenes.make_scene_c2(30_000, seed=5)
    cams = scenes.cameras_c2(3, w=320, h=180, f=233.0)                 # 12 tile rows, the last one ragged (180 = 11*16+4)
    sets = [settings_from_spec(c)[0] for c in cams]
    rv = _rv(sc)
    gen = torch.Generator().manual_seed(4)
    gc = torch.randn(3, 3, 180, 320, generator=gen).cuda()
    gd = torch.randn(3, 1, 180, 320, generator=gen).cuda() * 0.1
    cw, _, dw, aw, rw, stw = _fwd(sets, rv)
    gw = RZ.backward_views(stw, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"],
                           None, None, None, gc, None, gd, None)
    costs = RZ.tile_row_costs(stw)
    assert costs.shape == (3, 12) and int(costs.sum().item()) == stw.num_rendered
    parts = D.partition_units(costs.tolist(), 4, fixed_cost=0.1 * float(costs.sum()) / 3)
    assert any(len(p) > 1 for p in parts) and sum(len(p) for p in parts) > 3        # at least one view is cut
    acc = None
    total = 0
    vis = torch.zeros_like(rw)
    for units in parts:
        views, bands = D.split_units(units)
        out = RZ.forward_views([sets[v] for v in views], rv["means3D"], rv["opacities"], rv["colors_precomp"],
                               rv["scales"], rv["rotations"], bands=bands)
        cb, _, db, ab, rb, stb = out
        total += stb.num_rendered
        for i, (v, r0, r1) in enumerate(units):
            y0, y1 = r0 * 16, min(r1 * 16, 180)
            assert torch.equal(cb[i, :, y0:y1], cw[v, :, y0:y1]) and torch.equal(db[i, :, y0:y1], dw[v, :, y0:y1])
            assert torch.equal(ab[i, :, y0:y1], aw[v, :, y0:y1])
            outside = torch.ones(180, dtype=torch.bool, device="cuda")
            outside[y0:y1] = False
            assert (ab[i, 0, outside] == 0).all() and (cb[i, :, outside] == 0).all()           # background only
            assert ((rb[i] == 0) | (rb[i] == rw[v])).all()                     # radii: the view's, or 0 outside the band
            vis[v] |= rb[i]
        g = RZ.backward_views(stb, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"],
                              rv["rotations"], None, None, None, gc[views].contiguous(), None, gd[views].contiguous(),
                              None)
        part = {k: g[k] for k in ("means3D", "colors", "opacity", "scales", "rotations")}
        acc = part if acc is None else {k: acc[k] + part[k] for k in acc}
    assert total == stw.num_rendered and torch.equal(vis, rw)
    for k in acc:
        assert rel_err(acc[k], gw[k]) <= GRAD_TOL, k
    # the sharded step takes the same unit lists
    slab = D.GradSlab(rv["means3D"].shape[0], "cuda")
    D.sharded_step(sets, parts[0] + parts[1] + parts[2] + parts[3], rv, dict(color=gc, depth=gd, alpha=None), slab=slab)
    for k in acc:
        assert rel_err(slab.views[k].reshape(gw[k].shape), gw[k]) <= GRAD_TOL, k


def test_blend_loop_variants_agree_bit_for_bit():
    """The forward blend kernel exists in a one-entry-per-trip and a two-entries-per-trip variant (picked by launch
    size, common.cuh blend_ilp2).  Same arithmetic in the same order per pixel: images, n_contrib and final T must be
    identical (and so the gradients, up to the atomics' reordering)."""
    sc = scenes.make_scene_c2(30_000, seed=9)
    cams = scenes.cameras_c2(2, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c)[0] for c in cams]
    rv = _rv(sc)
    seg = sc.seg_colors.cuda()
    gen = torch.Generator().manual_seed(6)
    gc = torch.randn(2, 3, 180, 320, generator=gen).cuda()
    ge = torch.randn(2, 3, 180, 320, generator=gen).cuda()
    gd = torch.randn(2, 1, 180, 320, generator=gen).cuda() * 0.1
    res = {}
    for ilp in ("1", "2"):
        _lib.set_option("blend_ilp", int(ilp))                       # gsr_set_option: the library reads no environment
        assert _lib.get_option("blend_ilp") == int(ilp)
        for extra in (None, seg):
            c, e, d, a, r, st = _fwd(sets, rv, extra_colors=extra)
            ncon = [export_view(st, v)[3] for v in range(2)]
            g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"],
                                  rv["rotations"], None, None, extra, gc, None if extra is None else ge, gd, None)
            res[(ilp, extra is None)] = (c, e, d, a, ncon, g)
    _lib.set_option("blend_ilp", 0)
    for plain in (True, False):
        c1, e1, d1, a1, n1, g1 = res[("1", plain)]
        c2, e2, d2, a2, n2, g2 = res[("2", plain)]
        assert torch.equal(c1, c2) and torch.equal(d1, d2) and torch.equal(a1, a2)
        assert plain or torch.equal(e1, e2)
        assert all(torch.equal(x, y) for x, y in zip(n1, n2))
        for k in ("means3D", "colors", "opacity", "scales", "rotations", "means2D"):
            assert rel_err(g2[k], g1[k]) <= GRAD_TOL, k


def test_edge_cases_empty_culled_and_ragged():
    cs = scenes.camera_c1()
    st_, _ = settings_from_spec(cs, bg=(0.5, 0.25, 0.125))
    z = lambda *s: torch.zeros(*s, device="cuda")
    # P = 0
    color, _, depth, alpha, radii, st = RZ.forward_views([st_], z(0, 3), z(0, 1), z(0, 3), z(0, 3), z(0, 4))
    assert st.num_rendered == 0 and radii.shape == (1, 0)
    assert torch.allclose(color[0, :, 5, 7], torch.tensor([0.5, 0.25, 0.125], device="cuda"))
    assert float(depth.abs().max()) == 0 and float(alpha.abs().max()) == 0
    # every Gaussian behind the camera: culled, image = background, zero gradients
    sc = scenes.make_scene_c1(500)
    rv = _rv(sc)
    rv["means3D"] = rv["means3D"] + torch.tensor([0.0, 0.0, -50.0], device="cuda")
    color, _, depth, alpha, radii, st = _fwd([st_], rv)
    assert st.num_rendered == 0 and int(radii.abs().sum()) == 0
    g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                          None, None, torch.ones_like(color), None, None, None)
    assert all(float(g[k].abs().max()) == 0 for k in ("means3D", "opacity", "scales", "rotations", "colors"))
    # image size that is not a multiple of the tile size, Gaussians far larger than the image
    k = [[30.0, 0, 18.3], [0, 30.0, 11.1], [0, 0, 1]]
    cs2 = scenes.CameraSpec(37, 23, k, cs.w2c)
    st2, oc2 = settings_from_spec(cs2)
    sc = scenes.make_scene_c1(300, seed=9)
    rv = _rv(sc)
    rv["scales"] = rv["scales"] * 6.0
    color, _, depth, alpha, radii, st = _fwd([st2], rv)
    cpu = {k_: v.cpu() for k_, v in rv.items()}
    out = O.rasterize(oc2, cpu["means3D"], cpu["opacities"], colors_precomp=cpu["colors_precomp"], scales=cpu["scales"],
                      rotations=cpu["rotations"], collect_fragile=True)
    ok = ~out["fragile"]
    assert torch.equal(radii[0].cpu(), out["pre"]["radii"]) and st.num_rendered == out["lists"]["num_rendered"]
    assert float((color[0].cpu() - out["color"]).abs()[:, ok].max()) <= IMG_TOL


def test_drop_in_module_autograd_and_label_mask():
    """GaussianRasterizer(raster_settings)(**rendervar) exactly as REF/train.py:142 calls it, through autograd."""
    from diff_gaussian_rasterization import GaussianRasterizer as Renderer
    sc = scenes.make_scene_c1(3000, seed=4)
    cs = scenes.camera_c1()
    cam, _ = settings_from_spec(cs)
    params = {k: v.cuda().requires_grad_(True) for k, v in sc._asdict().items()}
    rendervar = {
        'means3D': params['means3D'], 'colors_precomp': params['rgb_colors'],
        'rotations': torch.nn.functional.normalize(params['unnorm_rotations']),
        'opacities': torch.sigmoid(params['logit_opacities']), 'scales': torch.exp(params['log_scales']),
        'means2D': torch.zeros_like(params['means3D'], requires_grad=True, device="cuda") + 0}
    rendervar['means2D'].retain_grad()
    im, radius, depth = Renderer(raster_settings=cam)(**rendervar)
    assert im.shape == (3, cs.h, cs.w) and radius.shape == (3000,) and depth.shape == (1, cs.h, cs.w)
    (im.mean() + 0.1 * depth.mean()).backward()
    for k in ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales"):
        assert params[k].grad is not None and torch.isfinite(params[k].grad).all() and float(params[k].grad.abs().sum()) > 0
    assert rendervar['means2D'].grad.shape == (3000, 3) and float(rendervar['means2D'].grad[:, 2].abs().max()) == 0
    # 4- and 5-tuples and the per-Gaussian label mask (R/.../__init__.py:159-170)
    label = torch.zeros(3000, device="cuda")
    label[1500:] = 1
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in rendervar.items() if k != 'means2D'}
    out4 = Renderer(raster_settings=cam, returns=4)(means2D=torch.zeros(3000, 3, device="cuda"), label=label, **leaves)
    assert len(out4) == 4
    out4[0].sum().backward()
    assert float(leaves["means3D"].grad[:1500].abs().max()) == 0 and float(leaves["means3D"].grad[1500:].abs().max()) > 0
    vis = Renderer(raster_settings=cam).markVisible(params['means3D'].detach())
    assert vis.dtype == torch.bool and bool(vis.all())


def test_semantic_feature_map_forward_and_backward():
    """The vendored variant's 32-channel feature map (forward.cu:381-383,405-406; backward.cu:594-611): bit-exact
    against the reference's own CUDA forward, <= 1e-5 against the oracle (features ride along as extra channels there),
    gradients (incl. dL/dfeature and the feature term in the geometry gradients) <= 1e-4 against the oracle's autograd;
    several views at once equal the views one at a time."""
    sc = scenes.make_scene_c1(8_000)
    cs = scenes.camera_c1()
    bg = (0.1, 0.2, 0.3)
    st_, oc = settings_from_spec(cs, bg=bg)
    rv = _rv(sc)
    P = rv["means3D"].shape[0]
    gen = torch.Generator().manual_seed(5)
    feat = torch.rand(P, 32, generator=gen)
    color, _, depth, alpha, radii, st = _fwd([st_], rv, semantic_feature=feat.cuda()[:, None, :])
    fmap = st.feature_map
    assert fmap.shape == (1, 32, cs.h, cs.w)
    if ref_cuda.available():
        rr = ref_cuda.RefRasterizer(st_)
        ro = rr.forward(rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"],
                        semantic=feat.cuda())
        assert torch.equal(color[0], ro["color"])
        assert torch.equal(fmap[0], ro["feature"])
    leaves = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in rv.items()}
    fleaf = feat.clone().requires_grad_(True)
    oc35 = oc._replace(bg=torch.cat([oc.bg.reshape(-1)[:3], oc.bg.reshape(-1)[:3], torch.zeros(29)]))
    out = O.rasterize(oc35, leaves["means3D"], leaves["opacities"], colors_precomp=leaves["colors_precomp"],
                      scales=leaves["scales"], rotations=leaves["rotations"], extra_colors=fleaf, collect_fragile=True)
    ok = ~out["fragile"]
    assert float((color[0].cpu() - out["color"][:3]).abs()[:, ok].max()) <= IMG_TOL
    assert float((fmap[0].cpu() - out["color"][3:]).abs()[:, ok].max()) <= IMG_TOL
    okf = ok.float()
    gc, gf, gd, ga = [torch.randn(c, cs.h, cs.w, generator=gen) * okf for c in (3, 32, 1, 1)]
    ((out["color"][:3] * gc).sum() + (out["color"][3:] * gf).sum() + (out["depth"] * gd).sum()
     + (out["alpha"] * ga).sum()).backward()
    g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                          None, None, gc.cuda()[None], None, gd.cuda()[None], ga.cuda()[None],
                          semantic_feature=feat.cuda(), dL_dfeature=gf.cuda()[None])
    assert rel_err(g["features"], fleaf.grad) <= GRAD_TOL
    assert rel_err(g["means3D"], leaves["means3D"].grad) <= GRAD_TOL
    assert rel_err(g["colors"], leaves["colors_precomp"].grad) <= GRAD_TOL
    assert rel_err(g["opacity"], leaves["opacities"].grad) <= GRAD_TOL
    assert rel_err(g["scales"], leaves["scales"].grad) <= GRAD_TOL
    assert rel_err(g["rotations"], leaves["rotations"].grad) <= GRAD_TOL
    assert rel_err(g["means2D"][0], O.means2D_grad_ndc(out["pre"])) <= GRAD_TOL
    # feature-only cotangent: the colour kernels contribute nothing, the geometry gradient is the feature term alone
    g0 = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                           None, None, None, None, None, None, semantic_feature=feat.cuda(), dL_dfeature=gf.cuda()[None])
    assert float(g0["colors"].abs().max()) == 0.0
    assert float(g0["means3D"].abs().max()) > 0.0

    # batched views == single views (forward exactly; feature gradients to summation order)
    sc2 = scenes.make_scene_c2(20_000, seed=4)
    cams = scenes.cameras_c2(3, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c, bg=bg)[0] for c in cams]
    rv2 = _rv(sc2)
    f2 = torch.rand(rv2["means3D"].shape[0], 32, generator=gen).cuda()
    *_, stb = _fwd(sets, rv2, semantic_feature=f2)
    gfs = torch.randn(3, 32, 180, 320, generator=gen).cuda()
    gb = RZ.backward_views(stb, rv2["means3D"], rv2["opacities"], rv2["colors_precomp"], rv2["scales"], rv2["rotations"],
                           None, None, None, None, None, None, None, semantic_feature=f2, dL_dfeature=gfs)
    acc_f = torch.zeros_like(gb["features"])
    acc_m = torch.zeros_like(gb["means3D"])
    for v, s in enumerate(sets):
        *_, s1 = _fwd([s], rv2, semantic_feature=f2)
        assert torch.equal(s1.feature_map[0], stb.feature_map[v])
        g1 = RZ.backward_views(s1, rv2["means3D"], rv2["opacities"], rv2["colors_precomp"], rv2["scales"],
                               rv2["rotations"], None, None, None, None, None, None, None, semantic_feature=f2,
                               dL_dfeature=gfs[v:v + 1])
        acc_f += g1["features"]
        acc_m += g1["means3D"]
    assert rel_err(gb["features"], acc_f) <= GRAD_TOL
    assert rel_err(gb["means3D"], acc_m) <= GRAD_TOL


def test_drop_in_module_semantic_feature_autograd():
    """GaussianRasterizer(...)(semantic_feature=[P,1,32]) returns the 5-tuple of the vendored variant
    (R/diff_gaussian_rasterization/__init__.py:108) and routes dL/dfeature_map to gradient slot 5 (:166)."""
    sc = scenes.make_scene_c1(3_000)
    st_, _ = settings_from_spec(scenes.camera_c1(), bg=(0.0, 0.0, 0.0))
    rv = {k: v.requires_grad_(True) for k, v in _rv(sc).items()}
    P = rv["means3D"].shape[0]
    feat = torch.rand(P, 1, 32, device="cuda", requires_grad=True)
    rast = RZ.GaussianRasterizer(st_, returns=5)
    means2D = torch.zeros_like(rv["means3D"], requires_grad=True)
    color, radii, fmap, depth, alpha = rast(means3D=rv["means3D"], means2D=means2D, opacities=rv["opacities"],
                                            semantic_feature=feat, colors_precomp=rv["colors_precomp"],
                                            scales=rv["scales"], rotations=rv["rotations"])
    assert fmap.shape == (32, st_.image_height, st_.image_width)
    (fmap.square().sum() + color.sum()).backward()
    assert feat.grad.shape == feat.shape and float(feat.grad.abs().max()) > 0
    assert torch.isfinite(rv["means3D"].grad).all() and torch.isfinite(means2D.grad).all()
    # no semantic_feature: empty feature map, as before
    out = rast(means3D=rv["means3D"], means2D=means2D, opacities=rv["opacities"], colors_precomp=rv["colors_precomp"],
               scales=rv["scales"], rotations=rv["rotations"])
    assert out[2].numel() == 0


def test_c5_stress_size_properties():
    """BASELINE.json configs[4]: 2 M Gaussians at 1920x1080 (two of the eight views, to bound the test's time).  Far
    beyond what the oracle can run, so the checks are size-independent properties: the tile ranges partition the
    instance list, keys are sorted inside every tile, a batch equals its views rendered alone bit for bit, the
    gradient is finite and linear in the cotangent."""
    sc = scenes.make_scene_c2(2_000_000, seed=5, scale=2.0)
    cams = scenes.cameras_c2(2, w=1920, h=1080, f=1400.0, radius=6.0)
    sets = [settings_from_spec(c)[0] for c in cams]
    rv = _rv(sc)
    color, _, depth, alpha, radii, st = _fwd(sets, rv)
    n, overflowed = RZ.read_counters(st)
    assert not overflowed and n == st.num_rendered and n > 10_000_000
    total = 0
    for v in range(2):
        keys, plist, ranges, ncon = export_view(st, v)
        r = ranges.long()
        lens = r[:, 1] - r[:, 0]
        assert int(lens.min()) >= 0
        nz = r[lens > 0]
        assert int(nz[0, 0]) == 0 and torch.equal(nz[1:, 0], nz[:-1, 1])         # ranges tile [0, R_v) without gaps
        assert keys.numel() == int(nz[-1, 1])
        assert bool((keys[1:] >= keys[:-1]).all())                               # (tile, depth) order
        tile_of = keys >> 32
        assert torch.equal(tile_of[nz[:, 0]], torch.nonzero(lens > 0).squeeze(1))
        assert int(ncon.max()) <= int(lens.max())
        assert int(plist.min()) >= 0 and int(plist.max()) < 2_000_000
        total += keys.numel()
    assert total == n
    c1, _, d1, a1, r1, st1 = _fwd([sets[1]], rv)
    assert torch.equal(color[1], c1[0]) and torch.equal(depth[1], d1[0]) and torch.equal(radii[1], r1[0])
    del st1, c1, d1, a1
    gen = torch.Generator().manual_seed(6)
    gc = torch.randn(2, 3, 1080, 1920, generator=gen).cuda()
    g1 = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                           None, None, gc, None, None, None, want_means2D=False)
    g2 = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                           None, None, 2.0 * gc, None, None, None, want_means2D=False)
    for k in ("means3D", "colors", "opacity", "scales", "rotations"):
        assert torch.isfinite(g1[k]).all(), k
        assert rel_err(g2[k], 2.0 * g1[k]) <= 1e-5, k
    assert float(g1["means3D"].abs().max()) > 0
