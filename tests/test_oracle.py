"""CPU tests of the oracle: golden vectors, analytic known answers, finite differences, invariants (no GPU needed)."""
import math
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import GOLDEN, O  # noqa: E402
from dynamic3dgaussians_b200 import scenes  # noqa: E402


def _cam(w=64, h=48, f=80.0, eye=(0.0, 0.0, -3.0), bg=(0.0, 0.0, 0.0), cx=None, cy=None):
    k = [[f, 0.0, w / 2 if cx is None else cx], [0.0, f, h / 2 if cy is None else cy], [0.0, 0.0, 1.0]]
    return O.camera_from_spec(w, h, k, scenes.look_at_w2c(eye, (0.0, 0.0, 0.0)), bg=bg)


def _one(mean=(0.0, 0.0, 0.0), scale=0.05, opacity=0.7, rgb=(0.2, 0.5, 0.9)):
    return dict(means3D=torch.tensor([mean], dtype=torch.float32), opacities=torch.tensor([[opacity]]),
                colors_precomp=torch.tensor([rgb]), scales=torch.full((1, 3), scale),
                rotations=torch.tensor([[1.0, 0.0, 0.0, 0.0]]))


def test_get_higher_msb_matches_reference_values():
    # rasterizer_impl.cu:35-50 ; SURVEY.md 8: 920 -> 10, 8160 -> 13, 256 -> 9
    assert O.get_higher_msb(920) == 10
    assert O.get_higher_msb(8160) == 13
    assert O.get_higher_msb(256) == 9
    assert O.get_higher_msb(1) == 1


def test_golden_vectors_reproduce():
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden
    out, g = make_golden.run_oracle()
    z = np.load(GOLDEN)
    assert np.array_equal(out["pre"]["radii"].numpy(), z["radii"])
    assert np.array_equal(out["lists"]["point_list"].astype(np.int32), z["point_list"])
    assert np.array_equal(out["lists"]["keys_sorted"], z["keys_sorted"])
    assert np.array_equal(out["lists"]["ranges"].astype(np.int32), z["ranges"])
    assert np.array_equal(out["n_contrib"].numpy().astype(np.int32), z["n_contrib"])
    assert np.abs(out["color"].detach().numpy() - z["color"]).max() <= 1e-6
    assert np.abs(out["depth"].detach().numpy() - z["depth"]).max() <= 1e-5
    for k in ("means3D", "opacities", "scales", "rotations", "colors_precomp", "seg"):
        ref = z["g_" + k]
        assert np.abs(g[k].numpy() - ref).max() <= 1e-4 * np.abs(ref).max() + 1e-7, k


def test_single_isotropic_gaussian_known_answer():
    """SURVEY.md 8c analytic case: centre pixel alpha = min(0.99, o); colour = c*alpha + bg*(1-alpha);
    radius = ceil(3*sqrt(sigma_px^2 + 0.3))."""
    cam = _cam(bg=(0.1, 0.2, 0.3))
    s = _one()
    out = O.rasterize(cam, **s)
    sigma_px = 0.05 * 80.0 / 3.0
    assert int(out["pre"]["radii"][0]) == math.ceil(3.0 * math.sqrt(sigma_px ** 2 + 0.3))
    px, py = out["pre"]["means2D"][0]
    assert abs(float(px) - 31.5) < 1e-4 and abs(float(py) - 23.5) < 1e-4     # ((ndc+1)*S-1)/2 with ndc = 0
    # nearest pixel centre is 0.5 px away in x and y
    cov = sigma_px ** 2 + 0.3
    a = 0.7 * math.exp(-0.5 * (0.25 + 0.25) / cov)
    col = out["color"][:, 24, 32]
    for ch, c in enumerate((0.2, 0.5, 0.9)):
        assert abs(float(col[ch]) - (c * a + (0.1, 0.2, 0.3)[ch] * (1 - a))) < 1e-5
    assert abs(float(out["alpha"][0, 24, 32]) - a) < 1e-6
    assert abs(float(out["depth"][0, 24, 32]) - 3.0 * a) < 1e-5
    assert int(out["n_contrib"][24, 32]) == 1


def test_opacity_one_hits_the_099_clamp():
    cam = _cam(w=33, h=33, f=40.0, cx=16.5, cy=16.5)   # pixel = c - 0.5: the mean lands exactly on pixel (16,16)
    out = O.rasterize(cam, **_one(opacity=1.0, scale=0.3))
    assert abs(float(out["alpha"][0, 16, 16]) - 0.99) < 1e-6


def test_depth_order_swap_and_tie_break():
    cam = _cam()
    two = dict(means3D=torch.tensor([[0.0, 0.0, 0.2], [0.0, 0.0, -0.2]]), opacities=torch.tensor([[0.6], [0.6]]),
               colors_precomp=torch.tensor([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0]]), scales=torch.full((2, 3), 0.05),
               rotations=torch.tensor([[1.0, 0, 0, 0], [1.0, 0, 0, 0]]))
    out = O.rasterize(cam, **two)
    t = (24 // 16) * 4 + (32 // 16)
    r = out["lists"]["ranges"][t]
    assert list(out["lists"]["point_list"][r[0]:r[1]]) == [1, 0]          # the nearer one (z=-0.2) first
    assert float(out["color"][1, 24, 32]) > float(out["color"][0, 24, 32])
    same = dict(two, means3D=torch.zeros(2, 3))
    out = O.rasterize(cam, **same)
    r = out["lists"]["ranges"][t]
    assert list(out["lists"]["point_list"][r[0]:r[1]]) == [0, 1]          # equal depth: ascending index (stable)


def test_tile_corner_touches_four_tiles_and_culls():
    cam = _cam(w=64, h=64, f=80.0)
    # project onto the tile corner (32,32): pixel = 31.5 + 80*x/3 -> x = 0.5*3/80
    x = 0.5 * 3.0 / 80.0
    out = O.rasterize(cam, **_one(mean=(x, x, 0.0), scale=0.02))
    assert int(out["pre"]["tiles_touched"][0]) == 4
    behind = O.rasterize(cam, **_one(mean=(0.0, 0.0, -5.0)))                # z_view <= 0 -> culled
    assert int(behind["pre"]["radii"][0]) == 0 and behind["lists"]["num_rendered"] == 0
    assert float(behind["color"].abs().max()) == 0.0


def test_termination_with_opaque_stack():
    cam = _cam()
    n = 12
    s = dict(means3D=torch.stack([torch.tensor([0.0, 0.0, 0.01 * i]) for i in range(n)]),
             opacities=torch.full((n, 1), 0.95), colors_precomp=torch.rand(n, 3, generator=torch.Generator().manual_seed(0)),
             scales=torch.full((n, 3), 0.2), rotations=torch.tensor([[1.0, 0, 0, 0]] * n))
    out = O.rasterize(cam, **s)
    nc = int(out["n_contrib"][24, 32])
    # T after k blends = 0.05^k (approx.): 0.05^3 = 1.25e-4 >= 1e-4, 0.05^4 < 1e-4 -> 3 contributors
    assert nc == 3
    assert float(out["final_T"][24, 32]) >= 1e-4
    assert abs(float(out["alpha"][0, 24, 32]) - (1.0 - float(out["final_T"][24, 32]))) < 1e-6


def test_invariants_on_random_scene():
    sc = scenes.make_scene_c1(400, seed=3)
    cs = scenes.camera_c1()
    cam = O.camera_from_spec(96, 80, [[100.0, 0, 50.0], [0, 100.0, 38.0], [0, 0, 1]], cs.w2c, bg=(0.3, 0.3, 0.3))
    rv = sc.rendervar()
    out = O.rasterize(cam, rv["means3D"], rv["opacities"], colors_precomp=rv["colors_precomp"], scales=rv["scales"],
                      rotations=rv["rotations"])
    assert float((out["alpha"][0] - (1 - out["final_T"])).abs().max()) < 2e-6     # alpha = 1 - T_final
    assert float(out["color"].max()) <= 1.0 + 1e-5 and float(out["color"].min()) >= 0.0
    keys = out["lists"]["keys_sorted"]
    assert np.all(keys[1:] >= keys[:-1])                                          # sortedness
    rg = out["lists"]["ranges"]
    assert int((rg[:, 1] - rg[:, 0]).sum()) == out["lists"]["num_rendered"]
    # permutation of the inputs permutes radii and leaves the image unchanged (no ties in this scene)
    perm = torch.randperm(400, generator=torch.Generator().manual_seed(5))
    out2 = O.rasterize(cam, rv["means3D"][perm], rv["opacities"][perm], colors_precomp=rv["colors_precomp"][perm],
                       scales=rv["scales"][perm], rotations=rv["rotations"][perm])
    assert torch.equal(out2["pre"]["radii"], out["pre"]["radii"][perm])
    assert torch.equal(out2["color"], out["color"])


def test_gradients_match_float64_finite_differences():
    torch.manual_seed(0)
    n = 6
    cam = _cam(w=32, h=32, f=40.0, bg=(0.2, 0.1, 0.4))
    base = dict(means3D=(torch.rand(n, 3, dtype=torch.float64) - 0.5) * 0.6,
                opacities=torch.rand(n, 1, dtype=torch.float64) * 0.6 + 0.2,
                colors_precomp=torch.rand(n, 3, dtype=torch.float64),
                scales=torch.rand(n, 3, dtype=torch.float64) * 0.15 + 0.1,
                rotations=torch.nn.functional.normalize(torch.randn(n, 4, dtype=torch.float64)))
    wc = torch.randn(3, 32, 32, dtype=torch.float64)
    wd = torch.randn(1, 32, 32, dtype=torch.float64)
    wa = torch.randn(1, 32, 32, dtype=torch.float64)

    def loss_of(inp):
        o = O.rasterize(cam, inp["means3D"], inp["opacities"], colors_precomp=inp["colors_precomp"],
                        scales=inp["scales"], rotations=inp["rotations"], dtype=torch.float64)
        return (o["color"] * wc).sum() + (o["depth"] * wd).sum() + (o["alpha"] * wa).sum()

    leaves = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    loss_of(leaves).backward()
    rng = np.random.default_rng(0)
    eps = 1e-6
    for name, t in base.items():
        for _ in range(4):
            idx = tuple(int(rng.integers(0, s)) for s in t.shape)
            plus = {k: v.clone() for k, v in base.items()}
            minus = {k: v.clone() for k, v in base.items()}
            plus[name][idx] += eps
            minus[name][idx] -= eps
            fd = float(loss_of(plus) - loss_of(minus)) / (2 * eps)
            an = float(leaves[name].grad[idx])
            assert abs(fd - an) <= 2e-4 * max(1.0, abs(fd)), (name, idx, fd, an)


def test_sh_forward_degree3_matches_direct_evaluation():
    torch.manual_seed(1)
    P = 50
    A = O._Arith(torch.float32)
    means = torch.randn(P, 3)
    campos = torch.tensor([0.1, -0.2, -3.0])
    shs = torch.randn(P, 16, 3) * 0.3
    rgb, clamped = O.eval_sh(A, 3, means, campos, shs)
    d = torch.nn.functional.normalize((means - campos).double(), dim=1)
    x, y, z = d[:, 0:1], d[:, 1:2], d[:, 2:3]
    sh = shs.double()
    C2, C3 = O.SH_C2, O.SH_C3
    res = O.SH_C0 * sh[:, 0] - O.SH_C1 * y * sh[:, 1] + O.SH_C1 * z * sh[:, 2] - O.SH_C1 * x * sh[:, 3]
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    res = res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6] + \
        C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8]
    res = res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10] + \
        C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12] + \
        C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14] + C3[6] * x * (xx - 3 * yy) * sh[:, 15]
    ref = torch.clamp(res + 0.5, min=0.0)
    assert float((rgb.double() - ref).abs().max()) < 1e-5
    assert torch.equal(clamped, (res + 0.5) < 0)
