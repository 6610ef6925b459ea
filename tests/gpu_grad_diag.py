"""Diagnostic (not a test): per-gradient error statistics of the product against (a) autograd through the oracle at C1
and (b) the patched reference backward (oracle/_ref/libref_rasterizer_patched.so) at C1 and at headline size.  Prints
max-norm relative error, the share of elements outside |a-b| <= 1e-4|b| + 1e-6 max|b|, and error quantiles.
    python tests/gpu_grad_diag.py [P_c2]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import O, RZ, elem_violations, final_T, rel_err, settings_from_spec  # noqa: E402
from dynamic3dgaussians_b200 import scenes  # noqa: E402
from oracle import ref_cuda  # noqa: E402


def stats(name, a, b):
    a = a.detach().double().cpu().reshape(-1)
    b = b.detach().double().cpu().reshape(-1)
    err = (a - b).abs()
    rel = err / (b.abs() + 1e-6 * b.abs().max() / 1e-4 + 1e-300)      # error in units of the criterion's allowance / 1e-4
    bad, n = elem_violations(a, b)
    q = torch.quantile(rel[torch.randperm(rel.numel())[:2_000_000]], torch.tensor([0.5, 0.99, 0.9999], dtype=torch.float64))
    print(f"  {name:12s} maxnorm {rel_err(a, b):.2e}  elem-viol {bad}/{n} ({bad / max(n, 1):.2e})  "
          f"q50 {q[0]:.1e} q99 {q[1]:.1e} q99.99 {q[2]:.1e} max {float(rel.max()):.1e}")


def vs_oracle():
    print("C1 10k vs autograd through the oracle (float32) and (float64)")
    sc = scenes.make_scene_c1(10_000)
    cs = scenes.camera_c1()
    st_, oc = settings_from_spec(cs, bg=(0.1, 0.2, 0.3))
    rv = {k: v.cuda() for k, v in sc.rendervar().items()}
    color, _, depth, alpha, radii, st = RZ.forward_views([st_], rv["means3D"], rv["opacities"], rv["colors_precomp"],
                                                         rv["scales"], rv["rotations"])
    for dt in (torch.float32, torch.float64):
        leaves = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in rv.items()}
        out = O.rasterize(oc, leaves["means3D"], leaves["opacities"], colors_precomp=leaves["colors_precomp"],
                          scales=leaves["scales"], rotations=leaves["rotations"], collect_fragile=True, dtype=dt)
        ok = ~out["fragile"]
        gen = torch.Generator().manual_seed(1)
        okf = ok.float()
        gc, gd, ga = [torch.randn(c, cs.h, cs.w, generator=gen) * okf for c in (3, 1, 1)]
        ((out["color"] * gc).sum() + (out["depth"] * gd).sum() + (out["alpha"] * ga).sum()).backward()
        g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"],
                              None, None, None, gc.cuda()[None], None, gd.cuda()[None], ga.cuda()[None])
        print(" oracle dtype", dt)
        for mine, theirs in (("means3D", "means3D"), ("colors", "colors_precomp"), ("opacity", "opacities"),
                             ("scales", "scales"), ("rotations", "rotations")):
            stats(mine, g[mine], leaves[theirs].grad)
        stats("means2D", g["means2D"][0], O.means2D_grad_ndc(out["pre"]))


def vs_patched(P, cams, sc, label):
    print(label)
    rv = {k: v.cuda() for k, v in sc.rendervar().items()}
    sets = [settings_from_spec(c, bg=(0.2, 0.3, 0.1))[0] for c in cams]
    V = len(sets)
    H, W = sets[0].image_height, sets[0].image_width
    color, _, depth, alpha, radii, st = RZ.forward_views(sets, rv["means3D"], rv["opacities"], rv["colors_precomp"],
                                                         rv["scales"], rv["rotations"])
    gen = torch.Generator().manual_seed(11)
    gc = torch.randn(V, 3, H, W, generator=gen).cuda()
    gd = (torch.randn(V, 1, H, W, generator=gen) * 0.1).cuda()
    ga = torch.randn(V, 1, H, W, generator=gen).cuda()
    g = RZ.backward_views(st, rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], None,
                          None, None, gc, None, gd, ga)
    fT = final_T(st)
    for direct in (False, True):
        print(" T_final route:", "handed over (alphas = -final_T)" if direct else "1 - alpha_out (as the reference does)")
        acc = None
        for v, s in enumerate(sets):
            rr = ref_cuda.RefRasterizer(s, patched=True)
            ro = rr.forward(rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
            print("  view", v, "color equal", torch.equal(ro["color"], color[v]), "alpha max diff",
                  float((ro["alpha"] - alpha[v]).abs().max()), "| (1 - final_T) - alpha | max",
                  float(((1.0 - fT[v]) - ro["alpha"][0]).abs().max()))
            rg = rr.backward(gc[v], gd[v], ga[v], alphas=(-fT[v]).contiguous() if direct else None)
            stats(f"means2D[{v}]", g["means2D"][v], rg["means2D"])
            part = {k: rg[k] for k in ("means3D", "opacity", "colors", "scales", "rotations")}
            acc = part if acc is None else {k: acc[k] + part[k] for k in acc}
        for k in acc:
            stats(k, g[k], acc[k].reshape(g[k].shape))


if __name__ == "__main__":
    vs_oracle()
    if ref_cuda.patched_available():
        vs_patched(10_000, [scenes.camera_c1()], scenes.make_scene_c1(10_000), "C1 10k vs patched reference backward")
        P = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
        vs_patched(P, scenes.cameras_c2(27)[4:4 + 3 * 9:9], scenes.make_scene_c2(P),
                   f"C2 {P} x 3 views vs patched reference backward")
