"""Workload for compute-sanitizer (tests/gpu_sanitize.sh): a small scene through EVERY kernel of the hot path -- several
views in one batch (deep tile lists: the blend rings wrap many times), fused extra colours, semantic features, forward and
backward, the kNN and the fused dynamic losses -- sized so that racecheck (10-100x slowdown) finishes in a minute.
    python tests/gpu_sanitize_case.py [P] [views]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import rasterizer as RZ, scenes  # noqa: E402
from dynamic3dgaussians_b200.camera import settings_from_spec  # noqa: E402


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 12_000
    V = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    dev = torch.device("cuda", 0)
    sc = scenes.make_scene_c2(P, seed=3)
    cams = scenes.cameras_c2(V, w=320, h=180, f=233.0)
    sets = [settings_from_spec(c, bg=(0.1, 0.2, 0.3), device=dev) for c in cams]
    rv = {k: v.to(dev).contiguous() for k, v in sc.rendervar().items()}
    seg = sc.seg_colors.to(dev).contiguous()
    args = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    g = torch.Generator().manual_seed(0)
    gc = torch.randn(V, 3, 180, 320, generator=g).to(dev)
    gd = torch.randn(V, 1, 180, 320, generator=g).to(dev) * 0.1
    # colour only, then colour + extra (both blend templates, both ILP variants of the forward)
    for ilp in (1, 2):
        RZ._lib.set_option("blend_ilp", ilp)
        for extra in (None, seg):
            *_, st = RZ.forward_views(sets, *args, extra_colors=extra)
            RZ.backward_views(st, *args, None, None, extra, gc, None if extra is None else gc, gd, gd,
                              want_means2D_color=extra is not None)
    RZ._lib.set_option("blend_ilp", 0)
    n = RZ.read_flags(st)["num_rendered"]
    # semantic features (one view)
    feat = torch.rand(P, 32, generator=g).to(dev)
    *_, st = RZ.forward_views(sets[:1], *args, semantic_feature=feat)
    RZ.backward_views(st, *args, None, None, None, gc[:1], None, gd[:1], None, semantic_feature=feat,
                      dL_dfeature=torch.randn(1, 32, 180, 320, generator=g).to(dev))
    # SH colours + cov3D is covered by pytest; kNN + dynamic losses
    from dynamic3dgaussians_b200.dyn_losses import build_neighbor_graph, rigidity_losses
    fg = rv["means3D"][: P // 2].contiguous().requires_grad_(True)
    rot = rv["rotations"][: P // 2].contiguous().requires_grad_(True)
    graph = build_neighbor_graph(fg.detach(), rot.detach(), 20)
    rigidity_losses(fg + 0.01, rot, graph).sum().backward()
    torch.cuda.synchronize()
    print(f"ok: {V} views, {P} Gaussians, {n} instances in the last batch")


if __name__ == "__main__":
    main()
