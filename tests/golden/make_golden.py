"""Generates tests/golden/small_scene.npz from the CPU oracle (oracle/eager_rasterizer.py).

The reference ships no fixtures for this path and its implementation is CUDA-only, so it cannot be run
in the GPU-less authoring container; the golden file therefore freezes the ORACLE's outputs on a seeded scene.  The
oracle itself is pinned against the reference CUDA build on the GPU box (tests/test_gpu_parity.py::test_vs_reference_*),
where radii, sorted lists, tile ranges, n_contrib and the images were found bit-identical.
Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from dynamic3dgaussians_b200 import scenes  # noqa: E402
from oracle import eager_rasterizer as O  # noqa: E402

P, W, H = 1500, 128, 96


def build_inputs():
    sc = scenes.make_scene_c1(P, seed=7)
    k = [[140.0, 0.0, W / 2 + 2.5], [0.0, 140.0, H / 2 - 1.75], [0.0, 0.0, 1.0]]
    cs = scenes.CameraSpec(W, H, k, scenes.look_at_w2c((0.5, -0.4, -2.6), (0.0, 0.0, 0.0)))
    gen = torch.Generator().manual_seed(11)
    cot = dict(color=torch.randn(3, H, W, generator=gen), extra=torch.randn(3, H, W, generator=gen),
               depth=torch.randn(1, H, W, generator=gen), alpha=torch.randn(1, H, W, generator=gen))
    return sc, cs, cot


def run_oracle():
    sc, cs, cot = build_inputs()
    cam = O.camera_from_spec(cs.w, cs.h, cs.k, cs.w2c, bg=(0.15, 0.25, 0.35))
    rv = sc.rendervar()
    leaves = {k: v.clone().requires_grad_(True) for k, v in rv.items()}
    seg = sc.seg_colors.clone().requires_grad_(True)
    out = O.rasterize(cam, leaves["means3D"], leaves["opacities"], colors_precomp=leaves["colors_precomp"],
                      scales=leaves["scales"], rotations=leaves["rotations"], extra_colors=seg, collect_fragile=True)
    ok = (~out["fragile"]).float()
    loss = (out["color"][:3] * cot["color"] * ok).sum() + (out["color"][3:] * cot["extra"] * ok).sum() + \
           (out["depth"] * cot["depth"] * ok).sum() + (out["alpha"] * cot["alpha"] * ok).sum()
    loss.backward()
    g = {k: v.grad for k, v in leaves.items()}
    g["seg"] = seg.grad
    g["means2D"] = O.means2D_grad_ndc(out["pre"])
    return out, g


if __name__ == "__main__":
    out, g = run_oracle()
    f32 = lambda t: t.detach().numpy().astype(np.float32)
    np.savez_compressed(
        os.path.join(os.path.dirname(os.path.abspath(__file__)), "small_scene.npz"),
        radii=out["pre"]["radii"].numpy().astype(np.int32),
        tiles_touched=out["pre"]["tiles_touched"].numpy().astype(np.int32),
        point_list=out["lists"]["point_list"].astype(np.int32),
        keys_sorted=out["lists"]["keys_sorted"].astype(np.uint64),
        ranges=out["lists"]["ranges"].astype(np.int32),
        n_contrib=out["n_contrib"].numpy().astype(np.int32),
        fragile=out["fragile"].numpy(),
        color=f32(out["color"]), depth=f32(out["depth"]), alpha=f32(out["alpha"]),
        **{"g_" + k: f32(v) for k, v in g.items()})
    print("num_rendered", out["lists"]["num_rendered"], "fragile", int(out["fragile"].sum()))
