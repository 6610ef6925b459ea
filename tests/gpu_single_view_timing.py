"""Stage times of a single C2 view, fused colour + seg forward + backward (diagnostic): python tests/gpu_single_view_timing.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import _lib, rasterizer as RZ  # noqa: E402
import bench  # noqa: E402


def main():
    dev = torch.device("cuda", 0)

    class A:  # noqa
        gaussians, views, workload = 300_000, 27, "c2"
    sc, sets = bench.build_workload(A, dev)
    rv = {k: v.to(dev).contiguous().requires_grad_(True) for k, v in sc.rendervar().items()}
    seg = sc.seg_colors.to(dev).contiguous()
    g = torch.Generator().manual_seed(0)
    wc = torch.randn(3, 360, 640, generator=g).to(dev)
    _lib.stage_timing(True)
    acc = {}
    views = [int(a) for a in sys.argv[1:]] or [0, 5, 10, 15, 20, 25]
    reps = 3
    for v in views:
        for it in range(reps + 1):
            color, extra, depth, alpha, radii = RZ.rasterize_views(
                [sets[v]], rv["means3D"], rv["opacities"], colors_precomp=rv["colors_precomp"], scales=rv["scales"],
                rotations=rv["rotations"], extra_colors=seg, capacity=5_000_000)
            ((color[0] + extra[0]) * wc).sum().backward()
            torch.cuda.synchronize()
            if it == 0:
                continue
            for k, t in _lib.read_stage_timing().items():
                acc[k] = acc.get(k, 0.0) + t / (reps * len(views))
    print({k: round(t * 1e3, 1) for k, t in acc.items()}, "us; sum", round(sum(acc.values()) * 1e3, 1))


if __name__ == "__main__":
    main()
