"""One C2 view, fused colour + seg forward and backward, a few times (diagnostic for ncu captures of the small-launch
regime the train loop and the 8-GPU shares run in): python tests/gpu_single_view.py [view] [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import rasterizer as RZ  # noqa: E402
import bench  # noqa: E402


def main():
    view = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dev = torch.device("cuda", 0)

    class A:  # noqa
        gaussians, views, workload = 300_000, 27, "c2"
    sc, sets = bench.build_workload(A, dev)
    rv = {k: v.to(dev).contiguous().requires_grad_(True) for k, v in sc.rendervar().items()}
    seg = sc.seg_colors.to(dev).contiguous()
    g = torch.Generator().manual_seed(0)
    wc = torch.randn(3, 360, 640, generator=g).to(dev)
    for _ in range(reps):
        color, extra, depth, alpha, radii = RZ.rasterize_views(
            [sets[view]], rv["means3D"], rv["opacities"], colors_precomp=rv["colors_precomp"], scales=rv["scales"],
            rotations=rv["rotations"], extra_colors=seg, capacity=5_000_000)
        ((color[0] + extra[0]) * wc).sum().backward()
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
