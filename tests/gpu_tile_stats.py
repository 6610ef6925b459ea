"""Tile-list statistics of C2 views (diagnostic, not a test): python tests/gpu_tile_stats.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import rasterizer as RZ  # noqa: E402
import bench  # noqa: E402
from helpers import export_view  # noqa: E402


def main():
    dev = torch.device("cuda", 0)

    class A:  # noqa
        gaussians, views, workload = 300_000, 27, "c2"
    sc, sets = bench.build_workload(A, dev)
    rv = {k: v.to(dev).contiguous() for k, v in sc.rendervar().items()}
    for v in (0, 10, 20):
        out = RZ.forward_views([sets[v]], rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
        st = out[5]
        keys, plist, ranges, ncon = export_view(st, 0)
        lens = (ranges[:, 1] - ranges[:, 0]).long()
        H, W = ncon.shape
        ty, tx = (H + 15) // 16, (W + 15) // 16
        pad = torch.zeros(ty * 16, tx * 16, dtype=torch.int64, device=dev)
        pad[:H, :W] = ncon.long()
        t = pad.view(ty, 16, tx, 16).permute(0, 2, 1, 3).reshape(ty * tx, 256)
        tmax = t.max(dim=1).values
        # 8x8 patches
        pt = pad.view(ty, 2, 8, tx, 2, 8).permute(0, 3, 1, 4, 2, 5).reshape(ty * tx, 4, 64)
        pmax = pt.max(dim=2).values            # [tiles, 4]
        s = torch.sort(lens, descending=True).values
        print(f"view {v}: instances {int(lens.sum())}, tiles {lens.numel()}, list len mean {lens.float().mean():.0f} "
              f"top {s[:8].tolist()} p50 {int(s[len(s) // 2])}")
        sm = torch.sort(tmax, descending=True).values
        print(f"   depth walked per tile (max n_contrib): mean {tmax.float().mean():.0f} top {sm[:8].tolist()} p50 {int(sm[len(sm) // 2])}")
        print(f"   sum over tiles of walked depth {int(tmax.sum())}; per-SM share (/148) {int(tmax.sum()) // 148}; "
              f"max tile / per-SM share = {float(sm[0]) / (float(tmax.sum()) / 148):.2f}")
        print(f"   pixel mean n_contrib {ncon.float().mean():.0f}; patch max mean {pmax.float().mean():.0f}")


if __name__ == "__main__":
    main()
