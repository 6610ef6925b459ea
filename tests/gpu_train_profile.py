"""Where one training iteration at C2 spends its GPU time (diagnostic, not a test): python tests/gpu_train_profile.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import _lib, scenes, train_loop as TL  # noqa: E402
from dynamic3dgaussians_b200 import rasterizer as RZ  # noqa: E402
import bench  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    sc = scenes.make_scene_c2(300_000)
    cams = scenes.cameras_c2(27)
    sets = [bench.make_settings(c, dev) for c in cams] if hasattr(bench, "make_settings") else None
    if sets is None:
        class A:  # noqa
            gaussians, views, workload = 300_000, 27, "c2"
        sc, sets = bench.build_workload(A, dev)
    rv = {k: v.to(dev).contiguous() for k, v in sc.rendervar().items()}
    seg = sc.seg_colors.to(dev)[:, 0].contiguous()
    g = torch.Generator().manual_seed(0)
    params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, 27, 3.3)
    opt = TL.initialize_optimizer(params, variables)
    data = [{'cam': s_, 'id': i, 'im': torch.rand(3, 360, 640, generator=g).to(dev),
             'seg': torch.rand(3, 360, 640, generator=g).to(dev)} for i, s_ in enumerate(sets)]
    cap = 5_000_000
    TL.train_timestep(params, variables, opt, data, 10, True, generator=g, do_densify=False, capacity=cap)
    torch.cuda.synchronize()
    _lib.stage_timing(True)
    acc = {}
    tot = 0.0
    n = 20
    for i in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        TL.train_timestep(params, variables, opt, [data[i % 27]], 1, True, generator=g, do_densify=False, capacity=cap)
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
        for k, v in _lib.read_stage_timing().items():
            acc[k] = acc.get(k, 0.0) + v / n
    print("ms per eager iteration (device span):", round(tot / n, 3))
    print("rasterizer stages:", {k: round(v, 4) for k, v in acc.items()}, "sum", round(sum(acc.values()), 3))
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        TL.train_timestep(params, variables, opt, data, 5, True, generator=g, do_densify=False, capacity=cap)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))


if __name__ == "__main__":
    main()
