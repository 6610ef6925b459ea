"""Per-CTA timeline of the blend kernels for one C2 view (diagnostic): which SM ran which tile, from when to when.
python tests/gpu_tile_trace.py [view | -n_views]
With a library built with -DGSR_TRACE_COUNTS (tests/build_variant.sh counts -DGSR_TRACE_COUNTS; GSR_LIB_PATH=...) the
backward also reports how many work-list entries were walked and how many of them some pixel of the patch blended
(C2, 27 views: 12.8 M walked, 78.9 % blended -- the bounding-box filter lets 21 % through to the 32-instruction skip path)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import _lib, rasterizer as RZ  # noqa: E402
import bench  # noqa: E402
from helpers import export_view  # noqa: E402


def report(name, tr, lens, walked):
    tr = tr.cpu().long()
    t0 = int(tr[:, 0].min())
    start, end, sm, vt = (tr[:, 0] - t0) / 1e3, (tr[:, 1] - t0) / 1e3, tr[:, 2], tr[:, 3]
    dur = end - start
    print(f"{name}: kernel span {float(end.max()):.1f} us; CTA duration mean {float(dur.mean()):.1f} max {float(dur.max()):.1f}; "
          f"latest start {float(start.max()):.1f}")
    top = torch.argsort(end, descending=True)[:10]
    for i in top.tolist():
        t = int(vt[i])
        print(f"   block {i:4d} tile {t:4d} sm {int(sm[i]):3d} start {float(start[i]):6.1f} end {float(end[i]):6.1f} dur {float(dur[i]):6.1f} "
              f"list {int(lens[t])} walked {int(walked[t])}")
    nsm = int(sm.max()) + 1
    busy_end = torch.zeros(nsm)
    for s_ in range(nsm):
        sel = sm == s_
        if sel.any():
            busy_end[s_] = end[sel].max()
    print(f"   per-SM last end: mean {float(busy_end.mean()):.1f} min {float(busy_end.min()):.1f} max {float(busy_end.max()):.1f}")
    # duration vs walked depth for heavy tiles
    w = walked[vt].float().clamp(min=1)
    heavy = torch.argsort(dur, descending=True)[:20]
    print("   ns per walked position (20 longest CTAs):", [round(float(dur[i] * 1e3 / w[i])) for i in heavy.tolist()])


def full(nv):
    """All nv views in one launch (the headline step): how long is the tail?"""
    dev = torch.device("cuda", 0)

    class A:  # noqa
        gaussians, views, workload = 300_000, 27, "c2"
    sc, sets = bench.build_workload(A, dev)
    rv = {k: t.to(dev).contiguous().requires_grad_(True) for k, t in sc.rendervar().items()}
    g = torch.Generator().manual_seed(0)
    wc = torch.randn(nv, 3, 360, 640, generator=g).to(dev)
    wd = torch.randn(nv, 1, 360, 640, generator=g).to(dev)
    VT = 23 * 40 * nv
    tf = torch.zeros(VT, 4, dtype=torch.int64, device=dev)
    tb = torch.zeros(VT, 4, dtype=torch.int64, device=dev)
    lib = _lib.load()
    for it in range(3):
        if it == 2:
            torch.cuda.synchronize()
            _lib.check(lib.gsr_debug_blend_trace(tf.data_ptr(), tb.data_ptr()), "trace")
        color, extra, depth, alpha, radii = RZ.rasterize_views(
            sets[:nv], rv["means3D"], rv["opacities"], colors_precomp=rv["colors_precomp"], scales=rv["scales"],
            rotations=rv["rotations"], capacity=48_000_000)
        ((color * wc).sum() + (depth * wd).sum()).backward()
    torch.cuda.synchronize()
    _lib.check(lib.gsr_debug_blend_trace(None, None), "trace off")
    for name, tr in (("forward", tf), ("backward", tb)):
        tr = tr.cpu().long()
        if name == "backward" and int((tr[:, 2] >> 16).sum()) > 0:     # a -DGSR_TRACE_COUNTS build
            n_list, n_done = int((tr[:, 2] >> 16).sum()), int((tr[:, 3] >> 32).sum())
            print(f"backward work lists: {n_list} entries walked, {n_done} blended by some pixel of the patch "
                  f"({100.0 * n_done / max(n_list, 1):.1f} %)")
            tr[:, 2] &= 0xffff
            tr[:, 3] &= 0xffffffff
        t0 = int(tr[:, 0].min())
        start, end, sm = (tr[:, 0] - t0) / 1e3, (tr[:, 1] - t0) / 1e3, tr[:, 2]
        dur = end - start
        nsm = int(sm.max()) + 1
        last = torch.tensor([float(end[sm == s_].max()) for s_ in range(nsm)])
        busy = torch.tensor([float(dur[sm == s_].sum()) for s_ in range(nsm)])
        print(f"{name}: {VT} CTAs, span {float(end.max()):.1f} us; per-SM last end mean {float(last.mean()):.1f} min {float(last.min()):.1f} "
              f"max {float(last.max()):.1f}; latest CTA start {float(start.max()):.1f}; CTA dur mean {float(dur.mean()):.1f} max {float(dur.max()):.1f}")
        top = torch.argsort(dur, descending=True)[:12]
        print("   longest CTAs (block, start, dur):", [(int(i), round(float(start[i])), round(float(dur[i]))) for i in top.tolist()])
        late = torch.argsort(end, descending=True)[:12]
        print("   last to finish (block, start, dur):", [(int(i), round(float(start[i])), round(float(dur[i]))) for i in late.tolist()])
        # work finished before time t: how much of the span runs with < 50% of the SMs busy
        ts = torch.linspace(0, float(end.max()), 200)
        act = torch.tensor([float(((start <= t) & (end > t)).sum()) for t in ts])
        print("   resident CTAs at 50/80/90/95/98% of the span:", [int(act[int(f * 199)]) for f in (0.5, 0.8, 0.9, 0.95, 0.98)])


def main():
    v = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    if v < 0:
        return full(-v)
    dev = torch.device("cuda", 0)

    class A:  # noqa
        gaussians, views, workload = 300_000, 27, "c2"
    sc, sets = bench.build_workload(A, dev)
    rv = {k: t.to(dev).contiguous().requires_grad_(True) for k, t in sc.rendervar().items()}
    seg = sc.seg_colors.to(dev).contiguous()
    g = torch.Generator().manual_seed(0)
    wc = torch.randn(3, 360, 640, generator=g).to(dev)
    VT = 23 * 40
    tf = torch.zeros(VT, 4, dtype=torch.int64, device=dev)
    tb = torch.zeros(VT, 4, dtype=torch.int64, device=dev)
    lib = _lib.load()
    for it in range(3):
        if it == 2:
            torch.cuda.synchronize()
            _lib.check(lib.gsr_debug_blend_trace(tf.data_ptr(), tb.data_ptr()), "trace")
        color, extra, depth, alpha, radii = RZ.rasterize_views(
            [sets[v]], rv["means3D"], rv["opacities"], colors_precomp=rv["colors_precomp"], scales=rv["scales"],
            rotations=rv["rotations"], extra_colors=seg, capacity=5_000_000)
        ((color[0] + extra[0]) * wc).sum().backward()
    torch.cuda.synchronize()
    _lib.check(lib.gsr_debug_blend_trace(None, None), "trace off")
    with torch.no_grad():
        out = RZ.forward_views([sets[v]], rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])
    keys, plist, ranges, ncon = export_view(out[5], 0)
    lens = (ranges[:, 1] - ranges[:, 0]).long().cpu()
    H, W = ncon.shape
    ty, tx = (H + 15) // 16, (W + 15) // 16
    pad = torch.zeros(ty * 16, tx * 16, dtype=torch.int64, device=dev)
    pad[:H, :W] = ncon.long()
    walked = pad.view(ty, 16, tx, 16).permute(0, 2, 1, 3).reshape(ty * tx, 256).max(dim=1).values.cpu()
    report("forward", tf, lens, walked)
    report("backward", tb, lens, walked)


if __name__ == "__main__":
    main()
