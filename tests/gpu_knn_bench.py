"""GPU diagnostic (not a pytest): times the kNN / dynamic-loss kernels against the reference's own kernels / eager code."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynamic3dgaussians_b200 import dyn_losses as DL, knn as KN, scenes  # noqa: E402
from oracle import ref_cuda  # noqa: E402


def timeit(fn, n=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    for P in (300_000, 2_000_000):
        sc = scenes.make_scene_c2(P)
        pts = sc.means3D.cuda()
        t3 = timeit(lambda: KN.knn(pts, 3))
        t20 = timeit(lambda: KN.knn(pts, 20), n=3)
        line = f"P={P}: distCUDA2 (k=3) {t3:.3f} ms = {P / t3 / 1e3:.1f} Mpts/s; k=20 {t20:.3f} ms"
        if ref_cuda.available():
            t0 = time.time()
            tr = timeit(lambda: ref_cuda.knn(pts), n=3)
            line += f"; reference distCUDA2 {tr:.3f} ms ({tr / t3:.1f}x slower)"
        print(line, flush=True)
    # dynamic losses at C3 size: F = 180k foreground Gaussians, K = 20
    sc = scenes.make_scene_c2(300_000)
    fg = sc.seg_colors[:, 0] > 0.5
    pts = sc.means3D[fg].cuda()
    rot = torch.nn.functional.normalize(sc.unnorm_rotations[fg]).cuda()
    g = DL.build_neighbor_graph(pts, rot, 20)
    p2 = (pts + 0.01 * torch.randn_like(pts)).requires_grad_(True)
    r2 = torch.nn.functional.normalize(rot + 0.01 * torch.randn_like(rot)).requires_grad_(True)
    w = torch.tensor([4.0, 4.0, 2.0], device="cuda")

    def ours():
        p2.grad = None
        r2.grad = None
        (DL.rigidity_losses(p2, r2, g) * w).sum().backward()

    from oracle import eager_dyn_losses as OD
    idx = g.idx.long()

    def eager():
        p2.grad = None
        r2.grad = None
        rel = OD.quat_mult(r2, g.prev_inv_rot)
        q = rel / rel.norm(dim=1, keepdim=True)
        r, x, y, z = q.unbind(1)
        Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y), 2 * (x * y + r * z),
                          1 - 2 * (x * x + z * z), 2 * (y * z - r * x), 2 * (x * z - r * y), 2 * (y * z + r * x),
                          1 - 2 * (x * x + y * y)], dim=1).reshape(-1, 3, 3)
        off = p2[idx] - p2[:, None]
        c = (Rm.transpose(2, 1)[:, None] @ off[:, :, :, None]).squeeze(-1)
        L = torch.stack([OD.weighted_l2_loss_v2(c, g.prev_offset, g.weight),
                         OD.weighted_l2_loss_v2(rel[idx], rel[:, None], g.weight),
                         OD.weighted_l2_loss_v1(torch.sqrt((off ** 2).sum(-1) + 1e-20), g.dist, g.weight)])
        (L * w).sum().backward()
        return L

    to, te = timeit(ours, 20), timeit(eager, 20)
    Lo = DL.rigidity_losses(p2, r2, g).detach().cpu()
    Le = eager().detach().cpu()
    print(f"dyn losses F={pts.shape[0]} K=20 fwd+bwd: fused {to:.3f} ms, eager torch on the same GPU {te:.3f} ms "
          f"({te / to:.1f}x); losses fused {Lo.tolist()} eager {Le.tolist()}", flush=True)


if __name__ == "__main__":
    main()
