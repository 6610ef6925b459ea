"""Rigidity / rotation / isometry losses: oracle pinned by float64 finite differences (CPU); CUDA kernel vs the oracle (GPU)."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402,F401
from oracle import eager_dyn_losses as OD  # noqa: E402
from oracle import eager_knn as OK  # noqa: E402


def _case(F=400, K=20, seed=0, dtype=torch.float32, motion=0.02):
    g = torch.Generator().manual_seed(seed)
    prev = torch.randn(F, 3, generator=g, dtype=torch.float64) * 0.3
    prev_rot = torch.nn.functional.normalize(torch.randn(F, 4, generator=g, dtype=torch.float64))
    _, idx, d2 = OK.knn_kdtree(prev.float().numpy(), K)
    idx = torch.from_numpy(idx.astype(np.int64))
    d2 = torch.from_numpy(d2).double()
    weight = torch.exp(-2000 * d2) + 0.05                # keep every edge alive so gradients are informative
    dist = torch.sqrt(d2)
    inv = prev_rot.clone()
    inv[:, 1:] *= -1
    off = prev[idx] - prev[:, None]
    pts = prev + motion * torch.randn(F, 3, generator=g, dtype=torch.float64)
    rot = torch.nn.functional.normalize(prev_rot + motion * torch.randn(F, 4, generator=g, dtype=torch.float64))
    c = lambda t: t.to(dtype)
    return dict(pts=c(pts), rot=c(rot), inv=c(inv), idx=idx, weight=c(weight), off=c(off), dist=c(dist))


def _oracle(c, pts=None, rot=None):
    return OD.rigidity_losses(c["pts"] if pts is None else pts, c["rot"] if rot is None else rot, c["inv"], c["idx"],
                              c["weight"], c["off"], c["dist"])


def test_oracle_zero_motion_gives_zero_losses():
    c = _case(motion=0.0, dtype=torch.float64)
    L = _oracle(c)
    assert float(L.abs().max()) < 1e-7                      # sqrt(1e-20) floor


def test_oracle_rigid_motion_is_free():
    """A global rigid motion leaves rigid / iso at zero (the point of the loss); rot measures neighbour consistency."""
    c = _case(motion=0.0, dtype=torch.float64)
    q = torch.nn.functional.normalize(torch.tensor([[0.9, 0.1, -0.3, 0.2]], dtype=torch.float64))
    Rm = OD.build_rotation(q)[0]
    pts = c["pts"] @ Rm.T + torch.tensor([0.3, -0.2, 0.5], dtype=torch.float64)
    prev_rot = c["inv"].clone()
    prev_rot[:, 1:] *= -1
    rot = OD.quat_mult(q.expand_as(prev_rot), prev_rot)    # every Gaussian rotated by q
    L = _oracle(c, pts=pts, rot=rot)
    assert float(L.abs().max()) < 1e-6


def test_oracle_gradients_match_finite_differences():
    c = _case(F=60, K=8, dtype=torch.float64, seed=3)
    pts = c["pts"].clone().requires_grad_(True)
    rot = c["rot"].clone().requires_grad_(True)
    wts = torch.tensor([4.0, 4.0, 2.0], dtype=torch.float64)     # REF/dense_train.py loss weights (rigid, rot, iso)
    (_oracle(c, pts, rot) * wts).sum().backward()
    eps = 1e-6
    rng = np.random.default_rng(0)
    for t, gname in ((pts, "pts"), (rot, "rot")):
        for _ in range(10):
            i, j = int(rng.integers(t.shape[0])), int(rng.integers(t.shape[1]))
            d = torch.zeros_like(t)
            d[i, j] = eps
            kw_p = {gname: (t + d).detach()}
            kw_m = {gname: (t - d).detach()}
            fd = float(((_oracle(c, **kw_p) - _oracle(c, **kw_m)) * wts).sum() / (2 * eps))
            assert abs(fd - float(t.grad[i, j])) <= 1e-6 * max(1.0, abs(fd)), (gname, i, j)


@pytest.mark.gpu
@pytest.mark.parametrize("F,K", [(400, 20), (5000, 20), (257, 3)])
def test_cuda_losses_and_gradients_match_oracle(F, K):
    from dynamic3dgaussians_b200 import dyn_losses as DL
    c = _case(F=F, K=K, seed=F)
    graph = DL.NeighborGraph(c["idx"].cuda(), c["weight"].cuda(), c["dist"].cuda(), c["inv"].cuda(), c["off"].cuda())
    pts = c["pts"].cuda().requires_grad_(True)
    rot = c["rot"].cuda().requires_grad_(True)
    wts = torch.tensor([4.0, 4.0, 2.0])
    L = DL.rigidity_losses(pts, rot, graph)
    (L * wts.cuda()).sum().backward()
    op = c["pts"].clone().requires_grad_(True)
    orot = c["rot"].clone().requires_grad_(True)
    Lo = _oracle(c, op, orot)
    (Lo * wts).sum().backward()
    assert float((L.cpu() - Lo).abs().max() / Lo.abs().max()) <= 1e-5
    assert helpers.rel_err(pts.grad, op.grad) <= 1e-4
    assert helpers.rel_err(rot.grad, orot.grad) <= 1e-4


@pytest.mark.gpu
def test_cuda_graph_builder_matches_reference_recipe():
    """build_neighbor_graph == initialize_post_first_timestep's recipe (REF/train.py:316-326) with an exact kNN."""
    from dynamic3dgaussians_b200 import dyn_losses as DL
    c = _case(F=3000, K=20, seed=11)
    pts, rot = c["pts"].cuda(), c["rot"].cuda()
    g = DL.build_neighbor_graph(pts, rot, 20)
    _, oi, od = OK.knn_kdtree(c["pts"].numpy(), 20)
    assert np.array_equal(g.dist.cpu().numpy(), np.sqrt(od))
    assert torch.allclose(g.weight.cpu(), torch.exp(-2000 * torch.from_numpy(od)), rtol=1e-5, atol=1e-30)
    L = DL.rigidity_losses(pts, rot, g)                     # no motion since the graph was refreshed: all ~0
    assert float(L.abs().max()) < 1e-6
    with pytest.raises(ValueError):
        DL.rigidity_losses(pts.cpu(), rot.cpu(), g)
