"""kNN (distCUDA2 / rigidity graph): oracle self-checks on the CPU, CUDA parity on the GPU (through the C ABI)."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402,F401  (puts the repo root on sys.path)
from oracle import eager_knn as OK  # noqa: E402
from oracle import ref_cuda  # noqa: E402


def _clouds():
    rng = np.random.default_rng(7)
    yield "gauss", rng.normal(size=(3000, 3)).astype(np.float32)
    yield "plane", np.concatenate([rng.uniform(-2, 2, size=(2000, 2)), np.zeros((2000, 1))], axis=1).astype(np.float32)
    c = rng.normal(size=(1500, 3)).astype(np.float32)
    c[::7] = c[1::7][: len(c[::7])]                      # exact duplicates: neighbours at distance 0
    yield "dups", c


def _check_indices(pts, idx, d2, what):
    """Every returned index is another point at exactly the returned distance (ties may permute equal distances)."""
    P, k = idx.shape
    assert (idx != np.arange(P)[:, None]).all(), what
    assert np.array_equal(OK.dist2_ref(pts[:, None, :], pts[idx]), d2), what
    assert (d2[:, 1:] >= d2[:, :-1]).all(), what


def test_oracle_kdtree_equals_bruteforce():
    for name, pts in _clouds():
        for k in (3, 20):
            a = OK.knn_bruteforce(pts, k)
            b = OK.knn_kdtree(pts, k)
            assert np.array_equal(a[2], b[2]), name
            assert np.array_equal(a[0], b[0]), name
            _check_indices(pts, b[1], b[2], name)


def test_oracle_known_answers():
    # unit lattice: every interior point has 6 neighbours at distance 1 -> mean of the 3 nearest squared distances = 1
    g = np.stack(np.meshgrid(*[np.arange(5)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    mean, idx, d2 = OK.knn_bruteforce(g, 3)
    assert np.all(mean == 1.0) and np.all(d2 == 1.0)
    # collinear points at 0, 1, 3, 7: nearest of the last is the third (16), then 36, 49
    p = np.array([[0, 0, 0], [1, 0, 0], [3, 0, 0], [7, 0, 0]], dtype=np.float32)
    mean, idx, d2 = OK.knn_bruteforce(p, 3)
    assert idx[3].tolist() == [2, 1, 0] and d2[3].tolist() == [16.0, 36.0, 49.0]
    assert mean[3] == np.float32((np.float32(16.0) + np.float32(36.0) + np.float32(49.0)) / np.float32(3.0))


@pytest.mark.gpu
@pytest.mark.parametrize("k", [3, 20, 5])
def test_cuda_knn_matches_oracle(k):
    from dynamic3dgaussians_b200 import knn as KN
    for name, pts in _clouds():
        mean, idx, d2 = KN.knn(torch.from_numpy(pts).cuda(), k)
        om, oi, od = OK.knn_kdtree(pts, k)
        assert np.array_equal(d2.cpu().numpy(), od), name            # same contraction -> identical floats
        assert np.array_equal(mean.cpu().numpy(), om), name
        _check_indices(pts, idx.cpu().numpy(), od, name)
        ext = OK.knn_kdtree(pts, k + 1)[2]                      # one more neighbour: a tie may straddle the k-th
        untied = ~(ext[:, 1:] == ext[:, :-1]).any(axis=1)
        assert np.array_equal(idx.cpu().numpy()[untied], oi[untied]), name


@pytest.mark.gpu
def test_cuda_knn_edge_cases():
    from dynamic3dgaussians_b200 import knn as KN
    m, i, d = KN.knn(torch.zeros(0, 3, device="cuda"), 3)
    assert m.numel() == 0 and i.shape == (0, 3)
    # fewer points than k+1: missing neighbours keep the reference's FLT_MAX / index 0 (K/simple_knn.cu:157-158)
    p = torch.tensor([[0.0, 0, 0], [2.0, 0, 0]], device="cuda")
    m, i, d = KN.knn(p, 3)
    assert d[0, 0].item() == 4.0 and i[0, 0].item() == 1 and d[0, 1].item() == torch.finfo(torch.float32).max
    with pytest.raises(ValueError):
        KN.knn(torch.zeros(4, 3), 3)                                 # CPU tensor: there is no CPU path
    # sizes around the leaf / super-box boundaries
    rng = np.random.default_rng(3)
    for P in (31, 32, 33, 1023, 1024, 1025, 4097):
        pts = rng.uniform(-1, 1, size=(P, 3)).astype(np.float32)
        mean, idx, d2 = KN.knn(torch.from_numpy(pts).cuda(), 3)
        om, oi, od = OK.knn_bruteforce(pts, 3)
        assert np.array_equal(d2.cpu().numpy(), od) and np.array_equal(idx.cpu().numpy(), oi), P


@pytest.mark.gpu
@pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref not built")
def test_cuda_distcuda2_equals_reference_kernel():
    """simple_knn._C.distCUDA2 against the reference's own kernel (oracle/_ref/libref_knn.so) at C2 size: mean squared
    distances bit-identical; indices identical wherever the three distances are distinct."""
    import simple_knn._C as SK
    from dynamic3dgaussians_b200 import scenes
    pts = scenes.make_scene_c2(300_000).means3D.cuda()
    mean, idx = SK.distCUDA2(pts)
    rmean, ridx = ref_cuda.knn(pts)
    torch.cuda.synchronize()
    assert torch.equal(mean, rmean)
    same = (idx == ridx).all(dim=1)
    assert same.float().mean().item() > 0.999
    assert torch.equal(SK.distCUDA2_dist(pts), mean)
