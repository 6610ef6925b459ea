"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU oracle of the exact k-nearest-neighbour query that SimpleKNN::knn computes (K/simple_knn.cu:131-190, K/ =
submodules_fsgs/simple-knn/): for every point the k nearest OTHER points by squared Euclidean distance (self is
excluded by position, so exact duplicates are neighbours at distance 0), ascending, and their mean
(`(best[0] + best[1] + best[2]) / 3.0f`, :186).  `distCUDA2` (K/spatial.cu:15-27) is k = 3.  The 20-NN rigidity graph of
the training scripts is built by Open3D's KD-tree in the reference (REF/helpers.py:136-146, open3d==0.16.0, absent here);
its published behaviour -- exact kNN, self dropped as the first hit -- is the same query with k = 20.

Two restatements: `knn_bruteforce` (numpy, O(P^2), small P) and `knn_kdtree` (scipy cKDTree in float64 for candidate
search, distances re-evaluated in float32 with the reference binary's contraction fma(dz,dz, fma(dx,dx, dy*dy)) which is
emulated exactly through float64: the product of two float32 is exact in float64 and one rounding to float32 follows
every fused step).  Parity: unpinned by the reference (it ships no test for this path); pinned here by brute force and,
on the GPU box, by the reference's own kernel compiled into oracle/_ref/libref_knn.so.
"""
from __future__ import annotations

import numpy as np


def _fma32(a, b, c):
    """float32 fma(a, b, c) via float64 (exact for float32 inputs up to one double rounding that cannot matter here:
    a*b is exact in float64; the sum is rounded to float64, then to float32)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def dist2_ref(q: np.ndarray, p: np.ndarray) -> np.ndarray:
    """Squared distance as the reference binary evaluates it (K/simple_knn.cu:134-135)."""
    d = (p.astype(np.float32) - q.astype(np.float32)).astype(np.float32)
    dx, dy, dz = d[..., 0], d[..., 1], d[..., 2]
    t = (dy * dy).astype(np.float32)
    t = _fma32(dx, dx, t)
    return _fma32(dz, dz, t)


def knn_bruteforce(points: np.ndarray, k: int):
    """(mean_dist2 [P], idx [P,k] int32, dist2 [P,k]) -- stable in the candidate index on exact ties."""
    pts = np.ascontiguousarray(points, dtype=np.float32)
    P = pts.shape[0]
    d2 = dist2_ref(pts[:, None, :], pts[None, :, :])
    d2[np.arange(P), np.arange(P)] = np.inf
    order = np.argsort(d2, axis=1, kind="stable")[:, :k]
    dk = np.take_along_axis(d2, order, axis=1).astype(np.float32)
    return _mean(dk), order.astype(np.int32), dk


def _mean(dk: np.ndarray) -> np.ndarray:
    s = np.zeros(dk.shape[0], dtype=np.float32)
    for j in range(dk.shape[1]):
        s = (s + dk[:, j]).astype(np.float32)
    return (s / np.float32(dk.shape[1])).astype(np.float32)


def knn_kdtree(points: np.ndarray, k: int, slack: int = 4):
    """Same result for large P: cKDTree proposes k+slack candidates in float64, float32 reference distances re-rank."""
    from scipy.spatial import cKDTree
    pts = np.ascontiguousarray(points, dtype=np.float32)
    P = pts.shape[0]
    kk = min(P, k + 1 + slack)
    _, cand = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=kk, workers=-1)
    cand = cand.reshape(P, kk)
    d2 = dist2_ref(pts[:, None, :], pts[cand])
    d2 = np.where(cand == np.arange(P)[:, None], np.inf, d2)
    key = np.argsort(d2, axis=1, kind="stable")
    order = np.take_along_axis(cand, key, axis=1)[:, :k]
    dk = np.take_along_axis(d2, key, axis=1)[:, :k].astype(np.float32)
    return _mean(dk), order.astype(np.int32), dk
