"""Host side of the fused rigidity / rotation / isometry losses (REF/train.py:253-273): one autograd.Function over
libgsr_b200.so's gsr_dyn_losses.  No CPU path."""
from __future__ import annotations

import torch

from . import _lib


def _chk(t, name, dtype=torch.float32):
    if not torch.is_tensor(t) or not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


class NeighborGraph:
    """The per-timestep constants of the losses: what REF/train.py keeps in `variables` (neighbor_indices, neighbor_weight,
    neighbor_dist :321-326; prev_inv_rot_fg, prev_offset :303-308)."""

    def __init__(self, neighbor_indices, neighbor_weight, neighbor_dist, prev_inv_rot, prev_offset):
        self.idx = _chk(neighbor_indices.to(torch.int32), "neighbor_indices", torch.int32)
        self.weight = _chk(neighbor_weight, "neighbor_weight")
        self.dist = _chk(neighbor_dist, "neighbor_dist")
        self.prev_inv_rot = _chk(prev_inv_rot, "prev_inv_rot")
        self.prev_offset = _chk(prev_offset, "prev_offset")
        F, K = self.idx.shape
        if self.weight.shape != (F, K) or self.dist.shape != (F, K) or self.prev_offset.shape != (F, K, 3) \
                or self.prev_inv_rot.shape != (F, 4):
            raise ValueError("inconsistent neighbour-graph shapes")
        self.F, self.K = F, K


def _run(pts, rot, graph: NeighborGraph, upstream):
    lib = _lib.load()
    dev = pts.device
    F, K = graph.F, graph.K
    if pts.shape != (F, 3) or rot.shape != (F, 4):
        raise ValueError("pts must be [F,3] and rot [F,4] for the graph's F")
    nbytes = lib.gsr_dyn_losses_temp_bytes(F)
    tmp = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    losses = torch.empty(3, dtype=torch.float32, device=dev)
    dp = dr = None
    if upstream is not None:
        dp = torch.empty((F, 3), dtype=torch.float32, device=dev)
        dr = torch.empty((F, 4), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.gsr_dyn_losses(F, K, pts.data_ptr(), rot.data_ptr(), graph.prev_inv_rot.data_ptr(),
                                      graph.idx.data_ptr(), graph.weight.data_ptr(), graph.prev_offset.data_ptr(),
                                      graph.dist.data_ptr(), _lib.ptr(upstream), losses.data_ptr(), _lib.ptr(dp),
                                      _lib.ptr(dr), tmp.data_ptr(), nbytes,
                                      torch.cuda.current_stream(dev).cuda_stream), "dyn_losses")
    return losses, dp, dr


class _DynLosses(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pts, rot, graph):
        pts_c, rot_c = _chk(pts.detach(), "pts"), _chk(rot.detach(), "rot")
        losses, _, _ = _run(pts_c, rot_c, graph, None)
        ctx.graph = graph
        ctx.save_for_backward(pts_c, rot_c)
        return losses

    @staticmethod
    def backward(ctx, g_losses):
        pts, rot = ctx.saved_tensors
        up = g_losses.to(torch.float32).contiguous()
        _, dp, dr = _run(pts, rot, ctx.graph, up)
        return dp, dr, None


def rigidity_losses(fg_pts: torch.Tensor, fg_rot: torch.Tensor, graph: NeighborGraph) -> torch.Tensor:
    """[3] tensor (rigid, rot, iso) of REF/train.py:263-273, differentiable w.r.t. fg_pts [F,3] and fg_rot [F,4]."""
    return _DynLosses.apply(fg_pts, fg_rot, graph)


def build_neighbor_graph(fg_pts: torch.Tensor, fg_rot: torch.Tensor, num_knn: int = 20) -> NeighborGraph:
    """initialize_post_first_timestep + initialize_per_timestep's graph constants (REF/train.py:294-326) on the GPU:
    the reference builds the 20-NN graph on the CPU with Open3D; here it is the CUDA kNN of this package."""
    from .knn import knn
    pts = _chk(fg_pts.detach(), "fg_pts")
    rot = _chk(fg_rot.detach(), "fg_rot")
    _, idx, d2 = knn(pts, num_knn)
    weight = torch.exp(-2000.0 * d2)
    dist = torch.sqrt(d2)
    return refresh_graph(NeighborGraph(idx, weight, dist, torch.zeros_like(rot), torch.zeros(
        (pts.shape[0], num_knn, 3), device=pts.device)), pts, rot)


def refresh_graph(graph: NeighborGraph, fg_pts: torch.Tensor, fg_rot: torch.Tensor) -> NeighborGraph:
    """Per-timestep update (REF/train.py:300-308): prev_inv_rot = conjugate of the current rotation, prev_offset from the
    current positions."""
    pts, rot = fg_pts.detach(), fg_rot.detach()
    inv = rot.clone()
    inv[:, 1:] = -inv[:, 1:]
    off = pts[graph.idx.long()] - pts[:, None]
    return NeighborGraph(graph.idx, graph.weight, graph.dist, inv.contiguous(), off.contiguous())
