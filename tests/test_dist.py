"""View-sharded data parallelism: host logic on CPU with the gloo backend, world_size 2."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402,F401
from dynamic3dgaussians_b200 import dist as D  # noqa: E402


def test_partition_views_balances_counts_and_costs():
    for world in (1, 2, 4, 8):
        parts = D.partition_views([1.0] * 27, world)
        sizes = sorted(len(p) for p in parts)
        assert sum(sizes) == 27 and sizes[-1] - sizes[0] <= 1
        assert sorted(v for p in parts for v in p) == list(range(27))
    # cost-aware: one heavy view gets a rank (almost) to itself
    costs = [10.0] + [1.0] * 9
    parts = D.partition_views(costs, 2)
    heavy = next(p for p in parts if 0 in p)
    assert sum(costs[v] for v in heavy) <= 10.0 + 1e-9
    assert D.partition_views(costs, 2) == parts                     # deterministic


def test_partition_units_covers_every_tile_row_once_and_balances():
    import random
    rnd = random.Random(0)
    V, rows = 27, 23
    rc = [[rnd.random() * (2.0 - (r - 11) ** 2 * 0.005) for r in range(rows)] for _ in range(V)]
    fixed = 0.1 * sum(map(sum, rc)) / V
    for world in (1, 2, 3, 4, 8):
        parts = D.partition_units(rc, world, fixed_cost=fixed)
        assert len(parts) == world and parts == D.partition_units(rc, world, fixed_cost=fixed)       # deterministic
        seen = {}
        for p in parts:
            for v, a, b in p:
                assert 0 <= a < b <= rows
                for r in range(a, b):
                    seen[(v, r)] = seen.get((v, r), 0) + 1
        assert len(seen) == V * rows and set(seen.values()) == {1}
        loads = [sum(sum(rc[v][a:b]) + fixed for v, a, b in p) for p in parts]
        assert max(loads) <= 1.05 * sum(loads) / world                  # whole views alone: 4 / 3.375 = 1.19 at world 8
        shared = sum(1 for v in range(V) if sum(1 for p in parts for u in p if u[0] == v) > 1)
        assert shared <= world - 1                                      # views stay whole wherever possible
    # more ranks than rows, empty input, all-zero costs
    parts = D.partition_units([[1.0] * 3] * 2, 8)
    assert sorted(u for p in parts for u in p) == [(v, r, r + 1) for v in range(2) for r in range(3)]
    assert D.partition_units([], 2) == [[], []]
    z = D.partition_units([[0.0] * 4] * 3, 2)
    assert sorted((v, r) for p in z for v, a, b in p for r in range(a, b)) == [(v, r) for v in range(3) for r in range(4)]


def test_rescale_row_costs_feedback_converges_on_a_systematic_model_error():
    import random
    rnd = random.Random(1)
    V, rows, world = 27, 23, 8
    scale = [0.5 + 1.5 * rnd.random() for _ in range(V)]
    model = [[scale[v] * (0.5 + rnd.random()) for _ in range(rows)] for v in range(V)]
    true = [[model[v][r] ** 0.7 + 0.2 for r in range(rows)] for v in range(V)]        # what the ranks would measure
    fixed = 0.1 * sum(map(sum, model)) / V

    def run(parts):
        return [sum(sum(true[v][a:b]) + 1.5 * fixed for v, a, b in p) for p in parts]

    costs, worst = model, []
    for _ in range(4):
        parts = D.partition_units(costs, world, fixed_cost=fixed)
        ms = run(parts)
        worst.append(max(ms) / (sum(ms) / world))
        costs = D.rescale_row_costs(costs, parts, ms, fixed)
        assert abs(sum(map(sum, costs)) - sum(map(sum, model))) < 0.05 * sum(map(sum, model))   # total roughly kept
    assert worst[0] > 1.2 and worst[2] < 1.05 and worst[3] < 1.05
    # whole-view units (ints) are accepted too, and a degenerate measurement changes nothing
    c2 = D.rescale_row_costs([[1.0, 1.0], [2.0, 2.0]], [[0], [1]], [1.0, 1.0])
    assert c2[0][0] > 1.0 and c2[1][0] < 2.0
    assert D.rescale_row_costs([[1.0]], [[0]], [0.0]) == [[1.0]]


def test_split_units_mixes_whole_views_and_bands():
    assert D.split_units([3, 5]) == ([3, 5], None)
    assert D.split_units([(3, 0, 7), 5, (6, 2, 4)]) == ([3, 5, 6], [(0, 7), None, (2, 4)])
    cot = dict(color=torch.arange(8.0).reshape(8, 1, 1, 1), depth=None)
    sel = D.select_views(cot, [(3, 0, 7), 5])
    assert sel["depth"] is None and sel["color"].flatten().tolist() == [3.0, 5.0]


def test_pack_unpack_roundtrip():
    P = 17
    g = dict(means3D=torch.randn(P, 3), colors=torch.randn(P, 3), opacity=torch.randn(P, 1), scales=torch.randn(P, 3),
             rotations=torch.randn(P, 4))
    slab = D.pack_grads(g, P)
    assert slab.shape == (P, D.GRAD_WIDTH) and slab.is_contiguous()
    u = D.unpack_grads(slab)
    for k in D.GRAD_KEYS:
        assert torch.equal(u[k], g[k])


def test_grad_slab_regions_are_contiguous_views_of_one_buffer():
    P = 11
    slab = D.GradSlab(P, "cpu")
    assert slab.flat.numel() == P * D.GRAD_WIDTH
    for k in D.GRAD_KEYS:
        v = slab.views[k]
        assert v.is_contiguous() and v.shape == (P, D.GradSlab.WIDTHS[k])
        v.fill_(float(D.GRAD_KEYS.index(k) + 1))
    assert slab.flat.sum().item() == P * (3 * 1 + 3 * 2 + 1 * 3 + 3 * 4 + 4 * 5)       # writes landed in the flat buffer
    rows = slab.as_rows()
    assert rows.shape == (P, D.GRAD_WIDTH) and rows[0].tolist() == [1, 1, 1, 2, 2, 2, 3, 4, 4, 4, 5, 5, 5, 5]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        V, P = 7, 50
        g = torch.Generator().manual_seed(0)
        per_view = torch.randn(V, P, D.GRAD_WIDTH, generator=g)       # stand-in for each view's per-Gaussian gradients
        parts = D.partition_views([float(v + 1) for v in range(V)], world)
        mine = parts[rank]
        slab = per_view[mine].sum(dim=0) if mine else torch.zeros(P, D.GRAD_WIDTH)
        D.allreduce_grads(slab)
        ok = torch.allclose(slab, per_view.sum(dim=0), atol=1e-5)
        # an async handle works too, and a rank with no views contributes zeros
        z = torch.zeros(P, D.GRAD_WIDTH) if rank else per_view[0].clone()
        h = D.allreduce_grads(z, async_op=True)
        h.wait()
        ok = ok and torch.allclose(z, per_view[0], atol=1e-6)
        # make_slab: without GPUs / NCCL there is no peer memory -- a plain GradSlab whose allreduce() goes through the
        # process group (the peer kernel itself: tests/test_gpu_peer.py on one GPU, tests/test_gpu_peer_multi.py on several)
        ms = D.make_slab(P, "cpu")
        ok = ok and type(ms) is D.GradSlab and ms.exchange == "nccl all_reduce"
        ms.views["means3D"].copy_(per_view[mine].sum(dim=0)[:, 0:3] if mine else torch.zeros(P, 3))
        ms.allreduce()
        ok = ok and torch.allclose(ms.views["means3D"], per_view.sum(dim=0)[:, 0:3], atol=1e-5)
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_allreduce_equals_single_process_sum():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert out[0] and out[1]


def test_allreduce_is_a_noop_without_a_process_group():
    t = torch.ones(3, D.GRAD_WIDTH)
    assert D.allreduce_grads(t) is None and torch.equal(t, torch.ones(3, D.GRAD_WIDTH))
    slab = D.make_slab(5, "cpu")                      # no process group: the plain slab, and its allreduce is a no-op too
    assert type(slab) is D.GradSlab and slab.allreduce() is None
