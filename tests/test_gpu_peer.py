"""GPU test (pytest -m gpu, ONE GPU) of the peer-memory exchange kernel gsr_peer_allreduce (csrc/peer_allreduce.cu): the
`world` ranks are played by `world` streams of one device, each with its own buffer and flag block, so the protocol
(arrival words, epochs, slices, rank-ordered sum, push) is exercised without a second GPU.  The kernels of the different
"ranks" wait for one another inside the kernel, so they must be co-resident: a few CTAs each.  The real thing -- buffers
in symmetric memory on 2-8 GPUs against ncclAllReduce -- is tests/gpu_peer_check.py under torchrun (profiles/scale_run.sh).
The reference is single-GPU: there is no reference counterpart; the check is against the same sum done by torch in the
same (rank) order, bit for bit."""
import ctypes as C
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dynamic3dgaussians_b200 import _lib  # noqa: E402

pytestmark = pytest.mark.gpu


def _setup(world, n, seed):
    g = torch.Generator().manual_seed(seed)
    bufs = [(torch.randn(n, generator=g) * (1.0 + r)).cuda() for r in range(world)]
    flags = [torch.zeros(_lib.PEER_FLAG_WORDS, dtype=torch.int32, device="cuda") for _ in range(world)]
    bp = (C.c_void_p * world)(*[b.data_ptr() for b in bufs])
    fp = (C.c_void_p * world)(*[f.data_ptr() for f in flags])
    streams = [torch.cuda.Stream() for _ in range(world)]
    return bufs, flags, bp, fp, streams


def _call_all(lib, mode, world, bp, fp, n, streams, blocks=4):
    torch.cuda.synchronize()
    for r in range(world):
        _lib.check(lib.gsr_peer_allreduce(mode, world, r, bp, fp, None, n, blocks, streams[r].cuda_stream), "peer")
    torch.cuda.synchronize()


def _rank_ordered_sum(bufs):
    acc = bufs[0].clone()
    for b in bufs[1:]:
        acc = acc + b
    return acc


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_allreduce_is_the_rank_ordered_sum_on_every_rank(world):
    lib = _lib.load()
    n = 4 * 100_003
    bufs, flags, bp, fp, streams = _setup(world, n, world)
    want = _rank_ordered_sum(bufs)
    _call_all(lib, _lib.PEER_ALLREDUCE, world, bp, fp, n, streams)
    for r in range(world):
        assert int(flags[r][_lib.PEER_ERROR_WORD]) == 0, "a wait timed out: the streams did not run side by side"
        assert torch.equal(bufs[r], want), r
    # again on the same buffers and flag blocks (epoch 2), with another grid size
    want2 = _rank_ordered_sum(bufs)
    _call_all(lib, _lib.PEER_ALLREDUCE, world, bp, fp, n, streams, blocks=7)
    for r in range(world):
        assert int(flags[r][_lib.PEER_ERROR_WORD]) == 0
        assert torch.equal(bufs[r], want2), r


def test_reduce_scatter_then_allgather_equal_allreduce():
    lib = _lib.load()
    world, n = 4, 4 * 50_001
    bufs, flags, bp, fp, streams = _setup(world, n, 9)
    before = [b.clone() for b in bufs]
    want = _rank_ordered_sum(bufs)
    per = -(-(n // 4) // world) * 4
    _call_all(lib, _lib.PEER_REDUCE_SCATTER, world, bp, fp, n, streams)
    for r in range(world):
        lo, hi = min(r * per, n), min((r + 1) * per, n)
        assert torch.equal(bufs[r][lo:hi], want[lo:hi]), r
        keep = torch.ones(n, dtype=torch.bool, device="cuda")
        keep[lo:hi] = False
        assert torch.equal(bufs[r][keep], before[r][keep]), r            # the other slices are untouched
    _call_all(lib, _lib.PEER_ALLGATHER, world, bp, fp, n, streams)
    for r in range(world):
        assert int(flags[r][_lib.PEER_ERROR_WORD]) == 0
        assert torch.equal(bufs[r], want), r


def test_bad_arguments_and_world_one():
    lib = _lib.load()
    n = 4 * 1000
    bufs, flags, bp, fp, streams = _setup(1, n, 1)
    before = bufs[0].clone()
    _call_all(lib, _lib.PEER_ALLREDUCE, 1, bp, fp, n, streams)
    assert torch.equal(bufs[0], before)
    s = torch.cuda.current_stream().cuda_stream
    assert lib.gsr_peer_allreduce(0, 9, 0, bp, fp, None, n, 0, s) == -1            # world > GSR_PEER_MAX_WORLD
    assert lib.gsr_peer_allreduce(0, 1, 1, bp, fp, None, n, 0, s) == -1            # rank outside the world
    assert lib.gsr_peer_allreduce(5, 1, 0, bp, fp, None, n, 0, s) == -1            # unknown mode
    assert lib.gsr_peer_allreduce(0, 1, 0, bp, fp, None, n + 2, 0, s) == -1        # not whole float4s
    assert lib.gsr_peer_allreduce(0, 1, 0, None, fp, None, n, 0, s) == -1
    assert lib.gsr_peer_allreduce(0, 1, 0, bp, fp, None, 0, 0, s) == 0
