"""GPU test (pytest -m gpu) of the gradient exchange on REAL peer memory: needs >= 2 GPUs on the box (skipped on the
single-GPU round-end box; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_peer_multi.py -m gpu` runs it).  Launches
tests/gpu_peer_check.py under torch.distributed.run on 2 ranks: dist.PeerBuffer (symmetric memory + gsr_peer_allreduce)
must give ncclAllReduce's sums (<= 1e-6 relative), the same bits on every rank, twice in a row, through the
reduce-scatter / all-gather modes, and replayed from a CUDA graph, with no wait timing out."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs of one node")
def test_peer_exchange_equals_nccl_on_two_gpus():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "gpu_peer_check.py"), str(14 * 100_000)]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    p2p = out["p2p"]
    assert "unavailable" not in p2p, p2p
    assert not p2p["gave_up"] and not p2p["gave_up_after_timing"]
    assert p2p["bit_identical_on_all_ranks"]
    for k in ("max_rel_err_vs_nccl", "second_call_max_rel_err", "reduce_scatter_allgather_max_rel_err"):
        assert p2p[k] <= 1e-6, (k, p2p[k])
    assert out["graph_replay_max_rel_err"] <= 1e-6 and not out["graph_gave_up"]
    mm = out.get("multimem", {})
    if "unavailable" not in mm and mm:
        assert not mm["gave_up"] and mm["bit_identical_on_all_ranks"] and mm["max_rel_err_vs_nccl"] <= 1e-6
