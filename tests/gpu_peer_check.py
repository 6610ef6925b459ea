"""Multi-GPU check (not collected by pytest; run under torchrun on 2-8 GPUs of one node, profiles/scale_run.sh):
the peer-memory exchange kernel (dist.PeerBuffer -> gsr_peer_allreduce) against ncclAllReduce on the gradient-slab size
of the bench workload (14 x 300000 floats): same sums (<= 1e-6 relative), bit-identical on all ranks, and device time of
both, back to back.  Rank 0 prints one JSON line.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tests/gpu_peer_check.py [n_floats]"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from dynamic3dgaussians_b200 import dist as D  # noqa: E402


def timed(fn, iters=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t.item()) * 1e3, 2)           # us, max over ranks


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 14 * 300_000
    out = {"world": world, "n_floats": n}
    g = torch.Generator().manual_seed(100 + rank)
    src = (torch.randn(n, generator=g) * (1 + rank)).cuda()
    ref = src.clone()
    dist.all_reduce(ref)

    def same_on_all_ranks(t):
        h = t.view(torch.int32).to(torch.int64).sum().reshape(1)
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        return all(int(x) == int(hs[0]) for x in hs)

    for name, mc in (("p2p", False), ("multimem", True)):
        try:
            pb = D.PeerBuffer(n, "cuda", multicast=mc)
        except Exception as e:                                   # noqa: BLE001
            out[name] = {"unavailable": f"{type(e).__name__}: {str(e)[:300]}"}
            break
        if mc and not pb._mc:
            out[name] = {"unavailable": "no multicast mapping (multicast_ptr == 0)"}
            continue
        r = {}
        pb.t.copy_(src)
        pb.allreduce()
        torch.cuda.synchronize()
        r["gave_up"] = pb.gave_up()
        r["max_rel_err_vs_nccl"] = float(((pb.t - ref).abs().max() / ref.abs().max()).item())
        r["bit_identical_on_all_ranks"] = same_on_all_ranks(pb.t)
        # twice in a row on fresh data (epochs), then the two half operations
        pb.t.copy_(src)
        pb.allreduce()
        r["second_call_max_rel_err"] = float(((pb.t - ref).abs().max() / ref.abs().max()).item())
        if not mc:
            pb.t.copy_(src)
            pb.reduce_scatter()
            pb.allgather()
            r["reduce_scatter_allgather_max_rel_err"] = float(((pb.t - ref).abs().max() / ref.abs().max()).item())
            r["rs_us"] = timed(pb.reduce_scatter)
            r["ag_us"] = timed(pb.allgather)
        r["allreduce_us"] = timed(pb.allreduce)
        for blocks in (32, 64, 148, 256):
            pb.blocks = blocks
            r[f"allreduce_us_blocks{blocks}"] = timed(pb.allreduce)
        r["gave_up_after_timing"] = pb.gave_up()
        out[name] = r
    t = src.clone()
    out["nccl_allreduce_us"] = timed(lambda: dist.all_reduce(t))
    # size sweep: the small end is the two flag round trips, the large end the NVLink rate
    sweep = {}
    for m in (4096, 262144, 1048576, 4200000, 16777216):
        row = {}
        try:
            pb = D.PeerBuffer(m, "cuda")
            row["allreduce_us"] = timed(pb.allreduce, 30)
            row["reduce_scatter_us"] = timed(pb.reduce_scatter, 30)
            row["allgather_us"] = timed(pb.allgather, 30)
            pm = D.PeerBuffer(m, "cuda", multicast=True)
            if pm._mc:
                row["multimem_us"] = timed(pm.allreduce, 30)
            del pb, pm
        except Exception as e:                                   # noqa: BLE001
            row["error"] = f"{type(e).__name__}: {str(e)[:200]}"
        tt = torch.zeros(m, device="cuda")
        row["nccl_us"] = timed(lambda: dist.all_reduce(tt), 30)
        sweep[str(m)] = row
    out["size_sweep_floats"] = sweep
    # inside a CUDA graph, as the step uses it
    try:
        pb = D.PeerBuffer(n, "cuda")
        pb.t.copy_(src)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            pb.allreduce()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            pb.allreduce()
        pb.t.copy_(src)
        gr.replay()
        torch.cuda.synchronize()
        out["graph_replay_max_rel_err"] = float(((pb.t - ref).abs().max() / ref.abs().max()).item())
        out["graph_replay_us"] = timed(gr.replay)
        out["graph_gave_up"] = pb.gave_up()
    except Exception as e:                                       # noqa: BLE001
        out["graph"] = f"{type(e).__name__}: {str(e)[:300]}"
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
