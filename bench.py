#!/usr/bin/env python
"""Benchmark of the hot path: fwd+bwd rasterization of a CMU-Panoptic-shaped timestep (BASELINE.json config C2:
300k Gaussians, 27 cameras of 640x360), reported as whole-job Mpix/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = forward + backward of ALL 27 views (sharded over the N ranks; per-Gaussian gradients all-reduced over NCCL
when N > 1).  `value` is timed with the inputs resident in HBM; `e2e` goes through the public autograd API with the
Gaussian parameters and target images coming from pinned host memory every step and the loss and gradients going
back.  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

P_GAUSS, N_VIEWS, IMG_W, IMG_H = 300_000, 27, 640, 360
METRIC = "fwd+bwd rasterized Mpix/s @300k Gaussians, 27x640x360 views"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--gaussians", type=int, default=P_GAUSS)
    ap.add_argument("--views", type=int, default=N_VIEWS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the train-loop it/s sample")
    ap.add_argument("--whole-views", action="store_true", help="N>1: shard whole views only (no tile-row bands)")
    ap.add_argument("--rebalance", type=int, default=0, help="N>1: feedback rounds of the cost-aware partition (dist.rescale_row_costs); at C2 the tile-list model is already within the run-to-run noise of the ranks")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from the host instead of one CUDA graph per step")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--exchange", default=os.environ.get("GSR_EXCHANGE", "peer"), choices=["peer", "peer-p2p", "peer-mc", "nccl"],
                    help="N>1: sum of the gradient slab over the ranks -- the library's kernel over peer memory (P2P, or "
                         "multimem reduce in the switch) or ncclAllReduce")
    ap.add_argument("--no-extras", action="store_true", help="skip the apples-to-apples / c4 / kNN / dyn-loss sub-lines")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4", "c5"],
                    help="c2 (default, the metric's configuration: 300k Gaussians, 27 views 640x360, strong scaling); "
                         "c4 (BASELINE.json configs[3]: the same scene in 4-return mode -- colour, depth AND alpha "
                         "cotangents, `label` gradient mask on the last 40 %% of the Gaussians -- sharded over the ranks); "
                         "c5 (configs[4]: 2M Gaussians, 8 views 1920x1080 PER GPU, weak scaling)")
    ap.add_argument("--emulate-world", type=int, default=0,
                    help="diagnostic, 1 GPU: run only the share rank --emulate-rank would get of an N-way partition "
                         "(no collective), to profile the small per-rank step; the line is marked and is not a result")
    ap.add_argument("--emulate-rank", type=int, default=0)
    return ap.parse_args()


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe).  NVML is used in-process
    (nvidia_ml_py) and initialised before the warm-up: spawning nvidia-smi next to the timed region stalls CUDA launches
    of this process for as long as its driver initialisation takes (seen as 7 ms of host time per step on some boxes).
    Falls back to an `nvidia-smi -lms` child started equally early."""

    def __init__(self, index: int):
        self.index = index
        self.samples = []          # (sm_mhz, max_mhz, reasons bitmask or list)
        self.active = False
        self.stop_flag = False
        self.thread = None
        self.nvml = None
        self.proc = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
        except Exception:
            self.nvml = None
            q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                 "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                 "clocks_event_reasons.sw_power_cap")
            try:
                self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={q}",
                                              "--format=csv,noheader,nounits", "-lms", "50"],
                                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.thread = threading.Thread(target=self._read_smi, daemon=True)
                self.thread.start()
            except Exception:
                self.proc = None

    def _poll_nvml(self):
        n = self.nvml
        while not self.stop_flag:
            if self.active:
                try:
                    sm = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                    try:
                        r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                    except Exception:
                        r = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                    self.samples.append((sm, self.max_mhz, r))
                except Exception:
                    pass
            time.sleep(0.01)

    def _read_smi(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            if not self.active:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                self.samples.append((float(f[0]), float(f[1]),
                                     [n for n, v in zip(names, f[3:7]) if v.lower().startswith("active")]))
            except ValueError:
                continue

    def start(self):
        self.samples = []
        self.active = True

    def stop(self):
        if self.nvml is None and self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml and nvidia-smi unavailable"]}
        if not self.samples:
            time.sleep(0.12)           # a very short timed region: take one sample right behind it
        self.active = False
        self.stop_flag = True
        if self.proc is not None:
            self.proc.terminate()
        reasons = set()
        if self.nvml is not None:
            n = self.nvml
            bits = {"hw_slowdown": getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8),
                    "hw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                    "sw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                    "sw_power_cap": getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)}
            for _, _, r in self.samples:
                for name, bit in bits.items():
                    if r & bit:
                        reasons.add(name)
        else:
            for _, _, r in self.samples:
                reasons.update(r)
        sm = sorted(x[0] for x in self.samples)
        mx = [x[1] for x in self.samples]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
    return world, rank, local


def max_over_ranks(ms: float, world: int) -> float:
    if world == 1:
        return ms
    import torch.distributed as dist
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def build_workload(args, device):
    from dynamic3dgaussians_b200 import scenes
    from dynamic3dgaussians_b200.camera import settings_from_spec
    if args.workload == "c5":
        # the same 8 cameras for every GPU, so that the per-GPU work is exactly fixed as N grows
        sc = scenes.make_scene_c2(args.gaussians, seed=5, scale=2.0)
        cams = scenes.cameras_c2(8, w=IMG_W, h=IMG_H, f=1400.0, radius=6.0) * (args.views // 8)
    else:
        sc = scenes.make_scene_c2(args.gaussians)
        cams = scenes.cameras_c2(args.views, w=IMG_W, h=IMG_H)
    sets = [settings_from_spec(c, device=device) for c in cams]
    return sc, sets


def kernels_per_step(L, with_exchange: bool, views: int = 0) -> int:
    """Launches of OUR kernels in one forward+backward (memsets / memcpys and NCCL's own kernels not counted); the names
    are those of the committed ncu launch list (profiles/r02_launches_c2.txt)."""
    fwd = (-(-max(views, 1) // 16)      # store_cams_kernel: 16 cameras per launch, as kernel arguments
           + 2                          # gaussian_static_kernel, preprocess_kernel (also takes the depth sort's digit histograms)
           + 1 + L.depth_passes         # radix_scan_hist_kernel + onesweep_pass_kernel per depth digit
           + 3                          # st_count_reduce_kernel, block_sums_scan_kernel, emit_st_instances_kernel
           + 2 + L.tile_passes          # sort_hist_from_keys_kernel, radix_scan_hist_kernel + onesweep passes (supertile bits)
           + 5                          # chunk_table, chunk_count, tile_offsets, tile_scan, tile_order
           + 2                          # pack_colors_kernel, scatter_records_kernel
           + 1)                         # blend_forward_kernel
    bwd = 1 + 1                         # blend_backward_kernel, preprocess_backward_kernel
    # launches of a few views (<= 148 * 8 * 4 CTAs) order the backward's CTAs by walked depth: one more kernel
    small = 0 < views * L.tiles_x * L.tiles_y <= 148 * 8 * 4
    return fwd + bwd + (1 if small else 0) + (1 if with_exchange else 0)      # peer_allreduce_kernel


def fp32_peak_tflops(lib, device) -> dict:
    """Measured dependent-free FMA throughput of the FP32 pipe (MEASURED_PEAKS.json has no FP32 entry): scalar FFMA and
    packed fp32x2 FFMA2 (what the blend kernels issue), 128 FMAs between loop branches.  The roofline fraction is quoted
    against the higher of the two; nominal = 148 SMs x 128 lanes x 2 x 1.965 GHz = 74.4 TFLOP/s."""
    from dynamic3dgaussians_b200 import _lib
    out = torch.empty(148 * 8 * 256, device=device)
    iters = 1 << 10
    s = torch.cuda.current_stream().cuda_stream
    res = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for name, packed, per_iter in (("ffma", 0, 256), ("ffma2", 1, 512)):
        for _ in range(2):
            _lib.check(lib.gsr_fp32_peak_probe(out.data_ptr(), iters, packed, s), "probe")
        best = 1e9
        for _ in range(5):
            e0.record()
            _lib.check(lib.gsr_fp32_peak_probe(out.data_ptr(), iters, packed, s), "probe")
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        res[name] = 148 * 8 * 256 * iters * per_iter / (best * 1e-3) / 1e12
    res["nominal"] = 148 * 128 * 2 * 1.965e9 / 1e12
    res["best"] = max(res["ffma"], res["ffma2"])
    return res


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_ours(args, world, rank, local):
    from dynamic3dgaussians_b200 import _lib, dist as D, rasterizer as RZ
    dev = torch.device("cuda", local if world > 1 else 0)
    lib = _lib.load()
    sc, sets = build_workload(args, dev)
    rv = {k: v.to(dev).contiguous() for k, v in sc.rendervar().items()}
    P, V = rv["means3D"].shape[0], len(sets)
    gen = torch.Generator().manual_seed(123)
    cot = dict(color=torch.randn(V, 3, IMG_H, IMG_W, generator=gen).to(dev),
               depth=torch.randn(V, 1, IMG_H, IMG_W, generator=gen).to(dev) * 0.1, alpha=None)
    label = None
    if args.workload == "c4":
        # 4-return callers (REF/cvpr_dyn.py:284) put a loss on alpha too and mask gradients by `label`
        cot["alpha"] = torch.randn(V, 1, IMG_H, IMG_W, generator=gen).to(dev)
        label = torch.zeros(P, device=dev)
        label[int(0.6 * P):] = 1.0
    parts = D.partition_views([1.0] * V, world)
    mine = parts[rank]

    cap = [None]
    rebalance_ms = []
    slab = D.make_slab(P, dev, None, args.exchange) if world > 1 else D.GradSlab(P, dev)     # (collective when peer memory)
    local = [D.select_views(cot, mine)]       # the loss side hands each rank the cotangents of its own views

    def step():
        return D.sharded_step(sets, mine, rv, local[0], capacity=cap[0], slab=slab, local_cotangents=True, label=label)

    # first pass: per-view instance counts -> cost-aware partition (deterministic on every rank)
    out, _ = step()
    torch.cuda.synchronize()
    emu = args.emulate_world if world == 1 else 0
    if world > 1 or emu:
        # tile-list lengths per (view, row of tiles) -> contiguous runs of equal cost: whole views where possible, bands
        # of tile rows where a cut falls inside a view (27 views do not divide over 8 ranks)
        import torch.distributed as dist
        ty = (IMG_H + 15) // 16
        costs = torch.zeros(V, ty, device=dev, dtype=torch.float64)
        if mine:
            costs[mine] = RZ.tile_row_costs(out[5])
        if world > 1:
            dist.all_reduce(costs)
        if args.whole_views:
            parts = D.partition_views(costs.sum(dim=1).tolist(), emu or world)
        else:
            # per-view work that does not shrink with the band (preprocess, depth sort, instance count): ~12 % of a view
            parts = D.partition_units(costs.tolist(), emu or world, fixed_cost=0.12 * float(costs.sum()) / V)
        mine = parts[args.emulate_rank if emu else rank]
        local[0] = D.select_views(cot, mine)
        # feedback: every rank times its own share (graph replay, no collective), the row costs are rescaled by the
        # measured / modelled ratio of the rank that ran them, and the line is cut again
        for _ in range(0 if (args.whole_views or emu) else args.rebalance):
            out, _ = step()
            c0 = int(out[5].num_rendered * 1.05) + 4096 if out is not None else 0
            g0 = D.GraphedStep(sets, mine, rv, local[0], capacity=c0, reduce=False, label=label)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.replay()
            torch.cuda.synchronize()
            ev0.record()
            for _k in range(5):
                g0.replay()
            ev1.record()
            torch.cuda.synchronize()
            t = torch.zeros(world, device=dev, dtype=torch.float64)
            t[rank] = ev0.elapsed_time(ev1) / 5
            dist.all_reduce(t)
            del g0
            fixed = 0.12 * float(costs.sum()) / V
            costs = torch.tensor(D.rescale_row_costs(costs.tolist(), parts, t.tolist(), fixed), device=dev,
                                 dtype=torch.float64)
            parts = D.partition_units(costs.tolist(), world, fixed_cost=fixed)
            mine = parts[rank]
            local[0] = D.select_views(cot, mine)
            rebalance_ms.append([round(x, 3) for x in t.tolist()])
    # one synchronising step with the final partition tells the instance count; from then on the binning workspace is
    # sized for 1.05x that count and no step touches the host (overflow is checked after the timed region)
    out, _ = step()
    if out is not None:
        cap[0] = int(out[5].num_rendered * 1.05) + 4096
    sampler = ClockSampler(dev.index or 0) if rank == 0 else None      # NVML initialised before the warm-up
    graphed = None
    if not args.no_graph:
        # the whole step (forward, backward, all-reduce) as ONE CUDA graph launch: same kernels, no host work per step
        graphed = D.GraphedStep(sets, mine, rv, local[0], capacity=cap[0] if cap[0] is not None else 0, label=label,
                                exchange=args.exchange)
        eager_step = step

        def step():                                            # noqa: F811
            return graphed.replay()
    for _ in range(args.warmup):
        step()
    barrier(world)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(world)
    e0.record()
    t_host = time.perf_counter()
    for _ in range(args.steps):
        out, _ = step()
    t_host = (time.perf_counter() - t_host) * 1e3 / max(args.steps, 1)       # host enqueue time per step (diagnostic)
    e1.record()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    clocks = sampler.stop() if rank == 0 else None

    exchange = None
    if world > 1:
        # what the exchange costs: the step's graph without it (each rank alone, max over ranks), and the exchange alone
        # on the slab (all ranks start together, so this is its floor -- in the step it also absorbs the ranks' skew)
        ex_slab = graphed.slab if graphed is not None else slab
        assert not (hasattr(ex_slab, "gave_up") and ex_slab.gave_up()), "a peer never arrived inside the timed region"
        g0 = D.GraphedStep(sets, mine, rv, local[0], capacity=cap[0] if cap[0] is not None else 0, reduce=False, label=label)
        x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.replay()
        barrier(world)
        x0.record()
        for _ in range(10):
            g0.replay()
        x1.record()
        torch.cuda.synchronize()
        no_ex = max_over_ranks(x0.elapsed_time(x1) / 10, world)
        del g0
        for _ in range(3):
            ex_slab.allreduce()
        barrier(world)
        x0.record()
        for _ in range(20):
            ex_slab.allreduce()
        x1.record()
        torch.cuda.synchronize()
        alone = max_over_ranks(x0.elapsed_time(x1) / 20, world)
        exchange = {"how": ex_slab.exchange, "bytes": int(ex_slab.flat.numel() * 4),
                    "alone_ms": round(alone, 4), "step_without_it_ms_max_over_ranks": round(no_ex, 4),
                    "exposed_in_step_ms": round(ms / max(args.steps, 1) - no_ex, 4)}
    if graphed is not None:
        assert not graphed.overflowed(), "instance capacity exceeded inside the timed region"
        step = eager_step
    # per-stage device times, averaged over a few extra synchronised eager steps (events recorded by the library)
    _lib.stage_timing(True)
    stages = {}
    nrep = 5
    for _ in range(nrep):
        out, _ = step()
        torch.cuda.synchronize()
        for k, v in _lib.read_stage_timing().items():
            stages[k] = stages.get(k, 0.0) + v / nrep
    _lib.stage_timing(False)
    st = out[5] if out is not None else None
    pairs = 0
    R = 0
    if st is not None:
        n_now, overflowed = RZ.read_counters(st)
        assert not overflowed, "instance capacity exceeded inside the timed region"
        st.num_rendered = n_now
        R1 = int(st.geom[12:16].view(torch.int32).item())          # counters[3]: supertile instances
        L = RZ.workspace_layout(st.P, st.V, st.W, st.H, st.channels, 0, st.capacity)
        ncon = st.image[L.n_contrib: L.n_contrib + st.V * st.W * st.H * 4].view(torch.int32)
        pairs = int(ncon.sum(dtype=torch.int64).item())
        R = st.num_rendered
        vis = float((st.radii > 0).float().mean().item())
    res = dict(ms=ms, host_ms=t_host, visible=vis if st is not None else 0.0, stages=stages, pairs=pairs, R=R, R1=R1 if st is not None else 0, clocks=clocks, views_per_rank=[round(sum(1.0 if isinstance(u, int) else (u[2] - u[1]) / float((IMG_H + 15) // 16) for u in p), 2) for p in parts], parts=parts, rebalance_ms=rebalance_ms, exchange=exchange,
               launches=kernels_per_step(L, world > 1, st.V) * args.steps if st is not None else 0, state=st, dev=dev,
               rv=rv, sets=sets, sc=sc, lib=lib, cot=cot)
    # correctness signal that survives sharding: the all-reduced gradient slab of the LAST (eager, stage-timed) step is the
    # same quantity for every N (same scene, same cotangents); the driver can compare these across N = 1, 2, 4, 8
    f64 = slab.flat.double()
    res["grad_check"] = {"l2": float(f64.norm().item()), "sum": float(f64.sum().item()),
                         "abs_sum": float(f64.abs().sum().item()), "n": int(f64.numel()),
                         "what": "float64 norms of the all-reduced [14 x P] gradient slab (means3D | colours | opacity | "
                                 "scales | rotations); equal across N up to summation order (<= 1e-5 relative)"}
    return res


def run_e2e(args, res, world, rank):
    """Same metric through the public autograd API with HOST inputs: every step copies the Gaussian parameter slab
    (pinned) and the uint8 target images (pinned) to the device, renders the rank's views, takes an L1 loss against the
    targets, back-propagates, all-reduces the gradients and reads the loss and the gradient slab back to the host.
    The copies run on their own streams (double-buffered inputs), overlapping the neighbouring steps' compute.

    At N > 1 the job still moves every input and result across PCIe ONCE per step: rank r uploads slice r of the
    parameter slab and the ranks all-gather it over NVLink (8 processes uploading the same 16.8 MB and downloading the
    same gradients measured 23 / 12 GB/s per rank, the host side of the box being shared), and rank r reads slice r of
    the all-reduced gradient slab back."""
    from dynamic3dgaussians_b200 import dist as D, rasterizer as RZ, train_ops as TO
    dev, sets, sc = res["dev"], res["sets"], res["sc"]
    parts = res["parts"]
    mine = parts[rank]
    my_views, my_bands = D.split_units(mine)
    my_sets = [sets[v] for v in my_views]
    # a band unit's loss covers its own rows only (the rest of that image belongs to another rank)
    row_mask = None
    if my_bands is not None:
        row_mask = torch.zeros(len(mine), 1, IMG_H, 1, device=dev)
        for i, b in enumerate(my_bands):
            r0, r1 = (0, IMG_H) if b is None else (b[0] * 16, min(b[1] * 16, IMG_H))
            row_mask[i, :, r0:r1] = 1.0
        row_mask *= 1.0 / (3 * IMG_W * row_mask.sum())
    row_weight = None if row_mask is None else row_mask.reshape(len(mine), IMG_H).contiguous()
    rv_host = sc.rendervar()
    P = rv_host["means3D"].shape[0]
    names = ("means3D", "colors_precomp", "opacities", "scales", "rotations")
    # the five parameter tensors travel as ONE pinned slab (14 floats per Gaussian), and so do their gradients: one
    # DMA each way per step and one all-reduce, instead of five of each
    shapes = [tuple(rv_host[k].shape) for k in names]
    sizes = [rv_host[k].numel() for k in names]
    import torch.distributed as tdist
    nw = world if (world > 1 and tdist.is_initialized()) else 1
    drank = tdist.get_rank() if nw > 1 else 0
    n_par = sum(sizes)
    chunk = -(-n_par // (4 * nw)) * 4                              # whole float4s: the peer kernel's slice of a rank
    lo, hi = drank * chunk, (drank + 1) * chunk
    par_full = torch.zeros(chunk * nw)
    par_full[:n_par] = torch.cat([rv_host[k].reshape(-1).float() for k in names])
    par_host = par_full[lo:hi].clone().pin_memory()                # this rank's slice of the slab
    tgt_host = torch.randint(0, 256, (len(mine), 3, IMG_H, IMG_W), dtype=torch.uint8,
                             generator=torch.Generator().manual_seed(5)).pin_memory()
    grad_host = torch.empty_like(par_host).pin_memory()
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
    h2d = par_host.numel() * 4 + tgt_host.numel()
    d2h = grad_host.numel() * 4 + 4

    states = []
    cap_e2e = int(res["R"] * 1.05) + 4096 if mine else None
    # A three-stage pipeline like a training input pipeline: the copy engine uploads step i+1's inputs (H2D stream)
    # while step i renders and back-propagates (compute stream) and step i-1's gradient slab and loss go back to the
    # host (D2H stream).  Every step still copies ITS inputs in and ITS results out inside the timed region.
    cur = torch.cuda.current_stream(dev)
    h2d_s, d2h_s = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    # N > 1: the parameter slab is gathered and the gradient slab reduce-scattered by the library's kernel over peer
    # memory (dist.PeerBuffer), NCCL when symmetric memory is not available or --exchange nccl
    peer = nw > 1 and args.exchange != "nccl"
    bufs = []
    for _ in range(2):
        ppar = pgrad = None
        if peer:
            try:
                ppar, pgrad = D.PeerBuffer(chunk * nw, dev), D.PeerBuffer(chunk * nw, dev)
            except Exception:                                        # noqa: BLE001 -- same outcome on every rank
                peer, ppar, pgrad = False, None, None
        bufs.append(dict(par=ppar.t if ppar else torch.zeros(chunk * nw, dtype=torch.float32, device=dev), ppar=ppar,
                         pgrad=pgrad, tgt=torch.empty(tgt_host.shape, dtype=torch.uint8, device=dev),
                         ready=torch.cuda.Event(), free=torch.cuda.Event()))
    for b in bufs:
        b["free"].record(cur)

    def prefetch(i, gate=None):
        b = bufs[i % 2]
        with torch.cuda.stream(h2d_s):
            h2d_s.wait_event(b["free"])
            if gate is not None:
                h2d_s.wait_event(gate)
            b["par"][lo:hi].copy_(par_host, non_blocking=True)
            b["tgt"].copy_(tgt_host, non_blocking=True)
            b["ready"].record(h2d_s)

    def gather(b):
        if b["ppar"] is not None:
            b["ppar"].allgather()                                           # in place: every rank now holds the slab
        elif nw > 1:
            tdist.all_gather_into_tensor(b["par"], b["par"][lo:hi])

    def compute(b):
        gather(b)
        return compute_rest(b)

    def compute_rest(b):
        m, c, o, s, r = b["leaves"]
        color, _, depth, alpha, radii = RZ.rasterize_views(my_sets, m, o, colors_precomp=c, scales=s, rotations=r,
                                                           capacity=cap_e2e, state_out=states, bands=my_bands)
        del states[:-1]
        # mean |color - target / 255| straight from the uint8 frames (one fused kernel each way; with bands, weighted by
        # the rows this rank owns)
        loss = TO.l1_loss_u8(color, b["tgt"], row_weight)
        loss.backward()
        torch.cat([t.grad.reshape(-1) for t in b["leaves"]], out=b["grad"][:n_par])
        if b["pgrad"] is not None:
            b["pgrad"].reduce_scatter()                                     # this rank reads back its own slice only
        else:
            D.allreduce_grads(b["grad"])
        return loss.detach().reshape(1), b["grad"][lo:hi]

    for b in bufs:
        # leaves share storage with the upload target
        b["leaves"] = [t.view(sh).detach().requires_grad_(True) for t, sh in zip(b["par"][:n_par].split(sizes), shapes)]
        b["grad"] = b["pgrad"].t if b["pgrad"] is not None else torch.zeros_like(b["par"])
        b["drained"] = torch.cuda.Event()
        b["drained"].record(cur)
    if not args.no_graph:
        # the compute of a step (render, loss, backward, all-reduce) is ONE graph per input buffer: at N>1 a rank's GPU
        # work is ~1 ms and the ~2.5 ms of Python/autograd dispatch per eager step would bound the pipeline
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for b in bufs:
                b["par"][lo:hi].copy_(par_host, non_blocking=True)
                b["tgt"].copy_(tgt_host, non_blocking=True)
            for _ in range(2):
                for t in bufs[0]["leaves"]:
                    t.grad = None
                compute(bufs[0])
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        pool = None
        for b in bufs:
            for t in b["leaves"]:
                t.grad = None
            if nw > 1:
                b["g_gather"] = torch.cuda.CUDAGraph()                 # the all-gather is a graph of its own: see run()
                with torch.cuda.graph(b["g_gather"], pool=pool):
                    gather(b)
                pool = b["g_gather"].pool() if pool is None else pool
            b["graph"] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(b["graph"], pool=pool):
                b["out"] = compute_rest(b)
            pool = b["graph"].pool() if pool is None else pool

    def download(b, done, lossd, grads, gate=None):
        with torch.cuda.stream(d2h_s):
            d2h_s.wait_event(done)
            if gate is not None:
                d2h_s.wait_event(gate)
            grad_host.copy_(grads, non_blocking=True)
            loss_host.copy_(lossd, non_blocking=True)
            b["drained"].record(d2h_s)
        if "graph" not in b:
            lossd.record_stream(d2h_s)

    def run(k):
        # N = 1: upload of step i+1 and download of step i-1 simply run next to step i.  N > 1: the exchange kernels'
        # barriers use system-scope fences, and a system-scope fence does not return while a DMA copy is in flight
        # (tests/gpu_fence_dma_probe.py: the exchange takes 30 us alone, 65 us next to a D2H and 143 us next to an H2D
        # copy; NCCL's kernels behave the same) -- copies that start when a step ends land exactly on the next step's
        # all-gather.  So the copies are released by an event recorded AFTER the all-gather of the running step: they
        # overlap the render, which issues no system-scope fence until its last kernel, a whole render later.
        prefetch(0)
        pending = None                               # the previous step's results, not yet on their way to the host
        for i in range(k):
            b = bufs[i % 2]
            cur.wait_event(b["ready"])
            cur.wait_event(b["drained"])             # this buffer's previous results have left for the host
            gate = None
            if nw > 1:
                if "g_gather" in b:
                    b["g_gather"].replay()
                else:
                    gather(b)
                gate = torch.cuda.Event()
                gate.record(cur)
            if pending is not None:
                download(*pending, gate=gate)
            if i + 1 < k:
                prefetch(i + 1, gate=gate)
            if "graph" in b:
                b["graph"].replay()
                lossd, grads = b["out"]
            else:
                for t in b["leaves"]:
                    t.grad = None
                lossd, grads = compute_rest(b)
            b["free"].record(cur)
            done = torch.cuda.Event()
            done.record(cur)
            pending = (b, done, lossd, grads)
            if nw == 1:                              # nothing to stay clear of: the download starts as soon as the step ends
                download(*pending)
                pending = None
        if pending is not None:
            download(*pending)
        cur.wait_stream(d2h_s)
        cur.wait_stream(h2d_s)

    run(max(2, args.warmup // 2))
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = max(3, args.steps // 2)
    e0.record()
    t_host = time.perf_counter()
    run(n)
    t_host = (time.perf_counter() - t_host) * 1e3 / n        # host time to ENQUEUE a step (diagnostic: must stay below ms/step)
    e1.record()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    if states and cap_e2e is not None:
        assert not RZ.read_counters(states[-1])[1], "instance capacity exceeded inside the e2e region"
    mpix = args.views * IMG_W * IMG_H * n / (ms * 1e-3) / 1e6
    loss_value = float(loss_host.item())
    if nw > 1:                                     # bytes per step of the whole job (sum over the ranks)
        t = torch.tensor([h2d, d2h], device=dev, dtype=torch.float64)
        tdist.all_reduce(t)
        h2d, d2h = int(t[0].item()), int(t[1].item())

    # the three legs of the pipeline timed one at a time (diagnostic, outside the timed region; this rank's numbers)
    def leg(fn, reps=5):
        torch.cuda.synchronize(dev)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        return round(e0.elapsed_time(e1) / reps, 3)

    def leg_h2d():
        prefetch(0)
        cur.wait_stream(h2d_s)

    def leg_compute():
        b = bufs[0]
        if "graph" in b:
            if "g_gather" in b:
                b["g_gather"].replay()
            b["graph"].replay()
        else:
            for t in b["leaves"]:
                t.grad = None
            compute(b)

    def leg_d2h():
        with torch.cuda.stream(d2h_s):
            d2h_s.wait_stream(cur)
            grad_host.copy_(bufs[0]["grad"][lo:hi], non_blocking=True)
        cur.wait_stream(d2h_s)

    legs = {"h2d_ms": leg(leg_h2d), "compute_ms": leg(leg_compute), "d2h_ms": leg(leg_d2h)}
    return {"value": round(mpix, 2), "unit": "Mpix/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "steps": n, "ms_per_step": round(ms / n, 3), "host_enqueue_ms_per_step_rank0": round(t_host, 3),
            "loss_rank0": loss_value, "legs_alone_rank0": legs,
            "pipeline": ("copies of step i+1 / i-1 run next to step i"
                         + ("; they are released after step i's all-gather, because a system-scope fence (every collective "
                            "kernel has them) does not return while a DMA copy is in flight" if nw > 1 else "")),
            "host_io": "bytes are job totals; parameter slab uploaded / gradient slab read back once per step "
                       + (f"(1/{nw} per rank; all-gather of the parameters and reduce-scatter of the gradients over NVLink by "
                          + ("the library's peer-memory kernel)" if peer else "NCCL)") if nw > 1 else "(one rank)")}


def run_train_loop(args, res):
    """train it/s (the second half of BASELINE.json's metric): the train.py-shaped loop of
    dynamic3dgaussians_b200/train_loop.py at C2 -- ONE random view per optimizer step (REF/train.py:422-438), fused
    colour + segmentation render, fused L1+SSIM losses against random target images, Adam over the reference's eight
    parameter groups.  Densification does not fire inside the sample (its schedule starts at iteration 500)."""
    from dynamic3dgaussians_b200 import train_loop as TL
    dev, sets, sc = res["dev"], res["sets"], res["sc"]
    g = torch.Generator().manual_seed(11)
    rv = res["rv"]
    seg = sc.seg_colors.to(dev)[:, 0].contiguous()
    params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, len(sets), 3.3)
    opt = TL.initialize_optimizer(params, variables)
    data = [{'cam': s_, 'id': i,
             'im': torch.rand(3, IMG_H, IMG_W, generator=g).to(dev), 'seg': torch.rand(3, IMG_H, IMG_W, generator=g).to(dev)}
            for i, s_ in enumerate(sets)]
    cap = int(2.5 * res["R"] / max(len(sets), 1)) + 65536          # per-view instance capacity: no host sync per step
    TL.train_timestep(params, variables, opt, data, 30, True, generator=g, do_densify=False, capacity=cap)
    torch.cuda.synchronize()
    n = 150
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    hist = TL.train_timestep(params, variables, opt, data, n, True, generator=g, do_densify=False, capacity=cap)
    e1.record()
    torch.cuda.synchronize()
    ms_eager = e0.elapsed_time(e1)
    out = {"config": "C2, 1 view 640x360 per step, colour+seg fused render, fused L1+SSIM losses, fused Adam (8 groups), no densify",
           "eager_it_per_s": round(n / (ms_eager * 1e-3), 1), "readme_reference_it_per_s": 100.0}
    # the same iteration as one CUDA graph per camera (what the loop uses once the point count is fixed)
    params, variables = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, len(sets), 3.3)
    opt = TL.initialize_optimizer(params, variables)
    trainer = TL.GraphedTrainer(params, variables, opt, data, True, cap)
    for _ in range(30):
        trainer.step(g)
    torch.cuda.synchronize()
    n = 300
    e0.record()
    for _ in range(n):
        loss = trainer.step(g)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    out.update({"it_per_s": round(n / (ms * 1e-3), 1), "ms_per_it": round(ms / n, 3), "iterations": n,
                "final_loss": round(float(loss), 5), "launch": "one CUDA graph per iteration"})
    # BASELINE.json configs[2]: a dynamic timestep (t >= 1) -- fixed point count, 20-NN rigidity / rotation / isometry
    # losses (fused kernels, GPU-built graph) + floor / background / colour-consistency terms on top of the above
    del trainer
    variables = TL.initialize_post_first_timestep(params, variables, opt)
    params, variables = TL.initialize_per_timestep(params, variables, opt)
    trainer = TL.GraphedTrainer(params, variables, opt, data, False, cap)
    for _ in range(30):
        trainer.step(g)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        loss = trainer.step(g)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    # BASELINE.json configs[1] says "train.py loop WITH densification": a sample that crosses two densification rounds
    # (iterations 450..649: rounds at 500 and 600, REF/external.py:248).  The point count changes, so this stretch runs
    # eagerly with the instance count read back every step (the reference's own host sync) and re-grown workspaces.
    dyn_loss_value = float(loss)
    del trainer
    params2, variables2 = TL.initialize_params(rv["means3D"].clone(), rv["colors_precomp"].clone(), seg, len(sets), 3.3)
    opt2 = TL.initialize_optimizer(params2, variables2)
    TL.train_timestep(params2, variables2, opt2, data, 20, True, generator=g, do_densify=False, capacity=None, start_iter=430)
    torch.cuda.synchronize()
    p_before = int(params2['means3D'].shape[0])
    nd = 200
    e0.record()
    TL.train_timestep(params2, variables2, opt2, data, nd, True, generator=g, do_densify=True, capacity=None, start_iter=450)
    e1.record()
    torch.cuda.synchronize()
    msd = e0.elapsed_time(e1)
    out["with_densification"] = {"it_per_s": round(nd / (msd * 1e-3), 1), "ms_per_it": round(msd / nd, 3),
                                 "iterations": "450..649 (densification rounds at 500 and 600)",
                                 "gaussians_before": p_before, "gaussians_after": int(params2['means3D'].shape[0]),
                                 "launch": "eager, instance count read back every step (workspaces follow the point count)"}
    del params2, variables2, opt2
    out["dynamic_timestep"] = {"it_per_s": round(n / (ms * 1e-3), 1), "ms_per_it": round(ms / n, 3),
                               "fg_gaussians": int(variables["is_fg"].sum()), "knn": 20,
                               "final_loss": round(dyn_loss_value, 5), "readme_reference_it_per_s": 50.0}
    return out


def _time_ms(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def run_extras(args, res):
    """Sub-lines of the N = 1 line (not the metric): the same step with the work the reference is FORCED to do, the C4
    variant, and the kNN / dynamic-loss kernels next to the reference's own kernel / the eager code they replace."""
    from dynamic3dgaussians_b200 import dist as D, rasterizer as RZ
    dev, sets, rv, cot, sc = res["dev"], res["sets"], res["rv"], res["cot"], res["sc"]
    P, V = rv["means3D"].shape[0], len(sets)
    cap = int(res["R"] * 1.05) + 4096
    out = {}
    pix = V * IMG_W * IMG_H
    args5 = (rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"])

    # (1) apples to apples: the vendored reference always blends 32 semantic channels (forward.cu:381-383) and scatters
    # their gradients (backward.cu:594-611); bench --impl reference feeds it zeros[P,32].  Here the repo arm does that
    # work too: semantic_feature = zeros[P,32], feature-map cotangent zeros.
    feat = torch.zeros(P, 32, device=dev)
    gfeat = torch.zeros(V, 32, IMG_H, IMG_W, device=dev)

    def with_features():
        *_, st = RZ.forward_views(sets, *args5, capacity=cap, semantic_feature=feat)
        RZ.backward_views(st, *args5, None, None, None, cot["color"], None, cot["depth"], None, want_means2D=False,
                          semantic_feature=feat, dL_dfeature=gfeat)
    ms = _time_ms(with_features, 5)
    out["apples_to_apples"] = {
        "ms_per_step": round(ms, 3), "value": round(pix / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
        "what": "the headline step PLUS the 32-channel semantic feature map forward and backward (semantic_feature = "
                "zeros[P,32], dL_dfeature_map = zeros): the algorithmic work of the reference arm, which cannot switch "
                "the feature channels off (SURVEY.md Q6); eager launches, device-timed"}
    del feat, gfeat

    # (2) C4 at N = 1 (BASELINE.json configs[3]); `--workload c4 --gpus N` runs the same thing sharded
    if args.workload == "c2":
        gen = torch.Generator().manual_seed(321)
        cot4 = dict(cot, alpha=torch.randn(V, 1, IMG_H, IMG_W, generator=gen).to(dev))
        label = torch.zeros(P, device=dev)
        label[int(0.6 * P):] = 1.0
        g4 = D.GraphedStep(sets, list(range(V)), rv, cot4, capacity=cap, label=label)
        ms = _time_ms(lambda: g4.replay(), 10)
        assert not g4.overflowed()
        masked = float(g4.slab.views["means3D"][: int(0.6 * P)].abs().max().item())
        out["c4"] = {"ms_per_step": round(ms, 3), "value": round(pix / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
                     "masked_gradient_abs_max": masked,
                     "what": "4-return mode: colour + depth + alpha cotangents, label mask = 1 on the last 40 % of the "
                             "Gaussians (gradients of the first 60 % exactly 0), one CUDA graph per step"}
        del g4, cot4

    # (3) kNN (distCUDA2, K = 3; the 20-NN rigidity graph) and the fused dynamic losses
    from dynamic3dgaussians_b200 import dyn_losses as DL, knn as KN
    pts = rv["means3D"]
    k3 = _time_ms(lambda: KN.knn(pts, 3), 10)
    k20 = _time_ms(lambda: KN.knn(pts, 20), 5)
    knn = {"points": P, "knn3_ms": round(k3, 3), "knn20_ms": round(k20, 3), "knn3_mpts_per_s": round(P / k3 / 1e3, 1)}
    try:
        from oracle import ref_cuda
        if os.path.exists(ref_cuda.KNN):
            knn["ref_knn3_ms"] = round(_time_ms(lambda: ref_cuda.knn(pts), 3, warm=1), 3)
            knn["ref"] = "the reference's simple_knn.cu compiled unmodified (oracle/_ref/libref_knn.so), same B200"
    except Exception as e:                                   # the comparator is optional, the product number is not
        knn["ref_error"] = str(e)[:200]
    out["knn"] = knn
    fg = sc.seg_colors[:, 0].to(dev) > 0.5
    fpts = pts[fg].contiguous()
    frot = rv["rotations"][fg].contiguous()
    graph = DL.build_neighbor_graph(fpts, frot, 20)
    gsd = torch.Generator().manual_seed(9)
    p2 = (fpts + 0.01 * torch.randn(fpts.shape, generator=gsd).to(dev)).requires_grad_(True)
    r2 = torch.nn.functional.normalize(frot + 0.01 * torch.randn(frot.shape, generator=gsd).to(dev)).requires_grad_(True)
    w = torch.tensor([4.0, 4.0, 2.0], device=dev)

    def fused():
        p2.grad = r2.grad = None
        (DL.rigidity_losses(p2, r2, graph) * w).sum().backward()

    idx = graph.idx.long()

    def eager():                                             # REF/train.py:253-273 in plain torch on the same GPU
        p2.grad = r2.grad = None
        q1, q2 = r2, graph.prev_inv_rot
        w1, x1, y1, z1 = q1.T
        w2, x2, y2, z2 = q2.T
        rel = torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                           w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2]).T
        q = rel / rel.norm(dim=1, keepdim=True)
        r, x, y, z = q.unbind(1)
        Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y), 2 * (x * y + r * z),
                          1 - 2 * (x * x + z * z), 2 * (y * z - r * x), 2 * (x * z - r * y), 2 * (y * z + r * x),
                          1 - 2 * (x * x + y * y)], dim=1).reshape(-1, 3, 3)
        off = p2[idx] - p2[:, None]
        c = (Rm.transpose(2, 1)[:, None] @ off[:, :, :, None]).squeeze(-1)
        l2 = lambda a, b, ww: torch.sqrt(((a - b) ** 2).sum(-1) * ww + 1e-20).mean()
        L = torch.stack([l2(c, graph.prev_offset, graph.weight), l2(rel[idx], rel[:, None], graph.weight),
                         torch.sqrt((torch.sqrt((off ** 2).sum(-1) + 1e-20) - graph.dist) ** 2 * graph.weight + 1e-20).mean()])
        (L * w).sum().backward()
    out["dyn_losses"] = {"fg_gaussians": int(fpts.shape[0]), "knn": 20, "fused_ms": round(_time_ms(fused, 20), 3),
                         "eager_torch_ms": round(_time_ms(eager, 10), 3),
                         "what": "rigidity + rotation + isometry losses, forward + backward; eager = the reference's "
                                 "expressions (REF/train.py:253-273) as plain torch ops on the same B200; the "
                                 "reference's README quotes ~20 ms / iteration (REF/README.md:90)"}
    return out


def cpu_baseline_port(sample_P=10_000):
    """The reference has NO CPU implementation of this path (it is CUDA-only), so the CPU baseline is the oracle port
    (oracle/eager_rasterizer.py) on config C1: 10k Gaussians, one 256x256 view, forward + autograd backward, all host
    threads.  Bounded: one un-timed warm-up of 1k Gaussians, then timed passes for ~12 s."""
    from dynamic3dgaussians_b200 import scenes
    from oracle import eager_rasterizer as O
    torch.set_num_threads(os.cpu_count() or 1)
    cs = scenes.camera_c1()
    cam = O.camera_from_spec(cs.w, cs.h, cs.k, cs.w2c)

    def once(P):
        sc = scenes.make_scene_c1(P)
        leaves = {k: v.clone().requires_grad_(True) for k, v in sc.rendervar().items()}
        t0 = time.time()
        out = O.rasterize(cam, leaves["means3D"], leaves["opacities"], colors_precomp=leaves["colors_precomp"],
                          scales=leaves["scales"], rotations=leaves["rotations"])
        (out["color"].sum() + out["depth"].sum()).backward()
        return time.time() - t0
    once(1000)
    # bounded sample: whole passes until ~12 s of CPU work are spent (at least 2, at most 12 passes)
    dts = []
    while len(dts) < 2 or (sum(dts) < 12.0 and len(dts) < 12):
        dts.append(once(sample_P))
    dt = sum(dts) / len(dts)
    return {"value": round(cs.w * cs.h / dt / 1e6, 5), "unit": "Mpix/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"config C1: {sample_P} Gaussians, 1 view {cs.w}x{cs.h}, fwd+bwd, mean of {len(dts)} passes, "
                      f"{sum(dts):.1f} s of CPU work "
                      f"(the reference rasterizer is CUDA-only; this is the PyTorch-eager oracle of its semantics)"}


def run_reference(args, world, rank):
    """Reference arm: the reference's own CUDA rasterizer (oracle/_ref, compiled unmodified from /root/reference),
    one view at a time as its API works, same scene / cameras / cotangents.  The reference has no CPU path; if the
    library is absent the CPU oracle port is timed instead."""
    if rank != 0:
        return None
    from oracle import ref_cuda
    # the reference has no multi-GPU path: under torchrun rank 0 alone runs it, on ONE GPU, and says so
    base = {"metric": METRIC, "unit": "Mpix/s", "n_gpus": 1, "requested_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "impl": "reference",
            "config": {"workload": f"C2: {args.gaussians} Gaussians, {args.views} views {IMG_W}x{IMG_H}, fwd+bwd, "
                                   "colour+depth cotangents"}}
    if not (torch.cuda.is_available() and ref_cuda.available()):
        cb = cpu_baseline_port()
        base.update(value=cb["value"], ms_per_step=None, cpu_baseline=cb,
                    e2e={"value": cb["value"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        base["config"]["workload"] = cb["sample"]
        return base
    dev = torch.device("cuda", 0)
    sc, sets = build_workload(args, dev)
    rv = {k: v.to(dev).contiguous() for k, v in sc.rendervar().items()}
    V = len(sets)
    gen = torch.Generator().manual_seed(123)
    gc = torch.randn(V, 3, IMG_H, IMG_W, generator=gen).to(dev)
    gd = torch.randn(V, 1, IMG_H, IMG_W, generator=gen).to(dev) * 0.1
    ga = torch.zeros(1, IMG_H, IMG_W, device=dev)
    rrs = [ref_cuda.RefRasterizer(s) for s in sets]
    sem = torch.zeros(rv["means3D"].shape[0], 32, device=dev)     # mandatory in this variant

    def step():
        R = 0
        for v, rr in enumerate(rrs):
            ro = rr.forward(rv["means3D"], rv["opacities"], rv["colors_precomp"], rv["scales"], rv["rotations"], sem)
            rr.backward(gc[v], gd[v], ga)
            R += ro["num_rendered"]
        return R
    for _ in range(max(1, args.warmup)):
        step()
    torch.cuda.synchronize()
    n = max(1, args.steps)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        R = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    mpix = V * IMG_W * IMG_H * n / (ms * 1e-3) / 1e6
    base.update(value=round(mpix, 2), steps=n, ms_per_step=round(ms / n, 3),
                cpu_baseline={"value": round(mpix, 2), "unit": "Mpix/s", "cores": 0, "kind": "reference",
                              "sample": "the reference's own CUDA kernels on the same B200 (it has no CPU "
                                        "implementation); 27 sequential single-view forward+backward calls through "
                                        "CudaRasterizer::Rasterizer, buffers pre-grown, semantic_feature = zeros[P,32]"},
                e2e={"value": round(mpix, 2), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                num_rendered=R)
    return base


def main():
    global IMG_W, IMG_H, METRIC
    args = parse()
    if args.workload == "c5":
        IMG_W, IMG_H = 1920, 1080
        args.gaussians = 2_000_000
        args.views = 8 * int(os.environ.get("WORLD_SIZE", "1"))
        args.whole_views = args.no_train = args.no_cpu_baseline = args.no_extras = True
        METRIC = "fwd+bwd rasterized Mpix/s @2M Gaussians, 8x1920x1080 views per GPU"
    if args.impl == "reference":
        world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
        line = run_reference(args, world, rank)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")
    world, rank, local = dist_setup(args)
    res = run_ours(args, world, rank, local)
    e2e = None if args.no_e2e else run_e2e(args, res, world, args.emulate_rank if args.emulate_world and world == 1 else rank)
    if rank == 0:
        steps, ms = args.steps, res["ms"]
        V = args.views
        mpix = V * IMG_W * IMG_H * steps / (ms * 1e-3) / 1e6
        stages = res["stages"]
        dom = max(stages, key=stages.get) if stages else None
        hbm_gbs, hbm_src = measured_peaks()
        fp32 = fp32_peak_tflops(res["lib"], res["dev"])
        pairs, R, P = res["pairs"], res["R"], args.gaussians
        st = res["state"]
        Vl = st.V if st is not None else 0
        # algorithmic work per launch of each stage on THIS rank (SURVEY.md 8d constants)
        from dynamic3dgaussians_b200 import rasterizer as RZ
        L = RZ.workspace_layout(P, max(Vl, 1), IMG_W, IMG_H, 3, 0, max(R, 1))
        R1 = res.get("R1", 0)
        work = {
            "blend_backward": ("fp32", 80.0 * pairs), "blend_forward": ("fp32", 27.0 * pairs),
            "supertile_sort": ("hbm", 16.0 * L.tile_passes * R1),
            # first pass reads every (key, index) pair and writes the kept ones (Gaussians that touch a tile); the other
            # passes move the kept pairs; the digit histograms come from the preprocess kernel (no extra read)
            "depth_sort": ("hbm", (8.0 + 8.0 * res.get("visible", 1.0) + 16.0 * (L.depth_passes - 1) * res.get("visible", 1.0)) * Vl * P),
            "preprocess": ("hbm", (44.0 + 56.0 * Vl) * P), "scatter_records": ("hbm", (4.0 + 8.0 + 48.0) * R1 + 48.0 * R),
            "emit_supertile_instances": ("hbm", 12.0 * Vl * P + 8.0 * R1), "tile_offsets": ("hbm", (4.0 + 4.0 + 8.0) * R1),
            "grad_zero": ("hbm", 48.0 * Vl * P), "preprocess_backward": ("hbm", (44.0 + 56.0) * P + 68.0 * Vl * P),
        }
        # dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed ncu --set full
        # capture of this same command (profiles/; refreshed whenever the kernel changes) -- never measured under ncu here
        traffic = None
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")) as f:
                tj = json.load(f)
            if world == 1 and dom in tj.get("bytes_per_launch", {}) and tj.get("views") == V and tj.get("gaussians") == P:
                traffic = tj["bytes_per_launch"][dom]
        except (OSError, ValueError):
            pass
        roof = None
        if dom in work:
            kind, amount = work[dom]
            t = stages[dom] * 1e-3
            if kind == "fp32":
                ach = amount / t / 1e12
                roof = {"kernel": dom, "bound": "fp32", "achieved": round(ach, 3), "peak": round(fp32["best"], 2),
                        "unit": "TFLOP/s", "frac": round(ach / fp32["best"], 4),
                        "frac_of_nominal": round(ach / fp32["nominal"], 4), "traffic_from_committed_ncu": traffic,
                        "peak_source": "measured in this run: dependent-free FMA loops, 128 FMAs per trip, scalar FFMA "
                                       f"{fp32['ffma']:.1f} and packed FFMA2 {fp32['ffma2']:.1f} TFLOP/s (the higher is "
                                       f"the denominator); nominal {fp32['nominal']:.1f}",
                        "note": "achieved = the reference algorithm's 80 flop per examined pair / stage time: credit "
                                "for pairs this kernel skips, not executed FMAs (ncu: issue-slot bound)"}
            else:
                ach = amount / t / 1e9
                roof = {"kernel": dom, "bound": "hbm", "achieved": round(ach, 1), "peak": hbm_gbs, "unit": "GB/s",
                        "frac": round(ach / hbm_gbs, 4), "traffic_from_committed_ncu": traffic, "peak_source": hbm_src}
        stage_roofline = {}
        for k, v in stages.items():
            if k in work and v > 0:
                kind, amount = work[k]
                if kind == "fp32":
                    a = amount / (v * 1e-3) / 1e12
                    stage_roofline[k] = {"bound": "fp32", "achieved": round(a, 2), "unit": "TFLOP/s",
                                         "frac": round(a / fp32["best"], 3)}
                else:
                    a = amount / (v * 1e-3) / 1e9
                    stage_roofline[k] = {"bound": "hbm", "achieved": round(a, 1), "unit": "GB/s",
                                         "frac": round(a / hbm_gbs, 3)}
        ideal_ms = sum((work[k][1] / (fp32["best"] * 1e12) if work[k][0] == "fp32" else work[k][1] / (hbm_gbs * 1e9)) * 1e3
                       for k in stages if k in work)
        line = {
            "metric": METRIC, "value": round(mpix, 2), "unit": "Mpix/s", "n_gpus": world, "steps": steps,
            "warmup": args.warmup, "ms_per_step": round(ms / steps, 3), "higher_is_better": True,
            "scaling": "weak" if args.workload == "c5" else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload.upper()}: {P} Gaussians, {V} views {IMG_W}x{IMG_H}, fwd+bwd, "
                                   + ("colour+depth+alpha cotangents, label mask on the last 40 %" if args.workload == "c4"
                                      else "colour+depth cotangents"),
                       "views_per_rank": res["views_per_rank"],
                       **({"rank_ms_before_each_rebalance": res["rebalance_ms"]} if res["rebalance_ms"] else {}), "parallelism": f"view-sharded dp{world}" + ("" if world == 1 or args.whole_views else " (cuts inside a view fall on tile-row bands)"),
                       "launch": "eager" if args.no_graph else "one CUDA graph per step (forward + backward + exchange)",
                       **({"exchange": res["exchange"]["how"]} if res.get("exchange") else {}),
                       "l2": "per-step working set (>1 GB of per-view state and instance records) exceeds the 126 MB L2",
                       "num_rendered_rank0": R, "visible_fraction_rank0": round(res.get("visible", 0.0), 4), "supertile_instances_rank0": res.get("R1", 0), "pairs_rank0": pairs},
            "clocks": res["clocks"], "e2e": e2e, "gpu_launches": res["launches"],
            "roofline": roof,
            **({"diagnostic": f"NOT a result: only rank {args.emulate_rank}'s share of a {args.emulate_world}-way partition "
                              "ran (value counts all views)"} if args.emulate_world else {}),
            "stage_ms_rank0": {k: round(v, 4) for k, v in stages.items()},
            "stage_roofline_rank0": stage_roofline,
            "step_vs_ideal_machine": {"ideal_ms": round(ideal_ms, 3), "frac": round(ideal_ms / (ms / steps), 3),
                                      "what": "sum over stages of (this build's algorithmic work / measured peak) "
                                              "divided by the measured step"},
            "grad_check": res.get("grad_check"),
            **({"exchange": res["exchange"]} if res.get("exchange") else {}),
            "host_enqueue_ms_per_step_rank0": round(res["host_ms"], 3),
            "train_it_per_s_equiv": round(V * steps / (ms * 1e-3), 1),
        }
        if world == 1 and not args.no_extras and not args.emulate_world:
            line["extras"] = run_extras(args, res)
        if world == 1 and not args.no_train:
            line["train"] = run_train_loop(args, res)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_port()
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
