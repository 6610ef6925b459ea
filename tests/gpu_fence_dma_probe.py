"""Diagnostic (not a test): does a concurrent DMA copy (H2D / D2H on another stream) slow the exchange kernel, whose
barriers use system-scope fences?  Two emulated ranks on ONE GPU (two streams, tests/test_gpu_peer.py) run the all-reduce
back to back while a third stream copies pinned host memory to / from the device.  python tests/gpu_fence_dma_probe.py"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from dynamic3dgaussians_b200 import _lib  # noqa: E402


def main():
    lib = _lib.load()
    world, n = 2, 14 * 300_000
    bufs = [torch.randn(n, device="cuda") for _ in range(world)]
    flags = [torch.zeros(_lib.PEER_FLAG_WORDS, dtype=torch.int32, device="cuda") for _ in range(world)]
    bp = (C.c_void_p * world)(*[b.data_ptr() for b in bufs])
    fp = (C.c_void_p * world)(*[f.data_ptr() for f in flags])
    streams = [torch.cuda.Stream() for _ in range(world)]
    copy_s = torch.cuda.Stream()
    host = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    dev = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")

    def exchanges(k):
        for _ in range(k):
            for r in range(world):
                _lib.check(lib.gsr_peer_allreduce(0, world, r, bp, fp, None, n, 16, streams[r].cuda_stream), "peer")

    def timed(copy):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if copy:
            with torch.cuda.stream(copy_s):
                for _ in range(40):
                    if copy == "h2d":
                        dev.copy_(host, non_blocking=True)
                    else:
                        host.copy_(dev, non_blocking=True)
        e0.record(streams[0])
        exchanges(50)
        e1.record(streams[0])
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 50 * 1e3

    exchanges(5)
    for mode in (None, "h2d", "d2h", None):
        print("concurrent copy:", mode, f"-> {timed(mode):.1f} us per exchange (two emulated ranks, 16 CTAs each)")
    # the same for a plain kernel without system-scope fences: a device-to-device copy kernel
    a = torch.empty(n, device="cuda")

    def timed_plain(copy):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if copy:
            with torch.cuda.stream(copy_s):
                for _ in range(40):
                    dev.copy_(host, non_blocking=True) if copy == "h2d" else host.copy_(dev, non_blocking=True)
        with torch.cuda.stream(streams[0]):
            e0.record(streams[0])
            for _ in range(50):
                a.mul_(1.0001)
            e1.record(streams[0])
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 50 * 1e3
    for mode in (None, "h2d", "d2h"):
        print("plain elementwise kernel, concurrent copy:", mode, f"-> {timed_plain(mode):.1f} us")


if __name__ == "__main__":
    main()
