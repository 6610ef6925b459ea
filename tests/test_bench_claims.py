"""CPU checks that what bench.py CLAIMS agrees with the evidence committed under profiles/ (no GPU needed):
the `gpu_launches` figure against the ncu launch list of the same command, the committed bench lines against the bench
contract's keys, and the DRAM-traffic table against the stage names the bench line uses."""
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import ROOT  # noqa: E402

sys.path.insert(0, ROOT)
import bench  # noqa: E402
from dynamic3dgaussians_b200 import rasterizer as RZ  # noqa: E402

PROFILES = os.path.join(ROOT, "profiles")


def _line(name):
    with open(os.path.join(PROFILES, name)) as f:
        return json.loads(f.read().strip().splitlines()[-1])


def test_launch_count_claim_equals_the_committed_ncu_launch_list():
    # the list was taken over 10 forward+backward passes of the bench command (2 steps + warm-up + the stage-timed passes)
    ours = {}
    for ln in open(os.path.join(PROFILES, "r02_launches_c2_final.txt")):
        m = re.match(r"(\S.*?)\s+n=\s*(\d+)\s+sum=", ln)
        if not m or m.group(1).startswith(("at::", "gsr::fp32_peak")):
            continue
        ours[m.group(1)] = int(m.group(2))
    assert ours, "no kernels parsed"
    passes = ours["preprocess_kernel"]
    per_step = sum(ours.values()) / passes
    L = RZ.workspace_layout(300_000, 27, 640, 360, 3, 0, 45_000_000)
    assert per_step == bench.kernels_per_step(L, False, 27)
    line = _line("r02_bench_n1_final.json")
    assert line["gpu_launches"] == bench.kernels_per_step(L, False, 27) * line["steps"]


def test_committed_bench_lines_carry_the_contract_keys():
    for name, n in (("r02_bench_n1_final.json", 1), ("r02_bench_n2_final.json", 2), ("r02_bench_n4_final.json", 4),
                    ("r02_bench_n8_final.json", 8)):
        d = _line(name)
        for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                  "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "grad_check"):
            assert k in d, (name, k)
        assert d["n_gpus"] == n and d["warmup"] >= 3 and d["gpu_launches"] > 0
        assert d["metric"] == bench.METRIC and "workload" in d["config"]
        assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
        e = d["e2e"]
        assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
        r = d["roofline"]
        assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3
        if n == 1:
            assert "cpu_baseline" in d and d["cpu_baseline"]["kind"] == "port" and "train" in d and "extras" in d
        else:
            assert "exchange" in d and d["exchange"]["bytes"] == 14 * 300_000 * 4
    # the same gradient slab for every N (summation order only)
    l2 = [_line(f"r02_bench_n{n}_final.json")["grad_check"]["l2"] for n in (1, 2, 4, 8)]
    assert max(l2) - min(l2) <= 1e-5 * max(l2)
    ref = _line("r02_bench_reference_arm_final.json")
    assert ref["impl"] == "reference" and ref["e2e"]["h2d_bytes_per_step"] == 0 and ref["metric"] == bench.METRIC


def test_traffic_table_names_bench_stages():
    tj = json.load(open(os.path.join(PROFILES, "ncu_traffic.json")))
    stages = set(_line("r02_bench_n1_final.json")["stage_ms_rank0"])
    assert set(tj["bytes_per_launch"]) <= stages
    assert tj["views"] == 27 and tj["gaussians"] == 300_000
