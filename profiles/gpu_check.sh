#!/bin/bash
# One GPU-box visit (1 GPU): parity tests of the rasterizer, a short bench line, and the ncu launch list of the same
# command.  usage: bash profiles/gpu_check.sh <tag> [pytest targets...]   -> gpurun_out/{pytest,bench,launches}_<tag>.*
tag=${1:-x}; shift
targets=${@:-tests/test_gpu_parity.py tests/test_gpu_branches.py}
mkdir -p gpurun_out
python -m pytest $targets -m gpu -x -q --no-header -p no:cacheprovider > gpurun_out/pytest_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_$tag.log
tail -2 gpurun_out/pytest_$tag.log
python bench.py --steps 20 --warmup 5 --no-train --no-extras --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -2 gpurun_out/bench_$tag.err
# rank 3's share of an 8-way partition, alone on one GPU: the per-rank floor the 8-GPU step is made of (diagnostic line)
python bench.py --steps 20 --warmup 5 --no-train --no-extras --no-cpu-baseline --no-e2e --emulate-world 8 --emulate-rank 3 > gpurun_out/bench_${tag}_emu8.json 2> gpurun_out/bench_${tag}_emu8.err
python - <<PY
import json
for n in ("gpurun_out/bench_$tag.json", "gpurun_out/bench_${tag}_emu8.json"):
    try:
        d = json.loads(open(n).read().strip().splitlines()[-1])
        print(n, d["value"], d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), json.dumps(d["stage_ms_rank0"]))
    except Exception as e:
        print(n, "failed", e)
PY
# A/B of the programmatic dependent launch chain (same library, option "pdl")
GSR_PDL=1 python bench.py --steps 20 --warmup 5 --no-train --no-extras --no-cpu-baseline --no-e2e > gpurun_out/bench_${tag}_pdl.json 2> gpurun_out/bench_${tag}_pdl.err
GSR_PDL=1 python bench.py --steps 20 --warmup 5 --no-train --no-extras --no-cpu-baseline --no-e2e --emulate-world 8 --emulate-rank 3 > gpurun_out/bench_${tag}_pdl_emu8.json 2> gpurun_out/bench_${tag}_pdl_emu8.err
python - <<PY
import json
for n in ("gpurun_out/bench_${tag}_pdl.json", "gpurun_out/bench_${tag}_pdl_emu8.json"):
    try:
        d = json.loads(open(n).read().strip().splitlines()[-1])
        print(n, d["value"], d["ms_per_step"], json.dumps(d["stage_ms_rank0"]))
    except Exception as e:
        print(n, "failed", e)
PY
EMU=""; [ -n "$NCU_EMU" ] && EMU="--emulate-world 8 --emulate-rank 3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-train --no-e2e --no-graph --no-extras $EMU > gpurun_out/ncu_$tag.log 2>&1
python profiles/launch_summary.py gpurun_out/launches_$tag.csv > gpurun_out/launches_$tag.txt
head -32 gpurun_out/launches_$tag.txt
# A/B: every library under dynamic3dgaussians_b200/variants/ runs the same short bench (tests/build_variant.sh)
for so in dynamic3dgaussians_b200/variants/*.so; do
  [ -f "$so" ] || continue
  n=$(basename $so .so)
  GSR_LIB_PATH=$PWD/$so python bench.py --steps 20 --warmup 5 --no-train --no-extras --no-cpu-baseline --no-e2e > gpurun_out/bench_${tag}_$n.json 2> gpurun_out/bench_${tag}_$n.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${tag}_$n.json").read().strip().splitlines()[-1])
    print("variant $n", d["value"], d["ms_per_step"], json.dumps(d["stage_ms_rank0"]))
except Exception as e:
    print("variant $n failed", e)
PY
done
