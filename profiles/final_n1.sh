#!/bin/bash
# Round-end evidence on ONE GPU: the whole GPU test-suite, smoke(), the default bench line, the reference arm, the C5
# workload, and the ncu launch list of the bench command.  usage: bash profiles/final_n1.sh <tag>
tag=${1:-final}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --no-header -p no:cacheprovider > gpurun_out/pytest_all_$tag.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all_$tag.log; tail -3 gpurun_out/pytest_all_$tag.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_$tag.log 2>&1; tail -1 gpurun_out/smoke_$tag.log
python bench.py > gpurun_out/bench_n1_$tag.json 2> gpurun_out/bench_n1_$tag.err; tail -2 gpurun_out/bench_n1_$tag.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm_$tag.json 2> gpurun_out/bench_reference_arm_$tag.err
python bench.py --workload c5 --no-train --no-extras --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/bench_c5_n1_$tag.json 2> gpurun_out/bench_c5_n1_$tag.err
python bench.py --workload c4 --no-train --no-extras --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/bench_c4_n1_$tag.json 2> gpurun_out/bench_c4_n1_$tag.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-train --no-e2e --no-graph --no-extras > gpurun_out/ncu_$tag.log 2>&1
python profiles/launch_summary.py gpurun_out/launches_$tag.csv > gpurun_out/launches_$tag.txt
python - <<PY
import json
for n in ("bench_n1_$tag", "bench_reference_arm_$tag", "bench_c5_n1_$tag", "bench_c4_n1_$tag"):
    try:
        d = json.loads(open(f"gpurun_out/{n}.json").read().strip().splitlines()[-1])
        print(n, d.get("value"), d.get("ms_per_step"), "e2e", (d.get("e2e") or {}).get("value"), "launches", d.get("gpu_launches"))
    except Exception as e:
        print(n, "failed", e)
PY
