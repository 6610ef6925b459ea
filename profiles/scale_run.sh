#!/bin/bash
# N-GPU bench line the way the driver launches it.  usage (on the box): bash profiles/scale_run.sh <N> <tag> [bench flags]
N=${1:-8}; tag=${2:-x}; shift 2
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-train "$@" > gpurun_out/bench_n${N}_$tag.json 2> gpurun_out/bench_n${N}_$tag.err
tail -3 gpurun_out/bench_n${N}_$tag.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_n${N}_$tag.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), json.dumps(d.get("stage_ms_rank0")))
for k in ("exchange", "grad_check"):
    if k in d: print(k, json.dumps(d[k]))
PY
# the exchange kernel alone against ncclAllReduce (PEER_CHECK=1)
if [ -n "$PEER_CHECK" ]; then
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      tests/gpu_peer_check.py > gpurun_out/peer_check_n${N}_$tag.json 2> gpurun_out/peer_check_n${N}_$tag.err
  cat gpurun_out/peer_check_n${N}_$tag.json
fi
