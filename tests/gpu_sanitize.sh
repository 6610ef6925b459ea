#!/bin/bash
# compute-sanitizer over the hot path (run on the GPU box):  bash tests/gpu_sanitize.sh [outdir]
# memcheck / racecheck / synccheck / initcheck on tests/gpu_sanitize_case.py; one summary per tool is written to
# <outdir>/sanitize_<tool>.txt (default gpurun_out/), the ones to keep go to profiles/.
# The barrier-free TMA ring of the blend kernels (a slot is re-issued to cp.async.bulk by the last warp out, behind
# __threadfence_block + fence.proxy.async) is what racecheck / synccheck are here for.
set -u
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
CS=${COMPUTE_SANITIZER:-/usr/local/cuda/bin/compute-sanitizer}
rc_all=0
for tool in memcheck racecheck synccheck initcheck; do
  size="12000 4"
  [ "$tool" = racecheck ] && size="6000 2"
  extra=""
  [ "$tool" = initcheck ] && extra="--track-unused-memory no"
  log="$OUT/sanitize_${tool}.txt"
  echo "# $CS --tool $tool $extra --print-limit 20 python tests/gpu_sanitize_case.py $size" > "$log"
  timeout 900 "$CS" --tool "$tool" $extra --print-limit 20 python tests/gpu_sanitize_case.py $size >> "$log" 2>&1
  rc=$?
  echo "# exit code $rc" >> "$log"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|^ok:" "$log" | tail -3
  [ $rc -ne 0 ] && rc_all=$rc
done
exit $rc_all
