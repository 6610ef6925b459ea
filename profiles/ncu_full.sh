#!/bin/bash
# One `ncu --set full` capture of every kernel of ONE forward+backward step of the bench workload (run on the GPU box, 1 GPU).
# usage: bash profiles/ncu_full.sh <tag>   -> gpurun_out/ncu_<tag>.ncu-rep, ncu_<tag>_raw.csv, ncu_<tag>_summary.txt
tag=${1:-x}
mkdir -p gpurun_out
K='regex:gaussian_static|preprocess_kernel|onesweep_pass|radix_scan_hist|emit_st|st_count|block_sums|sort_hist_from_keys|chunk_table|chunk_count|tile_offsets|tile_scan|tile_order|pack_colors|scatter_records|preprocess_backward|blend_forward|blend_backward'
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-train --no-e2e --no-graph --no-extras"
$CMD > gpurun_out/ncu_${tag}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "$K" --launch-skip 48 --launch-count 24 -f -o gpurun_out/ncu_$tag $CMD > gpurun_out/ncu_${tag}.log 2>&1
ncu -i gpurun_out/ncu_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${tag}_raw.csv 2>/dev/null
python profiles/ncu_summary.py gpurun_out/ncu_${tag}_raw.csv > gpurun_out/ncu_${tag}_summary.txt 2>&1
ls -la gpurun_out/ncu_$tag.ncu-rep; grep -c "^-----" gpurun_out/ncu_${tag}_summary.txt
