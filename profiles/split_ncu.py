"""Splits the condensed ncu summary of one step (profiles/ncu_full.sh -> gpurun_out/ncu_<tag>_summary.txt) into the committed
per-group files profiles/r02_ncu_{blend,binning,preprocess}.txt and refreshes profiles/ncu_traffic.json (DRAM bytes per
launch, summed over the kernels of a bench stage).  usage: python profiles/split_ncu.py gpurun_out/ncu_<tag>_summary.txt"""
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
txt = open(sys.argv[1]).read()
blocks = ['----- ' + b for b in txt.split('----- ')[1:]]
name = lambda b: b.split('\n')[0]
first = {}
keep = []
for b in blocks:                      # one step: stop at the first kernel that repeats beyond its count in a step
    keep.append(b)
blocks = keep[:24]
groups = {'binning': [], 'preprocess': [], 'blend': []}
for b in blocks:
    n = name(b)
    if 'blend_' in n:
        groups['blend'].append(b)
    elif 'preprocess' in n or 'gaussian_static' in n:
        groups['preprocess'].append(b)
    else:
        groups['binning'].append(b)
hdr = ("ncu --set full --clock-control none --import-source on, the kernels of one forward+backward step of `python bench.py --steps 1 "
       "--warmup 1 --no-cpu-baseline --no-train --no-e2e --no-graph --no-extras` (C2: 300k Gaussians, 27 views 640x360) on one B200, "
       "profiles/ncu_full.sh, condensed by profiles/ncu_summary.py.  FINAL build of round 2.  Times under ncu are cold-cache and "
       "serialised: compare shares, not absolutes.\n"
       "Launch order inside the step: store_cams x2, gaussian_static, preprocess, radix_scan_hist, onesweep x4 (depth digits 0-3), "
       "st_count_reduce, block_sums_scan, emit_st_instances, sort_hist_from_keys, chunk_table (side stream), radix_scan_hist, onesweep x1 "
       "(supertile bits), chunk_count, tile_offsets, tile_scan, tile_order (side stream), pack_colors, scatter_records, blend_forward, "
       "blend_backward, preprocess_backward.\n\n")
for g, bl in groups.items():
    open(os.path.join(HERE, f'r02_ncu_{g}.txt'), 'w').write(hdr + ''.join(bl))
unit = {'Mbyte': 1e6, 'Gbyte': 1e9, 'Kbyte': 1e3, 'byte': 1}
tr = {}
for b in blocks:
    k = re.search(r'(\w+_kernel)', name(b)).group(1)
    rd = re.search(r'dram__bytes_read.sum\s+([0-9.]+) (\w+)', b)
    wr = re.search(r'dram__bytes_write.sum\s+([0-9.]+) (\w+)', b)
    tr.setdefault(k, []).append(int(float(rd.group(1)) * unit[rd.group(2)] + float(wr.group(1)) * unit[wr.group(2)]))
one = sorted(tr['onesweep_pass_kernel'])
stage = {
    "blend_backward": tr['blend_backward_kernel'][0], "blend_forward": tr['blend_forward_kernel'][0],
    "scatter_records": tr['scatter_records_kernel'][0] + tr['pack_colors_kernel'][0],
    "preprocess": tr['preprocess_kernel'][0] + tr['gaussian_static_kernel'][0],
    "preprocess_backward": tr['preprocess_backward_kernel'][0],
    "emit_supertile_instances": tr['st_count_reduce_kernel'][0] + tr['block_sums_scan_kernel'][0]
    + tr['emit_st_instances_kernel'][0] + tr['sort_hist_from_keys_kernel'][0],
    "depth_sort": sum(one[:4]) + min(tr['radix_scan_hist_kernel']),          # the four depth passes move less than the instance pass
    "supertile_sort": one[4] + min(tr['radix_scan_hist_kernel']),
    "tile_offsets": tr['chunk_count_kernel'][0] + tr['tile_offsets_kernel'][0] + tr['tile_scan_kernel'][0]
    + tr['chunk_table_kernel'][0] + tr['tile_order_kernel'][0]}
json.dump({"source": "profiles/r02_ncu_{blend,binning,preprocess}.txt (ncu --set full --clock-control none --import-source on, python "
                     "bench.py --steps 1 --warmup 1 --no-graph, C2, one B200, profiles/ncu_full.sh, final build of round 2; "
                     "dram__bytes_read.sum + dram__bytes_write.sum per launch, summed over the kernels of a stage)",
           "views": 27, "gaussians": 300000, "bytes_per_launch": stage},
          open(os.path.join(HERE, 'ncu_traffic.json'), 'w'), indent=2)
for b in blocks:
    def g(k):
        m = re.search(k + r'\s+([0-9.]+)', b)
        return float(m.group(1)) if m else 0
    print(f"{name(b)[6:56]:50s} {g('gpu__time_duration.sum') * 1e3:7.1f}us issue {g('smsp__issue_active.avg.pct_of_peak_sustained_active'):5.1f} "
          f"dram% {g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} inst {g('smsp__inst_executed.sum'):.3g}")
