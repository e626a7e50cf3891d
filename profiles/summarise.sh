#!/bin/bash
# Turn what profiles/capture.sh brought back in gpurun_out/ into the tracked summaries (no GPU needed; run from the repo root
# with the same build as the capture, so that the SASS line attribution matches).
set -u
cd gpurun_out
cp bench_default.json ../profiles/r01_bench_default.json
cp bench_reference.json ../profiles/r01_bench_reference_arm.json
cp launches.csv ../profiles/r01_launches.csv
tail -1 breakdown.txt > ../profiles/r01_breakdown_default.txt
ncu -i box_full.ncu-rep --page raw --csv > box_raw.csv 2>/dev/null
ncu -i roi_fwd_full.ncu-rep --page raw --csv > roi_fwd_raw.csv 2>/dev/null
ncu -i roi_bwd_full.ncu-rep --page raw --csv > roi_bwd_raw.csv 2>/dev/null
ncu -i box_full.ncu-rep --page source --csv --kernel-name regex:anchor_target > at_src.csv 2>/dev/null
{ echo "# ncu --set full --clock-control none, one launch each inside: python bench.py --steps 2 --warmup 1 --no-cpu --no-extras (B200, round 1, profiles/capture.sh)"
  python ../profiles/ncu_summary.py box_raw.csv; } > ../profiles/r01_ncu_box_kernels.txt
{ echo "# ncu --set full: RoIAlign kernels, BASELINE configs[2] (profiles/roi_probe.py); columns: NCHW | NHWC (channels-last)"
  echo "# forward"; python ../profiles/ncu_summary.py roi_fwd_raw.csv
  echo "# backward (scatter; the dense zero fill is a separate memset)"; python ../profiles/ncu_summary.py roi_bwd_raw.csv; } > ../profiles/r01_ncu_roi.txt
python - <<'PY'
import csv, json
rows = list(csv.reader(open('box_raw.csv')))
h = rows[0]; ki = h.index('Kernel Name')
sel = [rows[0], rows[1]] + [r for r in rows[2:] if 'anchor_target_kernel' in r[ki]]
csv.writer(open('at_only.csv', 'w')).writerows(sel)
def mb(r, name):
    i = h.index(name); v = float(r[i]); u = rows[1][i]
    return int(v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u])
p = '../profiles/ncu_traffic.json'
d = json.load(open(p))
for kid, key in (("3", 'anchor_target_kernel'), ("2", 'colmax_kernel')):
    r = [x for x in rows[2:] if key in x[ki]][0]
    rd, wr = mb(r, 'dram__bytes_read.sum'), mb(r, 'dram__bytes_write.sum')
    d[kid] = {"kernel": r[ki][:60], "bytes": rd + wr, "read": rd, "write": wr, "capture": "gpurun_out/box_full.ncu-rep via profiles/capture.sh (r01)"}
json.dump(d, open(p, 'w'), indent=1)
PY
{ echo "# ncu --set full, anchor_target_kernel inside bench.py --steps 2 --warmup 1 --no-cpu --no-extras (B200, round 1, profiles/capture.sh)"
  python ../profiles/ncu_summary.py at_only.csv; echo
  echo "# executed warp instructions by CUDA line (profiles/sass_lines.py)"
  python ../profiles/sass_lines.py at_src.csv ../detectron-pytorch_b200/build/anchor_target.o anchor_target_kernelILi2ELb1ELi4 25 2>&1 | tail -26; } > ../profiles/r01_ncu_anchor_target.txt
