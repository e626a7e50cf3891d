#!/bin/bash
# gpurun_out/<tag>/ (profiles/capture_r02.sh) -> tracked summaries profiles/<tag>_* (no GPU needed; same build as the capture).
set -u
T=${1:-r02}
cd gpurun_out/$T
cp launches.csv ../../profiles/${T}_launches.csv
tail -1 breakdown.txt > ../../profiles/${T}_breakdown.txt
ncu -i box_full.ncu-rep --page raw --csv > box_raw.csv 2>/dev/null
ncu -i roi_full.ncu-rep --page raw --csv > roi_raw.csv 2>/dev/null
ncu -i box_full.ncu-rep --page source --csv --kernel-name regex:anchor_target > at_src.csv 2>/dev/null
{ echo "# ncu --set full --clock-control none, one launch each inside: python bench.py --steps 2 --warmup 1 --no-cpu --no-extras (B200, $T, profiles/capture_r02.sh)"
  python ../../profiles/ncu_summary.py box_raw.csv; } > ../../profiles/${T}_ncu_box_kernels.txt
python - "$T" <<'PY'
import csv, json, sys
T = sys.argv[1]
rows = list(csv.reader(open('roi_raw.csv')))
h = rows[0]; ki = h.index('Kernel Name')
# the SECOND (warm) launch of every distinct kernel name, in capture order
seen, pick = {}, []
for r in rows[2:]:
    seen[r[ki]] = seen.get(r[ki], 0) + 1
    if seen[r[ki]] == 2:
        pick.append(r)
csv.writer(open('roi_pick.csv', 'w')).writerows([rows[0], rows[1]] + pick)
rows = list(csv.reader(open('box_raw.csv')))
h = rows[0]; ki = h.index('Kernel Name')
csv.writer(open('at_only.csv', 'w')).writerows([rows[0], rows[1]] + [r for r in rows[2:] if 'anchor_target_kernel' in r[ki]])
def mb(r, name):
    i = h.index(name); v = float(r[i]); u = rows[1][i]
    return int(v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u])
p = '../../profiles/ncu_traffic.json'
d = json.load(open(p))
for kid, key in (("3", 'anchor_target_kernel'), ("2", 'colmax_kernel')):
    r = [x for x in rows[2:] if key in x[ki]][0]
    rd, wr = mb(r, 'dram__bytes_read.sum'), mb(r, 'dram__bytes_write.sum')
    d[kid] = {"kernel": r[ki][:60], "bytes": rd + wr, "read": rd, "write": wr, "capture": "gpurun_out/%s/box_full.ncu-rep via profiles/capture_r02.sh" % T}
json.dump(d, open(p, 'w'), indent=1)
PY
{ echo "# ncu --set full: RoIAlign kernels, BASELINE configs[2] (profiles/roi_ncu_probe.py), the second (warm) launch of every flavour;"
  echo "# one column per kernel, identified by the 'Kernel Name' row (template arguments included): NCHW forward | NCHW backward | channels-last forward | channels-last backward"
  python ../../profiles/ncu_summary.py roi_pick.csv; } > ../../profiles/${T}_ncu_roi.txt
K=$(grep -o "anchor_target_kernelILi[0-9]ELb[01]ELi[0-9]ELb0ELb0" ../../detectron-pytorch_b200/build/anchor_target.ptxas.log | head -1)
{ echo "# ncu --set full, anchor_target_kernel inside bench.py --steps 2 --warmup 1 --no-cpu --no-extras (B200, $T, profiles/capture_r02.sh)"
  python ../../profiles/ncu_summary.py at_only.csv; echo
  echo "# executed warp instructions by CUDA line (profiles/sass_lines.py)"
  python ../../profiles/sass_lines.py at_src.csv ../../detectron-pytorch_b200/build/anchor_target.o anchor_target_kernelILi2ELb1ELi4ELb0ELb0 30 2>&1 | tail -31; } > ../../profiles/${T}_ncu_anchor_target.txt
