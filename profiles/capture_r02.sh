#!/bin/bash
# Round-2 profile refresh (on the GPU box, from the repo root: bash profiles/capture_r02.sh [tag]).  Each ncu pass runs
# only after the same command has exited 0 without ncu.  1. breakdown (no profiler), 2. the ncu launch list of
# `bench.py --steps 2 --warmup 1 --no-cpu --no-extras`, 3. `--set full` of the box kernels inside that command,
# 4. `--set full` of every RoIAlign flavour (profiles/roi_ncu_probe.py: read by kernel name, not by skip count).
set -u
T=${1:-r02}
O=gpurun_out/$T
mkdir -p $O
python bench.py --steps 20 --warmup 3 --no-cpu --no-extras --breakdown > $O/bench_noextras.json 2> $O/breakdown.txt || exit 1
python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > $O/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:"anchor_target_kernel|colmax_kernel|label_weight_kernel|objectness_kernel|partition_kernel|nms_mask_kernel|nms_resolve_seg" \
    -s 14 -c 7 -o $O/box_full -f python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > $O/ncu_box.log 2>&1
python profiles/roi_ncu_probe.py > $O/roi_probe.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"roi_" -c 8 -o $O/roi_full -f \
    python profiles/roi_ncu_probe.py > $O/ncu_roi.log 2>&1
ls -la $O/*.ncu-rep $O/launches.csv
