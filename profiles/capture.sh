#!/bin/bash
# One-call profile refresh (run on the GPU box from the repo root: bash profiles/capture.sh).
# 1. the default bench line and the reference arm (no profiler), 2. per-kernel breakdown, 3. the ncu launch list of
# the same command, 4. one `--set full` capture of the box kernels and of RoIAlign forward.  Summaries are made from
# the .ncu-rep files afterwards with profiles/ncu_summary.py / sass_lines.py (no GPU needed).
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2>/dev/null
python bench.py --steps 20 --warmup 3 --no-cpu --no-extras --breakdown > /dev/null 2> gpurun_out/breakdown.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:"anchor_target_kernel|colmax_kernel|label_weight_kernel|objectness_kernel|partition_kernel|nms_mask_kernel|nms_resolve_seg" \
    -s 14 -c 7 -o gpurun_out/box_full -f python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_box.log 2>&1
# roi_probe launches: 14 x fwd NCHW, 28 x bwd NCHW, 14 x fwd NHWC, 28 x bwd NHWC -> the last NCHW and the first NHWC of each
ncu --set full --clock-control none --import-source on -k regex:roi_fwd_kernel -s 13 -c 2 -o gpurun_out/roi_fwd_full -f \
    python profiles/roi_probe.py > gpurun_out/ncu_roi.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:roi_bwd_kernel -s 27 -c 2 -o gpurun_out/roi_bwd_full -f \
    python profiles/roi_probe.py >> gpurun_out/ncu_roi.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/launches.csv
