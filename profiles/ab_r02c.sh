# scratch A/B job of round 2 (run on the GPU box from the repo root)
mkdir -p gpurun_out/j4
python -m pytest tests/test_gpu_boxes.py tests/test_gpu_pipeline.py tests/test_gpu_libs_mirror.py tests/test_gpu_fused.py -m gpu -x -q > gpurun_out/j4/gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/j4/gputests.log
for i in 1 2; do
python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --breakdown > gpurun_out/j4/new_$i.json 2> gpurun_out/j4/new_$i.err
done
grep -H breakdown gpurun_out/j4/*.err
python bench.py --no-cpu > gpurun_out/j4/bench_extras.json 2> gpurun_out/j4/bench_extras.err
python -c "
import json
d=json.loads(open('gpurun_out/j4/bench_extras.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['roofline']['kernel_ms']); print(json.dumps(d['g_sweep']))"
