# scratch A/B job of round 2 (run on the GPU box from the repo root)
mkdir -p gpurun_out/j11
python -m pytest tests/test_gpu_boxes.py tests/test_gpu_post.py tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/j11/gputests.log 2>&1; echo "tests rc=$?"; tail -1 gpurun_out/j11/gputests.log
for i in 1 2 3; do
python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --breakdown > gpurun_out/j11/new_$i.json 2> gpurun_out/j11/new_$i.err
done
grep -H breakdown gpurun_out/j11/*.err
python bench.py --no-cpu > gpurun_out/j11/ex.json 2> gpurun_out/j11/ex.err
python -c "
import json
d=json.loads(open('gpurun_out/j11/ex.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['roofline']['kernel_ms'], {k:v['anchor_target_kernel_ms'] for k,v in d['g_sweep'].items()})"
