# final checks of round 2, session 12 (on the GPU box, from the repo root): full GPU suite, the no-alloc test alone, default bench, reference arm
mkdir -p gpurun_out/final
python -m pytest tests -x -q -m gpu > gpurun_out/final/gpu_tests.txt 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/final/gpu_tests.txt
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/final/no_alloc_alone.txt 2>&1; echo "no_alloc alone rc=$?"; tail -1 gpurun_out/final/no_alloc_alone.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final/smoke.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/final/smoke.txt
python bench.py --impl reference > gpurun_out/final/bench_reference_arm.json 2> gpurun_out/final/bench_reference_arm.err; echo "ref rc=$?"
python bench.py > gpurun_out/final/bench_default.json 2> gpurun_out/final/bench_default.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/final/bench_default.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/final/bench_reference_arm.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'frac',d['roofline']['frac'],'kernel_ms',d['roofline']['kernel_ms'],'e2e',d['e2e']['value'],'ref',r['value'], 'ratio', d['e2e']['value']/r['value'])"
