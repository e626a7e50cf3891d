# multi-GPU lines of round 2, session 12 (on an N-GPU box: bash profiles/multi_r02c.sh N)
N=${1:-2}
mkdir -p gpurun_out/multi
if [ "$N" = "2" ]; then python -m pytest tests/test_gpu_parallel.py -x -q -m gpu > gpurun_out/multi/parallel_tests.txt 2>&1; echo "parallel tests rc=$?"; tail -1 gpurun_out/multi/parallel_tests.txt; fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu > gpurun_out/multi/bench_${N}gpu.json 2> gpurun_out/multi/bench_${N}gpu.err; echo "bench rc=$?"
python -c "
import json,sys
d=json.loads(open('gpurun_out/multi/bench_${N}gpu.json').read().strip().splitlines()[-1])
print('n',d['n_gpus'],'value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'strong',d.get('strong_64',{}).get('ms_per_step'), d.get('strong_64',{}).get('checksum'), 'resident', d.get('e2e_targets_resident',{}).get('value'))"
