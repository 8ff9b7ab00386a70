#!/bin/bash
# multi-GPU session: the driver's launch for N > 1 (strong scaling by default, weak beside it)
mkdir -p gpurun_out
N=${1:-2}
python -m pytest tests/test_gpu_configs.py -m gpu -x -q -k "replicas or sequential" 2>&1 | tail -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2i_bench_n$N.json 2> gpurun_out/r2i_bench_n$N.err; tail -3 gpurun_out/r2i_bench_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/r2i_bench_n$N.json'))
print('N', d['n_gpus'], d['scaling'], round(d['value']), 'e2e', round(d['e2e']['value']), d['config']['pairs_per_gpu_per_step'], 'weak:', d.get('weak_scaling'), d['h2d_copy'])"
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2i_bench_n1.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2i_bench_n1.json'))
print('N', d['n_gpus'], d['scaling'], round(d['value']), 'e2e', round(d['e2e']['value']), d['latency_b1'])"
if [ "$N" -ge 4 ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --kind modelnet --points 1024 --pairs 1024 --partial --noise 0.04 --steps 20 --warmup 5 --no-latency > gpurun_out/r2i_bench_modelnet_partial_n$N.json 2> gpurun_out/r2i_bench_modelnet_partial_n$N.err; tail -2 gpurun_out/r2i_bench_modelnet_partial_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/r2i_bench_modelnet_partial_n$N.json'))
print('modelnet partial N', d['n_gpus'], d['scaling'], round(d['value']), 'e2e', round(d['e2e']['value']), d['config']['pairs_per_gpu_per_step'], 'weak:', d.get('weak_scaling'))"
fi
