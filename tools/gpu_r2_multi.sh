#!/bin/bash
# Multi-GPU lines of round 2 (run with gpurun --gpus N): usage tools/gpu_r2_multi.sh N
# strong scaling of the KITTI config (64 pairs per step in total) with the weak number beside it, and at N = 8 the
# ModelNet-shaped partial-overlap batch (configs[3]).
N=${1:-2}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
run --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans > gpurun_out/bench_n${N}_strong.json 2> gpurun_out/bench_n${N}_strong.err; echo "strong rc=$?"
if [ "$N" = 8 ]; then
  run --kind modelnet --points 1024 --pairs 1024 --partial --noise 0.04 --steps 20 --warmup 5 --no-cpu-baseline --no-latency --no-raw-scans > gpurun_out/bench_modelnet_partial_n8.json 2> gpurun_out/bench_modelnet_partial_n8.err; echo "modelnet rc=$?"
fi
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_*n${N}*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['scaling'], round(d['value']), round(d['e2e']['value']), d.get('weak_scaling') or d.get('strong_scaling'), d['config'].get('streams_per_gpu'))
    except Exception as e: print(f, 'ERR', e)
PY
