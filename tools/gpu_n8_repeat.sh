#!/bin/bash
# the 8-GPU strong-scaling line twice in a row (run-to-run spread of a 6 ms timed region)
N=8
for r in 1 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$r bench.py --gpus $N --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null > gpurun_out/bench_n8_strong_rep$r.json
python -c "import json; d=json.loads(open('gpurun_out/bench_n8_strong_rep$r.json').read().strip().splitlines()[-1]); print('rep $r', round(d['value']), round(d['e2e']['value']), d['weak_scaling']['value'], d['weak_scaling']['e2e_value'], d['config']['streams_per_gpu'])"
done
