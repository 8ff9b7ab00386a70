#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2e_pytest_gpu.log
show='import json,sys; d=json.loads(sys.stdin.read()); print("pairs", d["config"]["pairs_per_gpu_per_step"], "streams", d["config"]["streams_per_gpu"], round(d["value"]), round(d["e2e"]["value"]))'
for p in 64 8; do for s in 1 2 3 4; do python bench.py --pairs $p --streams $s --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "$show"; done; done
python bench.py --kind modelnet --points 1024 --pairs 1024 --steps 10 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2e_bench_modelnet.json 2>gpurun_out/r2e_bench_modelnet.err
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_modelnet.json')); print('modelnet', round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()})"
python bench.py --kind modelnet --points 1024 --pairs 1024 --partial --noise 0.04 --steps 10 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/r2e_bench_modelnet_partial.json 2>gpurun_out/r2e_bench_modelnet_partial.err
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_modelnet_partial.json')); print('modelnet partial', round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()})"
