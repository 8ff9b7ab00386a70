#!/bin/bash
# round-2 GPU session D: full GPU test-suite, smoke, bench (N=1) with the new line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2d_pytest_gpu.log 2>&1; tail -5 gpurun_out/r2d_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2d_smoke.log 2>&1; tail -5 gpurun_out/r2d_smoke.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; tail -3 gpurun_out/r2d_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r2d_bench.json'))
print(round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()}, d['latency_b1'])
print(d['roofline']); print(d['roofline_e2e']); print(d['h2d_copy']); print(d.get('b1_baseline')); print(d.get('cpu_baseline'))
print({k: (round(v['frac'],4) if v.get('frac') else None) for k,v in d['roofline_stages'].items()})"
for p in 8 16 32; do python bench.py --pairs $p --steps 30 --warmup 5 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('pairs', d['config']['pairs_per_gpu_per_step'], 'streams', d['config']['streams_per_gpu'], round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()})"
python bench.py --pairs $p --streams 1 --steps 30 --warmup 5 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('pairs', d['config']['pairs_per_gpu_per_step'], 'streams', d['config']['streams_per_gpu'], round(d['value']), round(d['e2e']['value']))"
done
python bench.py --streams 2 --steps 20 --warmup 5 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('pairs', d['config']['pairs_per_gpu_per_step'], 'streams', d['config']['streams_per_gpu'], round(d['value']), round(d['e2e']['value']))"
