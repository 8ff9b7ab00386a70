#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2f_pytest_gpu.log 2>&1; tail -8 gpurun_out/r2f_pytest_gpu.log
python tools/seq_bench.py > gpurun_out/r2f_seq.jsonl 2> gpurun_out/r2f_seq.err; cat gpurun_out/r2f_seq.jsonl; tail -2 gpurun_out/r2f_seq.err
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; tail -2 gpurun_out/r2f_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r2f_bench.json'))
print(round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()}, d['latency_b1'], d.get('e2e_raw_scans',{}).get('value'))"
