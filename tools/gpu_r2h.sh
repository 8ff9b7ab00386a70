#!/bin/bash
mkdir -p gpurun_out
show='import json,sys; d=json.loads(sys.stdin.read()); print("pairs", d["config"]["pairs_per_gpu_per_step"], "streams", d["config"]["streams_per_gpu"], round(d["value"]), round(d["e2e"]["value"]))'
for p in 8 16 32; do for s in 2 3 4 6 8; do python bench.py --pairs $p --streams $s --steps 40 --warmup 8 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "$show"; done; done
python tools/sa_density_sweep.py > gpurun_out/r2h_sa_density.jsonl 2> gpurun_out/r2h_sa_density.err; cat gpurun_out/r2h_sa_density.jsonl; tail -2 gpurun_out/r2h_sa_density.err
python tools/op_sweep.py --ops dense --sizes 1024,4096,16384,65536 > gpurun_out/r2h_op_sweep_dense.jsonl 2> gpurun_out/r2h_op_sweep_dense.err; cat gpurun_out/r2h_op_sweep_dense.jsonl | cut -c1-600; tail -2 gpurun_out/r2h_op_sweep_dense.err
