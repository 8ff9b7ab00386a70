#!/bin/bash
# parity + FPS shape tuning
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for s in 32 16 8; do
  DCLR_FPS_SHAPE=$s python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-latency > gpurun_out/bench_fps$s.json 2> gpurun_out/bench_fps$s.err; echo "shape $s rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench_fps$s.json'));print($s, d['value'], d['e2e']['value'], d['stage_ms']['fps'])"
done
