#!/bin/bash
# round-2 GPU session A: multi-pick FPS -- parity, variants side by side, per-round trace, quick bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "fps" > gpurun_out/r2a_pytest_fps.log 2>&1; tail -3 gpurun_out/r2a_pytest_fps.log
python tools/fps_ab.py > gpurun_out/r2a_fps_ab.jsonl 2> gpurun_out/r2a_fps_ab.err; tail -40 gpurun_out/r2a_fps_ab.jsonl
for v in "8 4" "8 2" "16 4" "8 1" "8 0"; do set -- $v
  DCLR_LIB=$PWD/deepclr_b200/lib_trace.so DCLR_FPS_SHAPE=$1 DCLR_FPS_MP=$2 python tools/fps_trace.py > gpurun_out/r2a_fps_trace_$1x$2.txt 2>&1
  tail -4 gpurun_out/r2a_fps_trace_$1x$2.txt
done
nvcc -O3 -arch=sm_100a -o tools/micro/lat tools/micro/lat.cu && tools/micro/lat > gpurun_out/r2a_lat.txt 2>&1; cat gpurun_out/r2a_lat.txt
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-raw-scans > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
python -c "import json; d=json.load(open('gpurun_out/r2a_bench.json')); print(round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()}, d['latency_b1'])"
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest_gpu.log 2>&1; tail -3 gpurun_out/r2a_pytest_gpu.log
