#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "fps" > gpurun_out/r2g_pytest_fps.log 2>&1; tail -2 gpurun_out/r2g_pytest_fps.log
python tools/fps_ab.py 8x0 8x2 8x3 8x4 > gpurun_out/r2g_fps_ab.jsonl 2> gpurun_out/r2g_fps_ab.err; cat gpurun_out/r2g_fps_ab.jsonl
show='import json,sys; d=json.loads(sys.stdin.read()); print("pairs", d["config"]["pairs_per_gpu_per_step"], "streams", d["config"]["streams_per_gpu"], round(d["value"]), round(d["e2e"]["value"]), {k: round(v, 3) for k, v in d["stage_ms"].items()})'
for s in 2 3 4; do python bench.py --streams $s --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans 2>gpurun_out/r2g_err_$s.log | python -c "$show"; done
for s in 2 3; do python bench.py --pairs 8 --streams $s --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "$show"; done
python bench.py --kind modelnet --points 1024 --pairs 1024 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "$show"
python bench.py --kind modelnet --points 1024 --pairs 1024 --streams 1 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "$show"
