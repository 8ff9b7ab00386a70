#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "knn or golden or forward_matches" > gpurun_out/r2k_pytest.log 2>&1; tail -3 gpurun_out/r2k_pytest.log
show='import json,sys; d=json.loads(sys.stdin.read()); print("pairs", d["config"]["pairs_per_gpu_per_step"], "streams", d["config"]["streams_per_gpu"], round(d["value"]), round(d["e2e"]["value"]), {k: round(v, 3) for k, v in d["stage_ms"].items()})'
python bench.py --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "$show"
python bench.py --streams 1 --steps 30 --warmup 6 --no-cpu-baseline --no-latency --no-raw-scans 2>/dev/null | python -c "$show"
python bench.py --kind modelnet --points 1024 --pairs 1024 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "$show"
python bench.py --kind modelnet --points 1024 --pairs 1024 --streams 1 --steps 10 --warmup 3 --no-cpu-baseline --no-latency 2>/dev/null | python -c "$show"
