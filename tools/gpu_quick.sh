#!/bin/bash
# quick check: FPS-related parity + a short bench with per-stage times
python -m pytest tests -m gpu -x -q -k "${1:-fps}" 2>&1 | tail -2
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-latency | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), {k: round(v, 4) for k, v in d['stage_ms'].items()})"
